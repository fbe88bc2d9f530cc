// tex_ramp_probe.cu - experiment: read the hardware's interpolation weight directly from a ramp
// texture T[i] = i: tex(x, centre row) - floor = alpha as the filter sees it.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)
__global__ void sweep(cudaTextureObject_t tex, float x0, float dx, int n, float *out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) out[k] = tex2DLayered<float>(tex, x0 + dx * (float)k, 5.5f, 0);
}
int main() {
    const int w = 64, h = 16;
    std::vector<float> img((size_t)w * h);
    for (int y = 0; y < h; ++y) for (int x = 0; x < w; ++x) img[y * w + x] = (float)x;
    cudaArray_t arr;
    cudaChannelFormatDesc desc = cudaCreateChannelDesc(32, 0, 0, 0, cudaChannelFormatKindFloat);
    CK(cudaMalloc3DArray(&arr, &desc, make_cudaExtent(w, h, 1), cudaArrayLayered));
    cudaMemcpy3DParms cp = {};
    cp.srcPtr = make_cudaPitchedPtr(img.data(), (size_t)w * 4, w, h);
    cp.dstArray = arr; cp.extent = make_cudaExtent(w, h, 1); cp.kind = cudaMemcpyHostToDevice;
    CK(cudaMemcpy3D(&cp));
    cudaResourceDesc res = {}; res.resType = cudaResourceTypeArray; res.res.array.array = arr;
    cudaTextureDesc td = {}; td.addressMode[0] = td.addressMode[1] = cudaAddressModeWrap;
    td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeElementType;
    cudaTextureObject_t tex; CK(cudaCreateTextureObject(&tex, &res, &td, nullptr));
    const int n = 1 << 16;                      // x = 10.5 + k / 65536 * 1 : one texel interval in 2^-16 steps
    float *d; CK(cudaMalloc(&d, n * 4));
    for (float base : {10.5f, 37.5f}) {
        sweep<<<n / 256, 256>>>(tex, base, 1.0f / 65536.0f, n, d);
        std::vector<float> o(n);
        CK(cudaMemcpy(o.data(), d, n * 4, cudaMemcpyDeviceToHost));
        printf("base %.1f: first values:", base);
        for (int k = 0; k < 4; ++k) printf(" %.9g", o[k]);
        printf("\n step edges (k where the value changes, value after):");
        int edges = 0;
        for (int k = 1; k < n && edges < 6; ++k)
            if (o[k] != o[k - 1]) { printf(" [k=%d -> %.9g]", k, o[k]); ++edges; }
        int distinct = 1;
        for (int k = 1; k < n; ++k) if (o[k] != o[k - 1]) ++distinct;
        printf("\n distinct values over one texel interval: %d; last: %.9g\n", distinct, o[n - 1]);
    }
    return 0;
}
