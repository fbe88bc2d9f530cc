// tex_weight_probe.cu - experiment: the hardware's four 2-D bilinear weights, read through a
// one-hot texture (T = 1 at texel (10, 5)): sampling at (cx + a/256, cy + b/256) with the hot texel in
// corner c of the footprint returns that corner's weight.  Tests w_c = round(wa * wb / 256) / 256.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)
__global__ void grid(cudaTextureObject_t tex, float x0, float y0, float *out) {   // a = threadIdx.x/256, b = blockIdx.x/256
    const float a = (float)threadIdx.x / 256.0f, b = (float)blockIdx.x / 256.0f;
    out[blockIdx.x * 256 + threadIdx.x] = tex2DLayered<float>(tex, x0 + a, y0 + b, 0);
}
int main() {
    const int w = 64, h = 16;
    std::vector<float> img((size_t)w * h, 0.0f);
    img[5 * w + 10] = 1.0f;
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
    float *d; CK(cudaMalloc(&d, 256 * 256 * 4));
    std::vector<float> W[4];
    // corner 0: hot texel is T00 (i = 10, j = 5): x in [10.5, 11.5); corner 1 (T10): x in [9.5, 10.5) ...
    const float x0s[4] = {10.5f, 9.5f, 10.5f, 9.5f}, y0s[4] = {5.5f, 5.5f, 4.5f, 4.5f};
    for (int c = 0; c < 4; ++c) {
        grid<<<256, 256>>>(tex, x0s[c], y0s[c], d);
        W[c].resize(65536);
        CK(cudaMemcpy(W[c].data(), d, 65536 * 4, cudaMemcpyDeviceToHost));
    }
    long match[4] = {0, 0, 0, 0}, match_floor[4] = {0, 0, 0, 0}, sum256 = 0, ints = 0;
    for (int b = 0; b < 256; ++b) for (int a = 0; a < 256; ++a) {
        const int wa[4] = {256 - a, a, 256 - a, a}, wb[4] = {256 - b, 256 - b, b, b};
        int s = 0;
        bool all_int = true;
        for (int c = 0; c < 4; ++c) {
            const double v = (double)W[c][b * 256 + a] * 256.0;
            const int vi = (int)(v + 0.5);
            if (fabs(v - vi) > 1e-6) all_int = false;
            s += vi;
            if (vi == (wa[c] * wb[c] + 128) / 256) match[c]++;
            if (vi == (wa[c] * wb[c]) / 256) match_floor[c]++;
        }
        if (all_int) ++ints;
        if (s == 256) ++sum256;
    }
    printf("weights are multiples of 1/256 for %ld of 65536 (a, b); the four weights sum to 1 for %ld\n", ints, sum256);
    for (int c = 0; c < 4; ++c) printf("corner %d: round-half-up(wa*wb/256) matches %ld, floor matches %ld of 65536\n", c, match[c], match_floor[c]);
    for (int b : {1, 17, 100, 128, 200}) for (int a : {1, 17, 100, 128, 200}) {
        printf("a=%3d b=%3d: hw", a, b);
        for (int c = 0; c < 4; ++c) printf(" %8.3f", W[c][b * 256 + a] * 256.0);
        printf("   exact");
        const int wa[4] = {256 - a, a, 256 - a, a}, wb[4] = {256 - b, 256 - b, b, b};
        for (int c = 0; c < 4; ++c) printf(" %8.3f", wa[c] * wb[c] / 256.0);
        printf("\n");
    }
    return 0;
}
