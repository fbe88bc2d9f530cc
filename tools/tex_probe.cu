// tex_probe.cu - stand-alone experiment (not part of the product library):
//  1. peak bilinear fetch rate of a layered R32F texture (the sweep kernel's roofline),
//     for coherent and for scattered coordinates;
//  2. the same for an R16F copy of the image;
//  3. whether R16F filtering returns bit-identical values for integer-valued texels
//     (decoded 8-bit images), which would allow the narrower format at full resolution.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o build/tex_probe tools/tex_probe.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x)                                                                              \
    do {                                                                                   \
        cudaError_t e = (x);                                                               \
        if (e != cudaSuccess) {                                                            \
            printf("%s failed: %s (line %d)\n", #x, cudaGetErrorString(e), __LINE__);      \
            exit(1);                                                                       \
        }                                                                                  \
    } while (0)

struct Tex {
    cudaArray_t arr = nullptr;
    cudaTextureObject_t tex = 0;
};

template <class T> Tex make_tex(const std::vector<T> &host, int w, int h, int layers, cudaChannelFormatDesc desc) {
    Tex t;
    CK(cudaMalloc3DArray(&t.arr, &desc, make_cudaExtent(w, h, layers), cudaArrayLayered));
    cudaMemcpy3DParms cp = {};
    cp.srcPtr = make_cudaPitchedPtr((void *)host.data(), (size_t)w * sizeof(T), w, h);
    cp.dstArray = t.arr;
    cp.extent = make_cudaExtent(w, h, layers);
    cp.kind = cudaMemcpyHostToDevice;
    CK(cudaMemcpy3D(&cp));
    cudaResourceDesc res = {};
    res.resType = cudaResourceTypeArray;
    res.res.array.array = t.arr;
    cudaTextureDesc td = {};
    td.addressMode[0] = td.addressMode[1] = cudaAddressModeWrap;
    td.filterMode = cudaFilterModeLinear;
    td.readMode = cudaReadModeElementType;
    td.normalizedCoords = 0;
    CK(cudaCreateTextureObject(&t.tex, &res, &td, nullptr));
    return t;
}

// mode 0: a warp reads 32 horizontally adjacent pixels (+ fractional offset), moving down
// mode 1: lanes in groups of 8 read around 4 pixels, each lane offset by a few pixels
//         (like 8 plane hypotheses of one pixel), taps advancing by 2 pixels
// mode 2: every lane somewhere else (pseudo-random)
__global__ void fetch_kernel(cudaTextureObject_t tex, int w, int h, int layers, int iters, int mode, float *out) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const int warp = tid >> 5;
    float acc = 0.0f;
    unsigned s = tid * 2654435761u + 12345u;
    float bx, by;
    if (mode == 0) {
        bx = (float)((warp * 32) % w + lane) + 0.37f;
        by = (float)((warp * 7) % h) + 0.61f;
    } else if (mode == 1) {
        bx = (float)((warp * 8) % w + 2 * (lane >> 3)) + 1.7f * (float)(lane & 7) + 0.37f;
        by = (float)((warp * 5) % h) + 0.9f * (float)(lane & 7) + 0.61f;
    } else {
        bx = 0.37f;
        by = 0.61f;
    }
    const int layer = warp % layers;
    for (int i = 0; i < iters; ++i) {
        float x, y;
        if (mode == 2) {
            s = s * 1664525u + 1013904223u;
            x = bx + (float)(s >> 8 & 1023u);
            y = by + (float)(s >> 20 & 1023u);
        } else {
            x = bx + (float)(2 * (i % 6));
            y = by + (float)(2 * ((i / 6) % 6)) + (float)(i / 36);
        }
        acc += tex2DLayered<float>(tex, x, y, layer);
    }
    if (acc == 123.456f) out[tid] = acc;   // keep the loop
}

__global__ void compare_kernel(cudaTextureObject_t a, cudaTextureObject_t b, float scale_b, int w, int h, int n,
                               unsigned long long *mismatch, float *maxdiff) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= n) return;
    unsigned s = tid * 2654435761u + 777u;
    s = s * 1664525u + 1013904223u;
    const float x = (float)(s >> 8) * (1.0f / 16777216.0f) * (float)(w + 8) - 4.0f;
    s = s * 1664525u + 1013904223u;
    const float y = (float)(s >> 8) * (1.0f / 16777216.0f) * (float)(h + 8) - 4.0f;
    const float va = tex2DLayered<float>(a, x, y, 0);
    const float vb = tex2DLayered<float>(b, x, y, 0) * scale_b;
    if (__float_as_uint(va) != __float_as_uint(vb)) {
        atomicAdd(mismatch, 1ull);
        atomicMax((int *)maxdiff, __float_as_int(fabsf(va - vb)));
    }
}

static double time_fetch(cudaTextureObject_t tex, int w, int h, int layers, int mode) {
    const int threads = 256, blocks = 148 * 8 * 4, iters = 2048;
    float *out;
    CK(cudaMalloc(&out, (size_t)threads * blocks * sizeof(float)));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    fetch_kernel<<<blocks, threads>>>(tex, w, h, layers, 64, mode, out);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    fetch_kernel<<<blocks, threads>>>(tex, w, h, layers, iters, mode, out);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    CK(cudaFree(out));
    return (double)threads * blocks * iters / (ms * 1e-3) / 1e9;   // G fetches / s
}

int main() {
    const int w = 1600, h = 1200, layers = 11;
    std::vector<float> img((size_t)w * h * layers);
    std::vector<__half> img16(img.size());
    std::vector<unsigned char> img8(img.size());
    unsigned s = 1u;
    for (size_t i = 0; i < img.size(); ++i) {
        s = s * 1664525u + 1013904223u;
        const int v = (int)(s >> 24);   // integer grey level 0..255
        img[i] = (float)v;
        img16[i] = __float2half((float)v);
        img8[i] = (unsigned char)v;
    }
    Tex t32 = make_tex(img, w, h, layers, cudaCreateChannelDesc(32, 0, 0, 0, cudaChannelFormatKindFloat));
    Tex t16 = make_tex(img16, w, h, layers, cudaCreateChannelDescHalf());
    const char *names[3] = {"coherent rows", "8 hypotheses x 4 pixels", "scattered"};
    for (int mode = 0; mode < 3; ++mode) {
        printf("R32F %-26s %8.1f Gfetch/s\n", names[mode], time_fetch(t32.tex, w, h, layers, mode));
        printf("R16F %-26s %8.1f Gfetch/s\n", names[mode], time_fetch(t16.tex, w, h, layers, mode));
    }
    unsigned long long *mm;
    float *md;
    CK(cudaMalloc(&mm, 8));
    CK(cudaMalloc(&md, 4));
    CK(cudaMemset(mm, 0, 8));
    CK(cudaMemset(md, 0, 4));
    const int n = 1 << 26;
    compare_kernel<<<(n + 255) / 256, 256>>>(t32.tex, t16.tex, 1.0f, w, h, n, mm, md);
    CK(cudaDeviceSynchronize());
    unsigned long long hmm;
    float hmd;
    CK(cudaMemcpy(&hmm, mm, 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(&hmd, md, 4, cudaMemcpyDeviceToHost));
    printf("R16F vs R32F on integer texels: %llu of %d samples differ (max |diff| %g)\n", hmm, n, hmd);
    int dev_clock = 0;
    CK(cudaDeviceGetAttribute(&dev_clock, cudaDevAttrClockRate, 0));
    printf("SM clock (attr) %d kHz -> 4 fetch/clk/SM x 148 = %.1f Gfetch/s\n", dev_clock, 4.0 * 148 * dev_clock * 1e-6);
    return 0;
}
