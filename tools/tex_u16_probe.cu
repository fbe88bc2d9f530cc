// tex_u16_probe.cu - stand-alone experiment (not part of the product library):
// can a pyramid level that is an exact 2^n down-scale of 8-bit images (cv::resize INTER_LINEAR at factor 1/2, 1/4: every
// texel is the mean of a 2 x 2 block = k/4, k = 0..1020) be sampled from an R16 unorm layered texture (texel = k) instead of R32F?
//  1. bilinear fetch rate R8 vs R32F for the three access shapes of tools/tex_probe.cu;
//  2. does  s = (t * 255 + 49152) - 49152  (round to a multiple of 1/256: the unit's weights are
//     multiples of 1/256, the texels integers) recover the R32F filter result bit for bit?
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o build/tex_u16_probe tools/tex_u16_probe.cu
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

template <class T>
Tex make_tex(const std::vector<T> &host, int w, int h, int layers, cudaChannelFormatDesc desc, cudaTextureReadMode rm) {
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
    td.readMode = rm;
    td.normalizedCoords = 0;
    CK(cudaCreateTextureObject(&t.tex, &res, &td, nullptr));
    return t;
}

template <bool U8>
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
        float t = tex2DLayered<float>(tex, x, y, layer);
        if (U8) t = __fadd_rn(__fmaf_rn(t, 16383.75f, 12288.0f), -12288.0f);
        acc += t;
    }
    if (acc == 123.456f) out[tid] = acc;   // keep the loop
}

__global__ void compare_kernel(cudaTextureObject_t a, cudaTextureObject_t b, int w, int h, int n,
                               unsigned long long *mismatch, float *maxdiff, float *max_raw_err) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= n) return;
    unsigned s = tid * 2654435761u + 777u;
    s = s * 1664525u + 1013904223u;
    const float x = (float)(s >> 8) * (1.0f / 16777216.0f) * (float)(w + 8) - 4.0f;
    s = s * 1664525u + 1013904223u;
    const float y = (float)(s >> 8) * (1.0f / 16777216.0f) * (float)(h + 8) - 4.0f;
    const float va = tex2DLayered<float>(a, x, y, 0);
    const float raw = tex2DLayered<float>(b, x, y, 0);
    const float vb = __fadd_rn(__fmaf_rn(raw, 16383.75f, 12288.0f), -12288.0f);
    // distance of the raw sample from the exact grid value, in units of 1/1024 grey level (the grid of quarter-integer texels under 1/256 weights)
    atomicMax((int *)max_raw_err, __float_as_int(fabsf(raw * 16383.75f * 1024.0f - va * 1024.0f)));
    if (__float_as_uint(va) != __float_as_uint(vb)) {
        atomicAdd(mismatch, 1ull);
        atomicMax((int *)maxdiff, __float_as_int(fabsf(va - vb)));
    }
}

template <bool U8> static double time_fetch(cudaTextureObject_t tex, int w, int h, int layers, int mode) {
    const int threads = 256, blocks = 148 * 8 * 4, iters = 2048;
    float *out;
    CK(cudaMalloc(&out, (size_t)threads * blocks * sizeof(float)));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    fetch_kernel<U8><<<blocks, threads>>>(tex, w, h, layers, 64, mode, out);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    fetch_kernel<U8><<<blocks, threads>>>(tex, w, h, layers, iters, mode, out);
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
    std::vector<unsigned short> img8(img.size());
    unsigned s = 1u;
    for (size_t i = 0; i < img.size(); ++i) {
        int k = 0;   // sum of a 2 x 2 block of 8-bit grey levels
        for (int j = 0; j < 4; ++j) {
            s = s * 1664525u + 1013904223u;
            k += (int)(s >> 24);
        }
        img[i] = (float)k * 0.25f;
        img8[i] = (unsigned short)k;
    }
    Tex t32 = make_tex(img, w, h, layers, cudaCreateChannelDesc(32, 0, 0, 0, cudaChannelFormatKindFloat), cudaReadModeElementType);
    Tex t8 = make_tex(img8, w, h, layers, cudaCreateChannelDesc(16, 0, 0, 0, cudaChannelFormatKindUnsigned), cudaReadModeNormalizedFloat);
    const char *names[3] = {"coherent rows", "8 hypotheses x 4 pixels", "scattered"};
    for (int mode = 0; mode < 3; ++mode) {
        printf("R32F %-26s %8.1f Gfetch/s\n", names[mode], time_fetch<false>(t32.tex, w, h, layers, mode));
        printf("R16  %-26s %8.1f Gfetch/s (with the 2-op rounding)\n", names[mode], time_fetch<true>(t8.tex, w, h, layers, mode));
    }
    unsigned long long *mm;
    float *md;
    CK(cudaMalloc(&mm, 8));
    CK(cudaMalloc(&md, 8));
    CK(cudaMemset(mm, 0, 8));
    CK(cudaMemset(md, 0, 8));
    const int n = 1 << 27;
    compare_kernel<<<(n + 255) / 256, 256>>>(t32.tex, t8.tex, w, h, n, mm, md, md + 1);
    CK(cudaDeviceSynchronize());
    unsigned long long hmm;
    float hmd[2];
    CK(cudaMemcpy(&hmm, mm, 8, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(hmd, md, 8, cudaMemcpyDeviceToHost));
    printf("R16 unorm (rounded to 1/1024) vs R32F on quarter-integer texels: %llu of %d samples differ (max |diff| %g); "
           "raw sample off the grid by at most %g / 1024 grey levels\n", hmm, n, hmd[0], hmd[1]);
    return 0;
}
