// tex_lds_probe.cu - stand-alone experiment (not part of the product library): the sweep kernel is bound by
// the L1TEX data stage, which serves texture wavefronts AND shared-memory wavefronts (ncu: 83 % + 17 %).
// Per texture fetch the NCC loop reads one float2 of per-pixel tap data that is identical for the 8 lanes of a
// pixel group.  Which way of getting it costs the texture rate least?
//   mode 0: no tap data (ceiling)
//   mode 1: LDS.128 per 2 fetches, same address in the 8 lanes of a group (what the kernel does)
//   mode 2: one LDS.64 per 6 fetches (lane b of the group loads tap b) + 2 SHFL per fetch
//   mode 3: LDS.64 per fetch, same address in the 8 lanes of a group
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o build/tex_lds_probe tools/tex_lds_probe.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s failed: %s (line %d)\n", #x, cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

template <int MODE>
__global__ void __launch_bounds__(256, 3) fetch_kernel(cudaTextureObject_t tex, int w, int h, int layers, int iters, float *out) {
    __shared__ __align__(16) float2 taps[32][40];   // per pixel group: 36 taps (+ pad: 320 B = 80 words -> 16 mod 32)
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = tid >> 5;
    const int g = threadIdx.x >> 3, gl = threadIdx.x & 7;
    for (int i = gl; i < 40; i += 8) taps[g][i] = make_float2(1.0f + i * 0.001f, 0.5f + g * 0.002f);
    __syncthreads();
    const float bx = (float)((warp * 8) % w + 2 * (lane >> 3)) + 0.03f * (float)(lane & 7) + 0.37f;
    const float by = (float)((warp * 5) % h) + 0.02f * (float)(lane & 7) + 0.61f;
    const int layer = warp % layers;
    const unsigned gmask = 0xffu << (lane & 24);
    const int gbase = lane & 24;
    float acc = 0.0f, acc2 = 0.0f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll 1
        for (int a = 0; a < 6; ++a) {
            float2 mine = make_float2(0.f, 0.f);
            if (MODE == 2) mine = taps[g][a * 6 + min(gl, 5)];
#pragma unroll
            for (int b = 0; b < 6; ++b) {
                const float x = bx + (float)(2 * a), y = by + (float)(2 * b) + (float)(i & 7);
                float t = tex2DLayered<float>(tex, x, y, layer);
                t = __fadd_rn(__fmaf_rn(t, 255.0f, 49152.0f), -49152.0f);
                float2 tp;
                if (MODE == 0) tp = make_float2(1.0f, 0.5f);
                if (MODE == 1) {   // LDS.128 for taps b, b+1
                    const float4 q = *reinterpret_cast<const float4 *>(&taps[g][a * 6 + (b & ~1)]);
                    tp = (b & 1) ? make_float2(q.z, q.w) : make_float2(q.x, q.y);
                }
                if (MODE == 2) tp = make_float2(__shfl_sync(gmask, mine.x, gbase + b), __shfl_sync(gmask, mine.y, gbase + b));
                if (MODE == 3) tp = taps[g][a * 6 + b];
                const float ws = t * tp.x;
                acc += ws;
                acc2 = fmaf(t, tp.y, acc2);
            }
        }
    }
    if (acc + acc2 == 123.456f) out[tid] = acc;
}

template <int MODE> static double time_fetch(cudaTextureObject_t tex, int w, int h, int layers) {
    const int threads = 256, blocks = 148 * 3 * 8, iters = 48;
    float *out;
    CK(cudaMalloc(&out, (size_t)threads * blocks * sizeof(float)));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    fetch_kernel<MODE><<<blocks, threads>>>(tex, w, h, layers, 2, out);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    fetch_kernel<MODE><<<blocks, threads>>>(tex, w, h, layers, iters, out);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    CK(cudaFree(out));
    return (double)threads * blocks * iters * 36 / (ms * 1e-3) / 1e9;
}

int main() {
    const int w = 1600, h = 1200, layers = 11;
    std::vector<unsigned char> img8((size_t)w * h * layers);
    unsigned s = 1u;
    for (auto &v : img8) { s = s * 1664525u + 1013904223u; v = (unsigned char)(s >> 24); }
    cudaArray_t arr;
    cudaChannelFormatDesc desc = cudaCreateChannelDesc(8, 0, 0, 0, cudaChannelFormatKindUnsigned);
    CK(cudaMalloc3DArray(&arr, &desc, make_cudaExtent(w, h, layers), cudaArrayLayered));
    cudaMemcpy3DParms cp = {};
    cp.srcPtr = make_cudaPitchedPtr(img8.data(), w, w, h);
    cp.dstArray = arr; cp.extent = make_cudaExtent(w, h, layers); cp.kind = cudaMemcpyHostToDevice;
    CK(cudaMemcpy3D(&cp));
    cudaResourceDesc res = {}; res.resType = cudaResourceTypeArray; res.res.array.array = arr;
    cudaTextureDesc td = {}; td.addressMode[0] = td.addressMode[1] = cudaAddressModeWrap; td.filterMode = cudaFilterModeLinear;
    td.readMode = cudaReadModeNormalizedFloat;
    cudaTextureObject_t tex;
    CK(cudaCreateTextureObject(&tex, &res, &td, nullptr));
    for (int rep = 0; rep < 2; ++rep) {
        printf("mode 0 no tap data                  %8.1f Gfetch/s\n", time_fetch<0>(tex, w, h, layers));
        printf("mode 1 LDS.128 / 2 fetches (group)  %8.1f Gfetch/s\n", time_fetch<1>(tex, w, h, layers));
        printf("mode 2 LDS.64 / 6 fetches + 2 SHFL  %8.1f Gfetch/s\n", time_fetch<2>(tex, w, h, layers));
        printf("mode 3 LDS.64 / fetch (group)       %8.1f Gfetch/s\n", time_fetch<3>(tex, w, h, layers));
    }
    return 0;
}
