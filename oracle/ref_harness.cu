// ref_harness.cu - parity comparator: the REFERENCE's own CUDA kernels, compiled
// from where they lie (/root/reference/src/ACMMP.cu is #included below, never
// copied into this repository) for sm_100 with the reference's flags, behind a
// small C ABI so that tests/ and bench.py --impl reference can run them on the
// same inputs as the product library.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under acmmp_b200/ links, loads or calls
// this file or its output (oracle/_ref/libacmmp_ref.so).
//
// Differences to running the reference binary, all outside the kernels:
//  P1  `clock64()` in RandomInitialization (ACMMP.cu:624) is macro-replaced by a
//      caller-provided seed so runs are repeatable;
//  P2  the harness launches the reference's __global__ kernels itself in the
//      order of ACMMP::RunPatchMatch (ACMMP.cu:1415-1448) instead of calling
//      that member, which (because of the hard-coded `#define DEBUG`,
//      ACMMP.cu:3) would copy the whole plane map to the host and open
//      cv::imshow windows twice per run;
//  *   OpenCV C++ is not installed: oracle/ref_stub/opencv2 lets the headers
//      parse; none of it executes.
#include <cuda_runtime.h>
#include <curand_kernel.h>

__device__ unsigned long long g_acmmp_ref_seed;
#define clock64() (g_acmmp_ref_seed)
#include "ACMMP.cu"
#undef clock64

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "acmmp_b200.h"   // POD types shared with the product ABI (camera, params)

// declared in the reference's ACMMP.h, defined in ACMMP.cpp (not compiled here)
void CudaSafeCall(const cudaError_t error, const std::string &file, const int line) {
    if (error != cudaSuccess) fprintf(stderr, "ref cuda error %s:%d: %s\n", file.c_str(), line, cudaGetErrorString(error));
}
void CudaCheckError(const char *, const int) {}

static_assert(sizeof(Camera) == sizeof(acmmp_camera), "Camera layout");

struct ref_ctx {
    int n = 0, W = 0, H = 0;
    std::vector<Camera> cams;
    std::vector<cudaArray_t> arr, darr;
    cudaTextureObjects tex_h, dtex_h;
    cudaTextureObjects *tex_d = nullptr, *dtex_d = nullptr;
    Camera *cams_d = nullptr;
    float4 *planes = nullptr, *scaled = nullptr, *prior = nullptr, *seeded = nullptr;
    float *costs = nullptr, *pre_costs = nullptr;
    curandState *rng = nullptr;
    unsigned int *sel = nullptr, *masks = nullptr;
    size_t scaled_n = 0;
    PatchMatchParams params;
    std::string error;
    cudaEvent_t e0, e1;
    float last_ms = 0;
};

static int rfail(ref_ctx *c, const char *msg, cudaError_t e = cudaSuccess) {
    if (c) c->error = std::string(msg) + (e != cudaSuccess ? std::string(": ") + cudaGetErrorString(e) : "");
    return -1;
}
#define RCU(c, call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return rfail(c, #call, e__); } while (0)

static int make_tex(ref_ctx *c, const float *img, int w, int h, cudaArray_t *arr, cudaTextureObject_t *tex) {
    // same descriptor as ACMMP.cpp:646-663
    cudaChannelFormatDesc channelDesc = cudaCreateChannelDesc(32, 0, 0, 0, cudaChannelFormatKindFloat);
    RCU(c, cudaMallocArray(arr, &channelDesc, w, h));
    RCU(c, cudaMemcpy2DToArray(*arr, 0, 0, img, (size_t)w * sizeof(float), (size_t)w * sizeof(float), h, cudaMemcpyHostToDevice));
    cudaResourceDesc resDesc;
    memset(&resDesc, 0, sizeof resDesc);
    resDesc.resType = cudaResourceTypeArray;
    resDesc.res.array.array = *arr;
    cudaTextureDesc texDesc;
    memset(&texDesc, 0, sizeof texDesc);
    texDesc.addressMode[0] = cudaAddressModeWrap;
    texDesc.addressMode[1] = cudaAddressModeWrap;
    texDesc.filterMode = cudaFilterModeLinear;
    texDesc.readMode = cudaReadModeElementType;
    texDesc.normalizedCoords = 0;
    RCU(c, cudaCreateTextureObject(tex, &resDesc, &texDesc, NULL));
    return 0;
}

// unit-level probes: call the reference's __device__ functions directly
__global__ void ref_ncc_probe(cudaTextureObjects *tex, Camera *cams, const float4 *planes, int view, float *out,
                              PatchMatchParams params) {
    const int2 p = make_int2(blockIdx.x * blockDim.x + threadIdx.x, blockIdx.y * blockDim.y + threadIdx.y);
    const int width = cams[0].width, height = cams[0].height;
    if (p.x >= width || p.y >= height) return;
    const int center = p.y * width + p.x;
    out[center] = ComputeBilateralNCC(tex[0].images[0], cams[0], tex[0].images[view], cams[view], p, planes[center], params);
}
__global__ void ref_homography_probe(Camera *cams, const float4 *planes, int view, int count, float *out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    float H[9];
    ComputeHomography(cams[0], cams[view], planes[i], H);
    for (int k = 0; k < 9; ++k) out[i * 9 + k] = H[k];
}
__global__ void ref_geom_probe(cudaTextureObjects *dtex, Camera *cams, const float4 *planes, int view, float *out) {
    const int2 p = make_int2(blockIdx.x * blockDim.x + threadIdx.x, blockIdx.y * blockDim.y + threadIdx.y);
    const int width = cams[0].width, height = cams[0].height;
    if (p.x >= width || p.y >= height) return;
    const int center = p.y * width + p.x;
    out[center] = ComputeGeomConsistencyCost(dtex[0].images[view], cams[0], cams[view], planes[center], p);
}

extern "C" {

int ref_create(ref_ctx **out, int device) {
    if (cudaSetDevice(device) != cudaSuccess) return -1;
    ref_ctx *c = new ref_ctx();
    memset(&c->tex_h, 0, sizeof c->tex_h);
    memset(&c->dtex_h, 0, sizeof c->dtex_h);
    cudaEventCreate(&c->e0);
    cudaEventCreate(&c->e1);
    *out = c;
    return 0;
}

const char *ref_last_error(const ref_ctx *c) { return c ? c->error.c_str() : ""; }

static void ref_free_problem(ref_ctx *c) {
    for (int i = 0; i < (int)c->arr.size(); ++i) {
        if (c->tex_h.images[i]) cudaDestroyTextureObject(c->tex_h.images[i]);
        if (c->arr[i]) cudaFreeArray(c->arr[i]);
        c->tex_h.images[i] = 0;
    }
    for (int i = 0; i < (int)c->darr.size(); ++i) {
        if (c->dtex_h.images[i]) cudaDestroyTextureObject(c->dtex_h.images[i]);
        if (c->darr[i]) cudaFreeArray(c->darr[i]);
        c->dtex_h.images[i] = 0;
    }
    c->arr.clear();
    c->darr.clear();
    void *ptrs[] = {c->tex_d, c->dtex_d, c->cams_d, c->planes, c->scaled, c->prior, c->seeded, c->costs, c->pre_costs, c->rng, c->sel, c->masks};
    for (void *p : ptrs) if (p) cudaFree(p);
    c->tex_d = c->dtex_d = nullptr; c->cams_d = nullptr; c->planes = c->scaled = c->prior = c->seeded = nullptr;
    c->costs = c->pre_costs = nullptr; c->rng = nullptr; c->sel = c->masks = nullptr;
}

void ref_destroy(ref_ctx *c) {
    if (!c) return;
    ref_free_problem(c);
    cudaEventDestroy(c->e0);
    cudaEventDestroy(c->e1);
    delete c;
}

// = InputInitialization (upload half) + CudaSpaceInitialization (ACMMP.cpp:600-681)
int ref_set_problem(ref_ctx *c, int n, const acmmp_camera *cams, const float *const *images) {
    ref_free_problem(c);
    c->n = n;
    c->cams.resize(n);
    memcpy(c->cams.data(), cams, sizeof(Camera) * n);
    c->W = cams[0].width;
    c->H = cams[0].height;
    c->arr.assign(n, nullptr);
    for (int i = 0; i < n; ++i)
        if (make_tex(c, images[i], cams[i].width, cams[i].height, &c->arr[i], &c->tex_h.images[i])) return -1;
    RCU(c, cudaMalloc(&c->tex_d, sizeof(cudaTextureObjects)));
    RCU(c, cudaMemcpy(c->tex_d, &c->tex_h, sizeof(cudaTextureObjects), cudaMemcpyHostToDevice));
    RCU(c, cudaMalloc(&c->dtex_d, sizeof(cudaTextureObjects)));
    RCU(c, cudaMemcpy(c->dtex_d, &c->dtex_h, sizeof(cudaTextureObjects), cudaMemcpyHostToDevice));
    RCU(c, cudaMalloc(&c->cams_d, sizeof(Camera) * n));
    RCU(c, cudaMemcpy(c->cams_d, c->cams.data(), sizeof(Camera) * n, cudaMemcpyHostToDevice));
    const size_t P = (size_t)c->W * c->H;
    RCU(c, cudaMalloc(&c->planes, P * sizeof(float4)));
    RCU(c, cudaMalloc(&c->prior, P * sizeof(float4)));
    RCU(c, cudaMalloc(&c->seeded, P * sizeof(float4)));
    RCU(c, cudaMalloc(&c->costs, P * sizeof(float)));
    RCU(c, cudaMalloc(&c->pre_costs, P * sizeof(float)));
    RCU(c, cudaMalloc(&c->rng, P * sizeof(curandState)));
    RCU(c, cudaMalloc(&c->sel, P * sizeof(unsigned int)));
    RCU(c, cudaMalloc(&c->masks, P * sizeof(unsigned int)));
    RCU(c, cudaMemset(c->planes, 0, P * sizeof(float4)));
    RCU(c, cudaMemset(c->prior, 0, P * sizeof(float4)));
    RCU(c, cudaMemset(c->masks, 0, P * sizeof(unsigned int)));
    RCU(c, cudaMemset(c->rng, 0, P * sizeof(curandState)));
    c->params = PatchMatchParams();
    c->params.depth_min = cams[0].depth_min * 0.6f;
    c->params.depth_max = cams[0].depth_max * 1.2f;
    c->params.num_images = n;
    c->params.disparity_min = cams[0].K[0] * c->params.baseline / c->params.depth_max;
    c->params.disparity_max = cams[0].K[0] * c->params.baseline / c->params.depth_min;
    return 0;
}

int ref_set_flags(ref_ctx *c, int geom, int planar_prior, int multi_geometry, int hierarchy, int upsample, int seeded,
                  float scaled_cols, float scaled_rows) {
    c->params.geom_consistency = geom != 0;
    c->params.planar_prior = planar_prior != 0;
    c->params.multi_geometry = multi_geometry != 0;
    c->params.hierarchy = hierarchy != 0;
    c->params.upsample = upsample != 0;
    c->params.seeded = seeded != 0;
    c->params.scaled_cols = scaled_cols;
    c->params.scaled_rows = scaled_rows;
    return 0;
}

int ref_set_depth_map(ref_ctx *c, int slot, const float *depth, int w, int h) {
    if ((int)c->darr.size() < c->n) c->darr.resize(c->n, nullptr);
    if (c->dtex_h.images[slot]) { cudaDestroyTextureObject(c->dtex_h.images[slot]); cudaFreeArray(c->darr[slot]); }
    if (make_tex(c, depth, w, h, &c->darr[slot], &c->dtex_h.images[slot])) return -1;
    RCU(c, cudaMemcpy(c->dtex_d, &c->dtex_h, sizeof(cudaTextureObjects), cudaMemcpyHostToDevice));
    return 0;
}

int ref_set_planes(ref_ctx *c, int which, const float *planes4, int count) {
    // which: 0 planes, 1 prior, 2 seeded, 3 scaled (count = coarse pixels)
    const size_t P = (size_t)c->W * c->H;
    float4 *dst = which == 0 ? c->planes : which == 1 ? c->prior : which == 2 ? c->seeded : nullptr;
    if (which == 3) {
        if (c->scaled) cudaFree(c->scaled);
        RCU(c, cudaMalloc(&c->scaled, (size_t)count * sizeof(float4)));
        dst = c->scaled;
    } else if ((size_t)count != P) return rfail(c, "ref_set_planes: wrong count");
    RCU(c, cudaMemcpy(dst, planes4, (size_t)count * sizeof(float4), cudaMemcpyHostToDevice));
    return 0;
}

int ref_set_scalars(ref_ctx *c, int which, const void *data) {
    // which: 0 costs(f32) 1 pre_costs(f32) 2 selected_views(u32) 3 plane_masks(u32)
    const size_t P = (size_t)c->W * c->H;
    void *dst = which == 0 ? (void *)c->costs : which == 1 ? (void *)c->pre_costs : which == 2 ? (void *)c->sel : (void *)c->masks;
    RCU(c, cudaMemcpy(dst, data, P * 4, cudaMemcpyHostToDevice));
    return 0;
}

int ref_set_rng(ref_ctx *c, const uint32_t *rng6) {
    const size_t P = (size_t)c->W * c->H;
    std::vector<curandState> st(P);
    memset(st.data(), 0, P * sizeof(curandState));
    for (size_t i = 0; i < P; ++i) {
        st[i].d = rng6[i * 6];
        for (int k = 0; k < 5; ++k) st[i].v[k] = rng6[i * 6 + 1 + k];
    }
    RCU(c, cudaMemcpy(c->rng, st.data(), P * sizeof(curandState), cudaMemcpyHostToDevice));
    return 0;
}

int ref_get_state(ref_ctx *c, float *planes4, float *costs, uint32_t *sel, uint32_t *rng6, float *pre_costs) {
    const size_t P = (size_t)c->W * c->H;
    RCU(c, cudaDeviceSynchronize());
    if (planes4) RCU(c, cudaMemcpy(planes4, c->planes, P * sizeof(float4), cudaMemcpyDeviceToHost));
    if (costs) RCU(c, cudaMemcpy(costs, c->costs, P * sizeof(float), cudaMemcpyDeviceToHost));
    if (sel) RCU(c, cudaMemcpy(sel, c->sel, P * sizeof(unsigned int), cudaMemcpyDeviceToHost));
    if (pre_costs) RCU(c, cudaMemcpy(pre_costs, c->pre_costs, P * sizeof(float), cudaMemcpyDeviceToHost));
    if (rng6) {
        std::vector<curandState> st(P);
        RCU(c, cudaMemcpy(st.data(), c->rng, P * sizeof(curandState), cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < P; ++i) {
            rng6[i * 6] = st[i].d;
            for (int k = 0; k < 5; ++k) rng6[i * 6 + 1 + k] = st[i].v[k];
        }
    }
    return 0;
}

static void ref_grids(const ref_ctx *c, dim3 &g_init, dim3 &b_init, dim3 &g_cb, dim3 &b_cb) {
    // ACMMP.cu:1385-1404
    const int width = c->W, height = c->H;
    const int BLOCK_W = 32, BLOCK_H = BLOCK_W / 2;
    g_init = dim3((width + 16 - 1) / 16, (height + 16 - 1) / 16, 1);
    b_init = dim3(16, 16, 1);
    g_cb = dim3((width + BLOCK_W - 1) / BLOCK_W, ((height / 2) + BLOCK_H - 1) / BLOCK_H, 1);
    b_cb = dim3(BLOCK_W, BLOCK_H, 1);
}

int ref_run_init(ref_ctx *c, unsigned long long seed) {
    dim3 gi, bi, gc, bc;
    ref_grids(c, gi, bi, gc, bc);
    RCU(c, cudaMemcpyToSymbol(g_acmmp_ref_seed, &seed, sizeof seed));
    RandomInitialization<<<gi, bi>>>(c->tex_d, c->cams_d, c->planes, c->scaled, c->costs, c->pre_costs, c->rng, c->sel,
                                    c->prior, c->masks, c->seeded, c->params);
    RCU(c, cudaGetLastError());
    return 0;
}

int ref_run_sweep(ref_ctx *c, int iter, int color) {
    dim3 gi, bi, gc, bc;
    ref_grids(c, gi, bi, gc, bc);
    if (color == 0)
        BlackPixelUpdate<<<gc, bc>>>(c->tex_d, c->dtex_d, c->cams_d, c->planes, c->costs, c->pre_costs, c->rng, c->sel, c->prior, c->masks, c->params, iter);
    else
        RedPixelUpdate<<<gc, bc>>>(c->tex_d, c->dtex_d, c->cams_d, c->planes, c->costs, c->pre_costs, c->rng, c->sel, c->prior, c->masks, c->params, iter);
    RCU(c, cudaGetLastError());
    return 0;
}

int ref_run_tail(ref_ctx *c) {
    dim3 gi, bi, gc, bc;
    ref_grids(c, gi, bi, gc, bc);
    GetDepthandNormal<<<gi, bi>>>(c->cams_d, c->planes, c->params);
    BlackPixelFilter<<<gc, bc>>>(c->cams_d, c->planes, c->costs);
    RedPixelFilter<<<gc, bc>>>(c->cams_d, c->planes, c->costs);
    RCU(c, cudaGetLastError());
    return 0;
}

// kernel sequence of ACMMP::RunPatchMatch, with the reference's device-wide
// synchronisation after every launch (ACMMP.cu:1421-1448); timed with events.
int ref_run(ref_ctx *c, unsigned long long seed) {
    cudaEventRecord(c->e0);
    if (ref_run_init(c, seed)) return -1;
    RCU(c, cudaDeviceSynchronize());
    for (int i = 0; i < c->params.max_iterations; ++i) {
        if (ref_run_sweep(c, i, 0)) return -1;
        RCU(c, cudaDeviceSynchronize());
        if (ref_run_sweep(c, i, 1)) return -1;
        RCU(c, cudaDeviceSynchronize());
    }
    if (ref_run_tail(c)) return -1;
    RCU(c, cudaDeviceSynchronize());
    cudaEventRecord(c->e1);
    RCU(c, cudaEventSynchronize(c->e1));
    cudaEventElapsedTime(&c->last_ms, c->e0, c->e1);
    return 0;
}

float ref_last_ms(const ref_ctx *c) { return c->last_ms; }

int ref_ncc(ref_ctx *c, const float *planes4, int view, float *out) {
    const size_t P = (size_t)c->W * c->H;
    float4 *dp = nullptr;
    float *dout = nullptr;
    RCU(c, cudaMalloc(&dp, P * sizeof(float4)));
    RCU(c, cudaMalloc(&dout, P * sizeof(float)));
    RCU(c, cudaMemcpy(dp, planes4, P * sizeof(float4), cudaMemcpyHostToDevice));
    dim3 b(16, 16), g((c->W + 15) / 16, (c->H + 15) / 16);
    ref_ncc_probe<<<g, b>>>(c->tex_d, c->cams_d, dp, view, dout, c->params);
    RCU(c, cudaDeviceSynchronize());
    RCU(c, cudaMemcpy(out, dout, P * sizeof(float), cudaMemcpyDeviceToHost));
    cudaFree(dp);
    cudaFree(dout);
    return 0;
}

int ref_homography(ref_ctx *c, const float *planes4, int count, int view, float *out9) {
    float4 *dp = nullptr;
    float *dout = nullptr;
    RCU(c, cudaMalloc(&dp, (size_t)count * sizeof(float4)));
    RCU(c, cudaMalloc(&dout, (size_t)count * 9 * sizeof(float)));
    RCU(c, cudaMemcpy(dp, planes4, (size_t)count * sizeof(float4), cudaMemcpyHostToDevice));
    ref_homography_probe<<<(count + 127) / 128, 128>>>(c->cams_d, dp, view, count, dout);
    RCU(c, cudaDeviceSynchronize());
    RCU(c, cudaMemcpy(out9, dout, (size_t)count * 9 * sizeof(float), cudaMemcpyDeviceToHost));
    cudaFree(dp);
    cudaFree(dout);
    return 0;
}

int ref_geom_cost(ref_ctx *c, const float *planes4, int view, float *out) {
    const size_t P = (size_t)c->W * c->H;
    float4 *dp = nullptr;
    float *dout = nullptr;
    RCU(c, cudaMalloc(&dp, P * sizeof(float4)));
    RCU(c, cudaMalloc(&dout, P * sizeof(float)));
    RCU(c, cudaMemcpy(dp, planes4, P * sizeof(float4), cudaMemcpyHostToDevice));
    dim3 b(16, 16), g((c->W + 15) / 16, (c->H + 15) / 16);
    ref_geom_probe<<<g, b>>>(c->dtex_d, c->cams_d, dp, view, dout);
    RCU(c, cudaDeviceSynchronize());
    RCU(c, cudaMemcpy(out, dout, P * sizeof(float), cudaMemcpyDeviceToHost));
    cudaFree(dp);
    cudaFree(dout);
    return 0;
}

// JBU depth upsample with the reference's kernel (RunJBU, ACMMP.cpp:1022-1063)
int ref_jbu(const float *image, int w, int h, const float *depth, int sw, int sh, float *out) {
    int Imagescale = std::max(h / sh, w / sw);
    if (Imagescale == 1) return 1;
    ref_ctx tmp;
    cudaArray_t arr[2];
    JBUTexObj jt_h;
    if (make_tex(&tmp, image, w, h, &arr[0], &jt_h.imgs[0])) return -1;
    if (make_tex(&tmp, depth, sw, sh, &arr[1], &jt_h.imgs[1])) return -1;
    JBUParameters jp_h;
    jp_h.height = h; jp_h.width = w; jp_h.s_height = sh; jp_h.s_width = sw; jp_h.Imagescale = Imagescale;
    JBUParameters *jp_d; JBUTexObj *jt_d; float *depth_d;
    cudaMalloc(&jp_d, sizeof jp_h); cudaMemcpy(jp_d, &jp_h, sizeof jp_h, cudaMemcpyHostToDevice);
    cudaMalloc(&jt_d, sizeof jt_h); cudaMemcpy(jt_d, &jt_h, sizeof jt_h, cudaMemcpyHostToDevice);
    cudaMalloc(&depth_d, (size_t)w * h * sizeof(float));
    dim3 b(16, 16), g((w + 15) / 16, (h + 15) / 16);
    JBU_cu<<<g, b>>>(jp_d, jt_d, depth_d);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(out, depth_d, (size_t)w * h * sizeof(float), cudaMemcpyDeviceToHost);
    cudaFree(jp_d); cudaFree(jt_d); cudaFree(depth_d);
    for (int i = 0; i < 2; ++i) { cudaDestroyTextureObject(jt_h.imgs[i]); cudaFreeArray(arr[i]); }
    return e == cudaSuccess ? 0 : -1;
}

}  // extern "C"
