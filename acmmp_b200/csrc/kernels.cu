// kernels.cu - every device kernel of the ACMMP hot path + its launch wrapper.
// Compiled for sm_100a only, with the reference's arithmetic flags
// (-O3 --use_fast_math): see common.cuh for why.
#include <curand_kernel.h>

#include <cstring>

#include "common.cuh"
#include "geometry.cuh"
#include "kernels.h"
#include "ncc.cuh"
#include "sweep_kernel.cuh"

#include "init_kernel.cu.inc"
#include "sweep_kernel.cu.inc"
#include "prior_kernels.cu.inc"

namespace acmmp {

// ---------------------------------------------------------------------------
// camera-pair constants (the hypothesis-independent part of ComputeHomography,
// ACMMP.cu:264-290, plus what ComputeGeomConsistencyCost needs).  One thread
// per source view.  Operation order = the reference build's (each 3-term dot
// product is mul, fma, fma with the MIDDLE term's product first).
// ---------------------------------------------------------------------------
ACMMP_DEV float dot3_ref_order(float a0, float b0, float a1, float b1, float a2, float b2) {
    // a0*b0 + a1*b1 + a2*b2 as nvcc contracts it: fma(a2,b2, fma(a0,b0, a1*b1))
    return ffma(a2, b2, ffma(a0, b0, fmul(a1, b1)));
}

// cameras and depth-map pointers travel as a kernel argument (<= 33 x 108 bytes):
// no staging buffer, no host synchronisation per problem
struct SetupArgs {
    acmmp_camera cams[ACMMP_MAX_SRC_VIEWS + 1];
    const float *depth[ACMMP_MAX_SRC_VIEWS + 1];
    int layer[ACMMP_MAX_SRC_VIEWS + 1];   // layer of each view in the layered texture
};

__global__ void setup_pairs_kernel(const __grid_constant__ SetupArgs args, int n, int have_depth, float depth_min, float depth_max,
                                   RefConst *rc_out, PairConst *pairs_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const acmmp_camera *cams = args.cams;
    const float *const *depth_maps = have_depth ? args.depth : nullptr;
    const acmmp_camera &r = cams[0];
    if (i == 0) {
        RefConst rc;
        rc.K0 = r.K[0]; rc.K2 = r.K[2]; rc.K4 = r.K[4]; rc.K5 = r.K[5];
        rc.rK0 = mufu_rcp(r.K[0]);
        rc.rK4 = mufu_rcp(r.K[4]);
        rc.K0_over_K4 = fmul(r.K[0], rc.rK4);
        for (int k = 0; k < 9; ++k) { rc.R[k] = r.R[k]; rc.K[k] = r.K[k]; }
        for (int k = 0; k < 3; ++k) rc.t[k] = r.t[k];
        rc.C[0] = -dot3_ref_order(r.R[0], r.t[0], r.R[3], r.t[1], r.R[6], r.t[2]);
        rc.C[1] = -dot3_ref_order(r.R[1], r.t[0], r.R[4], r.t[1], r.R[7], r.t[2]);
        rc.C[2] = -dot3_ref_order(r.R[2], r.t[0], r.R[5], r.t[1], r.R[8], r.t[2]);
        rc.width = r.width; rc.height = r.height;
        rc.depth_min = depth_min; rc.depth_max = depth_max;
        *rc_out = rc;
    }
    if (i >= n - 1) return;
    const acmmp_camera &s = cams[i + 1];
    PairConst pc;
    // -C (un-negated camera centres, the negations cancel in C_ref - C_src)
    const float nrc0 = dot3_ref_order(r.R[0], r.t[0], r.R[3], r.t[1], r.R[6], r.t[2]);
    const float nrc1 = dot3_ref_order(r.R[1], r.t[0], r.R[4], r.t[1], r.R[7], r.t[2]);
    const float nrc2 = dot3_ref_order(r.R[2], r.t[0], r.R[5], r.t[1], r.R[8], r.t[2]);
    const float nsc0 = dot3_ref_order(s.R[0], s.t[0], s.R[3], s.t[1], s.R[6], s.t[2]);
    const float nsc1 = dot3_ref_order(s.R[1], s.t[0], s.R[4], s.t[1], s.R[7], s.t[2]);
    const float nsc2 = dot3_ref_order(s.R[2], s.t[0], s.R[5], s.t[1], s.R[8], s.t[2]);
    for (int a = 0; a < 3; ++a)
        for (int b = 0; b < 3; ++b)
            pc.Rrel[3 * a + b] = dot3_ref_order(s.R[3 * a + 0], r.R[3 * b + 0], s.R[3 * a + 1], r.R[3 * b + 1],
                                                s.R[3 * a + 2], r.R[3 * b + 2]);
    const float c0 = fsub(nsc0, nrc0), c1 = fsub(nsc1, nrc1), c2 = fsub(nsc2, nrc2);
    for (int a = 0; a < 3; ++a)
        pc.trel[a] = dot3_ref_order(s.R[3 * a + 0], c0, s.R[3 * a + 1], c1, s.R[3 * a + 2], c2);
    pc.K0s = s.K[0]; pc.K2s = s.K[2]; pc.K4s = s.K[4]; pc.K5s = s.K[5]; pc.K8s = s.K[8];
    pc.ws = (float)s.width; pc.hs = (float)s.height; pc.pad = 0.0f;
    pc.layer = args.layer[i + 1]; pc.pad2 = 0;
    for (int k = 0; k < 9; ++k) { pc.Rs[k] = s.R[k]; pc.Ks[k] = s.K[k]; }
    for (int k = 0; k < 3; ++k) pc.ts[k] = s.t[k];
    pc.Cs[0] = -nsc0; pc.Cs[1] = -nsc1; pc.Cs[2] = -nsc2;
    pc.depth = depth_maps ? depth_maps[i + 1] : nullptr;
    pc.dw = s.width; pc.dh = s.height;
    pairs_out[i] = pc;
}

// ---------------------------------------------------------------------------
// XORWOW state of curand_init(seed, subsequence = y, offset = x) for all pixels
// (ACMMP.cu:624).  One thread per (row, 32-pixel segment): the cuRAND device
// API performs the subsequence/offset skip-ahead once per segment (the
// reference pays it once per pixel), then the generator is stepped.
// ---------------------------------------------------------------------------
__global__ void rng_init_kernel(unsigned long long seed, int W, int H, RngPlanes out) {
    const int segs = (W + 31) / 32;
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= segs * H) return;
    const int y = t / segs, x0 = (t % segs) * 32;
    curandStateXORWOW_t st;
    curand_init(seed, (unsigned long long)y, (unsigned long long)x0, &st);
    Rng s;
    s.d = st.d; s.v0 = st.v[0]; s.v1 = st.v[1]; s.v2 = st.v[2]; s.v3 = st.v[3]; s.v4 = st.v[4];
    const int x1 = min(W, x0 + 32);
    for (int x = x0; x < x1; ++x) {
        out.store(y * W + x, s);
        rng_next(s);
    }
}

// ---------------------------------------------------------------------------
// tail: plane -> (world normal, depth); checkerboard median filter
// (GetDepthandNormal ACMMP.cu:1199-1212, CheckerboardFilter :1214-1328)
// ---------------------------------------------------------------------------
__global__ void depth_normal_kernel(float4 *planes, const RefConst *rcp) {
    const RefConst &rc = *rcp;
    const int px = blockIdx.x * blockDim.x + threadIdx.x;
    const int py = blockIdx.y * blockDim.y + threadIdx.y;
    if (px >= rc.width || py >= rc.height) return;
    const int c = py * rc.width + px;
    Plane p = make_plane(planes[c]);
    p.w = plane_depth(rc, p, px, py);
    p = normal_to_world(rc, p);
    planes[c] = to_float4(p);
}

__global__ void median_filter_kernel(float4 *planes, const float *costs, int W, int H, int color) {
    // thread (tx, ty) -> pixel of the requested colour: x = tx, y = 2*ty + parity
    const int px = blockIdx.x * blockDim.x + threadIdx.x;
    int py = (blockIdx.y * blockDim.y + threadIdx.y) * 2;
    py += ((px & 1) ^ color) ? 1 : 0;   // black: x,y same parity
    if (px >= W || py >= H) return;
    const int c = py * W + px;
    if (costs[c] < 0.001f) return;
    float f[21];
    int n = 0;
#define ACMMP_TAP(cond, dx, dy) if (cond) f[n++] = planes[c + (dy) * W + (dx)].w;
    f[n++] = planes[c].w;
    ACMMP_TAP(py > 0, 0, -1)
    ACMMP_TAP(py > 2, 0, -3)
    ACMMP_TAP(py > 4, 0, -5)
    ACMMP_TAP(py < H - 1, 0, 1)
    ACMMP_TAP(py < H - 3, 0, 3)
    ACMMP_TAP(py < H - 5, 0, 5)
    ACMMP_TAP(px > 0, -1, 0)
    ACMMP_TAP(px > 2, -3, 0)
    ACMMP_TAP(px > 4, -5, 0)
    ACMMP_TAP(px < W - 1, 1, 0)
    ACMMP_TAP(px < W - 3, 3, 0)
    ACMMP_TAP(px < W - 5, 5, 0)
    ACMMP_TAP(py > 0 && px < W - 2, 2, -1)
    ACMMP_TAP(py < H - 1 && px < W - 2, 2, 1)
    ACMMP_TAP(py > 0 && px > 1, -2, -1)
    ACMMP_TAP(py < H - 1 && px > 1, -2, 1)
    ACMMP_TAP(px > 0 && py > 2, -1, -2)
    ACMMP_TAP(px < W - 1 && py > 2, 1, -2)
    ACMMP_TAP(px > 0 && py < H - 2, -1, 2)
    ACMMP_TAP(px < W - 1 && py < H - 2, 1, 2)
#undef ACMMP_TAP
    for (int i = 1; i < n; i++) {
        float tmp = f[i];
        int j;
        for (j = i; j >= 1 && tmp < f[j - 1]; j--) f[j] = f[j - 1];
        f[j] = tmp;
    }
    const int m = n / 2;
    planes[c].w = (n % 2 == 0) ? (f[m - 1] + f[m]) / 2 : f[m];
}

// split float4 (nx,ny,nz,depth) into the .dmb plane layouts
__global__ void split_maps_kernel(const float4 *planes, int n, float *depth, float *normal3) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 p = planes[i];
    if (depth) depth[i] = p.w;
    if (normal3) {
        normal3[3 * i + 0] = p.x;
        normal3[3 * i + 1] = p.y;
        normal3[3 * i + 2] = p.z;
    }
}

// (normal3, w) -> float4
__global__ void pack_planes_kernel(float4 *planes, int n, const float *normal3, const float *w, int zero_normal) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 p;
    if (zero_normal) { p.x = 0.f; p.y = 0.f; p.z = 0.f; }
    else { p.x = normal3[3 * i]; p.y = normal3[3 * i + 1]; p.z = normal3[3 * i + 2]; }
    p.w = w[i];
    planes[i] = p;
}

// ---------------------------------------------------------------------------
// joint bilateral upsample of a depth map (JBU_cu, ACMMP.cu:1458-1516)
// ---------------------------------------------------------------------------
__global__ void jbu_depth_kernel(const float *image, int cols, int rows, const float *sdepth, int s_width,
                                 int s_height, int Imagescale, float *depth) {
    const int px = blockIdx.x * blockDim.x + threadIdx.x;
    const int py = blockIdx.y * blockDim.y + threadIdx.y;
    if (px >= cols || py >= rows) return;
    const int center = py * cols + px;
    const float scale = 1.0 * s_width / cols;
    const float sigmad = 0.50;
    const float sigmar = 25.5;
    const int WinWidth = Imagescale * Imagescale + 1;
    int num_neighbors = WinWidth / 2;
    const float o_y = py * scale;
    const float o_x = px * scale;
    const float refPix = image[center];
    int r_y = 0, r_ys = 0, r_x = 0, r_xs = 0;
    float sgauss = 0.0, rgauss = 0.0, totalgauss = 0.0;
    float total_val = 0.0, normalizing_factor = 0.0;
    float srcPix = 0, neighborPix = 0;
    for (int j = -num_neighbors; j <= num_neighbors; ++j) {
        r_y = o_y + j;
        r_y = (r_y > 0 ? (r_y < s_height ? r_y : s_height - 1) : 0);
        r_ys = py + j;
        r_ys = (r_ys > 0 ? (r_ys < rows ? r_ys : rows - 1) : 0);
        for (int i = -num_neighbors; i <= num_neighbors; ++i) {
            r_x = o_x + i;
            r_x = (r_x > 0 ? (r_x < s_width ? r_x : s_width - 1) : 0);
            srcPix = sdepth[r_y * s_width + r_x];
            r_xs = px + i;
            r_xs = (r_xs > 0 ? (r_xs < cols ? r_xs : cols - 1) : 0);
            neighborPix = image[r_ys * cols + r_xs];
            sgauss = jbu_spatial_gauss(o_x, o_y, r_x, r_y, sigmad);
            rgauss = jbu_range_gauss(fabs(refPix - neighborPix), sigmar);
            totalgauss = sgauss * rgauss;
            normalizing_factor += totalgauss;
            total_val += srcPix * totalgauss;
        }
    }
    depth[center] = total_val / normalizing_factor;
}

// ---------------------------------------------------------------------------
// bilinear resize of a float image with OpenCV INTER_LINEAR geometry
// (cv::resize as used by ACMMP.cpp:589 / acmmp_definitions.cpp:435):
// sample position (d + 0.5) * (src/dst) - 0.5 in double, clamped taps,
// horizontal pass then vertical pass in float.
// ---------------------------------------------------------------------------
__global__ void resize_linear_kernel(const float *src, int sw, int sh, float *dst, int dw, int dh) {
    const int dx = blockIdx.x * blockDim.x + threadIdx.x;
    const int dy = blockIdx.y * blockDim.y + threadIdx.y;
    if (dx >= dw || dy >= dh) return;
    const double scale_x = (double)sw / dw, scale_y = (double)sh / dh;
    float fx = (float)((dx + 0.5) * scale_x - 0.5);
    int sx = (int)floorf(fx);
    fx -= sx;
    if (sx < 0) { fx = 0; sx = 0; }
    if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
    float fy = (float)((dy + 0.5) * scale_y - 0.5);
    int sy = (int)floorf(fy);
    fy -= sy;
    if (sy < 0) { fy = 0; sy = 0; }
    if (sy >= sh - 1) { fy = 0; sy = sh - 1; }
    const int sx1 = min(sx + 1, sw - 1), sy1 = min(sy + 1, sh - 1);
    const float a0 = 1.f - fx, a1 = fx, b0 = 1.f - fy, b1 = fy;
    const float r0 = fadd(fmul(src[(size_t)sy * sw + sx], a0), fmul(src[(size_t)sy * sw + sx1], a1));
    const float r1 = fadd(fmul(src[(size_t)sy1 * sw + sx], a0), fmul(src[(size_t)sy1 * sw + sx1], a1));
    dst[(size_t)dy * dw + dx] = fadd(fmul(r0, b0), fmul(r1, b1));
}

// image -> padded layer staging buffer with replicated edges: sampling the
// (larger) layer with clamp addressing is then identical to clamp addressing on
// the image itself.
__global__ void pad_replicate_kernel(const float *src, int w, int h, float *dst, int W, int H) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= W || y >= H) return;
    dst[(size_t)y * W + x] = src[(size_t)min(y, h - 1) * w + min(x, w - 1)];
}

// ---------------------------------------------------------------------------
// roofline denominators: FP32 FMA and MUFU issue rate of this chip
// ---------------------------------------------------------------------------
__global__ void fp32_peak_kernel(float *out, int iters) {
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f;
    float a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f, a7 = a0 + 7.f;
    const float b = 0.999f, c = 1e-4f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            a0 = ffma(a0, b, c); a1 = ffma(a1, b, c); a2 = ffma(a2, b, c); a3 = ffma(a3, b, c);
            a4 = ffma(a4, b, c); a5 = ffma(a5, b, c); a6 = ffma(a6, b, c); a7 = ffma(a7, b, c);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}
__global__ void mufu_peak_kernel(float *out, int iters) {
    float a0 = 1.f + threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            a0 = mufu_rcp(a0); a1 = mufu_rcp(a1); a2 = mufu_rcp(a2); a3 = mufu_rcp(a3);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3;
}

// ===========================================================================
// launch wrappers
// ===========================================================================
static inline dim3 grid2(int w, int h, int bx, int by) { return dim3((w + bx - 1) / bx, (h + by - 1) / by, 1); }

cudaError_t launch_setup_pairs(const acmmp_camera *h_cams, int n, const float *const *h_depth_maps, const int *h_layers, float depth_min, float depth_max,
                               RefConst *d_rc, PairConst *d_pairs, cudaStream_t st) {
    SetupArgs args;
    memset(&args, 0, sizeof args);
    for (int i = 0; i < n; ++i) {
        args.cams[i] = h_cams[i];
        args.depth[i] = h_depth_maps ? h_depth_maps[i] : nullptr;
        args.layer[i] = h_layers ? h_layers[i] : i;
    }
    setup_pairs_kernel<<<1, 64, 0, st>>>(args, n, h_depth_maps != nullptr, depth_min, depth_max, d_rc, d_pairs);
    return cudaGetLastError();
}

cudaError_t launch_rng_init(unsigned long long seed, int W, int H, uint32_t *d_rng6, cudaStream_t st) {
    RngPlanes rp{d_rng6, (size_t)W * H};
    const int segs = (W + 31) / 32;
    const int total = segs * H;
    rng_init_kernel<<<(total + 127) / 128, 128, 0, st>>>(seed, W, H, rp);
    return cudaGetLastError();
}

size_t init_smem_bytes(int nsrc) {
    size_t per_group = sizeof(GroupSmemFixed) + (size_t)2 * nsrc * sizeof(float);
    per_group = (per_group + 15) & ~(size_t)15;
    size_t pairs = ((size_t)nsrc * sizeof(PairConst) + 15) & ~(size_t)15;
    size_t rc = (sizeof(RefConst) + 15) & ~(size_t)15;
    return pairs + rc + per_group * kInitGroups;
}

cudaError_t launch_init(const InitLaunch &L, cudaStream_t st) {
    InitArgs a;
    a.planes = L.planes; a.costs = L.costs; a.pre_costs = L.pre_costs; a.selected_views = L.selected_views;
    a.seeded_planes = L.seeded_planes; a.prior_planes = L.prior_planes; a.plane_masks = L.plane_masks;
    a.scaled_planes = L.scaled_planes;
    a.rng = RngPlanes{L.rng6, (size_t)L.W * L.H};
    a.ref_image = L.ref_image; a.tex = L.tex; a.pairs = L.pairs; a.rc = L.rc; a.nsrc = L.nsrc; a.mode = L.mode;
    a.hierarchy = L.hierarchy; a.top_k = L.top_k; a.sigma_spatial = L.sigma_spatial; a.sigma_color = L.sigma_color;
    a.scaled_cols = L.scaled_cols; a.scaled_rows = L.scaled_rows;
    const size_t smem = init_smem_bytes(L.nsrc);
    // the opt-in is a per-device attribute and contexts may live on several devices of one process:
    // set it on every large launch (a host-side table look-up next to a kernel launch)
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(L.tex_u8 ? init_kernel<true> : init_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    if (L.tex_u8) init_kernel<true><<<grid2(L.W, L.H, kInitTileW, kInitTileH), kInitThreads, smem, st>>>(a);
    else init_kernel<false><<<grid2(L.W, L.H, kInitTileW, kInitTileH), kInitThreads, smem, st>>>(a);
    return cudaGetLastError();
}

// ACMMP_SWEEP_MAX_BLOCKS=3 (compile time) keeps the three-CTA configuration everywhere (variant timing)
#ifndef ACMMP_SWEEP_MAX_BLOCKS
#define ACMMP_SWEEP_MAX_BLOCKS 4
#endif
cudaError_t launch_sweep(const SweepLaunch &L, cudaStream_t st) {
    SweepArgs a;
    a.planes_other = L.planes_other; a.costs_other = L.costs_other;
    a.planes_in = L.planes_in; a.costs_in = L.costs_in; a.planes_self = L.planes_self; a.costs_self = L.costs_self;
    a.planes = L.planes; a.costs = L.costs;
    a.selected_views = L.selected_views; a.pre_costs = L.pre_costs; a.prior_planes = L.prior_planes;
    a.plane_masks = L.plane_masks;
    a.rng = RngPlanes{L.rng6, (size_t)L.W * L.H};
    a.ref_image = L.ref_image; a.tex = L.tex; a.pairs = L.pairs; a.rc = L.rc; a.nsrc = L.nsrc; a.iter = L.iter; a.color = L.color;
    a.sigma_spatial = L.sigma_spatial; a.sigma_color = L.sigma_color;
    a.geom_consistency = L.geom_consistency; a.planar_prior = L.planar_prior; a.hierarchy = L.hierarchy;
    a.compat = L.compat;
    a.wave_lo = L.wave_lo; a.wave_hi = L.wave_hi; a.ref_grid_x = L.ref_grid_x;
    a.ncc_counter = L.ncc_counter;
    const bool boxes = L.tex_u8 && L.views_map != nullptr && L.layer_w >= kBoxW && L.layer_h >= kBoxH;
    if (boxes) memcpy(&a.views_map, L.views_map, sizeof a.views_map);
    else memset(&a.views_map, 0, sizeof a.views_map);
    a.layer_w = L.layer_w; a.layer_h = L.layer_h;
    const size_t smem = sweep_smem_bytes(L.nsrc, boxes);
    // four CTAs per SM when their shared memory fits (1 KB per CTA is reserved by the system), else three
    const bool four = !boxes && 4 * (smem + 1024) <= (size_t)228 * 1024 && ACMMP_SWEEP_MAX_BLOCKS >= 4;
#ifdef ACMMP_PAIR_MATRIX_ALWAYS
    const bool cold = true;
#else
    const bool cold = L.cold != 0;
#endif
    void (*kernel)(SweepArgs);
    if (boxes) kernel = sweep_kernel<true, true, 3, false>;
    else if (L.tex_u8) kernel = four ? (cold ? sweep_kernel<true, false, 4, true> : sweep_kernel<true, false, 4, false>)
                                     : (cold ? sweep_kernel<true, false, 3, true> : sweep_kernel<true, false, 3, false>);
    else kernel = four ? (cold ? sweep_kernel<false, false, 4, true> : sweep_kernel<false, false, 4, false>)
                       : (cold ? sweep_kernel<false, false, 3, true> : sweep_kernel<false, false, 3, false>);
    // per-device attribute: see launch_init
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    kernel<<<grid2(L.W, L.H, kTileW, kTileH), kSweepThreads, smem, st>>>(a);
    return cudaGetLastError();
}

__global__ void merge_color_kernel(float4 *__restrict__ planes_dst, float *__restrict__ costs_dst, const float4 *__restrict__ planes_src,
                                   const float *__restrict__ costs_src, int W, int H, int color) {
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    const int x = 2 * (blockIdx.x * blockDim.x + threadIdx.x) + ((y + color) & 1);
    if (x >= W || y >= H) return;
    const int c = y * W + x;
    planes_dst[c] = planes_src[c];
    costs_dst[c] = costs_src[c];
}
cudaError_t launch_merge_color(float4 *planes_dst, float *costs_dst, const float4 *planes_src, const float *costs_src, int W, int H, int color,
                               cudaStream_t st) {
    merge_color_kernel<<<grid2((W + 1) / 2, H, 32, 8), dim3(32, 8), 0, st>>>(planes_dst, costs_dst, planes_src, costs_src, W, H, color);
    return cudaGetLastError();
}

cudaError_t launch_depth_normal(float4 *planes, const RefConst *rc, int W, int H, cudaStream_t st) {
    depth_normal_kernel<<<grid2(W, H, 32, 8), dim3(32, 8), 0, st>>>(planes, rc);
    return cudaGetLastError();
}

cudaError_t launch_median_filter(float4 *planes, const float *costs, int W, int H, int color, cudaStream_t st) {
    median_filter_kernel<<<grid2(W, (H + 1) / 2, 32, 8), dim3(32, 8), 0, st>>>(planes, costs, W, H, color);
    return cudaGetLastError();
}

cudaError_t launch_split_maps(const float4 *planes, int n, float *depth, float *normal3, cudaStream_t st) {
    split_maps_kernel<<<(n + 255) / 256, 256, 0, st>>>(planes, n, depth, normal3);
    return cudaGetLastError();
}

cudaError_t launch_pack_planes(float4 *planes, int n, const float *normal3, const float *w, int zero_normal,
                               cudaStream_t st) {
    pack_planes_kernel<<<(n + 255) / 256, 256, 0, st>>>(planes, n, normal3, w, zero_normal);
    return cudaGetLastError();
}

cudaError_t launch_jbu_depth(const float *image, int w, int h, const float *sdepth, int sw, int sh, int scale,
                             float *out, cudaStream_t st) {
    jbu_depth_kernel<<<grid2(w, h, 32, 8), dim3(32, 8), 0, st>>>(image, w, h, sdepth, sw, sh, scale, out);
    return cudaGetLastError();
}

cudaError_t launch_resize_linear(const float *src, int sw, int sh, float *dst, int dw, int dh, cudaStream_t st) {
    resize_linear_kernel<<<grid2(dw, dh, 32, 8), dim3(32, 8), 0, st>>>(src, sw, sh, dst, dw, dh);
    return cudaGetLastError();
}

cudaError_t launch_pad_replicate(const float *src, int w, int h, float *dst, int W, int H, cudaStream_t st) {
    pad_replicate_kernel<<<grid2(W, H, 32, 8), dim3(32, 8), 0, st>>>(src, w, h, dst, W, H);
    return cudaGetLastError();
}

// 8-bit grey -> float (imread GRAYSCALE + convertTo CV_32F, ACMMP.cpp:539-541)
__global__ void u8_to_float_kernel(const uint8_t *__restrict__ src, size_t n, float *__restrict__ dst) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = (float)src[i];
}
cudaError_t launch_u8_to_float(const uint8_t *src, size_t n, float *dst, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    u8_to_float_kernel<<<(unsigned)blocks, 256, 0, st>>>(src, n, dst);
    return cudaGetLastError();
}

// integer grey levels?  (decoded 8-bit images that were never resized)
__global__ void check_u8_kernel(const float *__restrict__ src, size_t n, int *not_u8) {
    bool bad = false;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float v = src[i];
        bad |= !(v >= 0.0f && v <= 255.0f && v == truncf(v));
    }
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(not_u8, 1);
}
cudaError_t launch_check_u8(const float *src, size_t n, int *not_u8, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    check_u8_kernel<<<(unsigned)blocks, 256, 0, st>>>(src, n, not_u8);
    return cudaGetLastError();
}
__global__ void to_u8_replicate_kernel(const float *__restrict__ src, int w, int h, uint8_t *__restrict__ dst, int W, int H) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= W || y >= H) return;
    dst[(size_t)y * W + x] = (uint8_t)__float2int_rn(src[(size_t)min(y, h - 1) * w + min(x, w - 1)]);
}
cudaError_t launch_to_u8_replicate(const float *src, int w, int h, uint8_t *dst, int W, int H, cudaStream_t st) {
    to_u8_replicate_kernel<<<grid2(W, H, 32, 8), dim3(32, 8), 0, st>>>(src, w, h, dst, W, H);
    return cudaGetLastError();
}

// bilinear fetch rate of a layered R32F texture.  mode 0: a warp reads 32 horizontally
// adjacent pixels (best case); mode 1: the sweep kernel's shape - 8 lanes around each of
// 4 pixels, a few pixels apart, taps advancing by 2 pixels.
// u8: R8 unorm texels + the two-operation rounding of bilateral_ncc<true>.
__global__ void tex_peak_kernel(cudaTextureObject_t tex, int u8, int w, int h, int layers, int iters, int mode, float *out) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = tid >> 5;
    float bx, by;
    if (mode == 0) {
        bx = (float)((warp * 32) % w + lane) + 0.37f;
        by = (float)((warp * 7) % h) + 0.61f;
    } else {
        bx = (float)((warp * 8) % w + 2 * (lane >> 3)) + 1.7f * (float)(lane & 7) + 0.37f;
        by = (float)((warp * 5) % h) + 0.9f * (float)(lane & 7) + 0.61f;
    }
    const int layer = warp % layers;
    float acc = 0.0f;
    for (int i = 0; i < iters; ++i) {
        const float x = bx + (float)(2 * (i % 6));
        const float y = by + (float)(2 * ((i / 6) % 6)) + (float)(i / 36);
        float t = tex2DLayered<float>(tex, x, y, layer);
        if (u8) t = fadd(ffma(t, 255.0f, 49152.0f), -49152.0f);
        acc += t;
    }
    if (acc == 123.456f) out[tid] = acc;   // keeps the loop alive
}
cudaError_t launch_tex_peak(cudaTextureObject_t tex, int u8, int w, int h, int layers, int blocks, int threads, int iters, int mode,
                            float *scratch, cudaStream_t st) {
    tex_peak_kernel<<<blocks, threads, 0, st>>>(tex, u8, w, h, layers, iters, mode, scratch);
    return cudaGetLastError();
}

// one bilinear fetch per coordinate, through the same texture set-up as the problem's views
// (u8: the same rounding back onto the 1/256 grid as bilateral_ncc<true>, ncc.cuh)
__global__ void tex_sample_kernel(cudaTextureObject_t tex, int u8, const float *__restrict__ xs, const float *__restrict__ ys, int n,
                                  float *__restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float s = tex2DLayered<float>(tex, xs[i], ys[i], 0);
    if (u8) s = fadd(ffma(s, 255.0f, 49152.0f), -49152.0f);
    out[i] = s;
}
cudaError_t launch_tex_sample(cudaTextureObject_t tex, int u8, const float *xs, const float *ys, int n, float *out, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    tex_sample_kernel<<<(n + 255) / 256, 256, 0, st>>>(tex, u8, xs, ys, n, out);
    return cudaGetLastError();
}

cudaError_t launch_fp32_peak(float *scratch, int blocks, int threads, int iters, cudaStream_t st) {
    fp32_peak_kernel<<<blocks, threads, 0, st>>>(scratch, iters);
    return cudaGetLastError();
}
cudaError_t launch_mufu_peak(float *scratch, int blocks, int threads, int iters, cudaStream_t st) {
    mufu_peak_kernel<<<blocks, threads, 0, st>>>(scratch, iters);
    return cudaGetLastError();
}

size_t pair_const_size() { return sizeof(PairConst); }
size_t ref_const_size() { return sizeof(RefConst); }

}  // namespace acmmp
