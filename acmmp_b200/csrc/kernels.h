// kernels.h - host-visible launch interface of kernels.cu (internal, C++).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/acmmp_b200.h"

namespace acmmp {

// box staging of the sweep kernel (ncc.cuh): bytes per box row / rows per box
constexpr int kBoxW = 64, kBoxH = 32;

struct PairConst;
struct RefConst;

struct InitLaunch {
    float4 *planes;
    float *costs;
    float *pre_costs;
    uint32_t *selected_views;
    const float4 *seeded_planes;
    const float4 *prior_planes;
    const uint32_t *plane_masks;
    const float4 *scaled_planes;
    uint32_t *rng6;
    const float *ref_image;
    cudaTextureObject_t tex;
    int tex_u8;   // tex is an R8 unorm array of integer grey levels (ncc.cuh), else R32F
    const PairConst *pairs;
    const RefConst *rc;
    int W, H, nsrc, mode, hierarchy, top_k;
    float sigma_spatial, sigma_color, scaled_cols, scaled_rows;
};

struct SweepLaunch {
    const float4 *planes_other;
    const float *costs_other;
    const float4 *planes_in;
    const float *costs_in;
    const float4 *planes_self;
    const float *costs_self;
    float4 *planes;
    float *costs;
    uint32_t *selected_views;
    const float *pre_costs;
    const float4 *prior_planes;
    const uint32_t *plane_masks;
    uint32_t *rng6;
    const float *ref_image;
    cudaTextureObject_t tex;
    int tex_u8;
    const PairConst *pairs;
    const RefConst *rc;
    int W, H, nsrc, iter, color;
    float sigma_spatial, sigma_color;
    int geom_consistency, planar_prior, hierarchy, compat;
    int wave_lo, wave_hi, ref_grid_x;   // wave emulation (sweep_kernel.cuh); wave_hi == 0: off
    int cold;   // a half-sweep on a freshly randomised state: cost matrix by lane pairs (sweep_kernel<..., kPairMatrix>)
    unsigned long long *ncc_counter;   // nullable; 4096 slots
    // box staging (ncc.cuh): host copy of the tensor map over the R8 layers as linear bytes, or null; layer size
    const void *views_map;
    int layer_w, layer_h;
};

// h_cams / h_depth_maps are HOST arrays (of device pointers for the depth maps): passed as a kernel argument
cudaError_t launch_setup_pairs(const acmmp_camera *h_cams, int n, const float *const *h_depth_maps, const int *h_layers, float depth_min, float depth_max,
                               RefConst *d_rc, PairConst *d_pairs, cudaStream_t st);
cudaError_t launch_rng_init(unsigned long long seed, int W, int H, uint32_t *d_rng6, cudaStream_t st);
cudaError_t launch_init(const InitLaunch &L, cudaStream_t st);
cudaError_t launch_sweep(const SweepLaunch &L, cudaStream_t st);
// pixels of `color` ((x + y) & 1) of src -> dst (planes and costs): merges a half-finished ping-pong pair
cudaError_t launch_merge_color(float4 *planes_dst, float *costs_dst, const float4 *planes_src, const float *costs_src, int W, int H, int color,
                               cudaStream_t st);
cudaError_t launch_depth_normal(float4 *planes, const RefConst *rc, int W, int H, cudaStream_t st);
cudaError_t launch_median_filter(float4 *planes, const float *costs, int W, int H, int color, cudaStream_t st);
cudaError_t launch_split_maps(const float4 *planes, int n, float *depth, float *normal3, cudaStream_t st);
cudaError_t launch_pack_planes(float4 *planes, int n, const float *normal3, const float *w, int zero_normal,
                               cudaStream_t st);
cudaError_t launch_jbu_depth(const float *image, int w, int h, const float *sdepth, int sw, int sh, int scale,
                             float *out, cudaStream_t st);
cudaError_t launch_resize_linear(const float *src, int sw, int sh, float *dst, int dw, int dh, cudaStream_t st);
cudaError_t launch_pad_replicate(const float *src, int w, int h, float *dst, int W, int H, cudaStream_t st);
cudaError_t launch_u8_to_float(const uint8_t *src, size_t n, float *dst, cudaStream_t st);
// *not_u8 |= 1 when a pixel of src is not an integer grey level 0..255
cudaError_t launch_check_u8(const float *src, size_t n, int *not_u8, cudaStream_t st);
// float grey levels -> bytes, edges replicated up to W x H (the layer size)
cudaError_t launch_to_u8_replicate(const float *src, int w, int h, uint8_t *dst, int W, int H, cudaStream_t st);
cudaError_t launch_fp32_peak(float *scratch, int blocks, int threads, int iters, cudaStream_t st);
// seeded.cu (IEEE build): seeded hypotheses of one view at the current level from its 16-bit prior images
cudaError_t launch_prior_maps_to_planes(const acmmp_camera &cam, const uint16_t *depth16, const uint16_t *normal16, int prows,
                                        int pcols, int rows, int cols, float4 *planes, cudaStream_t st);
cudaError_t launch_tex_sample(cudaTextureObject_t tex, int u8, const float *xs, const float *ys, int n, float *out, cudaStream_t st);
cudaError_t launch_tex_peak(cudaTextureObject_t tex, int u8, int w, int h, int layers, int blocks, int threads, int iters, int mode,
                            float *scratch, cudaStream_t st);
cudaError_t launch_mufu_peak(float *scratch, int blocks, int threads, int iters, cudaStream_t st);
cudaError_t launch_support_cells(const float *costs, int W, int H, int2 *cells, cudaStream_t st);
cudaError_t launch_tri_raster(const int *tris, int ntri, const float4 *planes, const RefConst *rc, int W,
                              float4 *tri_planes, unsigned *labels, int *big_list, cudaStream_t st);
cudaError_t launch_prior_finalize(const unsigned *labels, const float4 *tri_planes, const RefConst *rc, int W, int H,
                                  float4 *prior_planes, unsigned *masks, cudaStream_t st);
size_t pair_const_size();
size_t ref_const_size();

}  // namespace acmmp
