/*
 * oracle.h - CPU restatement of the reference's per-view PatchMatch path.
 *
 * TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library; nothing under
 * acmmp_b200/ does.  It exists to check the CUDA product, never to stand in
 * for it.
 *
 * Parity status: the reference (rlav440/ACMMP) ships no tests, golden vectors
 * or fixtures for this path, and its implementation of the
 * path is CUDA-only, so this restatement cannot be pinned against reference
 * outputs inside the CPU-only container.  It is pinned on the GPU box instead:
 * tests/test_gpu_parity.py compares it with the reference's own kernels
 * (oracle/_ref, built from /root/reference/src/ACMMP.cu) and golden vectors
 * captured from those kernels are committed under tests/golden/.  The RNG part
 * is pinned against cuRAND's host tables (tests/golden/xorwow_*.npy).
 *
 * Arithmetic: plain IEEE float in the reference's expression order, no FMA
 * contraction (-ffp-contract=off), exact sqrt/exp/division.  The reference
 * build uses --use_fast_math approximations, so agreement is to ~1e-5 on
 * costs, not bit-exact; texture bilinear filtering is emulated with the
 * hardware's 8-bit interpolation weights.
 *
 * Every function cites the reference lines it restates.
 */
#ifndef ACMMP_ORACLE_H_
#define ACMMP_ORACLE_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {  /* = reference struct Camera, acmmp_definitions.h:47-55 */
    float K[9], R[9], t[3];
    int32_t height, width;
    float depth_min, depth_max;
} orc_camera;

typedef struct {
    int32_t n;                  /* views: 0 = reference */
    const orc_camera *cams;
    const float *const *images; /* per view, cams[i].width x cams[i].height */
    const float *const *depths; /* per view or NULL (geometric consistency) */
    float depth_min, depth_max; /* params.depth_min/max (already x0.6 / x1.2) */
    float sigma_spatial, sigma_color;
    int32_t top_k;
    int32_t geom, planar_prior, hierarchy;
    int32_t compat;             /* 0 = as compiled for sm_100, 1 = upstream */
    float scaled_cols, scaled_rows;
    /* per-pixel state of the reference view (W x H) */
    float *planes;              /* 4 floats */
    float *costs;
    float *pre_costs;
    uint32_t *sel;
    uint32_t *rng6;             /* d, v0..v4 */
    const float *prior_planes;  /* 4 floats */
    const uint32_t *masks;
    const float *scaled_planes; /* coarse: world normal + cost */
    const float *seeded;        /* 4 floats */
} orc_problem;

/* XORWOW, curand_init(seed, subsequence, offset) compatible */
void orc_rng_seed(uint64_t seed, uint64_t subsequence, uint64_t offset, uint32_t st[6]);
uint32_t orc_rng_next(uint32_t st[6]);
float orc_rng_uniform(uint32_t st[6]);
void orc_rng_init_image(uint64_t seed, int w, int h, uint32_t *rng6);

/* sampling */
float orc_tex_bilinear(const float *img, int w, int h, float x, float y);

/* geometry + cost */
void orc_homography(const orc_camera *ref, const orc_camera *src, const float plane[4], float H[9]);
float orc_ncc(const orc_problem *p, int px, int py, const float plane[4], int view);
float orc_geom_cost(const orc_problem *p, int px, int py, const float plane[4], int view);
float orc_plane_depth(const orc_camera *cam, const float plane[4], int px, int py);
float orc_plane_offset(const orc_camera *cam, int px, int py, float depth, const float n[3]);
float orc_initial_cost(const orc_problem *p, int px, int py, const float plane[4], uint32_t *sel,
                       float *view_costs /* n-1 or NULL */);

/* kernels; region = [x0,x1) x [y0,y1) of pixels to process */
enum { ORC_INIT_RANDOM = 0, ORC_INIT_SEEDED = 1, ORC_INIT_PRIOR = 2, ORC_INIT_UPSAMPLE = 3, ORC_INIT_WARM = 4 };
void orc_init(orc_problem *p, int mode, int x0, int y0, int x1, int y1);
void orc_sweep(orc_problem *p, int iter, int color, int x0, int y0, int x1, int y1);
void orc_depth_normal(orc_problem *p);
void orc_median_filter(orc_problem *p, int color);
void orc_run(orc_problem *p, int mode, int iterations);

/* JBU / resize */
int orc_jbu_depth(const float *image, int w, int h, const float *depth, int sw, int sh, float *out);
void orc_resize_linear(const float *src, int sw, int sh, float *dst, int dw, int dh);

/* planar prior around the triangulation (ACMMP.cpp:868-958, acmmp_definitions.cpp:332-374) */
int orc_support_points(const float *costs, int W, int H, int32_t *xy /* 2 * ceil(W/5)*ceil(H/5) */);
void orc_prior_from_triangles(const orc_camera *cam, const float *depth, int W, int H, const int32_t *tris, int ntri,
                              float depth_min, float depth_max, float *tri_planes, float *prior4, uint32_t *mask);

/* fusion (RunFusion, acmmp_definitions.cpp:828-1043); returns number of points.
 * pair_offsets: n+1 prefix offsets into pair_ids (source ids as indices).
 * points: capacity*9 floats (x,y,z,nx,ny,nz,c0,c1,c2) or NULL to count only. */
long orc_fusion(int n, const orc_camera *cams, const float *const *depths, const float *const *normals,
                const uint8_t *const *colors, const int *pair_offsets, const int *pair_ids,
                float consistency_scalar, int con_num_thresh, float *points, long capacity);

/* prior-aware fusion (RunPriorAwareFusion, acmmp_definitions.cpp:443-826): two (depth, normal) sets per
 * view; masks0 (nullable, per view nullable): 1 = ignore the pixel.  Same point layout as orc_fusion. */
long orc_prior_fusion(int n, const orc_camera *cams, const float *const *depths, const float *const *normals,
                      const float *const *depths2, const float *const *normals2, const uint8_t *const *colors,
                      const uint8_t *const *masks0, const int *pair_offsets, const int *pair_ids, float consistency_scalar,
                      int num_consistent_thresh, int single_match_penalty, float *points, long capacity);

/* seeded hypotheses from 16-bit prior images (pSampler::GetPriorPlaneEstimate, acmmp_definitions.cpp:98-177) */
void orc_prior_planes_from_maps(const orc_camera *cam, const uint16_t *depth16, const uint16_t *normal16, int prows, int pcols,
                                int rows, int cols, float *planes4);

#ifdef __cplusplus
}
#endif
#endif
