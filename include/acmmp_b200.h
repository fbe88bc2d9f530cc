/*
 * acmmp_b200.h - C ABI of the B200-native ACMMP per-view PatchMatch hot path.
 *
 * This is the drop-in boundary.  The reference has no FFI; its "operator API"
 * for this path is the C++ host class `ACMMP` (+ `JBU`/`RunJBU`) declared in
 * /root/reference/src/ACMMP.h:58-124,142-160,18-20 and driven by
 * ProcessProblem / JointBilateralUpsampling
 * (/root/reference/src/acmmp_definitions.cpp:245-440).  Every entry point below
 * names the reference member it replaces.  The C++ host classes in
 * acmmp_b200/host/acmmp.hpp (command line: cli/acmmp_main.cpp) and the Python
 * mirror in acmmp_b200/host.py are thin callers of this ABI.  No torch / C++ types cross this boundary.
 *
 * Conventions: every function returns 0 on success and a negative acmmp_status
 * on failure (the reference prints and exit()s, ACMMP.cpp:67-100; here the
 * caller decides).  acmmp_last_error() returns a static, human readable
 * message for the last failure on that context (or the global one for
 * context-less calls).  All host pointers are plain row-major arrays.
 * Pointers named d_* are DEVICE pointers on the context's device.
 */
#ifndef ACMMP_B200_H_
#define ACMMP_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ACMMP_MAX_SRC_VIEWS 32 /* 32-bit view mask, reference ACMMP.cu:76-84,437 */

typedef enum {
    ACMMP_OK = 0,
    ACMMP_ERR_INVALID = -1, /* bad argument / call order */
    ACMMP_ERR_CUDA = -2,    /* CUDA runtime failure (message in last_error) */
    ACMMP_ERR_NO_DEVICE = -3,
    ACMMP_ERR_STATE = -4    /* required input (images, depths, prior ...) missing */
} acmmp_status;

/* Bit-compatible with the reference `struct Camera`
 * (/root/reference/src/acmmp_definitions.h:47-55): 100 bytes. */
typedef struct {
    float K[9];
    float R[9];
    float t[3];
    int32_t height;
    int32_t width;
    float depth_min;
    float depth_max;
} acmmp_camera;

/* Mirror of the reference `struct PatchMatchParams`
 * (/root/reference/src/ACMMP.h:32-56).  Flags are int32 instead of bool so the
 * layout is the same from C, C++ and ctypes. */
typedef struct {
    int32_t max_iterations;   /* 2 */
    int32_t patch_size;       /* 11 */
    int32_t num_images;       /* 1 reference + sources */
    int32_t max_image_size;   /* 3200 */
    int32_t radius_increment; /* 2 */
    float sigma_spatial;      /* 5 */
    float sigma_color;        /* 3 */
    int32_t top_k;            /* 4 */
    float baseline;           /* 0.54 */
    float depth_min;
    float depth_max;
    float disparity_min;
    float disparity_max;
    float scaled_cols; /* float on purpose: ACMMP.h:47-48 */
    float scaled_rows;
    int32_t geom_consistency;
    int32_t planar_prior;
    int32_t multi_geometry;
    int32_t hierarchy;
    int32_t upsample;
    int32_t seeded;
} acmmp_params;

/* How the uninitialised `plane_hypotheses_now` of the reference
 * (ACMMP.cu:1149-1172) is given a meaning.  See DESIGN.md "UB pinning". */
typedef enum {
    /* what nvcc 12.9 -arch sm_100 made of it: the refinement starts from the best
     * neighbour's plane whenever that arm exists (the arg-min arm; where a planar
     * prior covers the pixel, the arm with the largest restricted cost), while cost /
     * depth / views are taken over only if the arm is adopted; where the compiled
     * code reads garbage registers (the arm does not exist) the current plane is
     * used instead.  Measured against the reference's kernels: DESIGN.md 4(2). */
    ACMMP_COMPAT_REF_SM100 = 0,
    /* upstream GhiXu/ACMMP semantics: refine the current plane in place. */
    ACMMP_COMPAT_UPSTREAM = 1
} acmmp_compat_mode;

/* Texel format of the views' layered texture (the reference: one R32F cudaArray per image,
 * ACMMP.cpp:638-663).  AUTO stores decoded 8-bit grey images that were not resized as R8 unorm
 * texels - a quarter of the texture footprint - whenever every image of the problem / pyramid
 * level holds integer grey levels 0..255, and R32F otherwise; the bilinear samples the kernels
 * see are bit-identical either way (acmmp_b200/csrc/ncc.cuh, tools/tex_u8_probe.cu). */
typedef enum {
    ACMMP_TEX_R32F = 0,
    ACMMP_TEX_AUTO = 1
} acmmp_tex_format;

typedef struct acmmp_ctx acmmp_ctx;

/* ---- lifetime ------------------------------------------------------------
 * replaces ACMMP::ACMMP / ~ACMMP (ACMMP.cpp:107-152) and cudaSetDevice(0)
 * (acmmp_definitions.cpp:253). */
int acmmp_create(acmmp_ctx **out, int device);
void acmmp_destroy(acmmp_ctx *ctx);
const char *acmmp_last_error(const acmmp_ctx *ctx);
/* version / build info string ("acmmp_b200 <ver> sm_100a"). */
const char *acmmp_version(void);

/* default parameter block = PatchMatchParams{} (ACMMP.h:32-56). */
void acmmp_default_params(acmmp_params *p);

/* ---- mode setters ----------------------------------------------------------
 * ACMMP::SetGeomConsistencyParams / SetHierarchyParams / SetPlanarPriorParams
 * (ACMMP.cpp:447-464). */
int acmmp_set_geom_consistency(acmmp_ctx *ctx, int multi_geometry);
int acmmp_set_hierarchy(acmmp_ctx *ctx);
int acmmp_set_planar_prior_params(acmmp_ctx *ctx);
int acmmp_set_compat_mode(acmmp_ctx *ctx, int mode);
/* Emulation of the reference launch's block waves (0 = off, the default: every same-colour read sees the
 * state before the half-sweep - deterministic, independent of the GPU).  Black/RedPixelUpdate read
 * same-colour neighbours that other blocks of the same launch rewrite (ACMMP.cu:895-988 vs :1165-1171).
 * The reference launches 32 x 32-pixel blocks (32 x 16 threads, ACMMP.cu:1415-1424) of which
 * `resident_blocks` fit on the GPU at a time (B200: 148 = one 512-thread, 128-register block per SM);
 * blocks of a later wave start when the earlier wave has stored its results, so the top 4 rows of such a
 * block see NEW planes / costs above them.  With resident_blocks > 0 a half-sweep runs as one sub-launch
 * per wave that reproduces this (tools/parity_study.py, profiles/r2*_parity_study.json: 99.5 % of the
 * pixels where the product differs from the reference build sit in those 4 rows).  Parity / diagnosis
 * mode: same work, a few more launches. */
int acmmp_set_wave_emulation(acmmp_ctx *ctx, int resident_blocks);
/* acmmp_tex_format for the following acmmp_set_problem calls (default ACMMP_TEX_R32F);
 * acmmp_texture_is_u8: 1 when the current problem's views sit in R8 texels. */
int acmmp_set_texture_format(acmmp_ctx *ctx, int format);
int acmmp_texture_is_u8(const acmmp_ctx *ctx);
int acmmp_get_params(const acmmp_ctx *ctx, acmmp_params *out);

/* ---- inputs ----------------------------------------------------------------
 * acmmp_set_problem: the upload half of ACMMP::InputInitialization +
 * CudaSpaceInitialization (ACMMP.cpp:525-681): cameras[0] / images[0] are the
 * reference view, 1..n-1 the sources, images already at this scale's size
 * (cams[i].width x cams[i].height float32, row stride = width).  Widens the
 * depth range by x0.6 / x1.2 exactly like ACMMP.cpp:600-601.  Allocates the
 * per-pixel state.  `images[i]` may be a HOST pointer (images_on_device = 0)
 * or a DEVICE pointer (1; used by the resident-pyramid path). */
int acmmp_set_problem(acmmp_ctx *ctx, int n, const acmmp_camera *cams,
                      const float *const *images, int images_on_device);

/* geometric-consistency depth maps of ref + sources (ACMMP.cpp:608-635,
 * 683-708): slot 0..n-1, size must equal that view's image size. */
int acmmp_set_depth_map(acmmp_ctx *ctx, int slot, const float *depth, int w,
                        int h, int on_device);

/* warm start of a geometric-consistency run from the view's own previous
 * result (ACMMP.cpp:710-742): depth, world-frame normals (3 floats / px),
 * costs. */
int acmmp_set_warm_start(acmmp_ctx *ctx, const float *depth,
                         const float *normal3, const float *cost);

/* hierarchy run inputs (ACMMP.cpp:745-808): coarse world-frame normals and
 * coarse costs (cw x ch) and the JBU-upsampled depth at this scale. */
int acmmp_set_hierarchy_inputs(acmmp_ctx *ctx, const float *coarse_normal3,
                               const float *coarse_cost, int cw, int ch,
                               const float *fine_depth);

/* fixed per-pixel plane hypotheses (nx,ny,nz,d in the reference camera frame)
 * = ACMMP::SetPlanarPrior, the reference's `seeded` branch (ACMMP.cpp:476-484,
 * ACMMP.cu:634-639).  This is the parity hook of SURVEY 8(c). */
int acmmp_set_seeded_planes(acmmp_ctx *ctx, const float *planes4);

/* per-pixel prior planes + triangle labels =
 * ACMMP::CudaPlanarPriorInitialization (ACMMP.cpp:811-831). */
int acmmp_set_planar_prior(acmmp_ctx *ctx, const float *planes4,
                           const uint32_t *mask);

/* planar-prior construction on the current result (GetSupportPoints ->
 * DelaunayTriangulation -> raster + GetPriorPlaneParams -> range check ->
 * CudaPlanarPriorInitialization; ACMMP.cpp:868-958,
 * acmmp_definitions.cpp:306-378).  Outputs are optional (may be NULL). */
int acmmp_build_planar_prior(acmmp_ctx *ctx, int *num_support,
                             int *num_triangles);
/* read back the per-pixel prior planes (4 f32) and triangle labels (u32) that
 * the reference keeps in prior_planes_host / plane_masks_host
 * (ACMMP.cpp:813-827).  NULL = skip. */
int acmmp_get_planar_prior(acmmp_ctx *ctx, float *planes4, uint32_t *mask);
/* 2-D Delaunay triangulation of n integer points (x,y pairs) = the host step of
 * ACMMP::DelaunayTriangulation (ACMMP.cpp:896-918, cv::Subdiv2D there).  Writes
 * up to `capacity` triangles (6 ints each: x,y of the 3 vertices); returns the
 * triangle count (>= 0) or a negative status.  Pure host code: no device
 * needed. */
int acmmp_delaunay(const int32_t *xy, int n, int32_t *tris, int capacity);
/* host threads one triangulation may use (the top levels of the divide-and-conquer recursion run
 * concurrently): 1 = sequential, 0 = default (a third of the host's cores, at most 8; environment
 * variable ACMMP_DELAUNAY_THREADS overrides).  Triangles and their order do not depend on it. */
int acmmp_set_delaunay_threads(int threads);

/* ---- allocation-free planar-prior raster ---------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the raster half
 * of acmmp_build_planar_prior - plane fit per triangle, triangle raster (later triangles overwrite), depth-range
 * check per pixel - with all inputs and outputs passed by the caller as DEVICE pointers:
 *   ref_cam            the reference view's camera at this scale (width / height = map size), host struct
 *   depth_min/max      the widened range (acmmp_get_params: depth_min, depth_max)
 *   d_planes4          w*h float4, depth of the finished run in .w (acmmp_get_result layout)
 *   d_tris             num_triangles x 6 int32 (x,y of the 3 vertices, acmmp_delaunay layout)
 *   d_prior_planes4    out: w*h float4 prior plane per pixel (0 where uncovered)
 *   d_masks            out: w*h u32 triangle label (index + 1; 0 = uncovered)
 *   stream             cudaStream_t (NULL = default stream); the call only enqueues work
 *   err, err_len       optional buffer for the error text (may be NULL)
 * The kernels' scratch (label map, triangle planes, long-triangle queue, camera constants) lives in a file-scope
 * static __device__ array: 10.2 MiB that every GPU which loads the library holds until exit.  The call performs no
 * cudaMalloc / cudaFree / cudaHostAlloc and calls no allocating library routine.  Limits: w*h <=
 * ACMMP_NOALLOC_MAX_PIXELS, num_triangles <= ACMMP_NOALLOC_MAX_TRIANGLES; larger inputs are rejected with
 * ACMMP_ERR_INVALID and a message.  ALL CALLS SHARE THE ARRAY, which keeps the last call's contents (what must start
 * at zero is cleared by every call): calls that overlap on the GPU - different streams or host threads - race;
 * serialise them.  Results are bit-identical to acmmp_build_planar_prior's for the same triangles. */
#define ACMMP_NOALLOC_MAX_PIXELS (1600 * 1200)
#define ACMMP_NOALLOC_MAX_TRIANGLES 153600 /* 2 x the 5x5 cells of 1600 x 1200 */
int acmmp_prior_raster_noalloc(int device, const acmmp_camera *ref_cam, float depth_min, float depth_max,
                               const float *d_planes4, const int32_t *d_tris, int num_triangles,
                               float *d_prior_planes4, uint32_t *d_masks, void *stream, char *err, size_t err_len);

/* ---- run -------------------------------------------------------------------
 * ACMMP::RunPatchMatch (ACMMP.cu:1378-1456).  `seed` replaces the reference's
 * clock64() seed (ACMMP.cu:624).  Runs on the context's stream; results are
 * in device memory until fetched. */
int acmmp_run(acmmp_ctx *ctx, uint64_t seed);

/* the same sequence in separately callable steps (parity tests) */
int acmmp_run_init(acmmp_ctx *ctx, uint64_t seed);        /* RandomInitialization */
int acmmp_run_sweep(acmmp_ctx *ctx, int iter, int color); /* 0 = black, 1 = red  */
int acmmp_run_tail(acmmp_ctx *ctx); /* GetDepthandNormal + Black/RedPixelFilter */
/* Diagnostics: one half-sweep whose NEIGHBOUR reads (the candidates of the 8 arms) come from the
 * given host snapshot (planes 4 f32, cost) while the pixel's own plane is the context's state, which
 * is updated in place.  Black/RedPixelUpdate read same-colour neighbours while other threads rewrite
 * them (ACMMP.cu:895-988 vs :1092,1165-1171): a reference pixel has seen the state before the launch,
 * the state after it, or a mix.  Feeding the reference's OUTPUT as the snapshot bounds the second
 * case; tools/parity_study.py uses it to attribute differences to that race. */
int acmmp_run_sweep_snapshot(acmmp_ctx *ctx, int iter, int color, const float *snap_planes4, const float *snap_cost);

/* ---- outputs ---------------------------------------------------------------
 * ACMMP::GetPlaneHypothesis / GetCost (ACMMP.cpp:848-856): planes4 is
 * (nx,ny,nz,w) per pixel - after acmmp_run / acmmp_run_tail the normal is in
 * the WORLD frame and w is the depth.  Any pointer may be NULL. */
int acmmp_get_result(acmmp_ctx *ctx, float *planes4, float *cost,
                     uint32_t *selected_views);
/* the same split into the .dmb layouts (depth h*w, normal h*w*3, cost h*w). */
int acmmp_get_maps(acmmp_ctx *ctx, float *depth, float *normal3, float *cost);
/* device pointer of the current float depth map (valid after run/tail) for the
 * NCCL exchange; w*h floats. */
int acmmp_get_depth_device(acmmp_ctx *ctx, const float **d_depth);

/* raw state access for step-exact parity tests: planes (4 f32), costs,
 * selected views, rng state (6 x u32: d, v0..v4).  NULL = skip. */
int acmmp_get_state(acmmp_ctx *ctx, float *planes4, float *cost,
                    uint32_t *selected_views, uint32_t *rng6);
int acmmp_set_state(acmmp_ctx *ctx, const float *planes4, const float *cost,
                    const uint32_t *selected_views, const uint32_t *rng6);

int acmmp_width(const acmmp_ctx *ctx);
int acmmp_height(const acmmp_ctx *ctx);
int acmmp_sync(acmmp_ctx *ctx);
/* kernels launched by this context since creation (bench `gpu_launches`). */
uint64_t acmmp_launch_count(const acmmp_ctx *ctx);
/* device time of the last acmmp_run in ms, split per kernel family:
 * out[0]=init out[1]=sweeps out[2]=tail out[3]=total (CUDA events). */
int acmmp_last_timing(acmmp_ctx *ctx, float out[4]);
/* sums of the same four device times (ms) over every acmmp_run since the last
 * collect (synchronises the context's stream); *runs = how many. */
int acmmp_timing_collect(acmmp_ctx *ctx, double out_ms[4], int *runs);

/* ---- JBU ---------------------------------------------------------------------
 * RunJBU / JBU::CudaRun / JBU_cu (ACMMP.cpp:1022-1087, ACMMP.cu:1458-1549):
 * joint bilateral upsample of a (sw x sh) depth map to (w x h) guided by the
 * (w x h) image.  Returns ACMMP_JBU_NOTHING_TO_DO (+1; error codes are negative) and
 * writes nothing when the integer scale is 1, like the reference's early-out
 * (ACMMP.cpp:1032-1035).  A whole scan whose levels have the same size is refused by
 * acmmp_scan_begin_level / acmmp_scan_run_dataflow (ACMMP_ERR_STATE): the reference then
 * keeps the previous scale's photometric depths.dmb, a case this schedule does not model. */
#define ACMMP_JBU_NOTHING_TO_DO 1
int acmmp_jbu_depth(int device, const float *image, int w, int h,
                    const float *depth, int sw, int sh, float *out);

/* ---- image pyramid -----------------------------------------------------------
 * cv::resize(..., INTER_LINEAR) of a float image as used by
 * ACMMP::InputInitialization (ACMMP.cpp:566-598) and
 * JointBilateralUpsampling (acmmp_definitions.cpp:428-435); host in / host
 * out, computed on the device. */
int acmmp_resize_linear(int device, const float *src, int sw, int sh,
                        float *dst, int dw, int dh);

/* ---- RNG ---------------------------------------------------------------------
 * XORWOW state of curand_init(seed, subsequence=y, offset=x) for every pixel
 * of a w x h image (ACMMP.cu:624), 6 x u32 per pixel (d, v0..v4); host out. */
int acmmp_rng_init_states(int device, uint64_t seed, int w, int h,
                          uint32_t *rng6);

/* ---- instrumentation ------------------------------------------------------------
 * NCC evaluations (one pixel x hypothesis x source view = 36 bilinear fetches)
 * issued since the last call: by the initial-cost kernel (exact, from the problem
 * size) and by the propagation kernel (counted on the device when enabled with
 * acmmp_count_ncc; off by default).  Used for the executed-work side of the roofline. */
int acmmp_count_ncc(acmmp_ctx *ctx, int enable);
int acmmp_ncc_count(acmmp_ctx *ctx, uint64_t *init_ncc, uint64_t *sweep_ncc);

/* ---- micro-benchmarks used by bench.py for the roofline denominators --------
 * measured FP32 FMA throughput (TFLOP/s), MUFU rate (Gop/s). */
int acmmp_measure_fp32_peak(int device, double *tflops, double *mufu_gops);
/* measured bilinear fetch rate (G fetches/s) of a layered R32F texture of one
 * DTU-shape problem: warp-coherent coordinates, and the propagation kernel's
 * access shape (8 hypotheses around each of 4 pixels per warp). */
int acmmp_measure_tex_peak(int device, double *coherent_gfetch, double *grouped_gfetch);
/* the same on R8 unorm texels incl. the rounding of the kU8 kernels (ACMMP_TEX_AUTO levels) */
int acmmp_measure_tex_peak_u8(int device, double *coherent_gfetch, double *grouped_gfetch);
/* n bilinear fetches at (xs[i], ys[i]) (unnormalised, texel centres at +0.5) from a host image
 * through the texture set-up of the problem's views (reference: ACMMP.cpp:646-663).  Known-answer
 * hook for the filter emulation of the test oracle (oracle/cost.c: orc_tex_bilinear). */
int acmmp_tex_sample(int device, const float *image, int width, int height, const float *xs, const float *ys, int n, float *out);
/* the same through an R8 unorm texture + the rounding of the kU8 kernels (ACMMP_TEX_AUTO): image must hold
 * integer grey levels 0..255 (ACMMP_ERR_INVALID otherwise); equals acmmp_tex_sample bit for bit. */
int acmmp_tex_sample_u8(int device, const float *image, int width, int height, const float *xs, const float *ys, int n, float *out);

/* ==== whole-scan schedule ======================================================
 * The stage graph of main() (/root/reference/src/main_ACMMP.cpp:89-176) over a
 * set of views, with everything resident in device memory: image pyramids of
 * ALL views, and the depth / normal / cost maps the reference hands from stage
 * to stage through <out>/2333_%08d/{depths,depths_geom,normals,costs}.dmb
 * (acmmp_definitions.cpp:392-401, ACMMP.cpp:608-635,710-808).
 *
 * Sharding: rank r of `world` owns the views v with v % world == r and runs
 * ProcessProblem (acmmp_definitions.cpp:245-403) for them only.  Only depth
 * maps cross views (ACMMP.cpp:608-635); after every stage the ranks all-gather
 * the exchange buffer: acmmp_scan_all_gather (NCCL inside the library, section
 * "multi-GPU" below), or a caller's own transport over acmmp_scan_exchange.
 * acmmp_scan_run_dataflow runs the whole sharded scan without the stage barrier.
 *
 * levels: ComputeMultiScaleSettings (acmmp_definitions.cpp:207-243); level 0 is
 * the coarsest.  Stages per level, in order:
 *   ACMMP_STAGE_PHOTO  geom=F planar=T (hierarchy=T on levels > 0, after the JBU
 *                      upsample of the previous level's depths_geom)
 *   ACMMP_STAGE_GEOM0  geom=T multi=F      ACMMP_STAGE_GEOM1  geom=T multi=T
 */
typedef struct acmmp_scan acmmp_scan;
enum { ACMMP_STAGE_PHOTO = 0, ACMMP_STAGE_GEOM0 = 1, ACMMP_STAGE_GEOM1 = 2 };

/* pairs: pair.txt as CSR (GenerateSampleList, acmmp_definitions.cpp:179-205;
 * sources with score <= 0 already dropped); ids are view indices 0..n-1.
 * cams: full-resolution cameras as read by ReadCamera (ACMMP.cpp:154-179) with
 * width/height = the image size. */
int acmmp_scan_create(acmmp_scan **out, int device, int n_views, const acmmp_camera *cams,
                      const int32_t *pair_offsets, const int32_t *pair_ids);
void acmmp_scan_destroy(acmmp_scan *s);
const char *acmmp_scan_last_error(const acmmp_scan *s);
/* full-resolution grey image of one view as float (imread GRAYSCALE +
 * convertTo CV_32F, ACMMP.cpp:539-541); host or device pointer. */
int acmmp_scan_set_image(acmmp_scan *s, int view, const float *gray, int on_device);
/* the same from the 8-bit image itself (what cv::imread(path, IMREAD_GRAYSCALE) returns, before the
 * reference's convertTo(CV_32F)): a quarter of the host-to-device bytes; the conversion runs on the
 * device.  Identical results to acmmp_scan_set_image of the converted image. */
int acmmp_scan_set_image_u8(acmmp_scan *s, int view, const uint8_t *gray, int on_device);
int acmmp_scan_set_shard(acmmp_scan *s, int rank, int world);
/* acmmp_tex_format of the level-wide texture (default ACMMP_TEX_AUTO: R8 texels on levels where no
 * view is resized and every image holds integer grey levels); call before acmmp_scan_begin_level.
 * acmmp_scan_level_is_u8: what the current level uses. */
int acmmp_scan_set_texture_format(acmmp_scan *s, int format);
/* Box staging of the sweep kernel on R8 levels (north_star: "source-view patches staged in shared memory via TMA
 * tiles"): per 16 x 4 tile and source view a 64 x 32-byte box of the view's layer goes to shared memory by
 * cp.async.bulk.tensor, placed where the tile's centre lands under its current plane; taps whose texels lie inside are
 * filtered from it with the texture unit's own 1/256 weights (bit-identical on 8-bit texels), every other tap goes
 * through the unit.  1 = on, 0 = off, -1 = the process default (environment ACMMP_BOX_STAGING, else the build's).
 * Results do not depend on the setting.
 * (Another scheduling knob without an entry point: environment ACMMP_COLD_SWEEPS = number of half-sweeps after a random
 * initialisation whose cost matrix is scored by lane pairs, default 2 on R8 levels and 3 on R32F ones - timing
 * experiments only, results do not depend on it either: tests/test_gpu_cold_sweeps.py.) */
int acmmp_set_box_staging(acmmp_ctx *ctx, int on);
int acmmp_scan_set_box_staging(acmmp_scan *s, int on);
int acmmp_scan_level_is_u8(const acmmp_scan *s);
/* Order of the second geometric-consistency pass.  1 = the reference's in-place
 * order dependence (ProcessProblem rewrites depths_geom.dmb, so geom-1 of view i reads
 * what earlier views of the same loop already wrote; ACMMP.cpp:608-635) - only
 * meaningful with world == 1.  0 = Jacobi (geom-1 reads the geom-0 maps of all
 * views).  -1 = default: 1 on an unsharded scan, 0 on a sharded one. */
int acmmp_scan_set_sequential_geom(acmmp_scan *s, int on);

/* ---- multi-GPU: NCCL inside the library ----------------------------------------------
 * The reference is single-GPU (acmmp_definitions.cpp:253).  A sharded scan (acmmp_scan_set_shard) exchanges the
 * depth maps of the owned views after every stage - the only data that crosses views (ACMMP.cpp:608-635) - with ONE
 * ncclAllGather on the scan's own stream: no host synchronisation, no staging copy.  libnccl.so.2 is loaded with
 * dlopen() on first use (the copy a torch process already holds, else the system's); acmmp_nccl_version() is 0 when
 * it is missing.  Communicators are created with maxCTAs = 4 (ncclCommInitRankConfig; ACMMP_NCCL_MAX_CTAS overrides,
 * 0 = NCCL's default): a collective spins on its SMs until the slowest rank arrives, and those SMs are lost to the
 * next task's kernels.
 *   one process per GPU:  rank 0 calls acmmp_nccl_get_unique_id and hands the 128 bytes to the other ranks by any
 *                         means (torch.distributed broadcast, MPI, a file); every rank then calls acmmp_scan_comm_init.
 *   one thread per GPU:   acmmp_scan_comm_init_all(scans, n) (ncclCommInitAll); each thread drives its own scan.
 * acmmp_scan_all_gather: call after every acmmp_scan_run_stage (a no-op on an unsharded scan).
 * acmmp_scan_collect_view: owner -> `root`, normals (3 f32 / px) + costs of one view into device buffers the root
 * provides (ncclSend / ncclRecv; every rank may call it, only owner and root act): brings the final maps to the rank
 * that fuses. */
#define ACMMP_NCCL_ID_BYTES 128
int acmmp_nccl_version(void);
int acmmp_nccl_get_unique_id(uint8_t id[ACMMP_NCCL_ID_BYTES]);
int acmmp_scan_comm_init(acmmp_scan *s, const uint8_t id[ACMMP_NCCL_ID_BYTES]);
int acmmp_scan_comm_init_all(acmmp_scan *const *scans, int n);
int acmmp_scan_comm_destroy(acmmp_scan *s);
int acmmp_scan_all_gather(acmmp_scan *s);
int acmmp_scan_collect_view(acmmp_scan *s, int view, int root, float *d_normal3, float *d_cost);
/* The whole scan - every level, every stage (main_ACMMP.cpp:89-176) - as ONE call, scheduled as a dataflow of
 * (stage, view) tasks instead of stage by stage (csrc/capi_dataflow.cu.inc).  A task needs only the previous stage's
 * maps of its own view and, for a geometric-consistency run, of its source views (ACMMP.cpp:608-635), so the tasks are
 * list-scheduled into rounds of `world` tasks that cross stage boundaries: 49 views on 8 GPUs take 37 rounds instead
 * of 6 stages x 7.  After a round every rank broadcasts the record [depth | normal | cost] it wrote (grouped
 * ncclBroadcast on a second, high-priority stream, overlapped with the next round's kernels); a task waits only for
 * the broadcast of the newest round it reads.  Collective: every rank calls it with the same seed after
 * acmmp_scan_comm_init.  Replaces the acmmp_scan_begin_level / acmmp_scan_run_stage / acmmp_scan_all_gather loop
 * (do not mix them within one scan pass; acmmp_scan_reset between passes).  Results: those of the staged schedule
 * with Jacobi geom-1 order, bit for bit.  Afterwards EVERY rank holds the final maps of EVERY view:
 * acmmp_scan_get_view / acmmp_scan_view_device / acmmp_scan_collect_view serve them locally.
 * acmmp_scan_keep_photometric(s, 1) before the run keeps a copy of every view's photometric (+ planar prior) depth
 * of the finest level - the reference's depths.dmb (acmmp_definitions.cpp:392-397), which later stages overwrite - for
 * acmmp_scan_get_photometric_depth (host buffer, the view's size at the finest level).
 * acmmp_scan_dataflow_stats: rounds and lag (rounds between a task and the newest round it reads) of the last run.
 * acmmp_plan_dataflow: the plan alone, on the host (no GPU): task = stage * n_views + view with
 * stage = level * 3 + {0 photometric+prior, 1 geom-0, 2 geom-1} (stage = level when photometric_only);
 * task_round / task_rank: n_levels * (photometric_only ? 1 : 3) * n_views entries each.  Within a round the tasks go
 * to the ranks with the least WORK so far (rounds are no barriers): level_cost[l] = relative cost of one RunPatchMatch
 * at level l (NULL: 4^l; the scan uses the level's mean pixel count), a photometric task counts twice. */
int acmmp_scan_run_dataflow(acmmp_scan *s, uint64_t seed);
int acmmp_scan_dataflow_stats(const acmmp_scan *s, int *rounds, int *lag);
/* sharded scans: ms after the start of the last run at which this rank had finished round k (CUDA events on the
 * compute stream); returns the number of rounds written (0 on an unsharded scan), < 0 on error */
int acmmp_scan_dataflow_trace(acmmp_scan *s, float *round_done_ms, int capacity);
/* RunPatchMatch kernel time (init, sweeps, tail, total; CUDA events on the launching stream) of `level` since the last
 * acmmp_scan_timing / acmmp_scan_level_stats of that level, under either schedule; ncc (may be NULL): the level's NCC
 * evaluations (init, sweeps) counted by a dataflow run under acmmp_scan_count_ncc; texels_u8 (may be NULL): the
 * level's texel format.  Waits for the stream. */
int acmmp_scan_level_stats(acmmp_scan *s, int level, double out_ms[4], uint64_t ncc[2], int *texels_u8);
int acmmp_scan_keep_photometric(acmmp_scan *s, int on);
int acmmp_scan_get_photometric_depth(acmmp_scan *s, int view, float *depth);
int acmmp_plan_dataflow(int n_views, const int32_t *pair_offsets, const int32_t *pair_ids, int n_levels, int photometric_only,
                        int world, const float *level_cost, int32_t *task_round, int32_t *task_rank, int32_t *n_rounds, int32_t *lag);
/* Reduced schedules (before the first level).  single_scale: every view at its (capped) full size from the
 * start, no pyramid (ComputeMultiScaleSettings, acmmp_definitions.cpp:207-243, switched off).
 * photometric_only: the photometric stage is ONE RunPatchMatch per view, without the planar-prior construction
 * and the second run (needs single_scale or a one-level scan).  Both set = BASELINE.json configs[1]. */
int acmmp_scan_set_schedule(acmmp_scan *s, int single_scale, int photometric_only);
int acmmp_scan_num_levels(const acmmp_scan *s);
/* builds the image pyramid level (InputInitialization's resize,
 * ACMMP.cpp:566-598) for all views, and for level > 0 runs
 * JointBilateralUpsampling (acmmp_definitions.cpp:405-440) for the owned views. */
int acmmp_scan_begin_level(acmmp_scan *s, int level);
int acmmp_scan_run_stage(acmmp_scan *s, int stage, uint64_t seed);
/* the depth maps written by the last stage / begin_level: one device buffer of
 * world * slots * stride floats, rank r's segment = [r * slots * stride,
 * (r+1) * slots * stride); view v lives in segment v % world, slot v / world. */
int acmmp_scan_exchange(acmmp_scan *s, float **d_buffer, int64_t *stride, int *slots);
/* size of view v at the current level */
int acmmp_scan_view_size(const acmmp_scan *s, int view, int *w, int *h);
/* current maps of an owned view (.dmb payloads; depth = the last written depth
 * map); host pointers, any may be NULL. */
int acmmp_scan_get_view(acmmp_scan *s, int view, float *depth, float *normal3, float *cost);
/* Seeded mode (the reference's `-p`: pSampler, /root/reference/src/acmmp_definitions.cpp:7-177,
 * ProcessProblem :274-280, main_ACMMP.cpp:107-121).  Hands over the 16-bit prior images of a view
 * (what cv::imread(path, -1) returns for <dense>priors/depths/%08d.png and .../normals/%08d.png:
 * prior_rows x prior_cols, the normal image's three channels in stored order, channel 0 = x).  The
 * photometric stage of the FIRST scale then starts from these planes instead of random ones.
 * Both pointers NULL: back to the unseeded schedule for this view.  Call before acmmp_scan_run_stage. */
int acmmp_scan_set_prior_maps(acmmp_scan *s, int view, const uint16_t *depth16, const uint16_t *normal16x3,
                              int prior_rows, int prior_cols);
/* The conversion on its own (pSampler::GetPriorPlaneEstimate): planes4 = rows x cols x (nx, ny, nz in
 * the camera frame, distance to the origin); cam = the view's camera at rows x cols. */
int acmmp_prior_planes_from_maps(int device, const acmmp_camera *cam, const uint16_t *depth16,
                                 const uint16_t *normal16x3, int prior_rows, int prior_cols, int rows, int cols,
                                 float *planes4);
const char *acmmp_seeded_last_error(void);

/* the same maps as device pointers (valid until the next stage / level of the scan), e.g. to
 * hand them to acmmp_fusion_set_view(..., on_device = 1) after acmmp_scan_sync */
int acmmp_scan_view_device(acmmp_scan *s, int view, const float **d_depth, const float **d_normal3, const float **d_cost);
/* device time spent in kernels, summed over all runs since the last call:
 * out_ms[0..3] = init, sweeps, tail, total; *runs = RunPatchMatch count;
 * *launches = kernel launches so far. */
int acmmp_scan_timing(acmmp_scan *s, double out_ms[4], int *runs, uint64_t *launches);
int acmmp_scan_sync(acmmp_scan *s);
/* host-side cost of the planar-prior triangulation since the last call: total host
 * ms spent in compaction + Delaunay (worker threads) and ms the issuing thread was
 * blocked waiting for one (0 = fully hidden behind device work). */
int acmmp_scan_host_stats(acmmp_scan *s, double *prior_host_ms, double *wait_ms);
/* acmmp_count_ncc / acmmp_ncc_count over every buffer set of the scan */
int acmmp_scan_count_ncc(acmmp_scan *s, int enable);
int acmmp_scan_ncc_count(acmmp_scan *s, uint64_t *init_ncc, uint64_t *sweep_ncc);
/* start the schedule over (level 0 next) keeping images and buffers */
int acmmp_scan_reset(acmmp_scan *s);
/* CUDA events on the scan's stream: record mark i (0..7); elapsed ms between
 * two recorded marks (waits for mark j). */
int acmmp_scan_mark(acmmp_scan *s, int i);
int acmmp_scan_elapsed(acmmp_scan *s, int i, int j, float *ms);

/* ==== depth-map fusion =========================================================
 * RunFusion (/root/reference/src/acmmp_definitions.cpp:828-1043) without the file
 * I/O: the caller hands over, per view, the depth map the reference reads from
 * depths[_geom].dmb, the normals of normals.dmb, the colour image rescaled to the
 * depth map (RescaleImageAndCamera, :864) as 3 bytes per pixel in stored channel
 * order (optional) and the optional depth mask (1 = ignore the pixel; the
 * reference derives it as `png < 128`, :889).  `cams` are the cameras rescaled to
 * the depth maps.  pair_offsets / pair_ids: the source views of every view as
 * indices (CSR), like acmmp_scan_create.  Point order and the per-view used_list
 * behaviour of the reference (:936) are preserved. */
typedef struct acmmp_fusion acmmp_fusion;
int acmmp_fusion_create(acmmp_fusion **out, int device, int n_views, const acmmp_camera *cams,
                        const int32_t *pair_offsets, const int32_t *pair_ids);
void acmmp_fusion_destroy(acmmp_fusion *f);
const char *acmmp_fusion_last_error(const acmmp_fusion *f);
/* host pointers, or device pointers on the fusion's device when on_device != 0 */
int acmmp_fusion_set_view(acmmp_fusion *f, int view, const float *depth, const float *normal3,
                          const uint8_t *color3, const uint8_t *mask, int on_device);
/* consistency_scalar / con_num_thresh: main_ACMMP.cpp:17-18 (0.3, 1).  May be
 * called again (the masks restart from the ones given to set_view). */
int acmmp_fusion_run(acmmp_fusion *f, float consistency_scalar, int con_num_thresh, int64_t *num_points);
/* points9: num_points x (x, y, z, nx, ny, nz, c0, c1, c2) floats in the order the
 * reference appends them to the PLY (view by view, row-major); per_view_counts:
 * n_views entries ("approved N pts", :1036).  Either may be NULL. */
int acmmp_fusion_get_points(acmmp_fusion *f, float *points9, int64_t capacity, int64_t *per_view_counts);
/* working mask of a view after the run (what later views see) */
int acmmp_fusion_get_mask(acmmp_fusion *f, int view, uint8_t *mask);
/* device ms of the last run: {score, greedy fixed point, compaction}; fixed-point rounds per view */
int acmmp_fusion_stats(acmmp_fusion *f, double ms[3], int32_t *rounds_per_view);

/* Prior-aware fusion (RunPriorAwareFusion, /root/reference/src/acmmp_definitions.cpp:443-826; the
 * reference's --multi_fusion / --force_fusion modes, main_ACMMP.cpp:178-189).  Every view carries a
 * second (depth, normal) set besides the one of acmmp_fusion_set_view: set 0 is what the reference
 * reads from its fusion folder, set 1 what it reads from the run's own output folder (:607-643).
 * A pixel tries both hypotheses against both maps of every source view; the one with more consistent
 * sources wins, a hypothesis passing alone needs single_match_penalty more consistent sources. */
int acmmp_fusion_set_view_second(acmmp_fusion *f, int view, const float *depth2, const float *normal3_2, int on_device);
int acmmp_fusion_run_prior_aware(acmmp_fusion *f, float consistency_scalar, int num_consistent_thresh,
                                 int single_match_penalty, int64_t *num_points);

/* ==== JPEG decode on the device (nvJPEG) ======================================
 * What the reference gets from cv::imread: mode 0 = IMREAD_GRAYSCALE (the luma
 * plane, ACMMP.cpp:534-541), mode 1 = IMREAD_COLOR (interleaved BGR,
 * acmmp_definitions.cpp:857).  `out` is a HOST buffer of width*height*(1|3)
 * bytes; with out == NULL only the size is returned. */
int acmmp_jpeg_info(const uint8_t *jpeg, size_t nbytes, int *width, int *height, int *components);
int acmmp_decode_jpeg(int device, const uint8_t *jpeg, size_t nbytes, int mode, uint8_t *out,
                      size_t capacity, int *width, int *height);
const char *acmmp_jpeg_last_error(void);

/* ==== input converter: view-pair scores ========================================
 * Replaces calc_score and the pair loop of the reference's converter
 * (/root/reference/python_scripts/colmap2mvsnet_acm.py:280-302, :399-413).
 * Image v owns the keypoints obs_points[obs_offsets[v] .. obs_offsets[v+1]): the index of the
 * keypoint's 3-D point in xyz (n_points x 3 doubles) or -1 when it has none (point3D_ids with the
 * sparse model's ids renumbered 0..n_points-1).  centers: n_images x 3 camera centres (-R^T t).
 * scores (n_images x n_images doubles, symmetric, zero diagonal): number of keypoints of image
 * min(i, j) whose point is also seen by the other image, or 0 when the angle at rank
 * int(quantile * n) of the sorted triangulation angles is below min_angle_deg (reference: 0.75, 1).
 * device_ms (nullable): device time of sort + scoring, inputs resident. */
int acmmp_pair_scores(int device, int n_images, const int64_t *obs_offsets, const int32_t *obs_points,
                      int64_t n_points, const double *xyz, const double *centers, double min_angle_deg,
                      double quantile, double *scores, double *device_ms);
const char *acmmp_pairs_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* ACMMP_B200_H_ */
