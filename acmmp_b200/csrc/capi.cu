// capi.cu - the C ABI of include/acmmp_b200.h: context, uploads, run sequence.
//
// Host-side counterpart of the reference's class ACMMP (ACMMP.cpp:107-152,
// 447-866) and ACMMP::RunPatchMatch (ACMMP.cu:1378-1456), restructured as a
// reusable context: buffers, textures and the stream live across problems (the
// reference builds and tears down everything per view per stage,
// acmmp_definitions.cpp:260-269), no cudaDeviceSynchronize between kernels
// (stream order suffices), nothing exits the process.
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <chrono>
#include <cstring>
#include <string>
#include <cuda.h>   // CUtensorMap + enums (types only; the encoder comes from cudaGetDriverEntryPoint)

#include <nvtx3/nvToolsExt.h>   // header-only; ranges cost nothing unless a tool is attached

#include <array>
#include <vector>

#include "../../include/acmmp_b200.h"
#include "kernels.h"
#include "prior.h"

using namespace acmmp;

namespace {

thread_local std::string g_global_error;

struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

// NVTX push/pop range over a scope on the issuing thread, named "acmmp:<task>:L<level>:v<view>" (no '/' or spaces:
// ncu's filter syntax uses them): `ncu --nvtx --nvtx-include "acmmp:geom1:L1:v1/" ...` profiles exactly the launches of
// that task; any NVTX-aware timeline tool shows the schedule
struct NvtxRange {
    explicit NvtxRange(const char *fmt, ...) {
        char buf[96];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof buf, fmt, ap);
        va_end(ap);
        nvtxRangePushA(buf);
    }
    NvtxRange(const NvtxRange &) = delete;
    NvtxRange &operator=(const NvtxRange &) = delete;
    ~NvtxRange() { nvtxRangePop(); }
};

// a device buffer that frees itself: the stand-alone operators return from the middle on errors
struct ScopedBuf : DevBuf {
    ScopedBuf() = default;
    ScopedBuf(const ScopedBuf &) = delete;
    ScopedBuf &operator=(const ScopedBuf &) = delete;
    ~ScopedBuf() { release(); }
};

// page-locked host staging (async D2H / H2D without implicit stream syncs)
struct PinnedBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaHostAlloc(&p, bytes, cudaHostAllocDefault);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release() {
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
    }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

// All views of a problem live in ONE layered cudaArray (layer i = view i) behind
// one texture object, so kernels sample any view through a warp-uniform handle.
#ifndef ACMMP_BOX_STAGING_DEFAULT
#define ACMMP_BOX_STAGING_DEFAULT 0
#endif

struct LayeredImages {
    cudaArray_t arr = nullptr;
    cudaTextureObject_t tex = 0;
    int w = 0, h = 0, layers = 0;
    bool u8 = false;   // R8 unorm texels (integer grey levels, read as normalised float) instead of R32F
    // R8 only: the same bytes pitch-linear (x, y, layer) under a TMA tensor map, for the sweep kernel's box staging
    unsigned char *linear = nullptr;
    CUtensorMap map;
    bool has_map = false;
};

// cuTensorMapEncodeTiled through the runtime (no link against libcuda); null when the driver lacks it
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                                  const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
        (void)cudaGetLastError();
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

}  // namespace

struct acmmp_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool owns_stream = true;
    // event quads (start, after init, after sweeps, after tail) of every acmmp_run since the
    // last acmmp_timing_collect; the newest quad also serves acmmp_last_timing
    std::vector<cudaEvent_t> evpool;
    std::vector<int> quad_tag;   // run_tag of every quad (the scan tags runs with their pyramid level)
    int run_tag = 0;
    size_t ev_used = 0;
    size_t last_quad = 0;
    bool have_quad = false;
    std::string error;
    acmmp_params params;
    int compat = ACMMP_COMPAT_REF_SM100;
    int split = -1;        // half-finished ping-pong pair: colour whose newest pixels sit in planes_snap (capi_run.cu.inc)
    int wave_blocks = 0;   // acmmp_set_wave_emulation: resident blocks of the reference launch (0 = off)
    int cold_left = 0;     // half-sweeps still counted as cold after a random initialisation (sweep_kernel<..., kPairMatrix>)
    int n = 0;   // images in the problem
    int W = 0, H = 0;
    std::vector<acmmp_camera> cams;
    LayeredImages views;
    // ACMMP_TEX_R32F: always float texels.  ACMMP_TEX_AUTO: R8 unorm texels when every image of the
    // problem holds integer grey levels 0..255 (bit-identical results, ncc.cuh), else float texels
    int tex_format = ACMMP_TEX_R32F;
    bool ext_tex_u8 = false;
    CUtensorMap ext_map;           // copy of the caller's tensor map over the R8 layers (box staging)
    bool ext_has_map = false;
    int ext_w = 0, ext_h = 0;
    int box_staging = -1;          // acmmp_set_box_staging: -1 = ACMMP_BOX_STAGING from the environment (default off)
    DevBuf stage_u8, u8_flag;
    // resident path: all views of a scan live in one layered array owned by the scan;
    // a problem then only names layers (no per-problem upload)
    cudaTextureObject_t ext_tex = 0;
    const float *ext_ref_image = nullptr;
    std::vector<int> layers;   // layer of view i in the texture in use
    std::vector<DevBuf> depth_maps;   // per slot
    std::vector<int> depth_set;
    DevBuf ref_image, d_cams, d_depth_ptrs, d_rc, d_pairs, stage_in, stage_pad;
    DevBuf planes, costs, pre_costs, sel, rng, planes_snap, costs_snap;
    DevBuf seeded, prior_planes, plane_masks, scaled_planes, tmp_a, tmp_b, tmp_c, tri_buf;
    // planar-prior construction: pinned staging of the support cells / triangles, the
    // event that marks the cells' arrival on the host, and the triangulation result
    PinnedBuf h_cells, h_tris;
    std::vector<int32_t> tri_scratch;
    cudaEvent_t ev_cells = nullptr;
    int prior_ns = 0, prior_nt = 0;
    double prior_host_ms = 0.0;   // host time spent in compaction + Delaunay (sum)
    bool have_seeded = false, have_prior = false, have_scaled = false, have_warm = false;
    bool pairs_dirty = true;
    bool tail_done = false;
    uint64_t launches = 0;
    uint64_t init_ncc = 0;       // NCC evaluations issued by init_kernel since the last acmmp_ncc_count
    int count_ncc = 0;           // sweep_kernel instrumentation on / off
    DevBuf ncc_counter;
    float timing[4] = {0, 0, 0, 0};
};

namespace {

int fail(acmmp_ctx *ctx, int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->error = buf;
    else g_global_error = buf;
    return code;
}

#define CU(ctx, call)                                                                          \
    do {                                                                                       \
        cudaError_t e__ = (call);                                                              \
        if (e__ != cudaSuccess)                                                                \
            return fail(ctx, ACMMP_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                        __FILE__, __LINE__);                                                   \
    } while (0)

// Every entry point starts here.  cudaGetLastError() drops a non-sticky error that OTHER code in the
// process may have left in this thread's runtime state (the launch wrappers read it right after
// their launch and would otherwise report somebody else's failure); sticky errors survive.
int bind_device(acmmp_ctx *ctx) {
    CU(ctx, cudaSetDevice(ctx->device));
    (void)cudaGetLastError();
    return 0;
}

void destroy_views(LayeredImages &v) {
    if (v.tex) cudaDestroyTextureObject(v.tex);
    if (v.arr) cudaFreeArray(v.arr);
    if (v.linear) cudaFree(v.linear);
    v = LayeredImages();
}

// layered float array + texture object with the reference's sampling state
// (ACMMP.cpp:646-663): linear filter, unnormalised coordinates, "wrap" (not
// honoured for unnormalised coordinates: the hardware clamps).
cudaError_t make_layered(LayeredImages &v, int w, int h, int layers, bool u8) {
    destroy_views(v);
    cudaChannelFormatDesc desc = u8 ? cudaCreateChannelDesc(8, 0, 0, 0, cudaChannelFormatKindUnsigned)
                                    : cudaCreateChannelDesc(32, 0, 0, 0, cudaChannelFormatKindFloat);
    cudaError_t e = cudaMalloc3DArray(&v.arr, &desc, make_cudaExtent(w, h, layers), cudaArrayLayered);
    if (e != cudaSuccess) return e;
    cudaResourceDesc res;
    memset(&res, 0, sizeof res);
    res.resType = cudaResourceTypeArray;
    res.res.array.array = v.arr;
    cudaTextureDesc td;
    memset(&td, 0, sizeof td);
    td.addressMode[0] = cudaAddressModeWrap;   // as ACMMP.cpp:655-661; clamps for unnormalised coordinates
    td.addressMode[1] = cudaAddressModeWrap;
    td.filterMode = cudaFilterModeLinear;
    td.readMode = u8 ? cudaReadModeNormalizedFloat : cudaReadModeElementType;
    td.normalizedCoords = 0;
    e = cudaCreateTextureObject(&v.tex, &res, &td, nullptr);
    if (e != cudaSuccess) return e;
    v.w = w;
    v.h = h;
    v.layers = layers;
    v.u8 = u8;
    // linear copy + tensor map (box kBoxW x kBoxH x 1); rows must be a multiple of 16 bytes
    if (u8 && w % 16 == 0 && w >= acmmp::kBoxW && h >= acmmp::kBoxH && encode_tiled_fn()) {
        if (cudaMalloc(&v.linear, (size_t)w * h * layers) != cudaSuccess) {
            (void)cudaGetLastError();
            v.linear = nullptr;
            return cudaSuccess;
        }
        const cuuint64_t dims[3] = {(cuuint64_t)w, (cuuint64_t)h, (cuuint64_t)layers};
        const cuuint64_t strides[2] = {(cuuint64_t)w, (cuuint64_t)w * h};
        const cuuint32_t box[3] = {acmmp::kBoxW, acmmp::kBoxH, 1}, estr[3] = {1, 1, 1};
        v.has_map = encode_tiled_fn()(&v.map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, v.linear, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                      CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
    }
    return cudaSuccess;
}

// bytes of one layer -> layer `layer` of an R8 array
cudaError_t copy_layer_u8(const LayeredImages &v, int layer, const uint8_t *d_src, cudaStream_t st) {
    if (v.linear) {
        cudaError_t e = cudaMemcpyAsync(v.linear + (size_t)layer * v.w * v.h, d_src, (size_t)v.w * v.h, cudaMemcpyDeviceToDevice, st);
        if (e != cudaSuccess) return e;
    }
    cudaMemcpy3DParms cp;
    memset(&cp, 0, sizeof cp);
    cp.srcPtr = make_cudaPitchedPtr(const_cast<uint8_t *>(d_src), (size_t)v.w, v.w, v.h);
    cp.dstArray = v.arr;
    cp.dstPos = make_cudaPos(0, 0, layer);
    cp.extent = make_cudaExtent(v.w, v.h, 1);
    cp.kind = cudaMemcpyDeviceToDevice;
    return cudaMemcpy3DAsync(&cp, st);
}

int ensure_views(acmmp_ctx *ctx, int w, int h, int layers, bool u8 = false) {
    LayeredImages &v = ctx->views;
    if (v.arr && v.w == w && v.h == h && v.layers >= layers && v.u8 == u8) return 0;
    if (u8) {
        CU(ctx, make_layered(v, w, h, layers, true));
        return 0;
    }
    destroy_views(v);
    cudaChannelFormatDesc desc = cudaCreateChannelDesc(32, 0, 0, 0, cudaChannelFormatKindFloat);
    CU(ctx, cudaMalloc3DArray(&v.arr, &desc, make_cudaExtent(w, h, layers), cudaArrayLayered));
    cudaResourceDesc res;
    memset(&res, 0, sizeof res);
    res.resType = cudaResourceTypeArray;
    res.res.array.array = v.arr;
    cudaTextureDesc td;
    memset(&td, 0, sizeof td);
    td.addressMode[0] = cudaAddressModeWrap;
    td.addressMode[1] = cudaAddressModeWrap;
    td.filterMode = cudaFilterModeLinear;
    td.readMode = cudaReadModeElementType;
    td.normalizedCoords = 0;
    CU(ctx, cudaCreateTextureObject(&v.tex, &res, &td, nullptr));
    v.w = w;
    v.h = h;
    v.layers = layers;
    v.u8 = false;
    return 0;
}

// ACMMP_TEX_AUTO: all views as bytes into an R8 array; *ok = 0 when some image does not hold
// integer grey levels (the caller then uploads float texels)
int upload_views_u8(acmmp_ctx *ctx, int n, const acmmp_camera *cams, const float *const *images, bool on_device, int maxw,
                    int maxh, int *ok) {
    *ok = 0;
    CU(ctx, ctx->u8_flag.ensure(sizeof(int)));
    CU(ctx, cudaMemsetAsync(ctx->u8_flag.p, 0, sizeof(int), ctx->stream));
    // every image is checked before the array changes format (a float problem keeps its array)
    std::vector<const float *> d_img(n);
    size_t total = 0;
    for (int i = 0; i < n; ++i) total += (size_t)cams[i].width * cams[i].height;
    if (!on_device) {
        CU(ctx, ctx->stage_in.ensure(total * sizeof(float)));
        size_t off = 0;
        for (int i = 0; i < n; ++i) {
            const size_t P = (size_t)cams[i].width * cams[i].height;
            CU(ctx, cudaMemcpyAsync(ctx->stage_in.as<float>() + off, images[i], P * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
            d_img[i] = ctx->stage_in.as<float>() + off;
            off += P;
        }
    } else {
        for (int i = 0; i < n; ++i) d_img[i] = images[i];
    }
    for (int i = 0; i < n; ++i) {
        CU(ctx, launch_check_u8(d_img[i], (size_t)cams[i].width * cams[i].height, ctx->u8_flag.as<int>(), ctx->stream));
        ctx->launches++;
    }
    int bad = 0;
    CU(ctx, cudaMemcpyAsync(&bad, ctx->u8_flag.p, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    if (bad) return 0;
    int rcode = ensure_views(ctx, maxw, maxh, n, true);
    if (rcode) return rcode;
    CU(ctx, ctx->stage_u8.ensure((size_t)maxw * maxh * n));
    for (int i = 0; i < n; ++i) {
        uint8_t *dst = ctx->stage_u8.as<uint8_t>() + (size_t)i * maxw * maxh;
        CU(ctx, launch_to_u8_replicate(d_img[i], cams[i].width, cams[i].height, dst, maxw, maxh, ctx->stream));
        ctx->launches++;
        CU(ctx, copy_layer_u8(ctx->views, i, dst, ctx->stream));
    }
    *ok = 1;
    return 0;
}

int upload_layer(acmmp_ctx *ctx, int layer, const float *img, int w, int h, bool on_device) {
    const LayeredImages &v = ctx->views;
    const float *d_src = img;
    if (!on_device) {
        CU(ctx, ctx->stage_in.ensure((size_t)w * h * sizeof(float)));
        CU(ctx, cudaMemcpyAsync(ctx->stage_in.p, img, (size_t)w * h * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
        d_src = ctx->stage_in.as<float>();
    }
    if (w != v.w || h != v.h) {   // smaller than the layer: replicate the edges
        CU(ctx, ctx->stage_pad.ensure((size_t)v.w * v.h * sizeof(float)));
        CU(ctx, launch_pad_replicate(d_src, w, h, ctx->stage_pad.as<float>(), v.w, v.h, ctx->stream));
        ctx->launches++;
        d_src = ctx->stage_pad.as<float>();
    }
    cudaMemcpy3DParms cp;
    memset(&cp, 0, sizeof cp);
    cp.srcPtr = make_cudaPitchedPtr(const_cast<float *>(d_src), (size_t)v.w * sizeof(float), v.w, v.h);
    cp.dstArray = v.arr;
    cp.dstPos = make_cudaPos(0, 0, layer);
    cp.extent = make_cudaExtent(v.w, v.h, 1);
    cp.kind = cudaMemcpyDeviceToDevice;
    CU(ctx, cudaMemcpy3DAsync(&cp, ctx->stream));
    // staging buffers are reused by the next layer
    if (!on_device || w != v.w || h != v.h) CU(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

int init_mode(const acmmp_params &p) {
    if (!p.geom_consistency && !p.hierarchy && !p.seeded) return 0;
    if (p.seeded) return 1;
    if (p.planar_prior) return 2;
    if (p.upsample) return 3;
    return 4;
}

int refresh_pairs(acmmp_ctx *ctx) {
    if (!ctx->pairs_dirty) return 0;
    const int n = ctx->n;
    std::vector<const float *> dptr(n, nullptr);
    for (int i = 0; i < n; ++i)
        if (i < (int)ctx->depth_maps.size() && ctx->depth_set[i]) dptr[i] = ctx->depth_maps[i].as<float>();
    CU(ctx, ctx->d_rc.ensure(ref_const_size()));
    CU(ctx, ctx->d_pairs.ensure(pair_const_size() * (n > 1 ? n - 1 : 1)));
    CU(ctx, launch_setup_pairs(ctx->cams.data(), n, dptr.data(), ctx->layers.data(), ctx->params.depth_min, ctx->params.depth_max,
                               ctx->d_rc.as<RefConst>(), ctx->d_pairs.as<PairConst>(), ctx->stream));
    ctx->launches++;
    ctx->pairs_dirty = false;
    return 0;
}

// per-problem state shared by the upload path (acmmp_set_problem) and the resident path
int problem_common(acmmp_ctx *ctx, int n, const acmmp_camera *cams) {
    ctx->W = cams[0].width;
    ctx->H = cams[0].height;
    const size_t P = (size_t)ctx->W * ctx->H;
    CU(ctx, ctx->planes.ensure(P * sizeof(float4)));
    CU(ctx, ctx->planes_snap.ensure(P * sizeof(float4)));
    CU(ctx, ctx->costs.ensure(P * sizeof(float)));
    CU(ctx, ctx->costs_snap.ensure(P * sizeof(float)));
    CU(ctx, ctx->pre_costs.ensure(P * sizeof(float)));
    CU(ctx, ctx->sel.ensure(P * sizeof(uint32_t)));
    CU(ctx, ctx->rng.ensure(P * 6 * sizeof(uint32_t)));
    // depth range widening + disparity range (ACMMP.cpp:600-606)
    ctx->params.depth_min = cams[0].depth_min * 0.6f;
    ctx->params.depth_max = cams[0].depth_max * 1.2f;
    ctx->params.num_images = n;
    ctx->params.disparity_min = cams[0].K[0] * ctx->params.baseline / ctx->params.depth_max;
    ctx->params.disparity_max = cams[0].K[0] * ctx->params.baseline / ctx->params.depth_min;
    ctx->depth_maps.resize(n);
    ctx->depth_set.assign(n, 0);
    ctx->have_seeded = ctx->have_prior = ctx->have_scaled = ctx->have_warm = false;
    ctx->params.seeded = 0;
    ctx->pairs_dirty = true;
    ctx->tail_done = false;
    return 0;
}

// resident path: the views already sit in layers of `tex` (owned by the caller);
// ref_image = row-major reference image on the device (borrowed)
int set_problem_resident(acmmp_ctx *ctx, int n, const acmmp_camera *cams, cudaTextureObject_t tex, bool tex_u8,
                         const float *ref_image, const int *layers, const LayeredImages *views = nullptr) {
    if (n < 2 || n - 1 > ACMMP_MAX_SRC_VIEWS) return fail(ctx, ACMMP_ERR_INVALID, "need 1 reference + 1..%d source views, got n=%d", ACMMP_MAX_SRC_VIEWS, n);
    if (bind_device(ctx)) return ACMMP_ERR_CUDA;
    ctx->n = n;
    ctx->cams.assign(cams, cams + n);
    ctx->ext_tex = tex;
    ctx->ext_tex_u8 = tex_u8;
    ctx->ext_has_map = views && views->has_map && tex_u8;
    if (ctx->ext_has_map) {
        ctx->ext_map = views->map;
        ctx->ext_w = views->w;
        ctx->ext_h = views->h;
    }
    ctx->ext_ref_image = ref_image;
    ctx->layers.assign(layers, layers + n);
    return problem_common(ctx, n, cams);
}

}  // namespace

extern "C" {

const char *acmmp_version(void) { return "acmmp_b200 0.1 sm_100a"; }

void acmmp_default_params(acmmp_params *p) {
    if (!p) return;
    memset(p, 0, sizeof *p);
    p->max_iterations = 2;
    p->patch_size = 11;
    p->num_images = 5;
    p->max_image_size = 3200;
    p->radius_increment = 2;
    p->sigma_spatial = 5.0f;
    p->sigma_color = 3.0f;
    p->top_k = 4;
    p->baseline = 0.54f;
    p->depth_min = 0.0f;
    p->depth_max = 1.0f;
    p->disparity_min = 0.0f;
    p->disparity_max = 1.0f;
}

int acmmp_create(acmmp_ctx **out, int device) {
    if (!out) return fail(nullptr, ACMMP_ERR_INVALID, "acmmp_create: null out pointer");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(nullptr, ACMMP_ERR_NO_DEVICE, "no CUDA device: %s", cudaGetErrorString(e));
    if (device < 0 || device >= count) return fail(nullptr, ACMMP_ERR_INVALID, "device %d out of range", device);
    acmmp_ctx *ctx = new acmmp_ctx();
    ctx->device = device;
    acmmp_default_params(&ctx->params);
    if (cudaSetDevice(device) != cudaSuccess || cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete ctx;
        return fail(nullptr, ACMMP_ERR_CUDA, "cannot create stream on device %d", device);
    }
    *out = ctx;
    return 0;
}

}  // extern "C"

namespace {
// a further context on the SAME stream as `first`: a second set of per-problem buffers so
// that the host can prepare / finish one problem while the device works on another, with
// all kernels still executing in issue order on one stream (no kernel concurrency, so
// per-kernel event times stay meaningful)
int create_sibling(acmmp_ctx **out, const acmmp_ctx *first) {
    acmmp_ctx *ctx = new acmmp_ctx();
    ctx->device = first->device;
    ctx->stream = first->stream;
    ctx->owns_stream = false;
    acmmp_default_params(&ctx->params);
    ctx->compat = first->compat;
    ctx->tex_format = first->tex_format;
    *out = ctx;
    return 0;
}
}  // namespace

extern "C" {

void acmmp_destroy(acmmp_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    destroy_views(ctx->views);
    for (auto &d : ctx->depth_maps) d.release();
    DevBuf *bufs[] = {&ctx->ref_image, &ctx->d_cams, &ctx->stage_in, &ctx->stage_pad, &ctx->d_depth_ptrs, &ctx->d_rc, &ctx->d_pairs,
                      &ctx->planes, &ctx->costs, &ctx->pre_costs, &ctx->sel, &ctx->rng, &ctx->planes_snap,
                      &ctx->costs_snap, &ctx->seeded, &ctx->prior_planes, &ctx->plane_masks, &ctx->scaled_planes,
                      &ctx->tmp_a, &ctx->tmp_b, &ctx->tmp_c, &ctx->tri_buf, &ctx->ncc_counter, &ctx->stage_u8, &ctx->u8_flag};
    for (auto *b : bufs) b->release();
    for (auto &ev : ctx->evpool)
        if (ev) cudaEventDestroy(ev);
    if (ctx->ev_cells) cudaEventDestroy(ctx->ev_cells);
    ctx->h_cells.release();
    ctx->h_tris.release();
    if (ctx->stream && ctx->owns_stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char *acmmp_last_error(const acmmp_ctx *ctx) { return ctx ? ctx->error.c_str() : g_global_error.c_str(); }

int acmmp_set_geom_consistency(acmmp_ctx *ctx, int multi_geometry) {
    if (!ctx) return ACMMP_ERR_INVALID;
    ctx->params.geom_consistency = 1;
    ctx->params.max_iterations = 2;
    if (multi_geometry) ctx->params.multi_geometry = 1;
    return 0;
}
int acmmp_set_hierarchy(acmmp_ctx *ctx) {
    if (!ctx) return ACMMP_ERR_INVALID;
    ctx->params.hierarchy = 1;
    return 0;
}
int acmmp_set_planar_prior_params(acmmp_ctx *ctx) {
    if (!ctx) return ACMMP_ERR_INVALID;
    ctx->params.planar_prior = 1;
    return 0;
}
int acmmp_set_compat_mode(acmmp_ctx *ctx, int mode) {
    if (!ctx) return ACMMP_ERR_INVALID;
    if (mode != ACMMP_COMPAT_REF_SM100 && mode != ACMMP_COMPAT_UPSTREAM)
        return fail(ctx, ACMMP_ERR_INVALID, "unknown compat mode %d", mode);
    ctx->compat = mode;
    return 0;
}
int acmmp_set_wave_emulation(acmmp_ctx *ctx, int resident_blocks) {
    if (!ctx) return ACMMP_ERR_INVALID;
    if (resident_blocks < 0) return fail(ctx, ACMMP_ERR_INVALID, "resident_blocks must be >= 0");
    ctx->wave_blocks = resident_blocks;
    return 0;
}
int acmmp_set_texture_format(acmmp_ctx *ctx, int format) {
    if (!ctx) return ACMMP_ERR_INVALID;
    if (format != ACMMP_TEX_R32F && format != ACMMP_TEX_AUTO)
        return fail(ctx, ACMMP_ERR_INVALID, "unknown texture format %d", format);
    ctx->tex_format = format;
    return 0;
}
int acmmp_texture_is_u8(const acmmp_ctx *ctx) {
    if (!ctx) return 0;
    return ctx->ext_tex ? (ctx->ext_tex_u8 ? 1 : 0) : (ctx->views.u8 ? 1 : 0);
}
int acmmp_get_params(const acmmp_ctx *ctx, acmmp_params *out) {
    if (!ctx || !out) return ACMMP_ERR_INVALID;
    *out = ctx->params;
    return 0;
}
int acmmp_width(const acmmp_ctx *ctx) { return ctx ? ctx->W : 0; }
int acmmp_height(const acmmp_ctx *ctx) { return ctx ? ctx->H : 0; }
uint64_t acmmp_launch_count(const acmmp_ctx *ctx) { return ctx ? ctx->launches : 0; }

int acmmp_sync(acmmp_ctx *ctx) {
    if (!ctx) return ACMMP_ERR_INVALID;
    if (bind_device(ctx)) return ACMMP_ERR_CUDA;
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

int acmmp_set_problem(acmmp_ctx *ctx, int n, const acmmp_camera *cams, const float *const *images,
                      int images_on_device) {
    if (!ctx) return ACMMP_ERR_INVALID;
    if (n < 2 || n - 1 > ACMMP_MAX_SRC_VIEWS)
        return fail(ctx, ACMMP_ERR_INVALID, "need 1 reference + 1..%d source views, got n=%d", ACMMP_MAX_SRC_VIEWS, n);
    if (!cams || !images) return fail(ctx, ACMMP_ERR_INVALID, "null cameras/images");
    for (int i = 0; i < n; ++i)
        if (cams[i].width <= 0 || cams[i].height <= 0 || !images[i])
            return fail(ctx, ACMMP_ERR_INVALID, "view %d: bad size %dx%d or null image", i, cams[i].width, cams[i].height);
    if (bind_device(ctx)) return ACMMP_ERR_CUDA;
    ctx->n = n;
    ctx->cams.assign(cams, cams + n);
    int maxw = 0, maxh = 0;
    for (int i = 0; i < n; ++i) {
        if (cams[i].width > maxw) maxw = cams[i].width;
        if (cams[i].height > maxh) maxh = cams[i].height;
    }
    int as_u8 = 0;
    if (ctx->tex_format == ACMMP_TEX_AUTO) {
        int rcode = upload_views_u8(ctx, n, cams, images, images_on_device != 0, maxw, maxh, &as_u8);
        if (rcode) return rcode;
    }
    if (!as_u8) {
        int rcode = ensure_views(ctx, maxw, maxh, n);
        if (rcode) return rcode;
        for (int i = 0; i < n; ++i) {
            rcode = upload_layer(ctx, i, images[i], cams[i].width, cams[i].height, images_on_device != 0);
            if (rcode) return rcode;
        }
    }
    ctx->ext_tex = 0;
    ctx->ext_tex_u8 = false;
    ctx->ext_has_map = false;
    ctx->ext_ref_image = nullptr;
    ctx->layers.resize(n);
    for (int i = 0; i < n; ++i) ctx->layers[i] = i;
    {
        const size_t P0 = (size_t)cams[0].width * cams[0].height;
        CU(ctx, ctx->ref_image.ensure(P0 * sizeof(float)));
        CU(ctx, cudaMemcpyAsync(ctx->ref_image.p, images[0], P0 * sizeof(float),
                                images_on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, ctx->stream));
    }
    {
        int rcode = problem_common(ctx, n, cams);
        if (rcode) return rcode;
    }
    // host images may be freed by the caller right after this returns
    if (!images_on_device) CU(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

int acmmp_set_depth_map(acmmp_ctx *ctx, int slot, const float *depth, int w, int h, int on_device) {
    if (!ctx || !depth) return ACMMP_ERR_INVALID;
    if (slot < 0 || slot >= ctx->n) return fail(ctx, ACMMP_ERR_INVALID, "depth slot %d out of range", slot);
    if (w != ctx->cams[slot].width || h != ctx->cams[slot].height)
        return fail(ctx, ACMMP_ERR_INVALID, "depth map %dx%d does not match view %d (%dx%d)", w, h, slot,
                    ctx->cams[slot].width, ctx->cams[slot].height);
    if (bind_device(ctx)) return ACMMP_ERR_CUDA;
    const size_t bytes = (size_t)w * h * sizeof(float);
    CU(ctx, ctx->depth_maps[slot].ensure(bytes));
    CU(ctx, cudaMemcpyAsync(ctx->depth_maps[slot].p, depth, bytes,
                            on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, ctx->stream));
    if (!on_device) CU(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->depth_set[slot] = 1;
    ctx->pairs_dirty = true;
    return 0;
}

int acmmp_set_warm_start(acmmp_ctx *ctx, const float *depth, const float *normal3, const float *cost) {
    if (!ctx || !depth || !normal3 || !cost) return ACMMP_ERR_INVALID;
    if (!ctx->W) return fail(ctx, ACMMP_ERR_STATE, "acmmp_set_problem first");
    if (bind_device(ctx)) return ACMMP_ERR_CUDA;
    const size_t P = (size_t)ctx->W * ctx->H;
    CU(ctx, ctx->tmp_a.ensure(P * 3 * sizeof(float)));
    CU(ctx, ctx->tmp_b.ensure(P * sizeof(float)));
    CU(ctx, cudaMemcpyAsync(ctx->tmp_a.p, normal3, P * 3 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, cudaMemcpyAsync(ctx->tmp_b.p, depth, P * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, cudaMemcpyAsync(ctx->costs.p, cost, P * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, launch_pack_planes(ctx->planes.as<float4>(), (int)P, ctx->tmp_a.as<float>(), ctx->tmp_b.as<float>(), 0, ctx->stream));
    ctx->launches++;
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->have_warm = true;
    return 0;
}

int acmmp_set_hierarchy_inputs(acmmp_ctx *ctx, const float *coarse_normal3, const float *coarse_cost, int cw, int ch,
                               const float *fine_depth) {
    if (!ctx || !coarse_normal3 || !coarse_cost || !fine_depth) return ACMMP_ERR_INVALID;
    if (!ctx->W) return fail(ctx, ACMMP_ERR_STATE, "acmmp_set_problem first");
    if (cw <= 0 || ch <= 0) return fail(ctx, ACMMP_ERR_INVALID, "bad coarse size %dx%d", cw, ch);
    if (bind_device(ctx)) return ACMMP_ERR_CUDA;
    const size_t P = (size_t)ctx->W * ctx->H, CP = (size_t)cw * ch;
    CU(ctx, ctx->tmp_a.ensure(CP * 3 * sizeof(float)));
    CU(ctx, ctx->tmp_b.ensure(CP * sizeof(float)));
    CU(ctx, ctx->tmp_c.ensure(P * sizeof(float)));
    CU(ctx, ctx->scaled_planes.ensure(CP * sizeof(float4)));
    CU(ctx, cudaMemcpyAsync(ctx->tmp_a.p, coarse_normal3, CP * 3 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, cudaMemcpyAsync(ctx->tmp_b.p, coarse_cost, CP * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, cudaMemcpyAsync(ctx->tmp_c.p, fine_depth, P * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    // coarse (world normal, COST in .w): ACMMP.cpp:777-792 with upsample = true.
    // The reference decides `upsample` by comparing the coarse width with the image
    // ROWS (ACMMP.cpp:766), i.e. it is true unless cw == H && ch == W.
    ctx->params.upsample = (cw != ctx->H || ch != ctx->W) ? 1 : 0;
    ctx->params.scaled_cols = (float)cw;
    ctx->params.scaled_rows = (float)ch;
    CU(ctx, launch_pack_planes(ctx->scaled_planes.as<float4>(), (int)CP, ctx->tmp_a.as<float>(), ctx->tmp_b.as<float>(), 0, ctx->stream));
    // fine planes: only .w (JBU depth) is defined by the reference (ACMMP.cpp:797-804
    // leaves x,y,z uninitialised); zeros are the defined stand-in.
    CU(ctx, launch_pack_planes(ctx->planes.as<float4>(), (int)P, nullptr, ctx->tmp_c.as<float>(), 1, ctx->stream));
    ctx->launches += 2;
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->have_scaled = true;
    return 0;
}

int acmmp_set_seeded_planes(acmmp_ctx *ctx, const float *planes4) {
    if (!ctx || !planes4) return ACMMP_ERR_INVALID;
    if (!ctx->W) return fail(ctx, ACMMP_ERR_STATE, "acmmp_set_problem first");
    if (bind_device(ctx)) return ACMMP_ERR_CUDA;
    const size_t P = (size_t)ctx->W * ctx->H;
    CU(ctx, ctx->seeded.ensure(P * sizeof(float4)));
    CU(ctx, cudaMemcpyAsync(ctx->seeded.p, planes4, P * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->have_seeded = true;
    ctx->params.seeded = 1;
    return 0;
}

int acmmp_set_planar_prior(acmmp_ctx *ctx, const float *planes4, const uint32_t *mask) {
    if (!ctx || !planes4 || !mask) return ACMMP_ERR_INVALID;
    if (!ctx->W) return fail(ctx, ACMMP_ERR_STATE, "acmmp_set_problem first");
    if (bind_device(ctx)) return ACMMP_ERR_CUDA;
    const size_t P = (size_t)ctx->W * ctx->H;
    CU(ctx, ctx->prior_planes.ensure(P * sizeof(float4)));
    CU(ctx, ctx->plane_masks.ensure(P * sizeof(uint32_t)));
    CU(ctx, cudaMemcpyAsync(ctx->prior_planes.p, planes4, P * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, cudaMemcpyAsync(ctx->plane_masks.p, mask, P * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->have_prior = true;
    return 0;
}

}  // extern "C"

#include "capi_run.cu.inc"
#include "capi_scan.cu.inc"
#include "capi_nccl.cu.inc"
#include "capi_dataflow.cu.inc"
