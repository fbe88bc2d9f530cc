// noalloc.cu - acmmp_prior_raster_noalloc: the planar-prior raster (tri_raster_kernel / big_raster_kernel /
// prior_finalize_kernel, prior_kernels.cu.inc) for callers that must not allocate: code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// The kernels need scratch: the per-pixel label map (atomicMax target, must start at zero), one fitted plane per
// triangle, the queue of long triangles and the camera constants.  acmmp_build_planar_prior keeps them in buffers of
// the context (cudaMalloc on first use); here they live in ONE file-scope `static __device__` array of a fixed
// maximum size, which every GPU that loads the library holds until the process exits.
//
// In host code the array's NAME is a host-side stand-in, not a device address (a kernel handed `g_scratch` faults or,
// where the GPU can reach host memory, silently uses the wrong memory): the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
//
// Every call shares the array, and it keeps the previous call's contents: the label map and the queue length are
// cleared on every call; calls that overlap on the GPU (two streams, two host threads) RACE - serialise them.
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "../../include/acmmp_b200.h"
#include "kernels.h"

using namespace acmmp;

namespace {

constexpr size_t kHeaderBytes = 4096;   // RefConst + one (unused) PairConst slot, each 256-byte aligned
constexpr size_t kLabelsBytes = (size_t)ACMMP_NOALLOC_MAX_PIXELS * sizeof(unsigned);
constexpr size_t kTriPlanesBytes = (size_t)ACMMP_NOALLOC_MAX_TRIANGLES * sizeof(float4);
constexpr size_t kBigListBytes = (((size_t)ACMMP_NOALLOC_MAX_TRIANGLES + 1) * sizeof(int) + 255) & ~(size_t)255;
constexpr size_t kScratchBytes = kHeaderBytes + kLabelsBytes + kTriPlanesBytes + kBigListBytes;
static_assert(kLabelsBytes % 256 == 0 && kTriPlanesBytes % 256 == 0, "sections stay 256-byte aligned");

// 1600 x 1200 pixels, 153,600 triangles: 10.2 MiB per GPU
static __device__ __align__(256) unsigned char g_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
unsigned char *g_scratch_ptr[kMaxDevices] = {nullptr};   // DEVICE address of g_scratch on device i
std::mutex g_setup_mutex;

int report(char *err, size_t err_len, int code, const char *fmt, ...) {
    if (err && err_len) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, err_len, fmt, ap);
        va_end(ap);
    }
    return code;
}

}  // namespace

extern "C" int acmmp_prior_raster_noalloc(int device, const acmmp_camera *ref_cam, float depth_min, float depth_max, const float *d_planes4,
                                          const int32_t *d_tris, int num_triangles, float *d_prior_planes4, uint32_t *d_masks, void *stream,
                                          char *err, size_t err_len) {
    if (err && err_len) err[0] = 0;
    if (!ref_cam || !d_planes4 || !d_prior_planes4 || !d_masks || (num_triangles > 0 && !d_tris))
        return report(err, err_len, ACMMP_ERR_INVALID, "acmmp_prior_raster_noalloc: null argument");
    const int W = ref_cam->width, H = ref_cam->height;
    if (W <= 0 || H <= 0 || num_triangles < 0) return report(err, err_len, ACMMP_ERR_INVALID, "acmmp_prior_raster_noalloc: bad size %d x %d, %d triangles", W, H, num_triangles);
    if ((long long)W * H > (long long)ACMMP_NOALLOC_MAX_PIXELS)
        return report(err, err_len, ACMMP_ERR_INVALID, "acmmp_prior_raster_noalloc: %d x %d = %lld pixels exceed the static scratch (ACMMP_NOALLOC_MAX_PIXELS = %d)", W, H,
                      (long long)W * H, ACMMP_NOALLOC_MAX_PIXELS);
    if (num_triangles > ACMMP_NOALLOC_MAX_TRIANGLES)
        return report(err, err_len, ACMMP_ERR_INVALID, "acmmp_prior_raster_noalloc: %d triangles exceed the static scratch (ACMMP_NOALLOC_MAX_TRIANGLES = %d)", num_triangles,
                      ACMMP_NOALLOC_MAX_TRIANGLES);
    if (device < 0 || device >= kMaxDevices) return report(err, err_len, ACMMP_ERR_INVALID, "acmmp_prior_raster_noalloc: device %d out of range", device);
    if (ref_const_size() > 2048 || pair_const_size() > 2048) return report(err, err_len, ACMMP_ERR_STATE, "acmmp_prior_raster_noalloc: header section too small");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return report(err, err_len, ACMMP_ERR_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    (void)cudaGetLastError();   // a stale non-sticky error of other code in this thread is not ours
    unsigned char *base = nullptr;
    {   // set-up, once per device, with that device current: the array's device address
        std::lock_guard<std::mutex> lock(g_setup_mutex);
        if (!g_scratch_ptr[device]) {
            void *p = nullptr;
            e = cudaGetSymbolAddress(&p, g_scratch);
            if (e != cudaSuccess || !p) return report(err, err_len, ACMMP_ERR_CUDA, "cudaGetSymbolAddress(g_scratch) on device %d: %s", device, cudaGetErrorString(e));
            g_scratch_ptr[device] = static_cast<unsigned char *>(p);
        }
        base = g_scratch_ptr[device];
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    RefConst *d_rc = reinterpret_cast<RefConst *>(base);
    PairConst *d_pairs = reinterpret_cast<PairConst *>(base + 2048);
    unsigned *labels = reinterpret_cast<unsigned *>(base + kHeaderBytes);
    float4 *tri_planes = reinterpret_cast<float4 *>(base + kHeaderBytes + kLabelsBytes);
    int *big_list = reinterpret_cast<int *>(base + kHeaderBytes + kLabelsBytes + kTriPlanesBytes);
    const size_t P = (size_t)W * H;
#define NA_CU(call)                                                                                                  \
    do {                                                                                                             \
        cudaError_t e__ = (call);                                                                                    \
        if (e__ != cudaSuccess) return report(err, err_len, ACMMP_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e__)); \
    } while (0)
    // camera constants of the reference view (n = 1: no source views, the PairConst slot stays unused)
    NA_CU(launch_setup_pairs(ref_cam, 1, nullptr, nullptr, depth_min, depth_max, d_rc, d_pairs, st));
    // the label map keeps the previous call's labels: clear (the queue length is cleared by launch_tri_raster)
    NA_CU(cudaMemsetAsync(labels, 0, P * sizeof(unsigned), st));
    NA_CU(launch_tri_raster(d_tris, num_triangles, reinterpret_cast<const float4 *>(d_planes4), d_rc, W, tri_planes, labels, big_list, st));
    NA_CU(launch_prior_finalize(labels, tri_planes, d_rc, W, H, reinterpret_cast<float4 *>(d_prior_planes4), d_masks, st));
#undef NA_CU
    return 0;
}
