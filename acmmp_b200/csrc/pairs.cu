// pairs.cu - view-pair scoring of the input converter, compiled
// WITHOUT --use_fast_math: the reference does this in numpy float64.
//
// Reference: calc_score, /root/reference/python_scripts/colmap2mvsnet_acm.py:280-302, called for
// every image pair i < j (:399-413).  Per pair it walks the keypoints of image i, keeps those whose
// 3-D point is also seen by image j (a Python `in` over a numpy array: O(n_i n_j) per pair), computes
// the triangulation angle of each, and scores the pair with the NUMBER of such keypoints - or 0
// when the angle at rank int(0.75 n) of the sorted list is below one degree.
//
// Here the loop is turned inside out.  Every observation (3-D point, image) becomes a 64-bit key
// point << 32 | image; one radix sort groups the observations of a point (its track), in image
// order.  One thread per distinct (point, image i) then walks the rest of the track - the images
// j > i that see the same point - and adds its multiplicity (a point can be observed by several
// keypoints of image i; the reference counts each) to count[i][j] and, when the angle is below the
// threshold, to below[i][j].  "Angle at rank k of the sorted list < threshold" is "more than k
// angles below the threshold", so no per-pair sort is needed: score = below > int(q n) ? 0 : n.
// Work: sum over points of (track length)^2 / 2 instead of V^2 n^2 / 2; integer atomics only, so the
// result does not depend on the order of execution.
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include <cub/device/device_radix_sort.cuh>

#include "../../include/acmmp_b200.h"

namespace {

thread_local std::string g_pairs_error;

int pfail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_pairs_error = buf;
    return code;
}

constexpr double kRadToDeg = 180.0 / 3.141592653589793;   // 180 / np.pi

__device__ __forceinline__ int key_image(uint64_t k) { return (int)(k & 0xffffffffu); }
__device__ __forceinline__ uint32_t key_point(uint64_t k) { return (uint32_t)(k >> 32); }

// obs -> keys; observations without a 3-D point (-1) get the all-ones key and sort to the end
__global__ void pair_keys_kernel(const int32_t *__restrict__ obs_points, const int32_t *__restrict__ obs_image, int64_t m,
                                 uint64_t *__restrict__ keys) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= m) return;
    const int32_t p = obs_points[e];
    keys[e] = p < 0 ? ~0ull : ((uint64_t)(uint32_t)p << 32) | (uint32_t)obs_image[e];
}

// triangulation angle in degrees at point p between camera centres a and b (calc_score :293)
__device__ __forceinline__ double triangulation_angle(const double *__restrict__ ca, const double *__restrict__ cb,
                                                      const double *__restrict__ p) {
    const double ax = ca[0] - p[0], ay = ca[1] - p[1], az = ca[2] - p[2];
    const double bx = cb[0] - p[0], by = cb[1] - p[1], bz = cb[2] - p[2];
    const double dot = ax * bx + ay * by + az * bz;
    const double na = sqrt(ax * ax + ay * ay + az * az);
    const double nb = sqrt(bx * bx + by * by + bz * bz);
    return kRadToDeg * acos(dot / na / nb);
}

__global__ void pair_accumulate_kernel(const uint64_t *__restrict__ keys, int64_t m, const double *__restrict__ xyz,
                                       const double *__restrict__ centers, int n_images, double min_angle,
                                       uint32_t *__restrict__ count, uint32_t *__restrict__ below) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= m) return;
    const uint64_t k = keys[e];
    if (k == ~0ull) return;
    if (e > 0 && keys[e - 1] == k) return;          // not the first keypoint of this (point, image)
    const uint32_t pt = key_point(k);
    const int i = key_image(k);
    int64_t f = e + 1;
    uint32_t mult = 1;
    while (f < m && keys[f] == k) { ++f; ++mult; }  // keypoints of image i on the same point
    const double *p = xyz + 3 * (size_t)pt;
    const double *ci = centers + 3 * (size_t)i;
    uint64_t prev = k;
    for (; f < m; ++f) {
        const uint64_t kf = keys[f];
        if (key_point(kf) != pt) break;             // also ends at the all-ones keys
        if (kf == prev) continue;                   // image j counts once however often it sees the point
        prev = kf;
        const int j = key_image(kf);
        const double theta = triangulation_angle(ci, centers + 3 * (size_t)j, p);
        const size_t cell = (size_t)i * n_images + j;
        atomicAdd(count + cell, mult);
        if (theta < min_angle) atomicAdd(below + cell, mult);
    }
}

__global__ void pair_finish_kernel(const uint32_t *__restrict__ count, const uint32_t *__restrict__ below, int n_images,
                                   double quantile, double *__restrict__ scores) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    if (j >= n_images || j <= i) return;
    const uint32_t n = count[(size_t)i * n_images + j];
    double s = (double)n;
    if (n > 0) {
        const uint32_t k = (uint32_t)((double)n * quantile);   // int(len(angles) * 0.75)
        if (below[(size_t)i * n_images + j] > k) s = 0.0;
    }
    scores[(size_t)i * n_images + j] = s;
    scores[(size_t)j * n_images + i] = s;
}

#define PCU(call)                                                                                      \
    do {                                                                                               \
        cudaError_t e__ = (call);                                                                      \
        if (e__ != cudaSuccess) {                                                                      \
            rc = pfail(ACMMP_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
            goto done;                                                                                 \
        }                                                                                              \
    } while (0)

}  // namespace

extern "C" {

const char *acmmp_pairs_last_error(void) { return g_pairs_error.c_str(); }

int acmmp_pair_scores(int device, int n_images, const int64_t *obs_offsets, const int32_t *obs_points, int64_t n_points,
                      const double *xyz, const double *centers, double min_angle_deg, double quantile, double *scores,
                      double *device_ms) {
    if (n_images <= 0 || !obs_offsets || !centers || !scores || n_points < 0 || (n_points > 0 && !xyz))
        return pfail(ACMMP_ERR_INVALID, "acmmp_pair_scores: bad argument");
    if (n_points > 0x7fffffffLL || n_images > 0x7fffffff / n_images)
        return pfail(ACMMP_ERR_INVALID, "acmmp_pair_scores: %lld points x %d images out of range", (long long)n_points, n_images);
    const int64_t m = obs_offsets[n_images];
    if (obs_offsets[0] != 0 || m < 0 || (m > 0 && !obs_points)) return pfail(ACMMP_ERR_INVALID, "acmmp_pair_scores: bad offsets");
    std::vector<int32_t> h_image((size_t)m);
    for (int v = 0; v < n_images; ++v) {
        if (obs_offsets[v + 1] < obs_offsets[v]) return pfail(ACMMP_ERR_INVALID, "acmmp_pair_scores: offsets decrease at image %d", v);
        for (int64_t e = obs_offsets[v]; e < obs_offsets[v + 1]; ++e) {
            if (obs_points[e] >= n_points) return pfail(ACMMP_ERR_INVALID, "acmmp_pair_scores: image %d refers to point %d of %lld", v, obs_points[e], (long long)n_points);
            h_image[(size_t)e] = v;
        }
    }
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return pfail(ACMMP_ERR_NO_DEVICE, "no CUDA device");
    if (device < 0 || device >= count) return pfail(ACMMP_ERR_INVALID, "device %d out of range", device);

    int rc = 0;
    const size_t cells = (size_t)n_images * n_images;
    int32_t *d_pts = nullptr, *d_img = nullptr;
    uint64_t *d_keys = nullptr, *d_sorted = nullptr;
    double *d_xyz = nullptr, *d_cen = nullptr, *d_scores = nullptr;
    uint32_t *d_count = nullptr;
    void *d_tmp = nullptr;
    size_t tmp_bytes = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    const size_t mm = (size_t)(m > 0 ? m : 1);
    PCU(cudaSetDevice(device));
    (void)cudaGetLastError();
    PCU(cudaMalloc(&d_pts, mm * sizeof(int32_t)));
    PCU(cudaMalloc(&d_img, mm * sizeof(int32_t)));
    PCU(cudaMalloc(&d_keys, mm * sizeof(uint64_t)));
    PCU(cudaMalloc(&d_sorted, mm * sizeof(uint64_t)));
    PCU(cudaMalloc(&d_xyz, (size_t)(n_points > 0 ? n_points : 1) * 3 * sizeof(double)));
    PCU(cudaMalloc(&d_cen, (size_t)n_images * 3 * sizeof(double)));
    PCU(cudaMalloc(&d_scores, cells * sizeof(double)));
    PCU(cudaMalloc(&d_count, 2 * cells * sizeof(uint32_t)));
    // all 64 key bits take part in the sort so that the all-ones keys end up last
    PCU(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, d_keys, d_sorted, (int64_t)mm, 0, 64, 0));
    PCU(cudaMalloc(&d_tmp, tmp_bytes > 0 ? tmp_bytes : 1));
    PCU(cudaEventCreate(&ev0));
    PCU(cudaEventCreate(&ev1));
    if (m > 0) {
        PCU(cudaMemcpy(d_pts, obs_points, (size_t)m * sizeof(int32_t), cudaMemcpyHostToDevice));
        PCU(cudaMemcpy(d_img, h_image.data(), (size_t)m * sizeof(int32_t), cudaMemcpyHostToDevice));
    }
    if (n_points > 0) PCU(cudaMemcpy(d_xyz, xyz, (size_t)n_points * 3 * sizeof(double), cudaMemcpyHostToDevice));
    PCU(cudaMemcpy(d_cen, centers, (size_t)n_images * 3 * sizeof(double), cudaMemcpyHostToDevice));
    PCU(cudaEventRecord(ev0, 0));
    PCU(cudaMemsetAsync(d_count, 0, 2 * cells * sizeof(uint32_t), 0));
    PCU(cudaMemsetAsync(d_scores, 0, cells * sizeof(double), 0));
    if (m > 0) {
        const unsigned blocks = (unsigned)((m + 255) / 256);
        pair_keys_kernel<<<blocks, 256>>>(d_pts, d_img, m, d_keys);
        PCU(cudaGetLastError());
        PCU(cub::DeviceRadixSort::SortKeys(d_tmp, tmp_bytes, d_keys, d_sorted, m, 0, 64, 0));
        pair_accumulate_kernel<<<blocks, 256>>>(d_sorted, m, d_xyz, d_cen, n_images, min_angle_deg, d_count, d_count + cells);
        PCU(cudaGetLastError());
    }
    pair_finish_kernel<<<dim3((unsigned)((n_images + 127) / 128), (unsigned)n_images), 128>>>(d_count, d_count + cells, n_images,
                                                                                              quantile, d_scores);
    PCU(cudaGetLastError());
    PCU(cudaEventRecord(ev1, 0));
    PCU(cudaMemcpy(scores, d_scores, cells * sizeof(double), cudaMemcpyDeviceToHost));
    if (device_ms) {
        float ms = 0.0f;
        PCU(cudaEventElapsedTime(&ms, ev0, ev1));
        *device_ms = ms;
    }
done:
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    cudaFree(d_tmp);
    cudaFree(d_count);
    cudaFree(d_scores);
    cudaFree(d_cen);
    cudaFree(d_xyz);
    cudaFree(d_sorted);
    cudaFree(d_keys);
    cudaFree(d_img);
    cudaFree(d_pts);
    return rc;
}

}  // extern "C"
