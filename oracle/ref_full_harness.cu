// ref_full_harness.cu - the reference PROGRAM as a library: its own kernels (ACMMP.cu), its own
// host class and helpers (ACMMP.cpp) and its own per-view driver, JBU step and fusion
// (acmmp_definitions.cpp: ProcessProblem, JointBilateralUpsampling, RunFusion), all compiled from
// where they lie under /root/reference/src (never copied into this repository), run in the order
// of its main() (main_ACMMP.cpp:64-199, restated below because that file needs Boost) on the
// reference's own on-disk layout, .dmb round trips included.
//
// TEST INFRASTRUCTURE ONLY (bench.py --impl reference, tests/): nothing under acmmp_b200/ links,
// loads or calls this file or its output (oracle/_ref/libacmmp_ref_full.so).
//
// Not the reference's: OpenCV (stand-in under oracle/ref_stub: images are read from a raw
// container instead of being JPEG-decoded; cv::Subdiv2D is the product's Delaunay triangulation
// through a hook - same triangles, possibly another order; cv::SVD::solveZ is a cross product;
// imshow / imwrite / line / merge do nothing), and `clock64()` in RandomInitialization
// (ACMMP.cu:624) is replaced by a caller-provided seed.  `#define DEBUG` (ACMMP.cu:3) stays, as
// shipped, in libacmmp_ref_full.so; libacmmp_ref_full_p2.so is the same program built from a copy of
// ACMMP.cu that the Makefile generates under oracle/_ref/gen_p2/ with that one line deleted (patch P2 of
// no per-run debug copies of the plane map, no leaked host buffer).
#include <cuda_runtime.h>
#include <curand_kernel.h>

// everything the reference's headers pull in, BEFORE `private` is lifted for the reference's own class (below)
#include <sys/stat.h>
#include <sys/types.h>

#include <algorithm>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <map>
#include <memory>
#include <sstream>
#include <string>
#include <vector>

#include "opencv2/opencv.hpp"

__device__ unsigned long long g_acmmp_ref_seed;
#define clock64() (g_acmmp_ref_seed)
// the pinning exports at the end of this file set up an ACMMP object by hand (cameras, costs_host): class ACMMP only
#define private public
#include "ACMMP.cu"
#undef private
#undef clock64
#include "ACMMP.cpp"

// ---- capture of what ProcessProblem hands to the planar-prior initialisation (acmmp_definitions.cpp:306-378) -------
// Function-like macros around two member calls of the reference's driver: the arguments pass through unchanged.
static bool g_capture = false;
static std::vector<float4> g_cap_planes;
static cv::Mat g_cap_mask, g_cap_depth;
static std::vector<int> g_cap_points, g_cap_tris;
static const std::vector<float4> &ref_capture_prior(const std::vector<float4> &planes, const cv::Mat_<float> &mask) {
    if (g_capture) {
        g_cap_planes = planes;
        g_cap_mask = mask.clone();
    }
    return planes;
}
static const cv::Mat_<float> &ref_capture_depths(const cv::Mat_<float> &depths) {
    if (g_capture && g_cap_depth.empty()) g_cap_depth = depths.clone();
    return depths;
}
#define CudaPlanarPriorInitialization(planes, mask) CudaPlanarPriorInitialization(ref_capture_prior(planes, mask), mask)
#define GetPriorPlaneParams(triangle, depths) GetPriorPlaneParams(triangle, ref_capture_depths(depths))
#include "acmmp_definitions.cpp"
#undef CudaPlanarPriorInitialization
#undef GetPriorPlaneParams

#include "../acmmp_b200/csrc/prior.h"

#include <chrono>
static double g_delaunay_seconds = 0.0;   // time spent in the stand-in for cv::Subdiv2D (reported by bench.py)
extern "C" double ref_full_delaunay_seconds(int reset) {
    const double v = g_delaunay_seconds;
    if (reset) g_delaunay_seconds = 0.0;
    return v;
}
static int delaunay_hook(const int *xy, int n, int *tris6, int capacity) {
    struct Timer {
        std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
        ~Timer() { g_delaunay_seconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); }
    } timer;
    std::vector<int32_t> tris;
    const int nt = acmmp::delaunay_triangulate(xy, n, tris);
    const int m = nt < capacity ? nt : capacity;
    memcpy(tris6, tris.data(), (size_t)m * 6 * sizeof(int));
    if (g_capture) {
        g_cap_points.assign(xy, xy + 2 * n);
        g_cap_tris.assign(tris.begin(), tris.begin() + (size_t)m * 6);
    }
    return m;
}
extern "C" int (*acmmp_ref_stub_delaunay_hook)(const int *, int, int *, int) = delaunay_hook;

extern "C" int ref_full_run(const char *dense_folder_c, const char *output_dir_c, unsigned long long seed, int fuse,
                            float consistency_scalar, int num_consistent_thresh, double *seconds_patchmatch, double *seconds_fusion) {
    std::string dense_folder(dense_folder_c), output_dir(output_dir_c);
    cudaMemcpyToSymbol(g_acmmp_ref_seed, &seed, sizeof seed);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    cudaEventRecord(e0, 0);
    // ---- main_ACMMP.cpp:64-176 -------------------------------------------------------------
    std::vector<Problem> problems;
    GenerateSampleList(dense_folder, problems);
    size_t num_images = problems.size();
    int max_num_downscale = ComputeMultiScaleSettings(dense_folder, problems);
    pSampler sample_handler(dense_folder, num_images);
    std::string output_folder = dense_folder + output_dir;
    mkdir(output_folder.c_str(), 0777);
    int flag = 0;
    int geom_iterations = 2;
    bool geom_consistency = false, planar_prior = false, hierarchy = false, multi_geometry = false;
    while (max_num_downscale >= 0) {
        for (size_t i = 0; i < num_images; ++i)
            if (problems[i].num_downscale >= 0) {
                problems[i].cur_image_size = problems[i].max_image_size / pow(2, problems[i].num_downscale);
                problems[i].num_downscale--;
            }
        if (flag == 0) {
            flag = 1;
        } else {
            for (size_t i = 0; i < num_images; ++i) JointBilateralUpsampling(dense_folder, output_folder, problems[i], problems[i].cur_image_size);
            hierarchy = true;
        }
        geom_consistency = false;
        planar_prior = true;
        for (size_t i = 0; i < num_images; ++i)
            ProcessProblem(sample_handler, output_folder, dense_folder, problems, i, geom_consistency, planar_prior, hierarchy, false, false);
        hierarchy = false;
        geom_consistency = true;
        planar_prior = false;
        for (int geom_iter = 0; geom_iter < geom_iterations; ++geom_iter) {
            multi_geometry = geom_iter != 0;
            for (size_t i = 0; i < num_images; ++i)
                ProcessProblem(sample_handler, output_folder, dense_folder, problems, i, geom_consistency, planar_prior, hierarchy, multi_geometry, false);
        }
        max_num_downscale--;
    }
    cudaDeviceSynchronize();
    cudaEventRecord(e1, 0);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (seconds_patchmatch) *seconds_patchmatch = ms * 1e-3;
    if (fuse) {
        cudaEventRecord(e0, 0);
        RunFusion(dense_folder, output_folder, problems, true, consistency_scalar, num_consistent_thresh, std::string("/images"), std::string(" "));
        cudaEventRecord(e1, 0);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        if (seconds_fusion) *seconds_fusion = ms * 1e-3;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    (void)cudaGetLastError();   // ~ACMMP frees members it never allocated (ACMMP.cpp:109-151): not the caller's error
    return 0;
}

// ---- pinning exports: the reference's OWN planar-prior functions (ACMMP.cpp:868-958) and what its driver builds --------
// (tests/test_prior_pinned.py, tests/test_gpu_prior_pinned.py: oracle/prior.c and the device construction against these)

namespace {
// an ACMMP object without its constructor's / destructor's device side: zeroed storage, never destroyed
ACMMP *bare_object(const acmmp_camera *cam) {
    void *mem = calloc(1, sizeof(ACMMP));
    ACMMP *a = new (mem) ACMMP();
    Camera c;
    memcpy(&c, cam, sizeof c);
    a->cameras.push_back(c);
    return a;
}
void drop_bare(ACMMP *a) {
    a->cameras.~vector();
    free(a);
}
}  // namespace

extern "C" int ref_full_support_points(const acmmp_camera *cam, const float *costs, int *xy, int capacity) {
    ACMMP *a = bare_object(cam);
    a->costs_host = const_cast<float *>(costs);
    std::vector<cv::Point> pts;
    a->GetSupportPoints(pts);
    a->costs_host = nullptr;
    drop_bare(a);
    const int n = (int)pts.size() < capacity ? (int)pts.size() : capacity;
    for (int i = 0; i < n; ++i) {
        xy[2 * i] = pts[i].x;
        xy[2 * i + 1] = pts[i].y;
    }
    return (int)pts.size();
}

// GetPriorPlaneParams of `n` triangles (6 ints each) on a depth map; cv::SVD::solveZ is the stand-in's (oracle/ref_stub)
extern "C" int ref_full_prior_planes(const acmmp_camera *cam, const float *depth, const int *tris6, int n, float *planes4) {
    ACMMP *a = bare_object(cam);
    cv::Mat_<float> d(cam->height, cam->width);
    memcpy(d.data, depth, (size_t)cam->width * cam->height * sizeof(float));
    for (int t = 0; t < n; ++t) {
        const int *v = tris6 + 6 * t;
        const float4 p = a->GetPriorPlaneParams(Triangle(cv::Point(v[0], v[1]), cv::Point(v[2], v[3]), cv::Point(v[4], v[5])), d);
        planes4[4 * t] = p.x; planes4[4 * t + 1] = p.y; planes4[4 * t + 2] = p.z; planes4[4 * t + 3] = p.w;
    }
    drop_bare(a);
    return 0;
}

extern "C" int ref_full_depths_from_planes(const acmmp_camera *cam, const float *planes4, const int *xy, int n, float *out) {
    ACMMP *a = bare_object(cam);
    for (int i = 0; i < n; ++i) out[i] = a->GetDepthFromPlaneParam(make_float4(planes4[4 * i], planes4[4 * i + 1], planes4[4 * i + 2], planes4[4 * i + 3]), xy[2 * i], xy[2 * i + 1]);
    drop_bare(a);
    return 0;
}

// one ProcessProblem(photometric + planar prior) of view `idx` at the first scale of `dense_folder`, capturing the depth map the
// prior was built from, the support points, the triangles, and the arguments of CudaPlanarPriorInitialization
extern "C" int ref_full_capture_prior(const char *dense_folder_c, const char *output_dir_c, int idx, unsigned long long seed, int *w, int *h,
                                      int *num_points, int *num_tris, int *num_planes) {
    std::string dense_folder(dense_folder_c), output_dir(output_dir_c);
    cudaMemcpyToSymbol(g_acmmp_ref_seed, &seed, sizeof seed);
    std::vector<Problem> problems;
    GenerateSampleList(dense_folder, problems);
    ComputeMultiScaleSettings(dense_folder, problems);
    for (auto &p : problems)
        if (p.num_downscale >= 0) {
            p.cur_image_size = p.max_image_size / pow(2, p.num_downscale);
            p.num_downscale--;
        }
    pSampler sample_handler(dense_folder, problems.size());
    std::string output_folder = dense_folder + output_dir;
    mkdir(output_folder.c_str(), 0777);
    g_cap_planes.clear();
    g_cap_points.clear();
    g_cap_tris.clear();
    g_cap_mask = cv::Mat();
    g_cap_depth = cv::Mat();
    g_capture = true;
    ProcessProblem(sample_handler, output_folder, dense_folder, problems, idx, false, true, false, false, false);
    g_capture = false;
    cudaDeviceSynchronize();
    (void)cudaGetLastError();   // see ref_full_run
    if (g_cap_mask.empty() || g_cap_depth.empty()) return -1;
    *w = g_cap_mask.cols;
    *h = g_cap_mask.rows;
    *num_points = (int)g_cap_points.size() / 2;
    *num_tris = (int)g_cap_tris.size() / 6;
    *num_planes = (int)g_cap_planes.size();
    return 0;
}

extern "C" int ref_full_captured(float *depth, int *points_xy, int *tris6, float *planes4, float *mask) {
    if (g_cap_mask.empty()) return -1;
    const size_t P = (size_t)g_cap_mask.rows * g_cap_mask.cols;
    if (depth) memcpy(depth, g_cap_depth.data, P * sizeof(float));
    if (mask) memcpy(mask, g_cap_mask.data, P * sizeof(float));
    if (points_xy) memcpy(points_xy, g_cap_points.data(), g_cap_points.size() * sizeof(int));
    if (tris6) memcpy(tris6, g_cap_tris.data(), g_cap_tris.size() * sizeof(int));
    if (planes4) memcpy(planes4, g_cap_planes.data(), g_cap_planes.size() * sizeof(float4));
    return 0;
}
