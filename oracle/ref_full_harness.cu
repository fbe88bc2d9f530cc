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

__device__ unsigned long long g_acmmp_ref_seed;
#define clock64() (g_acmmp_ref_seed)
#include "ACMMP.cu"
#undef clock64
#include "ACMMP.cpp"
#include "acmmp_definitions.cpp"

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
    return 0;
}
