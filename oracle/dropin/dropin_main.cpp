// dropin_main.cpp - main() of the reference (main_ACMMP.cpp:64-199) restated without Boost, as a library entry point:
// the same loop as oracle/ref_full_harness.cu, here driving the reference's UNCHANGED acmmp_definitions.cpp on top
// of oracle/dropin/ACMMP_dropin.cpp (class ACMMP over the C ABI of libacmmp_b200.so).  TEST INFRASTRUCTURE.
#include <cuda_runtime.h>
#include <curand_kernel.h>

#include "ACMMP.h"
#include "acmmp_definitions.cpp"   // the reference's per-view driver, JBU step, fusion and pSampler - as they lie

extern "C" void acmmp_dropin_set_seed(unsigned long long seed);

// cv::Subdiv2D of the stand-in OpenCV is not used by the drop-in (ACMMP::DelaunayTriangulation calls acmmp_delaunay)
extern "C" int (*acmmp_ref_stub_delaunay_hook)(const int *, int, int *, int) = nullptr;

extern "C" int dropin_run(const char *dense_folder_c, const char *output_dir_c, unsigned long long seed, int fuse, float consistency_scalar,
                          int num_consistent_thresh) {
    std::string dense_folder(dense_folder_c), output_dir(output_dir_c);
    acmmp_dropin_set_seed(seed);
    std::vector<Problem> problems;
    GenerateSampleList(dense_folder, problems);
    const size_t num_images = problems.size();
    int max_num_downscale = ComputeMultiScaleSettings(dense_folder, problems);
    pSampler sample_handler(dense_folder, num_images);
    std::string output_folder = dense_folder + output_dir;
    mkdir(output_folder.c_str(), 0777);
    bool first_scale = true;
    while (max_num_downscale >= 0) {
        for (size_t i = 0; i < num_images; ++i)
            if (problems[i].num_downscale >= 0) {
                problems[i].cur_image_size = problems[i].max_image_size / pow(2, problems[i].num_downscale);
                problems[i].num_downscale--;
            }
        bool hierarchy = false;
        if (!first_scale) {
            for (size_t i = 0; i < num_images; ++i) JointBilateralUpsampling(dense_folder, output_folder, problems[i], problems[i].cur_image_size);
            hierarchy = true;
        }
        first_scale = false;
        for (size_t i = 0; i < num_images; ++i)   // photometric + planar prior
            ProcessProblem(sample_handler, output_folder, dense_folder, problems, i, false, true, hierarchy, false, false);
        for (int geom_iter = 0; geom_iter < 2; ++geom_iter)
            for (size_t i = 0; i < num_images; ++i)
                ProcessProblem(sample_handler, output_folder, dense_folder, problems, i, true, false, false, geom_iter != 0, false);
        max_num_downscale--;
    }
    if (fuse) RunFusion(dense_folder, output_folder, problems, true, consistency_scalar, num_consistent_thresh, std::string("/images"), std::string(" "));
    return 0;
}
