// ref_host_harness.cpp - parity comparator for the HOST side of the reference: its own
// RunFusion (acmmp_definitions.cpp:828-1043) with ReadCamera / readDepthDmb / readNormalDmb /
// RescaleImageAndCamera / Get3DPointonWorld / ProjectonCamera / GetAngle /
// StoreColorPlyFileBinaryPointCloud (ACMMP.cpp), compiled from where the sources lie under
// /root/reference/src (never copied into this repository) against the stand-in OpenCV of
// oracle/ref_stub.  The test writes the reference's on-disk inputs (cams/*_cam.txt,
// images/*.jpg as a raw container, 2333_*/depths_geom.dmb, normals.dmb), calls RunFusion and
// reads ACMMP_model.ply back.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under acmmp_b200/ links, loads or calls this file or its
// output (oracle/_ref/libacmmp_ref_host.so).
#include "ACMMP.cpp"
#include "acmmp_definitions.cpp"

// no Delaunay behind cv::Subdiv2D in this library: the planar-prior path is not run through it
extern "C" int (*acmmp_ref_stub_delaunay_hook)(const int *, int, int *, int) = nullptr;

// members defined in the reference's ACMMP.cu (device side; not part of this library, never called)
void ACMMP::RunPatchMatch() {}
void JBU::CudaRun() {}

extern "C" int ref_host_run_fusion(const char *dense_folder, const char *out_folder, int n_views, const int *image_ids,
                                   const int *pair_offsets, const int *pair_image_ids, int geom_consistency,
                                   float consistency_scalar, int con_num_thresh) {
    std::vector<Problem> problems(n_views);
    for (int i = 0; i < n_views; ++i) {
        problems[i].ref_image_id = image_ids[i];
        for (int k = pair_offsets[i]; k < pair_offsets[i + 1]; ++k) problems[i].src_image_ids.push_back(pair_image_ids[k]);
    }
    std::string dense(dense_folder), out(out_folder);
    RunFusion(dense, out, problems, geom_consistency != 0, consistency_scalar, con_num_thresh, std::string("/images"), std::string(" "));
    return 0;
}

// RunPriorAwareFusion (acmmp_definitions.cpp:575-826): map set 0 is read from `fusion_folder`, set 1
// from `out_folder`; writes <out_folder>/ACMMP_prior_model.ply
extern "C" int ref_host_run_prior_fusion(const char *dense_folder, const char *out_folder, const char *fusion_folder, int n_views,
                                         const int *image_ids, const int *pair_offsets, const int *pair_image_ids,
                                         int geom_consistency, float consistency_scalar, int num_consistent_thresh,
                                         int single_match_penalty) {
    std::vector<Problem> problems(n_views);
    for (int i = 0; i < n_views; ++i) {
        problems[i].ref_image_id = image_ids[i];
        for (int k = pair_offsets[i]; k < pair_offsets[i + 1]; ++k) problems[i].src_image_ids.push_back(pair_image_ids[k]);
    }
    std::string dense(dense_folder), out(out_folder), fus(fusion_folder);
    RunPriorAwareFusion(dense, out, fus, problems, geom_consistency != 0, consistency_scalar, num_consistent_thresh,
                        single_match_penalty, std::string(" "));
    return 0;
}

// pSampler::GetPriorPlaneEstimate (acmmp_definitions.cpp:7-177): the seeded hypotheses of view `cam_num`
// from <dense_folder>priors/{depths,normals}/%08d.png (16-bit; the test writes the raw container).
// `cam` has the layout of the reference's Camera.  Returns -1 when pSampler finds no priors.
extern "C" int ref_host_prior_planes(const char *dense_folder, int ncams, int cam_num, const void *cam, int rows, int cols,
                                     float *planes4) {
    pSampler sampler(std::string(dense_folder), ncams);
    if (!sampler.confirm_using_prior()) return -1;
    Camera c;
    memcpy(&c, cam, sizeof c);
    std::unique_ptr<float4> planes = sampler.GetPriorPlaneEstimate(cam_num, c, rows, cols);
    memcpy(planes4, planes.get(), sizeof(float4) * (size_t)rows * cols);
    delete[] planes.release();
    return 0;
}
