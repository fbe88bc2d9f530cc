// acmmp.hpp - C++ host side above the C ABI (include/acmmp_b200.h): `class ACMMP` with the
// public methods of the reference's class (ACMMP.h:58-124), plus RAII owners of the
// whole-scan schedule and of the fusion.  Header-only, no OpenCV: cv::Mat_<T> becomes
// Map<T> (io.hpp).  Errors: the reference prints and exit()s inside CudaSafeCall
// (ACMMP.cpp:67-100); here every failing call throws std::runtime_error carrying
// acmmp_last_error(), and the CLI turns that into the reference's behaviour.
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/acmmp_b200.h"
#include "io.hpp"

namespace acmmp_b200 {

class ACMMP {
public:
    explicit ACMMP(int device = 0) {
        if (acmmp_create(&ctx_, device)) throw std::runtime_error(acmmp_last_error(nullptr));
    }
    ~ACMMP() { acmmp_destroy(ctx_); }
    ACMMP(const ACMMP &) = delete;
    ACMMP &operator=(const ACMMP &) = delete;

    void SetGeomConsistencyParams(bool multi_geometry) { check(acmmp_set_geom_consistency(ctx_, multi_geometry)); }   // ACMMP.cpp:447-454
    void SetHierarchyParams() { check(acmmp_set_hierarchy(ctx_)); }                                                    // :456-459
    void SetPlanarPriorParams() { check(acmmp_set_planar_prior_params(ctx_)); }                                        // :461-464

    // InputInitialization (:525-636) after the caller has read the files: grey float images
    // and cameras of the reference view followed by its sources (cameras carry width / height)
    void InputInitialization(const std::vector<Map<float>> &images, const std::vector<acmmp_camera> &cameras) {
        std::vector<const float *> ptrs;
        for (const auto &im : images) ptrs.push_back(im.data.data());
        check(acmmp_set_problem(ctx_, (int)images.size(), cameras.data(), ptrs.data(), 0));
        width_ = cameras[0].width;
        height_ = cameras[0].height;
    }
    void SetDepthMap(int slot, const Map<float> &depth) { check(acmmp_set_depth_map(ctx_, slot, depth.data.data(), depth.cols, depth.rows, 0)); }
    // CudaSpaceInitialization (:638-809): the .dmb payloads of the previous stage / scale
    void SetWarmStart(const Map<float> &depth, const Map<float> &normal, const Map<float> &cost) {
        check(acmmp_set_warm_start(ctx_, depth.data.data(), normal.data.data(), cost.data.data()));
    }
    void SetHierarchyInputs(const Map<float> &coarse_normal, const Map<float> &coarse_cost, const Map<float> &fine_depth) {
        check(acmmp_set_hierarchy_inputs(ctx_, coarse_normal.data.data(), coarse_cost.data.data(), coarse_cost.cols, coarse_cost.rows,
                                         fine_depth.data.data()));
    }
    // GetSupportPoints + DelaunayTriangulation + raster + GetPriorPlaneParams + CudaPlanarPriorInitialization
    void BuildPlanarPrior(int *num_support = nullptr, int *num_triangles = nullptr) { check(acmmp_build_planar_prior(ctx_, num_support, num_triangles)); }

    void RunPatchMatch(uint64_t seed) {   // ACMMP.cu:1378-1456; seed replaces clock64()
        check(acmmp_run(ctx_, seed));
        const size_t P = (size_t)width_ * height_;
        planes_.resize(4 * P);
        costs_.resize(P);
        check(acmmp_get_result(ctx_, planes_.data(), costs_.data(), nullptr));
    }
    int GetReferenceImageWidth() const { return width_; }
    int GetReferenceImageHeight() const { return height_; }
    // float4 (world normal, depth) of pixel `index`, ACMMP.cpp:848-851
    const float *GetPlaneHypothesis(int index) const { return &planes_[4 * (size_t)index]; }
    float GetCost(int index) const { return costs_[index]; }
    acmmp_ctx *handle() { return ctx_; }

private:
    void check(int rc) {
        if (rc) throw std::runtime_error(acmmp_last_error(ctx_));
    }
    acmmp_ctx *ctx_ = nullptr;
    int width_ = 0, height_ = 0;
    std::vector<float> planes_, costs_;
};

// RunJBU (ACMMP.cpp:1022-1087) without the file write; false when the integer scale is 1
inline bool RunJBU(const Map<float> &scaled_image_float, const Map<float> &src_depthmap, Map<float> &out, int device = 0) {
    out = Map<float>(scaled_image_float.rows, scaled_image_float.cols, 1);
    const int rc = acmmp_jbu_depth(device, scaled_image_float.data.data(), scaled_image_float.cols, scaled_image_float.rows,
                                   src_depthmap.data.data(), src_depthmap.cols, src_depthmap.rows, out.data.data());
    if (rc < 0) throw std::runtime_error(acmmp_last_error(nullptr));
    return rc == 0;
}

// the stage graph of main() (main_ACMMP.cpp:89-176), resident on the device
class Scan {
public:
    Scan(const std::vector<acmmp_camera> &cams, const std::vector<std::vector<int>> &pairs, int device = 0, int rank = 0, int world = 1) {
        std::vector<int32_t> offs(1, 0), ids;
        for (const auto &p : pairs) {
            ids.insert(ids.end(), p.begin(), p.end());
            offs.push_back((int32_t)ids.size());
        }
        if (acmmp_scan_create(&s_, device, (int)cams.size(), cams.data(), offs.data(), ids.data())) throw std::runtime_error(acmmp_last_error(nullptr));
        n_ = (int)cams.size();
        check(acmmp_scan_set_shard(s_, rank, world));
    }
    ~Scan() { acmmp_scan_destroy(s_); }
    Scan(const Scan &) = delete;
    Scan &operator=(const Scan &) = delete;
    void SetImage(int view, const Map<float> &gray) { check(acmmp_scan_set_image(s_, view, gray.data.data(), 0)); check(acmmp_scan_sync(s_)); }
    // the 8-bit grey image as cv::imread(path, IMREAD_GRAYSCALE) returns it; converted to float on the device
    void SetImageU8(int view, const Map<uint8_t> &gray) { check(acmmp_scan_set_image_u8(s_, view, gray.data.data(), 0)); check(acmmp_scan_sync(s_)); }
    // seeded mode (-p): the 16-bit prior images of a view in cv::imread(path, -1) order (pSampler, acmmp_definitions.cpp:98-131)
    void SetPriorMaps(int view, const Map<uint16_t> &depth16, const Map<uint16_t> &normal16) {
        if (depth16.channels != 1 || normal16.channels < 3 || normal16.rows != depth16.rows || normal16.cols != depth16.cols)
            throw std::runtime_error("prior images: need a 1-channel depth and a 3-channel normal image of the same size");
        if (normal16.channels == 3) {
            check(acmmp_scan_set_prior_maps(s_, view, depth16.data.data(), normal16.data.data(), depth16.rows, depth16.cols));
            return;
        }
        std::vector<uint16_t> n3((size_t)normal16.rows * normal16.cols * 3);   // drop the alpha channel
        for (size_t p = 0; p < (size_t)normal16.rows * normal16.cols; ++p)
            for (int k = 0; k < 3; ++k) n3[3 * p + k] = normal16.data[p * normal16.channels + k];
        check(acmmp_scan_set_prior_maps(s_, view, depth16.data.data(), n3.data(), depth16.rows, depth16.cols));
    }
    int NumLevels() const { return acmmp_scan_num_levels(s_); }
    void BeginLevel(int level) { check(acmmp_scan_begin_level(s_, level)); }
    void RunStage(int stage, uint64_t seed) { check(acmmp_scan_run_stage(s_, stage, seed)); }
    void Sync() { check(acmmp_scan_sync(s_)); }
    void ViewSize(int view, int &w, int &h) { check(acmmp_scan_view_size(s_, view, &w, &h)); }
    void GetView(int view, Map<float> *depth, Map<float> *normal, Map<float> *cost) {
        int w, h;
        ViewSize(view, w, h);
        if (depth) *depth = Map<float>(h, w, 1);
        if (normal) *normal = Map<float>(h, w, 3);
        if (cost) *cost = Map<float>(h, w, 1);
        check(acmmp_scan_get_view(s_, view, depth ? depth->data.data() : nullptr, normal ? normal->data.data() : nullptr,
                                  cost ? cost->data.data() : nullptr));
    }
    void ViewDevice(int view, const float **depth, const float **normal, const float **cost) { check(acmmp_scan_view_device(s_, view, depth, normal, cost)); }
    acmmp_scan *handle() { return s_; }

private:
    void check(int rc) {
        if (rc) throw std::runtime_error(acmmp_scan_last_error(s_));
    }
    acmmp_scan *s_ = nullptr;
    int n_ = 0;
};

// RunFusion (acmmp_definitions.cpp:828-1043) on the device
class Fusion {
public:
    Fusion(const std::vector<acmmp_camera> &cams, const std::vector<std::vector<int>> &pairs, int device = 0) {
        std::vector<int32_t> offs(1, 0), ids;
        for (const auto &p : pairs) {
            ids.insert(ids.end(), p.begin(), p.end());
            offs.push_back((int32_t)ids.size());
        }
        if (acmmp_fusion_create(&f_, device, (int)cams.size(), cams.data(), offs.data(), ids.data()))
            throw std::runtime_error(acmmp_fusion_last_error(nullptr));
    }
    ~Fusion() { acmmp_fusion_destroy(f_); }
    Fusion(const Fusion &) = delete;
    Fusion &operator=(const Fusion &) = delete;
    void SetView(int view, const float *depth, const float *normal, const uint8_t *color, const uint8_t *mask, bool on_device) {
        check(acmmp_fusion_set_view(f_, view, depth, normal, color, mask, on_device ? 1 : 0));
    }
    int64_t Run(float consistency_scalar, int con_num_thresh) {
        int64_t n = 0;
        check(acmmp_fusion_run(f_, consistency_scalar, con_num_thresh, &n));
        return n;
    }
    // second (depth, normal) set of a view + RunPriorAwareFusion (acmmp_definitions.cpp:575-826)
    void SetViewSecond(int view, const float *depth2, const float *normal2, bool on_device) {
        check(acmmp_fusion_set_view_second(f_, view, depth2, normal2, on_device ? 1 : 0));
    }
    int64_t RunPriorAware(float consistency_scalar, int num_consistent_thresh, int single_match_penalty) {
        int64_t n = 0;
        check(acmmp_fusion_run_prior_aware(f_, consistency_scalar, num_consistent_thresh, single_match_penalty, &n));
        return n;
    }
    void GetPoints(std::vector<float> &points9, std::vector<int64_t> &per_view, int64_t n, int n_views) {
        points9.resize((size_t)n * 9);
        per_view.resize(n_views);
        check(acmmp_fusion_get_points(f_, points9.data(), n, per_view.data()));
    }

private:
    void check(int rc) {
        if (rc) throw std::runtime_error(acmmp_fusion_last_error(f_));
    }
    acmmp_fusion *f_ = nullptr;
};

}  // namespace acmmp_b200
