// acmmp.hpp - C++ host side above the C ABI (include/acmmp_b200.h): `class ACMMP` with the
// public methods of the reference's class (ACMMP.h:58-124), plus RAII owners of the
// whole-scan schedule and of the fusion.  Header-only, no OpenCV: cv::Mat_<T> becomes
// Map<T> (io.hpp).  Errors: the reference prints and exit()s inside CudaSafeCall
// (ACMMP.cpp:67-100); here every failing call throws std::runtime_error carrying
// acmmp_last_error(), and the CLI turns that into the reference's behaviour.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/acmmp_b200.h"
#include "io.hpp"

namespace acmmp_b200 {

// `class ACMMP` with the reference's public methods and signatures (ACMMP.h:58-89).  The one systematic deviation:
// cv::Mat_<T> -> Map<T>, cv::Point / cv::Rect / float4 -> Point2i / Rect / Float4 (io.hpp), and RunPatchMatch takes the
// seed that replaces clock64() (ACMMP.cu:624).  Two extra overloads serve callers that already hold decoded images
// (InputInitialization(images, cameras)) and the device-side prior construction (BuildPlanarPrior).
class ACMMP {
public:
    explicit ACMMP(int device = 0) : device_(device) {
        if (acmmp_create(&ctx_, device)) throw std::runtime_error(acmmp_last_error(nullptr));
        acmmp_default_params(&params_);
    }
    ~ACMMP() { acmmp_destroy(ctx_); }
    ACMMP(const ACMMP &) = delete;
    ACMMP &operator=(const ACMMP &) = delete;

    void SetGeomConsistencyParams(bool multi_geometry = false) {   // ACMMP.cpp:447-454
        params_.geom_consistency = 1;
        params_.max_iterations = 2;
        if (multi_geometry) params_.multi_geometry = 1;
        check(acmmp_set_geom_consistency(ctx_, multi_geometry));
    }
    void SetHierarchyParams() {   // :456-459
        params_.hierarchy = 1;
        check(acmmp_set_hierarchy(ctx_));
    }
    void SetPlanarPriorParams() {   // :461-464
        params_.planar_prior = 1;
        check(acmmp_set_planar_prior_params(ctx_));
    }

    // InputInitialization (ACMMP.cpp:525-636): reads <dense_folder>/images/%08d.jpg (grey) and cams/%08d_cam.txt of the
    // reference view and its sources, limits every view to ITS problem's cur_image_size (cv::resize INTER_LINEAR on the
    // float image -> acmmp_resize_linear), widens the depth range, and in geometric-consistency mode reads the depth
    // maps of the previous stage from <output_folder>/2333_%08d/depths[_geom].dmb
    void InputInitialization(const std::string &dense_folder, const std::string &output_folder, const std::vector<Problem> &problem, const int idx) {
        images_.clear();
        cameras_.clear();
        depths_.clear();
        const Problem &local = problem[idx];
        std::vector<int> ids(1, local.ref_image_id);
        ids.insert(ids.end(), local.src_image_ids.begin(), local.src_image_ids.end());
        for (size_t i = 0; i < ids.size(); ++i) {
            // a source view is limited by ITS OWN problem's current size.  The reference indexes `problem` with the IMAGE ID
            // (:569-571), which is only right when ids are 0..n-1 in order; here the problem is looked up by its id
            int max_size = local.cur_image_size;
            if (i > 0)
                for (const Problem &q : problem)
                    if (q.ref_image_id == ids[i]) {
                        max_size = q.cur_image_size;
                        break;
                    }
            Map<float> image;
            acmmp_camera camera;
            LoadView(dense_folder, ids[i], max_size, image, camera);
            images_.push_back(std::move(image));
            cameras_.push_back(camera);
        }
        SetRangeParams();
        if (params_.geom_consistency) {
            const char *suffix = params_.multi_geometry ? "/depths_geom.dmb" : "/depths.dmb";
            for (int id : ids) {
                Map<float> depth;
                if (readDepthDmb(ResultFolder(output_folder, id) + suffix, depth)) throw std::runtime_error("cannot read the depth map of image " + std::to_string(id));
                depths_.push_back(std::move(depth));
            }
        }
    }
    // the same after the caller has read the files: grey float images and cameras of the reference view followed by
    // its sources (cameras carry width / height)
    void InputInitialization(const std::vector<Map<float>> &images, const std::vector<acmmp_camera> &cameras) {
        images_ = images;
        cameras_ = cameras;
        depths_.clear();
        SetRangeParams();
    }
    void SetDepthMap(int slot, const Map<float> &depth) {
        if ((int)depths_.size() <= slot) depths_.resize(slot + 1);
        depths_[slot] = depth;
    }

    // CudaSpaceInitialization (ACMMP.cpp:638-809): uploads the views; geometric consistency: depth maps + warm start
    // from depths[_geom].dmb / normals.dmb / costs.dmb; hierarchy: the coarse normals / costs and the JBU depth
    void CudaSpaceInitialization(const std::string &output_folder, const Problem &problem) {
        Upload();
        const std::string folder = ResultFolder(output_folder, problem.ref_image_id);
        if (params_.geom_consistency) {
            for (size_t i = 0; i < depths_.size(); ++i) check(acmmp_set_depth_map(ctx_, (int)i, depths_[i].data.data(), depths_[i].cols, depths_[i].rows, 0));
            Map<float> d, n, c;
            if (readDepthDmb(folder + (params_.multi_geometry ? "/depths_geom.dmb" : "/depths.dmb"), d) || readNormalDmb(folder + "/normals.dmb", n) ||
                readDepthDmb(folder + "/costs.dmb", c))
                throw std::runtime_error("cannot read the maps under " + folder);
            SetWarmStart(d, n, c);
        }
        if (params_.hierarchy) {
            Map<float> d, n, c;   // depths.dmb is at this scale already (RunJBU); normals / costs are the previous scale's
            if (readDepthDmb(folder + "/depths.dmb", d) || readNormalDmb(folder + "/normals.dmb", n) || readDepthDmb(folder + "/costs.dmb", c))
                throw std::runtime_error("cannot read the maps under " + folder);
            SetHierarchyInputs(n, c, d);
        }
    }
    // the same with the .dmb payloads in memory
    void CudaSpaceInitialization() { Upload(); }
    void SetWarmStart(const Map<float> &depth, const Map<float> &normal, const Map<float> &cost) {
        check(acmmp_set_warm_start(ctx_, depth.data.data(), normal.data.data(), cost.data.data()));
    }
    void SetHierarchyInputs(const Map<float> &coarse_normal, const Map<float> &coarse_cost, const Map<float> &fine_depth) {
        check(acmmp_set_hierarchy_inputs(ctx_, coarse_normal.data.data(), coarse_cost.data.data(), coarse_cost.cols, coarse_cost.rows,
                                         fine_depth.data.data()));
    }
    // SetPlanarPrior (:476-484): fixed per-pixel hypotheses of the seeded mode; takes ownership like the reference
    void SetPlanarPrior(std::unique_ptr<Float4[]> prior) {
        check(acmmp_set_seeded_planes(ctx_, reinterpret_cast<const float *>(prior.get())));
        params_.seeded = 1;
    }

    void RunPatchMatch(uint64_t seed = 0) {   // ACMMP.cu:1378-1456; seed replaces clock64()
        if (!uploaded_) Upload();
        check(acmmp_run(ctx_, seed));
        const size_t P = (size_t)cameras_[0].width * cameras_[0].height;
        planes_.resize(P);
        costs_.resize(P);
        check(acmmp_get_result(ctx_, reinterpret_cast<float *>(planes_.data()), costs_.data(), nullptr));
    }

    int GetReferenceImageWidth() const { return cameras_[0].width; }
    int GetReferenceImageHeight() const { return cameras_[0].height; }
    const Map<float> &GetReferenceImage() const { return images_[0]; }
    Float4 GetPlaneHypothesis(const int index) const { return planes_[index]; }   // world normal + depth after a run (:848-851)
    acmmp_camera GetCamera(const int index) const { return cameras_[index]; }
    float GetCost(const int index) const { return costs_[index]; }
    float GetMinDepth() const { return params_.depth_min; }
    float GetMaxDepth() const { return params_.depth_max; }

    // GetSupportPoints (:868-894): best pixel (cost < 0.1) of every 5 x 5 cell, cells in column-major order
    void GetSupportPoints(std::vector<Point2i> &support2DPoints) const {
        support2DPoints.clear();
        const int step = 5, width = GetReferenceImageWidth(), height = GetReferenceImageHeight();
        for (int col = 0; col < width; col += step)
            for (int row = 0; row < height; row += step) {
                float best = 2.0f;
                Point2i where;
                for (int c = col; c < std::min(width, col + step); ++c)
                    for (int r = row; r < std::min(height, row + step); ++r) {
                        const float cost = costs_[(size_t)r * width + c];
                        if (cost < 2.0f && best > cost) {
                            where = Point2i(c, r);
                            best = cost;
                        }
                    }
                if (best < 0.1f) support2DPoints.push_back(where);
            }
    }
    // DelaunayTriangulation (:896-918): cv::Subdiv2D -> the library's integer triangulation
    std::vector<Triangle> DelaunayTriangulation(const Rect boundRC, const std::vector<Point2i> &points) const {
        (void)boundRC;
        std::vector<Triangle> results;
        if (points.empty()) return results;
        std::vector<int32_t> xy(points.size() * 2), tris(points.size() * 12 + 12);
        for (size_t i = 0; i < points.size(); ++i) {
            xy[2 * i] = points[i].x;
            xy[2 * i + 1] = points[i].y;
        }
        const int nt = acmmp_delaunay(xy.data(), (int)points.size(), tris.data(), (int)tris.size() / 6);
        if (nt < 0) throw std::runtime_error("acmmp_delaunay failed");
        for (int t = 0; t < nt && t < (int)tris.size() / 6; ++t)
            results.push_back(Triangle(Point2i(tris[6 * t], tris[6 * t + 1]), Point2i(tris[6 * t + 2], tris[6 * t + 3]), Point2i(tris[6 * t + 4], tris[6 * t + 5])));
        return results;
    }
    // GetPriorPlaneParams (:920-953): plane n.X + d = 0, d > 0, through the three back-projected vertices
    Float4 GetPriorPlaneParams(const Triangle triangle, const Map<float> &depths) const {
        const acmmp_camera &cam = cameras_[0];
        const Point2i pts[3] = {triangle.pt1, triangle.pt2, triangle.pt3};
        double X[3][3];
        for (int k = 0; k < 3; ++k) {
            const float d = depths.at(pts[k].y, pts[k].x);
            X[k][0] = d * (pts[k].x - cam.K[2]) / cam.K[0];
            X[k][1] = d * (pts[k].y - cam.K[5]) / cam.K[4];
            X[k][2] = d;
        }
        const double ax = X[1][0] - X[0][0], ay = X[1][1] - X[0][1], az = X[1][2] - X[0][2];
        const double bx = X[2][0] - X[0][0], by = X[2][1] - X[0][1], bz = X[2][2] - X[0][2];
        const double nx = ay * bz - az * by, ny = az * bx - ax * bz, nz = ax * by - ay * bx;
        const double w = -(nx * X[0][0] + ny * X[0][1] + nz * X[0][2]);
        double nrm = std::sqrt(nx * nx + ny * ny + nz * nz);
        if (w < 0) nrm = -nrm;   // sign rule, :943-950
        Float4 out;
        out.x = (float)(nx / nrm);
        out.y = (float)(ny / nrm);
        out.z = (float)(nz / nrm);
        out.w = (float)(w / nrm);
        return out;
    }
    float GetDepthFromPlaneParam(const Float4 plane_hypothesis, const int x, const int y) const {   // :955-958
        const acmmp_camera &cam = cameras_[0];
        return -plane_hypothesis.w * cam.K[0] /
               ((x - cam.K[2]) * plane_hypothesis.x + (cam.K[0] / cam.K[4]) * (y - cam.K[5]) * plane_hypothesis.y + cam.K[0] * plane_hypothesis.z);
    }
    // CudaPlanarPriorInitialization (:811-831): per-triangle planes + per-pixel labels (label = triangle index + 1)
    void CudaPlanarPriorInitialization(const std::vector<Float4> &PlaneParams, const Map<float> &masks) {
        const int width = GetReferenceImageWidth(), height = GetReferenceImageHeight();
        std::vector<Float4> prior((size_t)width * height);
        std::vector<uint32_t> labels((size_t)width * height);
        for (int j = 0; j < height; ++j)
            for (int i = 0; i < width; ++i) {
                const size_t center = (size_t)j * width + i;
                labels[center] = (uint32_t)masks.at(j, i);
                if (masks.at(j, i) > 0) prior[center] = PlaneParams[(size_t)masks.at(j, i) - 1];
            }
        check(acmmp_set_planar_prior(ctx_, reinterpret_cast<const float *>(prior.data()), labels.data()));
    }
    // all of the above on the device (+ host Delaunay) in one call: what the resident schedule uses
    void BuildPlanarPrior(int *num_support = nullptr, int *num_triangles = nullptr) { check(acmmp_build_planar_prior(ctx_, num_support, num_triangles)); }

    void GetMaps(Map<float> &depth, Map<float> &normal, Map<float> &cost) {   // the three .dmb payloads (acmmp_definitions.cpp:283-295)
        const int width = GetReferenceImageWidth(), height = GetReferenceImageHeight();
        depth = Map<float>(height, width, 1);
        normal = Map<float>(height, width, 3);
        cost = Map<float>(height, width, 1);
        check(acmmp_get_maps(ctx_, depth.data.data(), normal.data.data(), cost.data.data()));
    }
    const acmmp_params &params() const { return params_; }
    acmmp_ctx *handle() { return ctx_; }

    static std::string ResultFolder(const std::string &output_folder, int id) {   // <output_folder>/2333_%08d
        char buf[64];
        snprintf(buf, sizeof buf, "/2333_%08d", id);
        return output_folder + buf;
    }

private:
    void check(int rc) {
        if (rc) throw std::runtime_error(acmmp_last_error(ctx_));
    }
    void SetRangeParams() {   // ACMMP.cpp:600-606
        params_.depth_min = cameras_[0].depth_min * 0.6f;
        params_.depth_max = cameras_[0].depth_max * 1.2f;
        params_.num_images = (int)images_.size();
        params_.disparity_min = cameras_[0].K[0] * params_.baseline / params_.depth_max;
        params_.disparity_max = cameras_[0].K[0] * params_.baseline / params_.depth_min;
        uploaded_ = false;
    }
    void Upload() {
        std::vector<const float *> ptrs;
        for (const auto &im : images_) ptrs.push_back(im.data.data());
        check(acmmp_set_problem(ctx_, (int)images_.size(), cameras_.data(), ptrs.data(), 0));
        for (size_t i = 0; i < depths_.size(); ++i)
            if (!depths_[i].empty() && !params_.geom_consistency) check(acmmp_set_depth_map(ctx_, (int)i, depths_[i].data.data(), depths_[i].cols, depths_[i].rows, 0));
        uploaded_ = true;
    }
    // one view as InputInitialization reads it (ACMMP.cpp:534-598)
    void LoadView(const std::string &dense_folder, int image_id, int max_image_size, Map<float> &image, acmmp_camera &camera) {
        char name[64];
        snprintf(name, sizeof name, "/%08d", image_id);
        Map<uint8_t> raw;
        std::string err;
        bool ok = false;
        for (const char *ext : {".jpg", ".pgm", ".ppm", ".raw"})
            if (ReadImageCached(dense_folder + "/images" + name + ext, raw, &err, false, device_)) {
                ok = true;
                break;
            }
        if (!ok) throw std::runtime_error(err.empty() ? "no image for id " + std::to_string(image_id) : err);
        image = ToGrayFloat(raw);
        if (!ReadCamera(dense_folder + "/cams" + name + "_cam.txt", camera)) throw std::runtime_error("cannot read the camera of image " + std::to_string(image_id));
        camera.height = image.rows;
        camera.width = image.cols;
        if (image.cols <= max_image_size && image.rows <= max_image_size) return;
        const float fx = static_cast<float>(max_image_size) / image.cols, fy = static_cast<float>(max_image_size) / image.rows;
        const float f = std::min(fx, fy);
        const int new_cols = (int)std::round(image.cols * f), new_rows = (int)std::round(image.rows * f);
        const float sx = new_cols / static_cast<float>(image.cols), sy = new_rows / static_cast<float>(image.rows);
        Map<float> scaled(new_rows, new_cols, 1);
        if (acmmp_resize_linear(device_, image.data.data(), image.cols, image.rows, scaled.data.data(), new_cols, new_rows)) throw std::runtime_error(acmmp_last_error(nullptr));
        image = std::move(scaled);
        camera.K[0] *= sx;
        camera.K[2] *= sx;
        camera.K[4] *= sy;
        camera.K[5] *= sy;
        camera.height = new_rows;
        camera.width = new_cols;
    }
    acmmp_ctx *ctx_ = nullptr;
    int device_ = 0;
    bool uploaded_ = false;
    acmmp_params params_;
    std::vector<Map<float>> images_, depths_;
    std::vector<acmmp_camera> cameras_;
    std::vector<Float4> planes_;
    std::vector<float> costs_;
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
    // sharded scans: the depth maps of the owned views -> every rank (ncclAllGather on the scan's stream; no-op when unsharded)
    void AllGather() { check(acmmp_scan_all_gather(s_)); }
    // the whole scan as a dataflow of (stage, view) tasks (all levels, all stages; collective on a sharded scan);
    // afterwards every rank serves every view's final maps
    void RunDataflow(uint64_t seed) { check(acmmp_scan_run_dataflow(s_, seed)); }
    void KeepPhotometric(bool on) { check(acmmp_scan_keep_photometric(s_, on ? 1 : 0)); }
    void GetPhotometricDepth(int view, Map<float> &depth) {
        int w = 0, h = 0;
        ViewSize(view, w, h);
        depth = Map<float>(h, w, 1);
        check(acmmp_scan_get_photometric_depth(s_, view, depth.data.data()));
    }
    // one process, one scan per GPU: communicators for scans[r] = rank r of scans.size() (ncclCommInitAll)
    template <class Ptr> static void CommInitAll(const std::vector<Ptr> &scans) {
        std::vector<acmmp_scan *> handles;
        for (const auto &p : scans) handles.push_back(p->handle());
        if (acmmp_scan_comm_init_all(handles.data(), (int)handles.size())) throw std::runtime_error(acmmp_scan_last_error(handles[0]));
    }
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
