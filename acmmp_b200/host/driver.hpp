// driver.hpp - C++ host mirror of the reference's per-view driver (acmmp_definitions.cpp): ComputeMultiScaleSettings
// (:207-243), ProcessProblem (:245-403) and JointBilateralUpsampling (:405-440), with the same names, argument
// meaning and on-disk hand-over (2333_%08d/{depths,depths_geom,normals,costs}.dmb between stages), written against
// `class ACMMP` of acmmp.hpp.  This is the reference's OWN flow - one problem at a time, files between stages - for
// callers that want exactly that; the resident schedule (class Scan) is the fast path.  OpenCV-free: Map<T>.
#pragma once
#include <sys/stat.h>

#include <cstdio>
#include <string>
#include <vector>

#include "acmmp.hpp"

namespace acmmp_b200 {

inline std::string ImagePath(const std::string &dense_folder, int id, const char *ext) {
    char buf[64];
    snprintf(buf, sizeof buf, "/images/%08d%s", id, ext);
    return dense_folder + buf;
}
inline bool ReadGreyImage(const std::string &dense_folder, int id, Map<uint8_t> &raw, int device) {
    std::string err;
    for (const char *ext : {".jpg", ".pgm", ".ppm", ".raw"})
        if (ReadImageCached(ImagePath(dense_folder, id, ext), raw, &err, false, device)) return true;
    return false;
}

// ComputeMultiScaleSettings (acmmp_definitions.cpp:207-243): halvings until the long side is <= 1000
inline int ComputeMultiScaleSettings(const std::string &dense_folder, std::vector<Problem> &problems, int device = 0) {
    int max_num_downscale = -1;
    const int size_bound = 1000;
    acmmp_params pmp;
    acmmp_default_params(&pmp);
    for (Problem &p : problems) {
        Map<uint8_t> raw;
        if (!ReadGreyImage(dense_folder, p.ref_image_id, raw, device)) throw std::runtime_error("no image for id " + std::to_string(p.ref_image_id));
        int max_size = std::max(raw.rows, raw.cols);
        if (max_size > pmp.max_image_size) max_size = pmp.max_image_size;
        p.max_image_size = max_size;
        int k = 0;
        while (max_size > size_bound) {
            max_size /= 2;
            k++;
        }
        if (k > max_num_downscale) max_num_downscale = k;
        p.num_downscale = k;
    }
    return max_num_downscale;
}

// the host construction of the planar prior inside ProcessProblem (acmmp_definitions.cpp:310-378), through the
// class's GetSupportPoints / DelaunayTriangulation / GetPriorPlaneParams / GetDepthFromPlaneParam
inline void BuildPlanarPriorOnHost(ACMMP &acmmp, const Map<float> &depths, std::vector<Float4> &planeParams_tri, Map<float> &mask_tri) {
    const int width = acmmp.GetReferenceImageWidth(), height = acmmp.GetReferenceImageHeight();
    const Rect imageRC(0, 0, width, height);
    std::vector<Point2i> support2DPoints;
    acmmp.GetSupportPoints(support2DPoints);
    const auto triangles = acmmp.DelaunayTriangulation(imageRC, support2DPoints);
    mask_tri = Map<float>(height, width, 1);
    planeParams_tri.clear();
    uint32_t idx = 0;
    for (const auto &triangle : triangles) {
        if (!(imageRC.contains(triangle.pt1) && imageRC.contains(triangle.pt2) && imageRC.contains(triangle.pt3))) continue;
        const float L01 = sqrt(pow(triangle.pt1.x - triangle.pt2.x, 2) + pow(triangle.pt1.y - triangle.pt2.y, 2));
        const float L02 = sqrt(pow(triangle.pt1.x - triangle.pt3.x, 2) + pow(triangle.pt1.y - triangle.pt3.y, 2));
        const float L12 = sqrt(pow(triangle.pt2.x - triangle.pt3.x, 2) + pow(triangle.pt2.y - triangle.pt3.y, 2));
        const float max_edge_length = std::max(L01, std::max(L02, L12));
        const float step = 1.0 / max_edge_length;
        // barycentric stepping; later triangles overwrite earlier ones (:346-352)
        for (float p = 0; p < 1.0; p += step)
            for (float q = 0; q < 1.0 - p; q += step) {
                const int x = p * triangle.pt1.x + q * triangle.pt2.x + (1.0 - p - q) * triangle.pt3.x;
                const int y = p * triangle.pt1.y + q * triangle.pt2.y + (1.0 - p - q) * triangle.pt3.y;
                mask_tri.at(y, x) = idx + 1.0;
            }
        planeParams_tri.push_back(acmmp.GetPriorPlaneParams(triangle, depths));
        idx++;
    }
    for (int i = 0; i < width; ++i)
        for (int j = 0; j < height; ++j)
            if (mask_tri.at(j, i) > 0) {
                const float d = acmmp.GetDepthFromPlaneParam(planeParams_tri[(size_t)mask_tri.at(j, i) - 1], i, j);
                if (!(d <= acmmp.GetMaxDepth() && d >= acmmp.GetMinDepth())) mask_tri.at(j, i) = 0;
            }
}

// ProcessProblem (acmmp_definitions.cpp:245-403).  `seed` replaces clock64(); `host_prior`: build the planar prior with
// the reference's host loop (BuildPlanarPriorOnHost) instead of the device construction (ACMMP::BuildPlanarPrior).
inline void ProcessProblem(const std::string &output_folder, const std::string &dense_folder, const std::vector<Problem> &problems, const int idx,
                           bool geom_consistency, bool planar_prior, bool hierarchy, bool multi_geometry = false, uint64_t seed = 0,
                           bool host_prior = true, int device = 0) {
    const Problem &problem = problems[idx];
    printf("Processing image %08d...\n", problem.ref_image_id);
    const std::string result_folder = ACMMP::ResultFolder(output_folder, problem.ref_image_id);
    mkdir(result_folder.c_str(), 0777);
    ACMMP acmmp(device);
    if (geom_consistency) acmmp.SetGeomConsistencyParams(multi_geometry);
    if (hierarchy) acmmp.SetHierarchyParams();
    acmmp.InputInitialization(dense_folder, output_folder, problems, idx);
    acmmp.CudaSpaceInitialization(output_folder, problem);
    acmmp.RunPatchMatch(seed);
    Map<float> depths, normals, costs;
    acmmp.GetMaps(depths, normals, costs);
    if (planar_prior) {
        printf("Run Planar Prior Assisted PatchMatch MVS ...\n");
        acmmp.SetPlanarPriorParams();
        if (host_prior) {
            std::vector<Float4> planeParams_tri;
            Map<float> mask_tri;
            BuildPlanarPriorOnHost(acmmp, depths, planeParams_tri, mask_tri);
            acmmp.CudaPlanarPriorInitialization(planeParams_tri, mask_tri);
        } else {
            acmmp.BuildPlanarPrior();
        }
        acmmp.RunPatchMatch(seed + 1);
        acmmp.GetMaps(depths, normals, costs);
    }
    if (writeDepthDmb(result_folder + (geom_consistency ? "/depths_geom.dmb" : "/depths.dmb"), depths) ||
        writeNormalDmb(result_folder + "/normals.dmb", normals) || writeDepthDmb(result_folder + "/costs.dmb", costs))
        throw std::runtime_error("cannot write the maps under " + result_folder);
    printf("Processing image %08d done!\n", problem.ref_image_id);
}

// JointBilateralUpsampling (acmmp_definitions.cpp:405-440): depths_geom.dmb of the previous scale -> depths.dmb at this
// scale, guided by the image resized to `acmmp_size`
inline void JointBilateralUpsampling(const std::string &dense_folder, const std::string &output_folder, const Problem &problem, int acmmp_size,
                                     int device = 0) {
    const std::string result_folder = ACMMP::ResultFolder(output_folder, problem.ref_image_id);
    Map<float> ref_depth;
    if (readDepthDmb(result_folder + "/depths_geom.dmb", ref_depth)) throw std::runtime_error("cannot read " + result_folder + "/depths_geom.dmb");
    Map<uint8_t> raw;
    if (!ReadGreyImage(dense_folder, problem.ref_image_id, raw, device)) throw std::runtime_error("no image for id " + std::to_string(problem.ref_image_id));
    const Map<float> image_float = ToGrayFloat(raw);
    const float factor_x = static_cast<float>(acmmp_size) / image_float.cols, factor_y = static_cast<float>(acmmp_size) / image_float.rows;
    const float factor = std::min(factor_x, factor_y);
    const int new_cols = (int)std::round(image_float.cols * factor), new_rows = (int)std::round(image_float.rows * factor);
    Map<float> scaled(new_rows, new_cols, 1);
    if (new_cols == image_float.cols && new_rows == image_float.rows) scaled = image_float;
    else if (acmmp_resize_linear(device, image_float.data.data(), image_float.cols, image_float.rows, scaled.data.data(), new_cols, new_rows))
        throw std::runtime_error(acmmp_last_error(nullptr));
    printf("Run JBU for image %d.jpg\n", problem.ref_image_id);
    Map<float> depthmap;
    if (!RunJBU(scaled, ref_depth, depthmap, device)) {   // integer scale 1: the reference returns without writing
        printf("Image.rows = Depthmap.rows\n");
        return;
    }
    mkdir(result_folder.c_str(), 0777);
    if (writeDepthDmb(result_folder + "/depths.dmb", depthmap)) throw std::runtime_error("cannot write " + result_folder + "/depths.dmb");
}

}  // namespace acmmp_b200
