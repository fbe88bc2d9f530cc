// fuse_common.hpp - the file-reading half of RunFusion / RunPriorAwareFusion (acmmp_definitions.cpp:575-666, 828-912)
// shared by the two command lines (acmmp_main: maps straight from the scan; fuse_data: maps from the output folder):
// colour image of every view at the size of its depth map, optional PNG masks, the two map sets of the prior-aware
// mode, then the fusion on the device and the PLY.
#pragma once
#include <cstdio>
#include <functional>
#include <string>
#include <vector>

#define ACMMP_IO_WITH_JPEG 1
#define ACMMP_IO_WITH_PNG 1
#include "../acmmp_b200/host/acmmp.hpp"

namespace acmmp_cli {
using namespace acmmp_b200;

inline std::string id_name(int id, const char *suffix) {
    char buf[64];
    snprintf(buf, sizeof buf, "%08d%s", id, suffix);
    return buf;
}

inline bool find_image(const std::string &folder, int id, Map<uint8_t> &im, std::string &err, bool color, int device) {
    for (const char *ext : {".jpg", ".pgm", ".ppm", ".raw"}) {   // the parity tests keep the reference's .jpg name for the raw container
        std::string e;
        if (ReadImage(folder + "/" + id_name(id, ext), im, &e, color, device)) return true;
        if (!e.empty() && e.find("cannot open") == std::string::npos) err = e;
    }
    if (err.empty()) err = "no image for id " + std::to_string(id) + " under " + folder;
    return false;
}

// <dense>/<mask_dir>/%08d.png as cv::imread(path, -1) returns it (acmmp_definitions.cpp:886-895), first channel, as bytes;
// 16-bit samples are clamped to 255 (only `< 128` is asked of them).  PGM / raw containers are accepted too.
inline bool find_mask(const std::string &folder, int id, Map<uint8_t> &m, std::string &err, int device) {
    Map<uint16_t> png;
    std::string e;
    if (ReadPng16(folder + "/" + id_name(id, ".png"), png, nullptr, &e)) {
        m = Map<uint8_t>(png.rows, png.cols, 1);
        for (size_t k = 0; k < (size_t)png.rows * png.cols; ++k) {
            const uint16_t v = png.data[k * png.channels];
            m.data[k] = (uint8_t)(v > 255 ? 255 : v);
        }
        return true;
    }
    if (e.find("cannot open") == std::string::npos) {
        err = e;
        return false;
    }
    Map<uint8_t> any;
    if (!find_image(folder, id, any, err, false, device)) return false;
    m = Map<uint8_t>(any.rows, any.cols, 1);
    for (size_t k = 0; k < (size_t)any.rows * any.cols; ++k) m.data[k] = any.data[k * any.channels];
    return true;
}

// RescaleImageAndCamera (ACMMP.cpp:179-203): the camera at the size of the depth map
inline acmmp_camera CameraAtMapSize(const acmmp_camera &cam, int cols, int rows) {
    acmmp_camera c = cam;
    if (cols == cam.width && rows == cam.height) return c;
    const float sx = cols / (float)cam.width, sy = rows / (float)cam.height;
    c.K[0] *= sx;
    c.K[2] *= sx;
    c.K[4] *= sy;
    c.K[5] *= sy;
    c.width = cols;
    c.height = rows;
    return c;
}

struct FuseOptions {
    std::string dense_folder, output_folder, fusion_folder, image_dir = "/images", mask_dir = " ";
    std::string set0_depth_name = "/depths_geom.dmb";   // map set 0 of the prior-aware mode (fusion folder)
    bool prior_aware = false;
    float consistency_scalar = 0.3f;
    int num_consistent_thresh = 1, single_match_penalty = 0, device = 0;
};

// maps(i, depth, normal): this run's maps of view i (host).  cams_out: cameras at the size of those maps.
// Returns the number of fused points, < 0 after printing an error.
inline long long FuseAndStore(const FuseOptions &o, const std::vector<Problem> &problems, const std::vector<acmmp_camera> &cams_out,
                              const std::vector<std::vector<int>> &pairs,
                              const std::function<bool(int, Map<float> &, Map<float> &)> &maps) {
    const int n = (int)problems.size();
    Fusion fusion(cams_out, pairs, o.device);
    for (int i = 0; i < n; ++i) {
        const int id = problems[i].ref_image_id;
        Map<uint8_t> color;   // colour read for the point colours (acmmp_definitions.cpp:857)
        std::string err;
        if (!find_image(o.dense_folder + o.image_dir, id, color, err, true, o.device)) {
            fprintf(stderr, "%s\n", err.c_str());
            return -1;
        }
        if (color.channels == 1) {   // grey input: replicate like imread(IMREAD_COLOR)
            Map<uint8_t> c3(color.rows, color.cols, 3);
            for (size_t k = 0; k < (size_t)color.rows * color.cols; ++k) c3.data[3 * k] = c3.data[3 * k + 1] = c3.data[3 * k + 2] = color.data[k];
            color = c3;
        }
        // RescaleImageAndCamera (ACMMP.cpp:179-203): the colour image at the size of the depth map
        if (color.cols != cams_out[i].width || color.rows != cams_out[i].height) color = ResizeLinear8(color, cams_out[i].width, cams_out[i].height);
        std::vector<uint8_t> mask;
        if (o.mask_dir != " ") {   // acmmp_definitions.cpp:884-899: %08d.png, resized to the depth map, mask = value < 128
            Map<uint8_t> m;
            if (!find_mask(o.dense_folder + "/" + o.mask_dir, id, m, err, o.device)) {
                fprintf(stderr, "mask of image %d: %s\n", id, err.c_str());
                return -1;
            }
            if (m.cols != cams_out[i].width || m.rows != cams_out[i].height) m = ResizeLinear8(m, cams_out[i].width, cams_out[i].height);
            mask.resize((size_t)m.rows * m.cols);
            for (size_t k = 0; k < mask.size(); ++k) mask[k] = m.data[k] < 128 ? 1 : 0;
        }
        Map<float> hd, hn;
        if (!maps(i, hd, hn) || hd.cols != cams_out[i].width || hd.rows != cams_out[i].height || hn.cols != hd.cols || hn.rows != hd.rows) {
            fprintf(stderr, "maps of image %d missing or of the wrong size\n", id);
            return -1;
        }
        if (o.prior_aware) {
            // set 0 = the maps of the fusion folder, set 1 = this run's (acmmp_definitions.cpp:607-643)
            const std::string folder = o.fusion_folder + "/2333_" + id_name(id, "");
            Map<float> d0, n0;
            if (readDepthDmb(folder + o.set0_depth_name, d0) || readNormalDmb(folder + "/normals.dmb", n0) || d0.cols != cams_out[i].width ||
                d0.rows != cams_out[i].height) {
                fprintf(stderr, "cannot read the maps of image %d under %s\n", id, folder.c_str());
                return -1;
            }
            fusion.SetView(i, d0.data.data(), n0.data.data(), color.data.data(), mask.empty() ? nullptr : mask.data(), false);
            fusion.SetViewSecond(i, hd.data.data(), hn.data.data(), false);
        } else {
            fusion.SetView(i, hd.data.data(), hn.data.data(), color.data.data(), mask.empty() ? nullptr : mask.data(), false);
        }
    }
    const int64_t count = o.prior_aware ? fusion.RunPriorAware(o.consistency_scalar, o.num_consistent_thresh, o.single_match_penalty)
                                        : fusion.Run(o.consistency_scalar, o.num_consistent_thresh);
    std::vector<float> pts;
    std::vector<int64_t> per_view;
    fusion.GetPoints(pts, per_view, count, n);
    for (int i = 0; i < n; ++i) printf("Fusing image %08d... approved %lld pts\n", i, (long long)per_view[i]);
    if (!StoreColorPlyFileBinaryPointCloud(o.output_folder + (o.prior_aware ? "/ACMMP_prior_model.ply" : "/ACMMP_model.ply"), pts.data(), (size_t)count)) {
        fprintf(stderr, "cannot write the PLY\n");
        return -1;
    }
    printf("store 3D points to ply file: %lld points\n", (long long)count);
    return (long long)count;
}

}  // namespace acmmp_cli
