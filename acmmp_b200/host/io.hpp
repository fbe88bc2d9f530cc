// io.hpp - the reference's on-disk formats either side of the path, without OpenCV:
// cam files, pair.txt, .dmb maps, the binary PLY, and grey / colour images.
// Reference: ReadCamera (ACMMP.cpp:149-177), GenerateSampleList
// (acmmp_definitions.cpp:179-205), read/write{Depth,Normal}Dmb (ACMMP.cpp:263-380),
// StoreColorPlyFileBinaryPointCloud (ACMMP.cpp:382-435).
#pragma once
#include <cfloat>
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <string>
#include <vector>

#include "../../include/acmmp_b200.h"

namespace acmmp_b200 {

// row-major map with interleaved channels: what cv::Mat_<T> holds in the reference
template <class T> struct Map {
    int rows = 0, cols = 0, channels = 1;
    std::vector<T> data;
    Map() {}
    Map(int r, int c, int ch = 1) : rows(r), cols(c), channels(ch), data((size_t)r * c * ch) {}
    bool empty() const { return data.empty(); }
    T &at(int r, int c, int k = 0) { return data[((size_t)r * cols + c) * channels + k]; }
    const T &at(int r, int c, int k = 0) const { return data[((size_t)r * cols + c) * channels + k]; }
};

struct Problem {   // acmmp_definitions.h:57-65
    int ref_image_id = 0;
    std::vector<int> src_image_ids;
};

// cams/%08d_cam.txt.  width / height are not in the file (the reference leaves them
// uninitialised until an image is loaded): they stay 0 here.
inline bool ReadCamera(const std::string &cam_path, acmmp_camera &camera) {
    std::ifstream file(cam_path);
    if (!file) return false;
    camera = acmmp_camera();
    std::string line;
    file >> line;
    for (int i = 0; i < 3; ++i) file >> camera.R[3 * i + 0] >> camera.R[3 * i + 1] >> camera.R[3 * i + 2] >> camera.t[i];
    float tmp[4];
    file >> tmp[0] >> tmp[1] >> tmp[2] >> tmp[3];
    file >> line;
    for (int i = 0; i < 3; ++i) file >> camera.K[3 * i + 0] >> camera.K[3 * i + 1] >> camera.K[3 * i + 2];
    float depth_num, interval;
    file >> camera.depth_min >> interval >> depth_num >> camera.depth_max;
    return !file.fail();
}

// pair.txt: sources with a score <= 0 are dropped, as in the reference
inline bool GenerateSampleList(const std::string &dense_folder, std::vector<Problem> &problems) {
    std::ifstream file(dense_folder + "/pair.txt");
    problems.clear();
    int num_images = 0;
    if (!(file >> num_images)) return false;
    for (int i = 0; i < num_images; ++i) {
        Problem problem;
        int num_src_images = 0;
        file >> problem.ref_image_id >> num_src_images;
        for (int j = 0; j < num_src_images; ++j) {
            int id;
            float score;
            file >> id >> score;
            if (score <= 0.0f) continue;
            problem.src_image_ids.push_back(id);
        }
        problems.push_back(problem);
    }
    return !file.fail();
}

// .dmb: int32 type (1 = float), h, w, channels, then floats
inline int writeDmb(const std::string &file_path, const float *data, int h, int w, int nb) {
    FILE *f = fopen(file_path.c_str(), "wb");
    if (!f) return -1;
    const int32_t hdr[4] = {1, h, w, nb};
    fwrite(hdr, sizeof(int32_t), 4, f);
    fwrite(data, sizeof(float), (size_t)h * w * nb, f);
    fclose(f);
    return 0;
}
inline int readDmb(const std::string &file_path, Map<float> &out) {
    FILE *f = fopen(file_path.c_str(), "rb");
    if (!f) return -1;
    int32_t hdr[4] = {-1, 0, 0, 0};
    if (fread(hdr, sizeof(int32_t), 4, f) != 4 || hdr[0] != 1) {
        fclose(f);
        return -1;
    }
    out = Map<float>(hdr[1], hdr[2], hdr[3]);
    const size_t n = fread(out.data.data(), sizeof(float), out.data.size(), f);
    fclose(f);
    return n == out.data.size() ? 0 : -1;
}
inline int writeDepthDmb(const std::string &p, const Map<float> &depth) { return writeDmb(p, depth.data.data(), depth.rows, depth.cols, 1); }
inline int writeNormalDmb(const std::string &p, const Map<float> &normal) { return writeDmb(p, normal.data.data(), normal.rows, normal.cols, 3); }
inline int readDepthDmb(const std::string &p, Map<float> &depth) { return readDmb(p, depth); }
inline int readNormalDmb(const std::string &p, Map<float> &normal) { return readDmb(p, normal); }

// Images.  OpenCV's codecs are not available to C++ in this build: binary PGM / PPM (P5 / P6,
// maxval 255) and the raw container the parity harness uses ("ACMRAW1 rows cols channels\n" +
// bytes, channels in stored = BGR order) are read here; JPEG streams are decoded on the device by
// nvJPEG through the library when the including program defines ACMMP_IO_WITH_JPEG (and links
// libacmmp_b200): `color` selects cv::imread's IMREAD_COLOR (BGR) or IMREAD_GRAYSCALE (luma).
inline bool ReadImage(const std::string &path, Map<uint8_t> &out, std::string *err = nullptr, bool color = true, int device = 0) {
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) {
        if (err) *err = "cannot open " + path;
        return false;
    }
    (void)color;
    (void)device;
#ifdef ACMMP_IO_WITH_JPEG
    {
        unsigned char soi[2] = {0, 0};
        if (fread(soi, 1, 2, f) == 2 && soi[0] == 0xFF && soi[1] == 0xD8) {
            fseek(f, 0, SEEK_END);
            const long n = ftell(f);
            fseek(f, 0, SEEK_SET);
            std::vector<uint8_t> bytes((size_t)n);
            const bool got = fread(bytes.data(), 1, bytes.size(), f) == bytes.size();
            fclose(f);
            int w = 0, h = 0;
            if (!got || acmmp_decode_jpeg(device, bytes.data(), bytes.size(), color ? 1 : 0, nullptr, 0, &w, &h)) {
                if (err) *err = path + ": " + (got ? acmmp_jpeg_last_error() : "read error");
                return false;
            }
            out = Map<uint8_t>(h, w, color ? 3 : 1);
            if (acmmp_decode_jpeg(device, bytes.data(), bytes.size(), color ? 1 : 0, out.data.data(), out.data.size(), &w, &h)) {
                if (err) *err = path + ": " + acmmp_jpeg_last_error();
                return false;
            }
            return true;
        }
        fseek(f, 0, SEEK_SET);
    }
#endif
    char magic[16] = {0};
    bool ok = false;
    if (fscanf(f, "%15s", magic) == 1) {
        const std::string m(magic);
        int a = 0, b = 0, c = 0;
        if (m == "ACMRAW1" && fscanf(f, "%d %d %d", &a, &b, &c) == 3 && fgetc(f) == '\n') {
            out = Map<uint8_t>(a, b, c);
            ok = fread(out.data.data(), 1, out.data.size(), f) == out.data.size();
        } else if ((m == "P5" || m == "P6") && fscanf(f, "%d %d %d", &a, &b, &c) == 3 && c == 255 && fgetc(f) != EOF) {
            out = Map<uint8_t>(b, a, m == "P6" ? 3 : 1);
            ok = fread(out.data.data(), 1, out.data.size(), f) == out.data.size();
            if (ok && m == "P6")   // PPM is RGB; keep OpenCV's BGR convention in memory
                for (size_t i = 0; i < out.data.size(); i += 3) std::swap(out.data[i], out.data[i + 2]);
        } else if (err) {
            *err = path + ": not a PGM / PPM / ACMRAW1 file (JPEG decoding is not built in)";
        }
    }
    fclose(f);
    if (!ok && err && err->empty()) *err = path + ": truncated image";
    return ok;
}

// grey float image as cv::imread(IMREAD_GRAYSCALE) + convertTo(CV_32FC1) gives it (ACMMP.cpp:534-541):
// BGR -> Y with OpenCV's fixed-point weights
inline Map<float> ToGrayFloat(const Map<uint8_t> &im) {
    Map<float> g(im.rows, im.cols, 1);
    const size_t n = (size_t)im.rows * im.cols;
    for (size_t i = 0; i < n; ++i) {
        if (im.channels == 1) g.data[i] = (float)im.data[i];
        else g.data[i] = (float)((im.data[3 * i] * 1868 + im.data[3 * i + 1] * 9617 + im.data[3 * i + 2] * 4899 + 8192) >> 14);
    }
    return g;
}

// binary little-endian PLY, vertex = x y z nx ny nz (float) red green blue (uchar); points9 as
// acmmp_fusion_get_points returns them (colour = stored BGR bytes): the reference writes
// red = c2, green = c1, blue = c0 (ACMMP.cpp:411-430)
inline bool StoreColorPlyFileBinaryPointCloud(const std::string &ply_path, const float *points9, size_t n) {
    FILE *f = fopen(ply_path.c_str(), "wb");
    if (!f) return false;
    fprintf(f, "ply\nformat binary_little_endian 1.0\nelement vertex %zu\n", n);
    fprintf(f, "property float x\nproperty float y\nproperty float z\nproperty float nx\nproperty float ny\nproperty float nz\n");
    fprintf(f, "property uchar red\nproperty uchar green\nproperty uchar blue\nend_header\n");
    std::vector<unsigned char> buf;
    buf.reserve(27 * 4096);
    for (size_t i = 0; i < n; ++i) {
        const float *p = points9 + 9 * i;
        float rec[6] = {p[0], p[1], p[2], p[3], p[4], p[5]};
        if (!(rec[0] < FLT_MAX && rec[0] > -FLT_MAX) || !(rec[1] < FLT_MAX && rec[1] > -FLT_MAX) || !(rec[2] < FLT_MAX && rec[2] >= -FLT_MAX))
            rec[0] = rec[1] = rec[2] = 0.0f;
        const unsigned char *b = reinterpret_cast<const unsigned char *>(rec);
        buf.insert(buf.end(), b, b + 24);
        buf.push_back((unsigned char)(int)p[8]);
        buf.push_back((unsigned char)(int)p[7]);
        buf.push_back((unsigned char)(int)p[6]);
        if (buf.size() >= 27 * 4096) {
            fwrite(buf.data(), 1, buf.size(), f);
            buf.clear();
        }
    }
    fwrite(buf.data(), 1, buf.size(), f);
    fclose(f);
    return true;
}

}  // namespace acmmp_b200
