// io.hpp - the reference's on-disk formats either side of the path, without OpenCV:
// cam files, pair.txt, .dmb maps, the binary PLY, and grey / colour images.
// Reference: ReadCamera (ACMMP.cpp:149-177), GenerateSampleList
// (acmmp_definitions.cpp:179-205), read/write{Depth,Normal}Dmb (ACMMP.cpp:263-380),
// StoreColorPlyFileBinaryPointCloud (ACMMP.cpp:382-435).
#pragma once
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <list>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include <sys/stat.h>

#include "../../include/acmmp_b200.h"
#ifdef ACMMP_IO_WITH_PNG
#include <zlib.h>   // link with -lz
#endif

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

struct Problem {   // acmmp_definitions.h:57-63
    int ref_image_id = 0;
    std::vector<int> src_image_ids;
    int max_image_size = 6400;
    int num_downscale = 0;
    int cur_image_size = 6400;
};

// stand-ins for the cv:: / CUDA vector types in the reference's signatures (ACMMP.h:60-89, acmmp_definitions.h:65-68)
struct Point2i {   // cv::Point
    int x = 0, y = 0;
    Point2i() {}
    Point2i(int x_, int y_) : x(x_), y(y_) {}
};
struct Rect {   // cv::Rect
    int x = 0, y = 0, width = 0, height = 0;
    Rect() {}
    Rect(int x_, int y_, int w_, int h_) : x(x_), y(y_), width(w_), height(h_) {}
    bool contains(const Point2i &p) const { return p.x >= x && p.x < x + width && p.y >= y && p.y < y + height; }
};
struct Triangle {   // acmmp_definitions.h:65-68
    Point2i pt1, pt2, pt3;
    Triangle(const Point2i &a, const Point2i &b, const Point2i &c) : pt1(a), pt2(b), pt3(c) {}
};
struct Float4 {   // CUDA's float4: plane hypothesis (nx, ny, nz, d)
    float x = 0, y = 0, z = 0, w = 0;
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
// the header is not trusted: a map is at most 2^15 x 2^15 with 1 or 3 channels (ACMMP.cpp:263-347 reads
// whatever the header says); `want_channels` > 0 additionally fixes the channel count
inline int readDmb(const std::string &file_path, Map<float> &out, int want_channels = 0) {
    FILE *f = fopen(file_path.c_str(), "rb");
    if (!f) return -1;
    int32_t hdr[4] = {-1, 0, 0, 0};
    if (fread(hdr, sizeof(int32_t), 4, f) != 4 || hdr[0] != 1 || hdr[1] <= 0 || hdr[2] <= 0 || hdr[1] > 32768 || hdr[2] > 32768 ||
        (hdr[3] != 1 && hdr[3] != 3) || (want_channels > 0 && hdr[3] != want_channels)) {
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
inline int readDepthDmb(const std::string &p, Map<float> &depth) { return readDmb(p, depth, 1); }
inline int readNormalDmb(const std::string &p, Map<float> &normal) { return readDmb(p, normal, 3); }

#ifdef ACMMP_IO_WITH_PNG
// PNG (8 / 16 bit; grey, grey + alpha, RGB, RGBA; not interlaced, no palette) decoded with zlib into
// 16-bit samples in cv::imread(path, -1) order: colour channels as B, G, R (, A).  Used for the 16-bit
// prior images of the seeded mode (acmmp_definitions.cpp:104-105) and for 8-bit masks (:886-902).
inline bool ReadPng16(const std::string &path, Map<uint16_t> &out, int *bit_depth = nullptr, std::string *err = nullptr) {
    auto fail = [&](const std::string &why) {
        if (err) *err = path + ": " + why;
        return false;
    };
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) return fail("cannot open");
    fseek(f, 0, SEEK_END);
    const long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::vector<uint8_t> buf((size_t)(n > 0 ? n : 0));
    const bool got = n > 0 && fread(buf.data(), 1, buf.size(), f) == buf.size();
    fclose(f);
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    if (!got || buf.size() < 8 || memcmp(buf.data(), sig, 8) != 0) return fail("not a PNG file");
    auto be32 = [&](size_t at) { return ((uint32_t)buf[at] << 24) | ((uint32_t)buf[at + 1] << 16) | ((uint32_t)buf[at + 2] << 8) | buf[at + 3]; };
    uint32_t w = 0, h = 0;
    int depth = 0, ctype = -1;
    std::vector<uint8_t> idat;
    for (size_t pos = 8; pos + 12 <= buf.size();) {
        const uint32_t len = be32(pos);
        if (pos + 12 + (size_t)len > buf.size()) return fail("truncated chunk");
        const std::string type(reinterpret_cast<const char *>(&buf[pos + 4]), 4);
        const uint8_t *d = &buf[pos + 8];
        if (type == "IHDR") {
            if (len < 13) return fail("bad IHDR");
            w = be32(pos + 8);
            h = be32(pos + 12);
            depth = d[8];
            ctype = d[9];
            if (d[12] != 0) return fail("interlaced PNG not supported");
        } else if (type == "IDAT") {
            idat.insert(idat.end(), d, d + len);
        } else if (type == "IEND") {
            break;
        }
        pos += 12 + (size_t)len;
    }
    const int cn = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (!w || !h || !cn || (depth != 8 && depth != 16)) return fail("unsupported PNG (need 8/16-bit grey or RGB, no palette)");
    const size_t bpp = (size_t)cn * depth / 8, stride = bpp * w;
    std::vector<uint8_t> raw((stride + 1) * h);
    uLongf raw_len = (uLongf)raw.size();
    if (uncompress(raw.data(), &raw_len, idat.data(), (uLong)idat.size()) != Z_OK || raw_len != raw.size()) return fail("bad zlib stream");
    std::vector<uint8_t> prev(stride, 0), cur(stride);
    out = Map<uint16_t>((int)h, (int)w, cn);
    for (uint32_t y = 0; y < h; ++y) {
        const uint8_t *line = &raw[(stride + 1) * y];
        const int filter = line[0];
        for (size_t x = 0; x < stride; ++x) {
            const int a = x >= bpp ? cur[x - bpp] : 0, b = prev[x], c = x >= bpp ? prev[x - bpp] : 0;
            int pred = 0;
            if (filter == 1) pred = a;
            else if (filter == 2) pred = b;
            else if (filter == 3) pred = (a + b) >> 1;
            else if (filter == 4) {
                const int pp = a + b - c, pa = abs(pp - a), pb = abs(pp - b), pc = abs(pp - c);
                pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
            } else if (filter != 0) return fail("bad filter type");
            cur[x] = (uint8_t)(line[1 + x] + pred);
        }
        for (uint32_t x = 0; x < w; ++x)
            for (int k = 0; k < cn; ++k) {
                const int src = (cn >= 3 && k < 3) ? 2 - k : k;   // RGB(A) in the file -> BGR(A) like cv::imread
                const uint8_t *sp = &cur[(size_t)x * bpp + (size_t)src * depth / 8];
                out.data[((size_t)y * w + x) * cn + k] = depth == 16 ? (uint16_t)((sp[0] << 8) | sp[1]) : sp[0];
            }
        prev.swap(cur);
    }
    if (bit_depth) *bit_depth = depth;
    return true;
}
#endif

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

// Content cache of decoded images.  The reference decodes every view of every problem again at every stage
// (InputInitialization, ACMMP.cpp:534-564: 49 x 11 x 6 cv::imread calls on the headline scan); its per-view flow
// (host/driver.hpp: ProcessProblem, JointBilateralUpsampling, ComputeMultiScaleSettings) goes through this cache
// instead: one decode per file and colour mode, keyed by path, validated against the file's size and modification
// time, least-recently-used entries dropped beyond the byte budget (ACMMP_IMAGE_CACHE_MB, default 4096; 0 = off).
// The resident schedule (class Scan) needs none: it holds every decoded view in HBM for the whole scan.
class ImageCache {
  public:
    struct Stats { uint64_t hits = 0, misses = 0, evictions = 0; size_t bytes = 0; };
    static ImageCache &Instance() {
        static ImageCache c;
        return c;
    }
    bool Read(const std::string &path, Map<uint8_t> &out, std::string *err, bool color, int device) {
        struct stat st;
        if (budget_ == 0 || stat(path.c_str(), &st) != 0) return ReadImage(path, out, err, color, device);
        const std::string key = path + (color ? "#c" : "#g");
        {
            std::lock_guard<std::mutex> lock(mu_);
            auto it = index_.find(key);
            if (it != index_.end()) {
                Entry &e = *it->second;
                if (e.size == st.st_size && e.mtime_s == st.st_mtim.tv_sec && e.mtime_ns == st.st_mtim.tv_nsec) {
                    lru_.splice(lru_.begin(), lru_, it->second);
                    out = e.image;
                    stats_.hits++;
                    return true;
                }
                stats_.bytes -= e.image.data.size();   // the file changed: decode again
                lru_.erase(it->second);
                index_.erase(it);
            }
        }
        if (!ReadImage(path, out, err, color, device)) return false;
        std::lock_guard<std::mutex> lock(mu_);
        stats_.misses++;
        if (out.data.size() > budget_ || index_.count(key)) return true;
        lru_.push_front(Entry{key, out, st.st_size, st.st_mtim.tv_sec, st.st_mtim.tv_nsec});
        index_[key] = lru_.begin();
        stats_.bytes += out.data.size();
        while (stats_.bytes > budget_ && lru_.size() > 1) {
            stats_.bytes -= lru_.back().image.data.size();
            index_.erase(lru_.back().key);
            lru_.pop_back();
            stats_.evictions++;
        }
        return true;
    }
    Stats stats() {
        std::lock_guard<std::mutex> lock(mu_);
        return stats_;
    }
    void SetBudget(size_t bytes) {
        std::lock_guard<std::mutex> lock(mu_);
        budget_ = bytes;
        while (stats_.bytes > budget_ && !lru_.empty()) {
            stats_.bytes -= lru_.back().image.data.size();
            index_.erase(lru_.back().key);
            lru_.pop_back();
            stats_.evictions++;
        }
    }

  private:
    struct Entry {
        std::string key;
        Map<uint8_t> image;
        off_t size;
        time_t mtime_s;
        long mtime_ns;
    };
    ImageCache() {
        const char *e = getenv("ACMMP_IMAGE_CACHE_MB");
        budget_ = (size_t)(e ? atoll(e) : 4096) << 20;
    }
    std::mutex mu_;
    std::list<Entry> lru_;
    std::map<std::string, std::list<Entry>::iterator> index_;
    size_t budget_ = 0;
    Stats stats_;
};
inline bool ReadImageCached(const std::string &path, Map<uint8_t> &out, std::string *err = nullptr, bool color = true, int device = 0) {
    return ImageCache::Instance().Read(path, out, err, color, device);
}

// cv::resize(src, dst, Size(dw, dh), 0, 0, INTER_LINEAR) on 8-bit images, as RunFusion applies it to the colour
// image (RescaleImageAndCamera, ACMMP.cpp:179-203) and to the mask (acmmp_definitions.cpp:897) when their size
// differs from the depth map's.  OpenCV's integer scheme: 11-bit coefficients saturate_cast<short>(w * 2048),
// horizontal pass in int, vertical pass ((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2; an exact
// 2 x 2 reduction takes the INTER_AREA fast path (a + b + c + d + 2) >> 2 instead (imgproc/resize.cpp).
inline Map<uint8_t> ResizeLinear8(const Map<uint8_t> &src, int dw, int dh) {
    const int cn = src.channels, sw = src.cols, sh = src.rows;
    Map<uint8_t> dst(dh, dw, cn);
    if (sw == dw && sh == dh) {
        dst.data = src.data;
        return dst;
    }
    if (sw == 2 * dw && sh == 2 * dh) {
        for (int y = 0; y < dh; ++y)
            for (int x = 0; x < dw; ++x)
                for (int k = 0; k < cn; ++k)
                    dst.at(y, x, k) = (uint8_t)((src.at(2 * y, 2 * x, k) + src.at(2 * y, 2 * x + 1, k) + src.at(2 * y + 1, 2 * x, k) +
                                                 src.at(2 * y + 1, 2 * x + 1, k) + 2) >> 2);
        return dst;
    }
    const double scale_x = (double)sw / dw, scale_y = (double)sh / dh;
    auto coef = [](float f) {   // saturate_cast<short>(f * INTER_RESIZE_COEF_SCALE): round half to even
        const long v = lrintf(f * 2048.0f);
        return (int)(v < -32768 ? -32768 : v > 32767 ? 32767 : v);
    };
    std::vector<int> xofs(dw), yofs(dh), alpha(2 * (size_t)dw), beta(2 * (size_t)dh);
    for (int dx = 0; dx < dw; ++dx) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = (int)floorf(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        xofs[dx] = sx;
        alpha[2 * dx] = coef(1.f - fx);
        alpha[2 * dx + 1] = coef(fx);
    }
    for (int dy = 0; dy < dh; ++dy) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = (int)floorf(fy);
        fy -= sy;
        yofs[dy] = sy;
        beta[2 * dy] = coef(1.f - fy);
        beta[2 * dy + 1] = coef(fy);
    }
    std::vector<int> row0((size_t)dw * cn), row1((size_t)dw * cn);
    auto hpass = [&](int sy, std::vector<int> &row) {
        sy = sy < 0 ? 0 : (sy > sh - 1 ? sh - 1 : sy);
        for (int dx = 0; dx < dw; ++dx) {
            const int sx = xofs[dx], sx1 = sx + 1 < sw ? sx + 1 : sw - 1;
            for (int k = 0; k < cn; ++k) row[(size_t)dx * cn + k] = src.at(sy, sx, k) * alpha[2 * dx] + src.at(sy, sx1, k) * alpha[2 * dx + 1];
        }
    };
    for (int dy = 0; dy < dh; ++dy) {
        hpass(yofs[dy], row0);
        hpass(yofs[dy] + 1, row1);
        const int b0 = beta[2 * dy], b1 = beta[2 * dy + 1];
        for (size_t i = 0; i < (size_t)dw * cn; ++i)
            dst.data[(size_t)dy * dw * cn + i] = (uint8_t)((((b0 * (row0[i] >> 4)) >> 16) + ((b1 * (row1[i] >> 4)) >> 16) + 2) >> 2);
    }
    return dst;
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

// the same before the convertTo: the 8-bit grey image (handed to acmmp_scan_set_image_u8, which converts on the device)
inline Map<uint8_t> ToGray8(const Map<uint8_t> &im) {
    if (im.channels == 1) return im;
    Map<uint8_t> g(im.rows, im.cols, 1);
    const size_t n = (size_t)im.rows * im.cols;
    for (size_t i = 0; i < n; ++i)
        g.data[i] = (uint8_t)((im.data[3 * i] * 1868 + im.data[3 * i + 1] * 9617 + im.data[3 * i + 2] * 4899 + 8192) >> 14);
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
