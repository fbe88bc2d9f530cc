// Stand-in for the OpenCV C++ headers (OpenCV C++ is not installed in this image), used
// ONLY to compile the reference's sources where they lie under /root/reference/src for the
// parity comparators under oracle/_ref:
//   * libacmmp_ref.so      (oracle/ref_harness.cu): the reference's CUDA kernels - none of
//     this header executes there;
//   * libacmmp_ref_host.so (oracle/ref_host_harness.cpp): the reference's host code
//     (RunFusion and its helpers) - uses the functional subset below: Mat / Mat_ storage and
//     element access, zeros, clone, convertTo, imread (a raw container, see imread below),
//     resize (bilinear / nearest), SVD::solveZ for 3 x 4 systems, Subdiv2D through a hook.
//     Display, drawing and imwrite are inert (debug windows / debug images only).
// Test infrastructure - not part of the product.
#pragma once
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

typedef unsigned char uchar;

#define CV_8U 0
#define CV_8UC1 0
#define CV_8UC3 16
#define CV_16U 2
#define CV_16UC1 2
#define CV_16UC3 18
#define CV_32F 5
#define CV_32FC1 5
#define CV_32FC3 21
#define CV_MAT_DEPTH(t) ((t) & 7)
#define CV_MAT_CN(t) ((((t) >> 3) & 63) + 1)

namespace cv {

struct Point { int x, y; Point() : x(0), y(0) {} Point(int a, int b) : x(a), y(b) {} };
struct Point2f { float x, y; Point2f() : x(0), y(0) {} Point2f(float a, float b) : x(a), y(b) {} };
struct Size { int width, height; Size() : width(0), height(0) {} Size(int w, int h) : width(w), height(h) {} };
struct Rect {
    int x, y, width, height;
    Rect() : x(0), y(0), width(0), height(0) {}
    Rect(int a, int b, int c, int d) : x(a), y(b), width(c), height(d) {}
    bool contains(const Point &p) const { return p.x >= x && p.x < x + width && p.y >= y && p.y < y + height; }
};
struct Scalar { double val[4]; Scalar(double a = 0, double b = 0, double c = 0, double d = 0) { val[0] = a; val[1] = b; val[2] = c; val[3] = d; } };

template <class T, int N> struct Vec {
    T val[N];
    Vec() { for (int i = 0; i < N; ++i) val[i] = T(); }
    Vec(T a, T b, T c) { static_assert(N == 3, "3 components"); val[0] = a; val[1] = b; val[2] = c; }
    T &operator[](int i) { return val[i]; }
    const T &operator[](int i) const { return val[i]; }
    Vec operator+(const Vec &o) const { Vec r; for (int i = 0; i < N; ++i) r.val[i] = val[i] + o.val[i]; return r; }
    Vec operator-(const Vec &o) const { Vec r; for (int i = 0; i < N; ++i) r.val[i] = val[i] - o.val[i]; return r; }
    Vec &operator/=(double d) { for (int i = 0; i < N; ++i) val[i] = (T)(val[i] / d); return *this; }
    Vec operator*(double d) const { Vec r; for (int i = 0; i < N; ++i) r.val[i] = (T)(val[i] * d); return r; }
    Vec operator/(double d) const { Vec r; for (int i = 0; i < N; ++i) r.val[i] = (T)(val[i] / d); return r; }
};
typedef Vec<float, 3> Vec3f;
typedef Vec<float, 6> Vec6f;
typedef Vec<unsigned char, 3> Vec3b;

template <class T> struct TypeOf;
template <> struct TypeOf<float> { enum { value = CV_32FC1 }; };
template <> struct TypeOf<uchar> { enum { value = CV_8UC1 }; };
template <> struct TypeOf<Vec3f> { enum { value = CV_32FC3 }; };
template <> struct TypeOf<Vec3b> { enum { value = CV_8UC3 }; };

enum { IMREAD_GRAYSCALE = 0, IMREAD_COLOR = 1, INTER_NEAREST = 0, INTER_LINEAR = 1, WINDOW_NORMAL = 0 };

struct Mat {
    int rows = 0, cols = 0;
    int type_ = CV_8UC1;
    unsigned char *data = nullptr;
    size_t step[2] = {0, 0};
    std::shared_ptr<std::vector<unsigned char>> store;

    static size_t elem_size(int type) { return (size_t)(CV_MAT_DEPTH(type) == CV_32F ? 4 : CV_MAT_DEPTH(type) == CV_16U ? 2 : 1) * CV_MAT_CN(type); }
    void create(int r, int c, int type) {
        rows = r; cols = c; type_ = type;
        store = std::make_shared<std::vector<unsigned char>>((size_t)r * c * elem_size(type), 0);
        data = store->data();
        step[1] = elem_size(type);
        step[0] = step[1] * c;
    }
    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(int r, int c, int type, const Scalar &s) {
        create(r, c, type);
        const int cn = CV_MAT_CN(type);
        for (size_t i = 0; i < (size_t)r * c; ++i)
            for (int k = 0; k < cn; ++k) {
                if (CV_MAT_DEPTH(type) == CV_32F) reinterpret_cast<float *>(data)[i * cn + k] = (float)s.val[k];
                else data[i * cn + k] = (unsigned char)s.val[k];
            }
    }
    static Mat zeros(int r, int c, int type) { return Mat(r, c, type); }
    int type() const { return type_; }
    int channels() const { return CV_MAT_CN(type_); }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    size_t total() const { return (size_t)rows * cols; }
    Mat clone() const {
        Mat m;
        if (!empty()) {
            m.create(rows, cols, type_);
            memcpy(m.data, data, store->size());
        }
        return m;
    }
    template <class T> T &at(int r, int c) { return *reinterpret_cast<T *>(data + (size_t)r * step[0] + (size_t)c * step[1]); }
    template <class T> const T &at(int r, int c) const { return *reinterpret_cast<const T *>(data + (size_t)r * step[0] + (size_t)c * step[1]); }
    template <class T> T *ptr(int r = 0) { return reinterpret_cast<T *>(data + (size_t)r * step[0]); }
    template <class T> const T *ptr(int r = 0) const { return reinterpret_cast<const T *>(data + (size_t)r * step[0]); }
    // value conversion with saturation for 8-bit targets, same channel count
    void convertTo(Mat &dst, int rtype, double alpha = 1.0, double beta = 0.0) const {
        const int cn = channels();
        const int dtype = (CV_MAT_DEPTH(rtype)) | ((cn - 1) << 3);
        Mat out(rows, cols, dtype);
        const size_t n = total() * cn;
        if (CV_MAT_DEPTH(type_) == CV_16U && CV_MAT_DEPTH(rtype) == CV_32F) {
            // 16-bit -> float with scale: OpenCV's cvtScale works in float for this pair (src * (float)alpha + (float)beta)
            const float fa = (float)alpha, fb = (float)beta;
            for (size_t i = 0; i < n; ++i) {
                const float prod = (float)reinterpret_cast<const unsigned short *>(data)[i] * fa;
                reinterpret_cast<float *>(out.data)[i] = prod + fb;
            }
            dst = out;
            return;
        }
        for (size_t i = 0; i < n; ++i) {
            const double v = (CV_MAT_DEPTH(type_) == CV_32F ? (double)reinterpret_cast<const float *>(data)[i] : (double)data[i]) * alpha + beta;
            if (CV_MAT_DEPTH(rtype) == CV_32F) reinterpret_cast<float *>(out.data)[i] = (float)v;
            else out.data[i] = (unsigned char)(v < 0 ? 0 : v > 255 ? 255 : std::lrint(v));
        }
        dst = out;
    }
};

template <class T> struct Mat_ : public Mat {
    Mat_() { type_ = TypeOf<T>::value; }
    Mat_(int r, int c) { create(r, c, TypeOf<T>::value); }
    Mat_(const Mat &m) { assign(m); }
    Mat_ &operator=(const Mat &m) { assign(m); return *this; }
    void assign(const Mat &m) {
        if (m.empty() || m.type_ == TypeOf<T>::value) {
            static_cast<Mat &>(*this) = m;
            if (m.empty()) type_ = TypeOf<T>::value;
        } else {
            Mat tmp;
            m.convertTo(tmp, CV_MAT_DEPTH(TypeOf<T>::value));
            static_cast<Mat &>(*this) = tmp;
        }
    }
    Mat_ clone() const { return Mat_(Mat::clone()); }
    T &operator()(int r, int c) { return this->template at<T>(r, c); }
    const T &operator()(int r, int c) const { return this->template at<T>(r, c); }
};

// element-wise "a < v" -> 255 / 0 and integer division, as used for the depth masks
inline Mat operator<(const Mat &a, double v) {
    Mat out(a.rows, a.cols, CV_8UC1);
    for (size_t i = 0; i < a.total(); ++i)
        out.data[i] = ((CV_MAT_DEPTH(a.type_) == CV_32F ? (double)reinterpret_cast<const float *>(a.data)[i * a.channels()] : (double)a.data[i * a.channels()]) < v) ? 255 : 0;
    return out;
}
inline Mat operator/(const Mat &a, double v) {
    Mat out;
    a.convertTo(out, CV_MAT_DEPTH(a.type_), 1.0 / v);
    return out;
}

// imread: OpenCV's codecs are not available.  The tests write images as a raw container
// under the file name the reference expects: "ACMRAW1\n" rows cols channels "\n" + bytes;
// "ACMRAW2" holds 16-bit little-endian samples (the prior PNGs of pSampler), returned as they are
// for flags < 0 (IMREAD_UNCHANGED) and reduced to 8 bits otherwise.
inline Mat imread(const std::string &path, int flags = IMREAD_COLOR) {
    Mat out;
    FILE *f = fopen(path.c_str(), "rb");
    if (!f) return out;
    char magic[16] = {0};
    int r = 0, c = 0, cn = 0;
    if (fscanf(f, "%15s %d %d %d", magic, &r, &c, &cn) != 4 || fgetc(f) != '\n') {
        fclose(f);
        return out;
    }
    if (std::string(magic) == "ACMRAW2") {
        Mat raw(r, c, cn == 3 ? CV_16UC3 : CV_16UC1);
        if (fread(raw.data, 2, (size_t)r * c * cn, f) == (size_t)r * c * cn) {
            if (flags < 0) {
                out = raw;
            } else {
                out.create(r, c, cn == 3 ? CV_8UC3 : CV_8UC1);
                for (size_t i = 0; i < (size_t)r * c * cn; ++i) out.data[i] = (unsigned char)(reinterpret_cast<unsigned short *>(raw.data)[i] >> 8);
            }
        }
    } else if (std::string(magic) == "ACMRAW1") {
        Mat raw(r, c, cn == 3 ? CV_8UC3 : CV_8UC1);
        if (fread(raw.data, 1, (size_t)r * c * cn, f) == (size_t)r * c * cn) {
            if (flags == IMREAD_GRAYSCALE && cn == 3) {
                out.create(r, c, CV_8UC1);
                for (size_t i = 0; i < (size_t)r * c; ++i)   // BGR -> grey, OpenCV's integer weights
                    out.data[i] = (unsigned char)((raw.data[3 * i] * 1868 + raw.data[3 * i + 1] * 9617 + raw.data[3 * i + 2] * 4899 + 8192) >> 14);
            } else if (flags == IMREAD_COLOR && cn == 1) {
                out.create(r, c, CV_8UC3);
                for (size_t i = 0; i < (size_t)r * c; ++i) out.data[3 * i] = out.data[3 * i + 1] = out.data[3 * i + 2] = raw.data[i];
            } else {
                out = raw;
            }
        }
    }
    fclose(f);
    return out;
}
inline bool imwrite(const std::string &, const Mat &) { return true; }

// bilinear (float and 8-bit) or nearest resize with OpenCV's pixel-centre convention
inline void resize(const Mat &src, Mat &dst, Size size, double = 0, double = 0, int interp = INTER_LINEAR) {
    Mat out(size.height, size.width, src.type_);
    const int cn = src.channels();
    const bool f32 = CV_MAT_DEPTH(src.type_) == CV_32F;
    const double sx = (double)src.cols / size.width, sy = (double)src.rows / size.height;
    for (int y = 0; y < size.height; ++y) {
        float fy = (float)((y + 0.5) * sy - 0.5);
        int y0 = (int)std::floor(fy);
        fy -= y0;
        if (y0 < 0) { y0 = 0; fy = 0; }
        if (y0 >= src.rows - 1) { y0 = src.rows - 1; fy = 0; }
        const int y1 = y0 + 1 < src.rows ? y0 + 1 : src.rows - 1;
        for (int x = 0; x < size.width; ++x) {
            float fx = (float)((x + 0.5) * sx - 0.5);
            int x0 = (int)std::floor(fx);
            fx -= x0;
            if (x0 < 0) { x0 = 0; fx = 0; }
            if (x0 >= src.cols - 1) { x0 = src.cols - 1; fx = 0; }
            const int x1 = x0 + 1 < src.cols ? x0 + 1 : src.cols - 1;
            for (int k = 0; k < cn; ++k) {
                auto get = [&](int yy, int xx) -> float {
                    return f32 ? reinterpret_cast<const float *>(src.data + yy * src.step[0])[xx * cn + k]
                               : (float)(src.data + yy * src.step[0])[xx * cn + k];
                };
                float v;
                if (interp == INTER_NEAREST) v = get((int)(y * sy), (int)(x * sx));
                else v = (get(y0, x0) * (1.f - fx) + get(y0, x1) * fx) * (1.f - fy) + (get(y1, x0) * (1.f - fx) + get(y1, x1) * fx) * fy;
                if (f32) reinterpret_cast<float *>(out.data + y * out.step[0])[x * cn + k] = v;
                else (out.data + y * out.step[0])[x * cn + k] = (unsigned char)std::lrint(v);
            }
        }
    }
    dst = out;
}

inline void imshow(const std::string &, const Mat &) {}
inline int waitKey(int = 0) { return 0; }
inline void namedWindow(const std::string &, int = 0) {}
inline void resizeWindow(const std::string &, int, int) {}
inline void line(Mat &, Point, Point, const Scalar &, int = 1) {}
inline void merge(const std::vector<Mat> &, Mat &) {}
inline void merge(const Mat *, size_t, Mat &) {}

// Subdiv2D: the including harness may provide a Delaunay triangulation through this hook
// (oracle/ref_full_harness.cu passes the product's integer Delaunay, acmmp_b200/csrc/prior.cpp);
// without a hook the triangle list is empty.  Unlike OpenCV there are no virtual outer vertices,
// i.e. exactly the triangles that survive the reference's imageRC.contains() filter.
}  // namespace cv
extern "C" int (*acmmp_ref_stub_delaunay_hook)(const int *xy, int n, int *tris6, int capacity);
namespace cv {
struct Subdiv2D {
    std::vector<int> pts;
    Subdiv2D() {}
    explicit Subdiv2D(Rect) {}
    int insert(Point2f p) { pts.push_back((int)p.x); pts.push_back((int)p.y); return (int)pts.size() / 2 - 1; }
    void getTriangleList(std::vector<Vec6f> &out) const {
        out.clear();
        if (!acmmp_ref_stub_delaunay_hook || pts.size() < 6) return;
        std::vector<int> tris(pts.size() * 6 + 6);   // < 2n triangles, 6 ints each
        const int nt = acmmp_ref_stub_delaunay_hook(pts.data(), (int)pts.size() / 2, tris.data(), (int)tris.size() / 6);
        for (int t = 0; t < nt; ++t) {
            Vec6f v;
            for (int k = 0; k < 6; ++k) v[k] = (float)tris[6 * t + k];
            out.push_back(v);
        }
    }
};
// solveZ for the one shape the reference uses (3 x 4, ACMMP.cpp:922-944): the unit null vector of
// [X 1] = generalised cross product of the three rows (the sign is normalised by the caller)
struct SVD {
    static void solveZ(const Mat &A, Mat &B) {
        double a[3][4];
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 4; ++c) a[r][c] = A.at<float>(r, c);
        auto det3 = [&](int c0, int c1, int c2) {
            return a[0][c0] * (a[1][c1] * a[2][c2] - a[1][c2] * a[2][c1]) - a[0][c1] * (a[1][c0] * a[2][c2] - a[1][c2] * a[2][c0]) +
                   a[0][c2] * (a[1][c0] * a[2][c1] - a[1][c1] * a[2][c0]);
        };
        double n[4] = {det3(1, 2, 3), -det3(0, 2, 3), det3(0, 1, 3), -det3(0, 1, 2)};
        double len = std::sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2] + n[3] * n[3]);
        if (len == 0) len = 1;
        if (B.empty()) B.create(4, 1, CV_32FC1);
        for (int k = 0; k < 4; ++k) B.at<float>(k, 0) = (float)(n[k] / len);
    }
};

}  // namespace cv
