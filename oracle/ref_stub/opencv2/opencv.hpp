// Minimal stand-in for the OpenCV C++ headers, used ONLY to let nvcc parse the
// reference's ACMMP.h / ACMMP.cu where they lie under /root/reference/src when
// building the parity comparator oracle/_ref/libacmmp_ref.so.  OpenCV C++ is
// not installed in this image.  Nothing here is called at run time: the harness
// (oracle/ref_harness.cu) launches the reference __global__ kernels directly
// and never executes the reference's cv:: debug-visualisation code.
// Test infrastructure - not part of the product.
#pragma once
#include <string>
#include <vector>
#include <cfloat>
#include <cmath>
#include <cstdio>
#define CV_8U 0
#define CV_32F 5
#define CV_32FC1 5
#define CV_32FC3 21
namespace cv {
struct Point { int x, y; Point() : x(0), y(0) {} Point(int a, int b) : x(a), y(b) {} };
struct Rect { int x, y, width, height; };
template <class T, int N> struct Vec { T val[N]; T &operator[](int i) { return val[i]; } const T &operator[](int i) const { return val[i]; } };
typedef Vec<float, 3> Vec3f;
typedef Vec<float, 6> Vec6f;
typedef Vec<unsigned char, 3> Vec3b;
struct Mat {
    int rows = 0, cols = 0;
    unsigned char *data = nullptr;
    static Mat zeros(int, int, int) { return Mat(); }
    void convertTo(Mat &, int, double = 1.0, double = 0.0) const {}
};
template <class T> struct Mat_ : public Mat {
    Mat_() {}
    Mat_(const Mat &) {}
    T &operator()(int, int) { static T dummy; return dummy; }
};
inline void imshow(const std::string &, const Mat &) {}
inline int waitKey(int = 0) { return 0; }
}  // namespace cv
