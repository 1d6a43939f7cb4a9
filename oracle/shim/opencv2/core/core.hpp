// Minimal cv:: stand-in used ONLY to compile the reference's CPU sources
// (/root/reference/src/{Solver,SGM,BM,cost}.cpp) as the parity oracle, and to
// compile-check our own Solver/SGM/BM shells, on a box that has no C++ OpenCV.
// TEST INFRASTRUCTURE — never linked into the product library.
//
// It provides exactly the subset of the OpenCV API those translation units touch:
// a ref-counted Mat (create/release/ptr/at/ROI/copyTo), Point/Size/Rect/Vec3b,
// MIN/MAX, cvtColor(GRAY2BGR) and an 8-bit GaussianBlur that reproduces OpenCV's
// fixed-point path bit for bit (checked against cv2 4.13 in tests/test_oracle_cpu.py).
#pragma once

#include <algorithm>
#include <cassert>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

typedef unsigned char uchar;

#ifndef MIN
#define MIN(a, b) ((a) > (b) ? (b) : (a))
#endif
#ifndef MAX
#define MAX(a, b) ((a) < (b) ? (b) : (a))
#endif

#define CV_8U 0
#define CV_32F 5
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn) - 1) << 3))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_GRAY2BGR 8

namespace cv {

template <typename T>
struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
};
typedef Point_<int> Point;

template <typename T>
struct Point3_ {
    T x, y, z;
    Point3_() : x(0), y(0), z(0) {}
    Point3_(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
};
typedef Point3_<double> Point3d;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};

struct Rect {
    int x, y, width, height;
    Rect() : x(0), y(0), width(0), height(0) {}
    Rect(int x_, int y_, int w, int h) : x(x_), y(y_), width(w), height(h) {}
};

struct Vec3b {
    uchar val[3];
    uchar& operator[](int i) { return val[i]; }
    const uchar& operator[](int i) const { return val[i]; }
};

class Mat {
public:
    int rows, cols;
    uchar* data;
    size_t step;  // bytes per row

    Mat() : rows(0), cols(0), data(nullptr), step(0), type_(0) {}
    Mat(int r, int c, int type) : rows(0), cols(0), data(nullptr), step(0), type_(0) { create(r, c, type); }
    // Header over caller-owned memory (no ownership taken), like cv::Mat(rows, cols, type, ptr, step).
    Mat(int r, int c, int type, void* ext, size_t ext_step = 0)
        : rows(r), cols(c), data(static_cast<uchar*>(ext)), step(0), type_(type) {
        step = ext_step ? ext_step : static_cast<size_t>(c) * elemSize();
    }

    int type() const { return type_; }
    int depth() const { return type_ & 7; }
    int channels() const { return (type_ >> 3) + 1; }
    size_t elemSize() const { return static_cast<size_t>(channels()) * (depth() == CV_32F ? 4 : 1); }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    bool isContinuous() const { return step == static_cast<size_t>(cols) * elemSize(); }

    void create(int r, int c, int type) {
        if (data && r == rows && c == cols && type == type_) return;
        release();
        rows = r;
        cols = c;
        type_ = type;
        step = static_cast<size_t>(c) * elemSize();
        size_t bytes = step * static_cast<size_t>(r);
        store_.reset(new std::vector<uchar>(bytes ? bytes : 1));
        data = store_->data();
    }

    void release() {
        store_.reset();
        data = nullptr;
        rows = cols = 0;
        step = 0;
    }

    template <typename T>
    T* ptr(int r = 0) { return reinterpret_cast<T*>(data + step * static_cast<size_t>(r)); }
    template <typename T>
    const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(data + step * static_cast<size_t>(r)); }
    template <typename T>
    T& at(int r, int c) { return ptr<T>(r)[c]; }
    template <typename T>
    const T& at(int r, int c) const { return ptr<T>(r)[c]; }

    static Mat zeros(int r, int c, int type) {
        Mat m(r, c, type);
        std::memset(m.data, 0, m.step * static_cast<size_t>(r));
        return m;
    }

    Mat operator()(const Rect& roi) const {
        Mat m;
        m.rows = roi.height;
        m.cols = roi.width;
        m.type_ = type_;
        m.step = step;
        m.store_ = store_;
        m.data = data + step * static_cast<size_t>(roi.y) + elemSize() * static_cast<size_t>(roi.x);
        return m;
    }

    void copyTo(Mat& dst) const {
        if (dst.data == nullptr || dst.rows != rows || dst.cols != cols || dst.type_ != type_) dst.create(rows, cols, type_);
        const size_t row_bytes = static_cast<size_t>(cols) * elemSize();
        for (int r = 0; r < rows; ++r) std::memcpy(dst.data + dst.step * r, data + step * r, row_bytes);
    }

    Mat clone() const {
        Mat m;
        copyTo(m);
        return m;
    }

private:
    int type_;
    std::shared_ptr<std::vector<uchar>> store_;
};

inline void cvtColor(const Mat& src, Mat& dst, int code) {
    assert(code == CV_GRAY2BGR && src.type() == CV_8UC1);
    (void)code;
    Mat out(src.rows, src.cols, CV_8UC3);
    for (int r = 0; r < src.rows; ++r) {
        const uchar* s = src.ptr<uchar>(r);
        uchar* d = out.ptr<uchar>(r);
        for (int c = 0; c < src.cols; ++c) d[3 * c] = d[3 * c + 1] = d[3 * c + 2] = s[c];
    }
    dst = out;
}

// OpenCV's CV_8U GaussianBlur for small kernels runs in fixed point: the 1-D kernels
// are quantised to 8 fractional bits (sum == 256), the horizontal pass keeps 8.8,
// the vertical pass 16.16, and the result is (v + 2^15) >> 16; borders are
// BORDER_REFLECT_101.  Only ksize 3x3 is needed by the reference
// (src/Solver.cpp:124-125: Size(3,3), sigmaX=2, sigmaY=1 -> [82,92,82] x [70,116,70]).
inline void gaussian_kernel3_q8(double sigma, int k[3]) {
    const double e = std::exp(-1.0 / (2.0 * sigma * sigma));
    const double s = 1.0 + 2.0 * e;
    const int side = static_cast<int>(std::lrint(256.0 * e / s));
    k[0] = side;
    k[2] = side;
    k[1] = 256 - 2 * side;
}

inline void GaussianBlur(const Mat& src, Mat& dst, Size ksize, double sigmaX, double sigmaY = 0) {
    assert(src.type() == CV_8UC1 && ksize.width == 3 && ksize.height == 3);
    (void)ksize;
    if (sigmaY <= 0) sigmaY = sigmaX;
    int kx[3], ky[3];
    gaussian_kernel3_q8(sigmaX, kx);
    gaussian_kernel3_q8(sigmaY, ky);
    const int H = src.rows, W = src.cols;
    Mat out(H, W, CV_8UC1);
    std::vector<int> hbuf(static_cast<size_t>(H) * W);
    auto refl = [](int p, int n) { return n == 1 ? 0 : (p < 0 ? -p : (p >= n ? 2 * n - 2 - p : p)); };
    for (int r = 0; r < H; ++r) {
        const uchar* s = src.ptr<uchar>(r);
        for (int c = 0; c < W; ++c)
            hbuf[static_cast<size_t>(r) * W + c] = kx[0] * s[refl(c - 1, W)] + kx[1] * s[c] + kx[2] * s[refl(c + 1, W)];
    }
    for (int r = 0; r < H; ++r) {
        const int* a = &hbuf[static_cast<size_t>(refl(r - 1, H)) * W];
        const int* b = &hbuf[static_cast<size_t>(r) * W];
        const int* c2 = &hbuf[static_cast<size_t>(refl(r + 1, H)) * W];
        uchar* d = out.ptr<uchar>(r);
        for (int c = 0; c < W; ++c) d[c] = static_cast<uchar>((ky[0] * a[c] + ky[1] * b[c] + ky[2] * c2[c] + 32768) >> 16);
    }
    dst = out;
}

}  // namespace cv
