// cost.cpp — single-pixel cost probes of inc/cost.h (reference: src/cost.cpp:4-144).
//
// Host utilities for inspecting one window at one disparity; the solvers never call them (the dense costs are device
// kernels: stereo_matching_b200/csrc/cost_kernels.cu).  Semantics follow the reference exactly: window coordinates are
// clamped to the image ("replicate" border), the right column is the clamped left column minus disp / scale clamped
// at 0, SAD / SSD return sum / win_h / win_w with two float divisions, CT packs (v | bit) << 1 and skips taps whose
// right column falls off the image, CT_pts packs (v << 1) | bit for the same (u, v) in both images.
#include "../inc/cost.h"

namespace {

inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// Visits the taps of a win_h x win_w window centred on `c`, row-major, with clamped coordinates:
// f(di, dj, row pointer of the left image, row pointer of the right image, clamped left column).
template <class F>
inline void for_each_tap(const Mat &img_l, const Mat &img_r, const Point &c, int win_h, int win_w, F f)
{
    const int rh = win_h / 2, rw = win_w / 2;
    for (int di = -rh; di <= rh; ++di) {
        const int y = clampi(c.y + di, 0, img_l.rows - 1);
        const uchar *row_l = img_l.ptr<uchar>(y);
        const uchar *row_r = img_r.ptr<uchar>(y);
        for (int dj = -rw; dj <= rw; ++dj) f(di, dj, row_l, row_r, clampi(c.x + dj, 0, img_l.cols - 1));
    }
}

}  // namespace

float SAD(const Mat &img_l, const Mat &img_r, const Point &l_pt, int disp, int win_h, int win_w, float* /*weight*/, int scale)
{
    float acc = 0;  // float accumulator like the reference; the integer sums stay exact (<= 2^24)
    const int shift = disp / scale;
    for_each_tap(img_l, img_r, l_pt, win_h, win_w, [&](int, int, const uchar *l, const uchar *r, int xl) {
        const int xr = xl - shift > 0 ? xl - shift : 0;
        acc += abs(l[xl] - r[xr]);
    });
    return acc / win_h / win_w;
}

float SSD(const Mat &img_l, const Mat &img_r, const Point &l_pt, int disp, int win_h, int win_w, float* /*weight*/, int scale)
{
    float acc = 0;
    const int shift = disp / scale;
    for_each_tap(img_l, img_r, l_pt, win_h, win_w, [&](int, int, const uchar *l, const uchar *r, int xl) {
        const int xr = xl - shift > 0 ? xl - shift : 0;
        const int e = l[xl] - r[xr];
        acc += e * e;
    });
    return acc / win_h / win_w;
}

int hamming_cost(uint64_t ct_l, uint64_t ct_r)
{
    return __builtin_popcountll(ct_l ^ ct_r);  // the reference counts the set bits one by one (src/cost.cpp:132-144)
}

int CT(const Mat &img_l, const Mat &img_r, const Point &l_pt, int disp, int win_h, int win_w, float* /*weight*/, int scale)
{
    const uchar centre_l = img_l.at<uchar>(l_pt.y, l_pt.x);
    const uchar centre_r = img_r.at<uchar>(l_pt.y, l_pt.x - disp > 0 ? l_pt.x - disp : 0);  // not divided by scale (src/cost.cpp:69)
    const int shift = disp / scale;
    uint64_t bits_l = 0, bits_r = 0;
    for_each_tap(img_l, img_r, l_pt, win_h, win_w, [&](int di, int dj, const uchar *l, const uchar *r, int xl) {
        if (di == 0 && dj == 0) return;
        const int xr = xl - shift;
        if (xr < 0) return;  // tap dropped from both strings
        bits_l = (bits_l | (uint64_t)(l[xl] > centre_l)) << 1;
        bits_r = (bits_r | (uint64_t)(r[xr] > centre_r)) << 1;
    });
    return hamming_cost(bits_l, bits_r);
}

void CT_pts(const Mat &img_l, const Mat &img_r, int u, int v, int win_h, int win_w, float* /*weight*/,
            uint64_t *cost_table_l, uint64_t *cost_table_r)
{
    const uchar centre_l = img_l.at<uchar>(v, u), centre_r = img_r.at<uchar>(v, u);
    uint64_t bits_l = 0, bits_r = 0;
    for_each_tap(img_l, img_r, Point(u, v), win_h, win_w, [&](int di, int dj, const uchar *l, const uchar *r, int x) {
        if (di == 0 && dj == 0) return;
        bits_l = (bits_l << 1) | (uint64_t)(l[x] > centre_l);
        bits_r = (bits_r << 1) | (uint64_t)(r[x] > centre_r);
    });
    cost_table_l[(size_t)v * img_l.cols + u] = bits_l;
    cost_table_r[(size_t)v * img_l.cols + u] = bits_r;
}
