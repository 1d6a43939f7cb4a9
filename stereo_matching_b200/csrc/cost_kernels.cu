// cost_kernels.cu — cost-volume construction for sm_100a: decimation, Gaussian pre-blur, census
// transform, Hamming volume fused with the reference's recursive horizontal cost filter, the recursive
// vertical filter, and the SAD volume.
//
// Replaces Solver::build_cost_table / build_dsi_from_table(_beta) / cost_horizontal_filter /
// cost_vertical_filter (src/Solver.cpp:120-368), CT_pts / hamming_cost / SAD (src/cost.cpp) of the reference.
#include <cstdlib>

#include "common.cuh"
#include "kernels.h"

namespace sgmb200 {

namespace {

// ---------------------------------------------------------------------------------------------
// src/SGM.cpp:40-56 (rows i*scale) / src/BM.cpp:17-33 (rows i): nearest-neighbour decimation into a
// dense [2][ih*iw] buffer; with scale == 1 this is just the pitched -> dense copy.
__global__ void subsample_kernel(const uint8_t* __restrict__ src, size_t src_stride, int col_step, int row_step,
                                 uint8_t* __restrict__ dst, int ih, int iw) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    if (j >= iw || i >= ih) return;
    dst[(size_t)i * iw + j] = src[(size_t)(i * row_step) * src_stride + (size_t)j * col_step];
}

// ---------------------------------------------------------------------------------------------
// src/Solver.cpp:124-125: cv::GaussianBlur(img, out, Size(3,3), 2, 1) on CV_8U.  OpenCV runs this in
// fixed point: kx = [82,92,82]/256, ky = [70,116,70]/256, (v + 2^15) >> 16, BORDER_REFLECT_101.
__device__ __forceinline__ int reflect101(int p, int n) {
    if (n == 1) return 0;
    if (p < 0) return -p;
    if (p >= n) return 2 * n - 2 - p;
    return p;
}
__global__ void blur3_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int h, int w) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    if (j >= w) return;
    const size_t img = (size_t)blockIdx.z * h * w;
    const uint8_t* s = src + img;
    const int jm = reflect101(j - 1, w), jp = reflect101(j + 1, w);
    const int rows[3] = {reflect101(i - 1, h), i, reflect101(i + 1, h)};
    int hsum[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const uint8_t* row = s + (size_t)rows[r] * w;
        hsum[r] = 82 * (int)__ldg(row + jm) + 92 * (int)__ldg(row + j) + 82 * (int)__ldg(row + jp);
    }
    dst[img + (size_t)i * w + j] = (uint8_t)((70 * hsum[0] + 116 * hsum[1] + 70 * hsum[2] + 32768) >> 16);
}

// ---------------------------------------------------------------------------------------------
// src/cost.cpp:99-129 (CT_pts): (2*rh+1) x (2*rw+1) census, replicate-clamped coordinates, centre skipped,
// bit = neighbour > centre appended as (v << 1) | bit in row-major order.  The image tile plus halo is
// staged in shared memory so the 62 compares per pixel read on-chip data.
constexpr int CT_TX = 32, CT_TY = 8, CT_MAX_RH = 3, CT_MAX_RW = 4;
__global__ void census_kernel(const uint8_t* __restrict__ img, uint64_t* __restrict__ table, int h, int w, int rh, int rw) {
    __shared__ uint8_t tile[CT_TY + 2 * CT_MAX_RH][CT_TX + 2 * CT_MAX_RW];
    const size_t base = (size_t)blockIdx.z * h * w;
    const uint8_t* s = img + base;
    const int x0 = blockIdx.x * CT_TX, y0 = blockIdx.y * CT_TY;
    const int tw = CT_TX + 2 * rw, th = CT_TY + 2 * rh;
    for (int t = threadIdx.y * CT_TX + threadIdx.x; t < tw * th; t += CT_TX * CT_TY) {
        const int ty = t / tw, tx = t % tw;
        const int y = min(max(y0 + ty - rh, 0), h - 1);
        const int x = min(max(x0 + tx - rw, 0), w - 1);
        tile[ty][tx] = __ldg(s + (size_t)y * w + x);
    }
    __syncthreads();
    const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
    if (x >= w || y >= h) return;
    const uint8_t ctr = tile[threadIdx.y + rh][threadIdx.x + rw];
    uint64_t v = 0;
    for (int a = 0; a <= 2 * rh; ++a)
        for (int b = 0; b <= 2 * rw; ++b) {
            if (a == rh && b == rw) continue;
            v = (v << 1) | (uint64_t)(tile[threadIdx.y + a][threadIdx.x + b] > ctr);
        }
    table[base + (size_t)y * w + x] = v;
}

// x / WIN for the small constant windows, correctly rounded like __fdiv_rn but in four instructions:
//   q0 = RN(x * y), y = RN(1 / WIN);  r = RN(x - WIN * q0) (exact through the FMA);  q = RN(q0 + r * y).
// tests/cpp/verify_div_const.c checks this against x / b for EVERY finite float and b in {1..5}: the only input
// whose bits differ is x = -0 (the FMAs give +0), which the sign transfer at the end repairs.
template <int WIN>
__device__ __forceinline__ float div_win(float x) {
    if constexpr (WIN == 1) return x;
    constexpr float b = (float)WIN, y = 1.0f / (float)WIN;
    const float q0 = __fmul_rn(x, y);
    const float r = __fmaf_rn(-b, q0, x);
    return copysignf(__fmaf_rn(r, y, q0), x);
}

// ---------------------------------------------------------------------------------------------
// The reference's cost filters run IN PLACE (src/Solver.cpp:319,325 and :357,363): once the running sum
// starts subtracting "the value WIN columns back", that column has already been overwritten with its
// filtered value, so the filter is a float recurrence along the line, not a box mean.
// RecBox reproduces it as a stream: values are fed in line order; it tracks the final value of every
// position (raw for the first FIRST positions, filtered afterwards) in a FIRST+1 deep shift register, whose
// oldest entry is exactly what the reference subtracts.  Operation order per step: add, subtract, divide.
template <int WIN>
struct RecBox {
    static constexpr int HALF = WIN / 2;
    static constexpr int FIRST = WIN - HALF - 1;  // first filtered position
    static constexpr int R = FIRST + 1;
    float ring[R];
    float sum;
    __device__ __forceinline__ void reset() {
        sum = 0.f;
#pragma unroll
        for (int r = 0; r < R; ++r) ring[r] = 0.f;
    }
    __device__ __forceinline__ void push(float v) {
#pragma unroll
        for (int r = 0; r + 1 < R; ++r) ring[r] = ring[r + 1];
        ring[R - 1] = v;
    }
    // Feeds the raw value of position p (0-based, in order).  Returns true when a filtered value for position
    // `out_pos` is produced.  `len` is the line length.
    __device__ __forceinline__ bool feed(int p, float c, int len, int& out_pos, float& out_val) {
        sum = __fadd_rn(sum, c);
        if (p >= WIN) sum = __fsub_rn(sum, ring[0]);
        if (p < FIRST) push(c);  // leading positions stay raw: their final value is the raw one
        if (p >= WIN - 1) {
            const int pos = FIRST + (p - (WIN - 1));
            const int last = FIRST + (len - 2 * HALF) - 1;
            if (pos <= last) {
                out_val = div_win<WIN>(sum);
                out_pos = pos;
                push(out_val);
                return true;
            }
        }
        return false;
    }
    __device__ static __forceinline__ int last_filtered(int len) { return FIRST + (len - 2 * HALF) - 1; }
};

// src/Solver.cpp:143-248 (Hamming volume of one view, sky override) fused with :296-330 (horizontal filter): one block per
// (row, view), one thread per disparity.  The two census
// rows are staged in shared memory once, padded by D entries of edge replication on the side the disparity shift runs
// off (src/Solver.cpp:185 max(j - d, 0), :239 min(j + d, W - 1)), so the inner loop is two LDS.64, XOR, POPC and the
// three-term recurrence; the steady state (columns WIN .. W-HALF-1) has no conditionals and keeps the last outputs in a
// statically indexed circular buffer.
template <int WIN, bool SKY>
__global__ void __launch_bounds__(256) hamming_hfilter_row_kernel(const uint64_t* __restrict__ ct_l, const uint64_t* __restrict__ ct_r,
                                                                   const uint8_t* __restrict__ sky_l, const uint8_t* __restrict__ sky_r,
                                                                   float* __restrict__ cost, int h, int w, int d, int scale) {
    using RB = RecBox<WIN>;
    constexpr int R = RB::R;
    extern __shared__ __align__(16) unsigned char hsm[];
    const int i = blockIdx.x, view = blockIdx.y, k = threadIdx.x;
    // view 0: A = left row (index j), B = right row padded on the left (index j - shift + d)
    // view 1: A = right row (index j), B = left row padded on the right (index j + shift)
    uint64_t* rowA = reinterpret_cast<uint64_t*>(hsm);
    uint64_t* rowB = rowA + w;
    uint8_t* skyrow = reinterpret_cast<uint8_t*>(rowB + w + d);
    const uint64_t* srcA = (view == 0 ? ct_l : ct_r) + (size_t)i * w;
    const uint64_t* srcB = (view == 0 ? ct_r : ct_l) + (size_t)i * w;
    for (int t = k; t < w; t += blockDim.x) rowA[t] = __ldg(srcA + t);
    for (int t = k; t < w + d; t += blockDim.x) {
        const int src = view == 0 ? max(t - d, 0) : min(t, w - 1);
        rowB[t] = __ldg(srcB + src);
    }
    if (SKY) {
        const uint8_t* sky = view == 0 ? sky_l : sky_r;  // one of the two may be absent
        for (int t = k; t < w; t += blockDim.x) skyrow[t] = sky ? __ldg(sky + (size_t)i * w + t) : (uint8_t)0;
    }
    __syncthreads();
    const int shift = k / scale;
    const uint64_t* pb = rowB + (view == 0 ? d - shift : shift);  // pb[j] is the partner of rowA[j]
    float* out = cost + ((size_t)view * h + i) * (size_t)w * d + k;
    const float sky_cost = k == 0 ? 0.0f : 999999.0f;  // src/Solver.cpp:165-179
    auto raw = [&](int j) -> float {
        const float c = (float)__popcll(rowA[j] ^ pb[j]);
        if (SKY) return skyrow[j] == 255 ? sky_cost : c;
        return c;
    };

    const int last = RB::last_filtered(w);
    const int fast_end = last + 1;  // columns > last keep their raw value (and, for even WIN, the final step has no output)
    RB box;
    box.reset();
    int j = 0;
    // ---- lead-in: the first WIN columns through the general feed ----
    for (; j < WIN && j < w; ++j) {
        const float c = raw(j);
        if (j < RB::FIRST || j > last) out[(size_t)j * d] = c;
        int pos;
        float val;
        if (box.feed(j, c, w, pos, val)) out[(size_t)pos * d] = val;
    }
    // ---- steady state: sum += c[j]; sum -= (value WIN columns back, already filtered); out[j - back] = sum / WIN ----
    constexpr int BACK = WIN - 1 - RB::FIRST;  // distance from the newest input to the output position
    float circ[R];
#pragma unroll
    for (int r = 0; r < R; ++r) circ[r] = box.ring[r];
    float sum = box.sum;
    float* po = out + (size_t)(j - BACK) * d;
    for (; j + R <= fast_end; j += R) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            sum = __fadd_rn(sum, raw(j + r));
            sum = __fsub_rn(sum, circ[r]);
            const float o = div_win<WIN>(sum);
            circ[r] = o;
            *po = o;
            po += d;
        }
    }
    // ---- tail: back to the general feed (re-aligned shift register) ----
#pragma unroll
    for (int r = 0; r < R; ++r) box.ring[r] = circ[r];
    box.sum = sum;
    for (; j < w; ++j) {
        const float c = raw(j);
        if (j > last) out[(size_t)j * d] = c;
        int pos;
        float val;
        if (box.feed(j, c, w, pos, val)) out[(size_t)pos * d] = val;
    }
}

// src/Solver.cpp:333-368: the same recurrence down the rows, in place.  One thread per (column, d); rows are
// read in batches of U so the loads of a batch are in flight together while the float chain advances.
template <int WIN>
__global__ void __launch_bounds__(256) vfilter_kernel(float* cost, int h, int w, int d) {
    const size_t line = (size_t)w * d;
    const size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= line) return;
    float* base = cost + (size_t)blockIdx.y * h * line + g;
    constexpr int U = 8;
    RecBox<WIN> box;
    box.reset();
    for (int i0 = 0; i0 < h; i0 += U) {
        float c[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (i0 + u < h) c[u] = base[(size_t)(i0 + u) * line];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (i0 + u < h) {
                int pos;
                float val;
                if (box.feed(i0 + u, c[u], h, pos, val)) base[(size_t)pos * line] = val;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// src/cost.cpp:4-30 (SAD) for every (pixel, d): with replicate-clamped window coordinates the sum separates
// into a vertical sum per (clamped column, d) and a horizontal box over clamped columns.  One block handles a
// row segment of SAD_TJ columns for all d: vertical sums (<= 9*255, uint16) go to shared memory, the horizontal
// box slides over them there.  Integer sums are exact; the float result is sum / win_h / win_w like the reference.
constexpr int SAD_TJ = 32;
__global__ void __launch_bounds__(256) sad_volume_kernel(const uint8_t* __restrict__ img_l, const uint8_t* __restrict__ img_r,
                                                         float* __restrict__ cost, int h, int w, int d, int win_h, int win_w,
                                                         int scale) {
    extern __shared__ uint16_t vs[];  // [(SAD_TJ + 2*rw)][d]
    const int rh = win_h / 2, rw = win_w / 2;
    const int i = blockIdx.y, j0 = blockIdx.x * SAD_TJ, view = blockIdx.z;
    const int ncol = SAD_TJ + 2 * rw;
    for (int t = threadIdx.x; t < ncol * d; t += blockDim.x) {
        const int cx = t / d, k = t % d;
        const int x = min(max(j0 + cx - rw, 0), w - 1);  // clamped window column (the reference clamps x_l / x_r)
        int xl, xr;
        if (view == 0) {
            xl = x;
            xr = max(x - k / scale, 0);
        } else {
            xr = x;
            xl = min(x + k / scale, w - 1);
        }
        int acc = 0;
        for (int a = -rh; a <= rh; ++a) {
            const int y = min(max(i + a, 0), h - 1);
            acc += abs((int)__ldg(img_l + (size_t)y * w + xl) - (int)__ldg(img_r + (size_t)y * w + xr));
        }
        vs[t] = (uint16_t)acc;
    }
    __syncthreads();
    float* out = cost + ((size_t)view * h + i) * (size_t)w * d;
    for (int t = threadIdx.x; t < SAD_TJ * d; t += blockDim.x) {
        const int cj = t / d, k = t % d;
        const int j = j0 + cj;
        if (j >= w) continue;
        int acc = 0;
        for (int b = 0; b <= 2 * rw; ++b) acc += vs[(cj + b) * d + k];
        out[(size_t)j * d + k] = __fdiv_rn(__fdiv_rn((float)acc, (float)win_h), (float)win_w);
    }
}


// src/cost.cpp:33-59 (SSD) for every (pixel, d): the same separable scheme as sad_volume_kernel with squared differences.
// The reference accumulates in float; every partial sum is an integer below 2^24 as long as win_h * win_w * 255^2 < 2^24
// (windows up to 258 pixels; the launcher rejects larger ones), so integer accumulation gives the same float, and the
// result is sum / win_h / win_w with the reference's two divisions.
__global__ void __launch_bounds__(256) ssd_volume_kernel(const uint8_t* __restrict__ img_l, const uint8_t* __restrict__ img_r,
                                                         float* __restrict__ cost, int h, int w, int d, int win_h, int win_w, int scale) {
    extern __shared__ unsigned vs32[];  // [(SAD_TJ + 2*rw)][d] vertical sums
    const int rh = win_h / 2, rw = win_w / 2;
    const int i = blockIdx.y, j0 = blockIdx.x * SAD_TJ;
    const int ncol = SAD_TJ + 2 * rw;
    for (int t = threadIdx.x; t < ncol * d; t += blockDim.x) {
        const int cx = t / d, k = t % d;
        const int xl = min(max(j0 + cx - rw, 0), w - 1);  // clamped window column (src/cost.cpp:45-46)
        const int xr = max(xl - k / scale, 0);            // :47
        unsigned acc = 0;
        for (int a = -rh; a <= rh; ++a) {
            const int y = min(max(i + a, 0), h - 1);
            const int df = (int)__ldg(img_l + (size_t)y * w + xl) - (int)__ldg(img_r + (size_t)y * w + xr);
            acc += (unsigned)(df * df);
        }
        vs32[t] = acc;
    }
    __syncthreads();
    float* out = cost + (size_t)i * w * d;
    for (int t = threadIdx.x; t < SAD_TJ * d; t += blockDim.x) {
        const int cj = t / d, k = t % d;
        const int j = j0 + cj;
        if (j >= w) continue;
        unsigned acc = 0;
        for (int b = 0; b <= 2 * rw; ++b) acc += vs32[(cj + b) * d + k];
        out[(size_t)j * d + k] = __fdiv_rn(__fdiv_rn((float)acc, (float)win_h), (float)win_w);
    }
}

// src/cost.cpp:62-96 (CT) for every (pixel, d): a census string of the left window around (x, y) against one of the right
// window, recomputed per disparity, with the function's own conventions: the right centre pixel is R(y, max(x - d, 0)) —
// d NOT divided by scale (:69) — while the taps use x_r = x_l - d / scale and are DROPPED from both strings when x_r < 0
// (:85-88); the strings are packed (v | bit) << 1, which the Hamming distance does not see.  So the cost is the number of
// kept taps whose two comparisons disagree.  One block = (row, 32 columns); the 2*rh+1 rows of both images that the
// windows touch are staged in shared memory, the right one D-1+rw columns further to the left.
constexpr int CT_VOL_TJ = 32;
__global__ void __launch_bounds__(256) ct_volume_kernel(const uint8_t* __restrict__ img_l, const uint8_t* __restrict__ img_r,
                                                        float* __restrict__ cost, int h, int w, int d, int win_h, int win_w, int scale) {
    extern __shared__ unsigned char ctsm[];
    const int rh = win_h / 2, rw = win_w / 2, nrow = 2 * rh + 1;
    const int i = blockIdx.y, j0 = blockIdx.x * CT_VOL_TJ;
    const int lw = CT_VOL_TJ + 2 * rw;              // left tile: columns j0-rw .. j0+31+rw (clamped on load)
    const int rshift = rw + d - 1;                  // right tile starts at column j0 - rshift
    const int rwid = CT_VOL_TJ + rw + rshift;       // .. up to j0+31+rw
    unsigned char* tl = ctsm;                       // [nrow][lw]
    unsigned char* tr = ctsm + (size_t)nrow * lw;   // [nrow][rwid]
    for (int t = threadIdx.x; t < nrow * lw; t += blockDim.x) {
        const int r = t / lw, c = t % lw;
        const int y = min(max(i + r - rh, 0), h - 1), x = min(max(j0 + c - rw, 0), w - 1);
        tl[t] = __ldg(img_l + (size_t)y * w + x);
    }
    for (int t = threadIdx.x; t < nrow * rwid; t += blockDim.x) {
        const int r = t / rwid, c = t % rwid;
        const int y = min(max(i + r - rh, 0), h - 1), x = min(max(j0 + c - rshift, 0), w - 1);  // x < 0 is never used (dropped taps)
        tr[t] = __ldg(img_r + (size_t)y * w + x);
    }
    __syncthreads();
    float* out = cost + (size_t)i * w * d;
    for (int t = threadIdx.x; t < CT_VOL_TJ * d; t += blockDim.x) {
        const int cj = t / d, k = t % d;
        const int x = j0 + cj;
        if (x >= w) continue;
        const int ks = k / scale;
        const unsigned char cl = tl[rh * lw + cj + rw];                              // L(y, x)
        const unsigned char cr = tr[rh * rwid + (max(x - k, 0) - (j0 - rshift))];    // R(y, max(x - d, 0))
        int acc = 0;
        for (int a = 0; a < nrow; ++a) {
            for (int b = -rw; b <= rw; ++b) {
                if (a == rh && b == 0) continue;
                const int xl = min(max(x + b, 0), w - 1);
                const int xr = xl - ks;
                if (xr < 0) continue;
                const bool bl = tl[a * lw + (xl - (j0 - rw))] > cl;
                const bool br = tr[a * rwid + (xr - (j0 - rshift))] > cr;
                acc += bl != br ? 1 : 0;
            }
        }
        out[(size_t)x * d + k] = (float)acc;
    }
}

// Banded form: a block owns SADB_XC window columns (SADB_XC - 2*rw output columns) x all D disparities and walks down a
// band of rows.  Thread (k, g) keeps the vertical sums V(x, k) = sum_a |own(y+a, x) - other(y+a, x -/+ k)| of its
// columns in registers and slides them one row at a time (+ the entering row, - the leaving row; the image rows of the
// window sit in a small circular shared-memory buffer with the reference's clamping already applied), publishes them
// as uint16 and then slides a horizontal box over them for its output columns.  Every (row, x, k) absolute
// difference is computed twice per band instead of win_h * win_w times.  The two float divisions of src/cost.cpp:29
// come from a table indexed by the integer sum (built on the host with the same IEEE divisions) when one is given.
constexpr int SADB_XC = 32;

template <int D>
__global__ void __launch_bounds__(256) sad_volume_band_kernel(const uint8_t* __restrict__ img_l, const uint8_t* __restrict__ img_r,
                                                              const float* __restrict__ lut, float* __restrict__ cost, int h, int w,
                                                              int win_h, int win_w, int scale, int band) {
    constexpr int XC = SADB_XC, G = 256 / D, NV = XC / G;
    static_assert(256 % D == 0 && XC % G == 0, "D must be a power of two in 8..256");
    extern __shared__ __align__(16) unsigned char sad_sm[];
    const int rh = win_h / 2, rw = win_w / 2, TJ = XC - 2 * rw, NR = 2 * rh + 2;
    const int OW = XC + (D - 1) / scale + 1;
    const int RS = (XC + OW + 3) & ~3;                       // bytes per buffered image row: [own tile | other tile]
    uint16_t* vs = reinterpret_cast<uint16_t*>(sad_sm);      // [XC][D]
    unsigned char* rows = sad_sm + (size_t)XC * D * 2;       // [NR][RS]
    const int view = blockIdx.z, j0 = blockIdx.x * TJ, i0 = blockIdx.y * band, i1 = min(i0 + band, h);
    const uint8_t* own = view ? img_r : img_l;
    const uint8_t* other = view ? img_l : img_r;
    const int obase = view ? j0 - rw : j0 - rw - (D - 1) / scale;
    const int t = threadIdx.x, k = t % D, g = t / D, ks = k / scale;

    // index of the other image's pixel for each of this thread's window columns (clamped like src/cost.cpp:14-25)
    int oidx[NV];
#pragma unroll
    for (int n = 0; n < NV; ++n) {
        const int xcl = min(max(j0 + g + n * G - rw, 0), w - 1);
        oidx[n] = (view ? min(xcl + ks, w - 1) : max(xcl - ks, 0)) - obase + XC;
    }
    auto load_row = [&](int yy) {  // yy may lie outside the image: rows are replicated
        const int y = min(max(yy, 0), h - 1);
        unsigned char* dst = rows + (size_t)((yy % NR + NR) % NR) * RS;
        for (int q = t; q < XC + OW; q += 256) {
            const int x = q < XC ? j0 + q - rw : obase + (q - XC);
            dst[q] = (q < XC ? own : other)[(size_t)y * w + min(max(x, 0), w - 1)];
        }
    };
    auto row_ptr = [&](int yy) { return rows + (size_t)((yy % NR + NR) % NR) * RS; };

    int V[NV];
#pragma unroll
    for (int n = 0; n < NV; ++n) V[n] = 0;
    for (int yy = i0 - rh; yy < i0 + rh; ++yy) load_row(yy);
    __syncthreads();
    for (int yy = i0 - rh; yy < i0 + rh; ++yy) {
        const unsigned char* rp = row_ptr(yy);
#pragma unroll
        for (int n = 0; n < NV; ++n) V[n] += abs((int)rp[g + n * G] - (int)rp[oidx[n]]);
    }
    const int cpg = (TJ + G - 1) / G, c0 = g * cpg, c1 = min(min(c0 + cpg, TJ), w - j0);
    for (int i = i0; i < i1; ++i) {
        load_row(i + rh);
        __syncthreads();  // the entering row is in place (and every thread is done with the previous vs)
        {
            const unsigned char* rin = row_ptr(i + rh);
            const unsigned char* rout = row_ptr(i - rh - 1);
            const bool sub = i > i0;
#pragma unroll
            for (int n = 0; n < NV; ++n) {
                V[n] += abs((int)rin[g + n * G] - (int)rin[oidx[n]]);
                if (sub) V[n] -= abs((int)rout[g + n * G] - (int)rout[oidx[n]]);
                vs[(g + n * G) * D + k] = (uint16_t)V[n];
            }
        }
        __syncthreads();  // vertical sums of row i are published
        if (c0 < c1) {
            int acc = 0;
            for (int b = 0; b <= 2 * rw; ++b) acc += vs[(c0 + b) * D + k];
            float* out = cost + (((size_t)view * h + i) * w + j0 + c0) * D + k;
            for (int cj = c0;; ++cj) {
                *out = lut ? __ldg(lut + acc) : __fdiv_rn(__fdiv_rn((float)acc, (float)win_h), (float)win_w);
                if (cj + 1 >= c1) break;
                acc += (int)vs[(cj + 2 * rw + 1) * D + k] - (int)vs[cj * D + k];
                out += D;
            }
        }
    }
}

}  // namespace

void launch_subsample(const uint8_t* src, size_t src_stride, int col_step, int row_step, uint8_t* dst, int ih, int iw,
                      cudaStream_t st, LaunchCounter& lc) {
    dim3 grid((iw + 255) / 256, ih);
    subsample_kernel<<<grid, 256, 0, st>>>(src, src_stride, col_step, row_step, dst, ih, iw);
    lc.note("subsample_kernel");
}

void launch_blur3(const uint8_t* src, uint8_t* dst, int h, int w, cudaStream_t st, LaunchCounter& lc) {
    dim3 grid((w + 255) / 256, h, 2);
    blur3_kernel<<<grid, 256, 0, st>>>(src, dst, h, w);
    lc.note("blur3_kernel");
}

void launch_census(const uint8_t* img, uint64_t* table, int h, int w, int win_h, int win_w, cudaStream_t st,
                   LaunchCounter& lc) {
    const int rh = win_h / 2, rw = win_w / 2;
    if (rh > CT_MAX_RH || rw > CT_MAX_RW || (2 * rh + 1) * (2 * rw + 1) - 1 > 64) {
        lc.err = cudaErrorInvalidValue;
        return;
    }
    dim3 grid((w + CT_TX - 1) / CT_TX, (h + CT_TY - 1) / CT_TY, 2);
    census_kernel<<<grid, dim3(CT_TX, CT_TY), 0, st>>>(img, table, h, w, rh, rw);
    lc.note("census_kernel");
}

static size_t hamming_row_smem_bytes(int w, int d, bool sky) { return (size_t)(2 * w + d) * sizeof(uint64_t) + (sky ? (size_t)w : 0); }

template <int WIN>
static void launch_hh(const uint64_t* ct_l, const uint64_t* ct_r, const uint8_t* sky_l, const uint8_t* sky_r, float* cost, int views,
                      int h, int w, int d, int scale, cudaStream_t st) {
    const bool sky = sky_l || sky_r;
    const size_t smem = hamming_row_smem_bytes(w, d, sky);  // sgm_create rejects rows that do not fit (hamming_rows_fit)
    dim3 grid(h, views);
    if (sky) {
        if (smem > 48 * 1024) cudaFuncSetAttribute(hamming_hfilter_row_kernel<WIN, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        hamming_hfilter_row_kernel<WIN, true><<<grid, d, smem, st>>>(ct_l, ct_r, sky_l, sky_r, cost, h, w, d, scale);
    } else {
        if (smem > 48 * 1024) cudaFuncSetAttribute(hamming_hfilter_row_kernel<WIN, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        hamming_hfilter_row_kernel<WIN, false><<<grid, d, smem, st>>>(ct_l, ct_r, sky_l, sky_r, cost, h, w, d, scale);
    }
}

bool hamming_rows_fit(int w, int d) { return hamming_row_smem_bytes(w, d, true) <= 200 * 1024; }

void launch_hamming_hfilter(const uint64_t* ct_l, const uint64_t* ct_r, const uint8_t* sky_l, const uint8_t* sky_r,
                            float* cost, int views, int h, int w, int d, int scale, int win, cudaStream_t st,
                            LaunchCounter& lc) {
    switch (win) {
        case 5: launch_hh<5>(ct_l, ct_r, sky_l, sky_r, cost, views, h, w, d, scale, st); break;
        case 4: launch_hh<4>(ct_l, ct_r, sky_l, sky_r, cost, views, h, w, d, scale, st); break;
        case 3: launch_hh<3>(ct_l, ct_r, sky_l, sky_r, cost, views, h, w, d, scale, st); break;
        case 2: launch_hh<2>(ct_l, ct_r, sky_l, sky_r, cost, views, h, w, d, scale, st); break;
        case 1: launch_hh<1>(ct_l, ct_r, sky_l, sky_r, cost, views, h, w, d, scale, st); break;
        default: lc.err = cudaErrorInvalidValue; return;
    }
    lc.note("hamming_hfilter_kernel");
}

// Integer mode: the raw Hamming volume of src/Solver.cpp:143-248 (no sky mask, no cost filter) as
// s16x2 words, two disparities {2k, 2k+1} per 32-bit word, the layout S16x2Arith works on.  Block = (row, view); both
// census rows are staged in shared memory like hamming_hfilter_row_kernel does; a thread produces two adjacent words
// (four disparities) per step and stores them as one 8-byte access, fully coalesced along d.
//
// Sky masks (src/Solver.cpp:165-179: masked pixels cost 0 at d = 0 and 999999 elsewhere): 999999 does not fit 16 bits, so masked
// pixels get sky_big = 2 * P2 instead, which provably gives the same results.  With the cost filter off a masked pixel's path
// values are L(0) <= P2 and L(d > 0) = fill + t, t in [0, P2]; every later min(L(d), L(d +- 1) + P1, m + P2) with m = L(0)
// takes m + P2 (or L(0) + P1 at d = 1) whether the fill is 999999 or any value >= 2 * P2, so the path values of all other
// pixels are unchanged; at the masked pixel itself S(0) <= 8 * P2 < 8 * fill <= S(d > 0), so the winner is d = 0 and
// S(0) / S(second) <= 0.5 < UNIQUE_RATIO keeps it valid, as in the reference.  Only the cost / path-sum TAPS of masked pixels
// differ from the float volumes (fill + t instead of 999999 + t).
__global__ void __launch_bounds__(256) hamming_s16x2_kernel(const uint64_t* __restrict__ ct_l, const uint64_t* __restrict__ ct_r,
                                                            const uint8_t* __restrict__ sky_l, const uint8_t* __restrict__ sky_r,
                                                            unsigned sky_big, uint2* __restrict__ cost, int h, int w, int d, int scale) {
    extern __shared__ __align__(16) unsigned char hsm[];
    const int i = blockIdx.x, view = blockIdx.y;
    const uint8_t* sky = view == 0 ? sky_l : sky_r;
    if (sky) sky += (size_t)i * w;
    uint64_t* rowA = reinterpret_cast<uint64_t*>(hsm);
    uint64_t* rowB = rowA + w;  // view 0: right row padded on the left by d entries; view 1: left row padded on the right
    const uint64_t* srcA = (view == 0 ? ct_l : ct_r) + (size_t)i * w;
    const uint64_t* srcB = (view == 0 ? ct_r : ct_l) + (size_t)i * w;
    for (int t = threadIdx.x; t < w; t += blockDim.x) rowA[t] = __ldg(srcA + t);
    for (int t = threadIdx.x; t < w + d; t += blockDim.x) rowB[t] = __ldg(srcB + (view == 0 ? max(t - d, 0) : min(t, w - 1)));
    __syncthreads();
    const int q = d / 4;                      // 8-byte stores per pixel; a power of two <= 64
    const int g = threadIdx.x & (q - 1);      // this thread's group of four disparities 4g .. 4g+3
    const int p = threadIdx.x / q, np = blockDim.x / q;
    uint2* out = cost + ((size_t)view * h + i) * (size_t)w * q + g;
    const uint2 sky_words = make_uint2((g == 0 ? 0u : sky_big) | (sky_big << 16), sky_big * 0x10001u);  // d = 0 costs nothing
    auto emit = [&](int j, uint64_t a, uint64_t b0, uint64_t b1, uint64_t b2, uint64_t b3) {
        const unsigned c0 = __popcll(a ^ b0), c1 = __popcll(a ^ b1), c2 = __popcll(a ^ b2), c3 = __popcll(a ^ b3);
        const bool masked = sky && __ldg(sky + j) == 255;
        out[(size_t)j * q] = masked ? sky_words : make_uint2(c0 | (c1 << 16), c2 | (c3 << 16));
    };
    if (scale != 1) {  // general form: four look-ups per step
        for (int j = p; j < w; j += np) {
            uint64_t b[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int shift = (4 * g + e) / scale;
                b[e] = rowB[view == 0 ? j + d - shift : j + shift];
            }
            emit(j, rowA[j], b[0], b[1], b[2], b[3]);
        }
        return;
    }
    // scale 1: the thread walks a contiguous run of columns; moving one column on shifts the four partners by one entry,
    // so only one of them is new (a shift register instead of four shared-memory reads per step)
    const int seg = (w + np - 1) / np, j0 = p * seg, j1 = min(w, j0 + seg);
    if (j0 >= j1) return;
    if (view == 0) {  // partner of disparity k at column j: rowB[j + d - k]
        const uint64_t* pb = rowB + d - 4 * g;
        uint64_t b1 = pb[j0 - 1], b2 = pb[j0 - 2], b3 = pb[j0 - 3];
        for (int j = j0; j < j1; ++j) {
            const uint64_t b0 = pb[j];
            emit(j, rowA[j], b0, b1, b2, b3);
            b3 = b2, b2 = b1, b1 = b0;
        }
    } else {          // rowB[j + k]
        const uint64_t* pb = rowB + 4 * g;
        uint64_t b0 = pb[j0], b1 = pb[j0 + 1], b2 = pb[j0 + 2];
        for (int j = j0; j < j1; ++j) {
            const uint64_t b3 = pb[j + 3];
            emit(j, rowA[j], b0, b1, b2, b3);
            b0 = b1, b1 = b2, b2 = b3;
        }
    }
}

void launch_hamming_s16x2(const uint64_t* ct_l, const uint64_t* ct_r, const uint8_t* sky_l, const uint8_t* sky_r, unsigned sky_big, void* cost,
                          int views, int h, int w, int d, int scale, cudaStream_t st, LaunchCounter& lc) {
    const size_t smem = (size_t)(2 * w + d) * sizeof(uint64_t);
    if (smem > 200 * 1024 || d % 4) {
        lc.err = cudaErrorInvalidValue;
        lc.failed = "hamming_s16x2_kernel (row too wide)";
        return;
    }
    if (smem > 48 * 1024) cudaFuncSetAttribute(hamming_s16x2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    hamming_s16x2_kernel<<<dim3(h, views), 256, smem, st>>>(ct_l, ct_r, sky_l, sky_r, sky_big, static_cast<uint2*>(cost), h, w, d, scale);
    lc.note("hamming_s16x2_kernel");
}

void launch_vfilter(float* cost, int views, int h, int w, int d, int win, cudaStream_t st, LaunchCounter& lc) {
    dim3 grid((unsigned)(((size_t)w * d + 255) / 256), views);
    switch (win) {
        case 5: vfilter_kernel<5><<<grid, 256, 0, st>>>(cost, h, w, d); break;
        case 4: vfilter_kernel<4><<<grid, 256, 0, st>>>(cost, h, w, d); break;
        case 3: vfilter_kernel<3><<<grid, 256, 0, st>>>(cost, h, w, d); break;
        case 2: vfilter_kernel<2><<<grid, 256, 0, st>>>(cost, h, w, d); break;
        case 1: vfilter_kernel<1><<<grid, 256, 0, st>>>(cost, h, w, d); break;
        default: lc.err = cudaErrorInvalidValue; return;
    }
    lc.note("vfilter_kernel");
}

template <int D>
static void launch_sad_band(const uint8_t* img_l, const uint8_t* img_r, const float* lut, float* cost, int views, int h, int w, int win_h,
                            int win_w, int scale, cudaStream_t st) {
    const int rh = win_h / 2, rw = win_w / 2, tj = SADB_XC - 2 * rw;
    const int ow = SADB_XC + (D - 1) / scale + 1, rs = (SADB_XC + ow + 3) & ~3;
    const size_t smem = (size_t)SADB_XC * D * 2 + (size_t)(2 * rh + 2) * rs;
    const int band = h >= 256 ? 48 : 32;
    dim3 grid((w + tj - 1) / tj, (h + band - 1) / band, views);
    sad_volume_band_kernel<D><<<grid, 256, smem, st>>>(img_l, img_r, lut, cost, h, w, win_h, win_w, scale, band);
}

// lut (optional, device): lut[s] = (float(s) / win_h) / win_w for s = 0 .. 255 * win_h * win_w
void launch_sad_volume(const uint8_t* img_l, const uint8_t* img_r, const float* lut, float* cost, int views, int h, int w, int d,
                       int win_h, int win_w, int scale, cudaStream_t st, LaunchCounter& lc) {
    bool legacy = false;  // the per-row form below serves disparity counts / windows the banded kernel is not built for
#ifdef SGM_B200_TUNING
    legacy = std::getenv("SGM_B200_SAD_LEGACY") != nullptr;
#endif
    const int rw = win_w / 2;
    const bool band_ok = !legacy && rw <= 12 && win_h <= 63 && (d == 8 || d == 16 || d == 32 || d == 64 || d == 128 || d == 256);
    if (band_ok) {
        switch (d) {
            case 8: launch_sad_band<8>(img_l, img_r, lut, cost, views, h, w, win_h, win_w, scale, st); break;
            case 16: launch_sad_band<16>(img_l, img_r, lut, cost, views, h, w, win_h, win_w, scale, st); break;
            case 32: launch_sad_band<32>(img_l, img_r, lut, cost, views, h, w, win_h, win_w, scale, st); break;
            case 64: launch_sad_band<64>(img_l, img_r, lut, cost, views, h, w, win_h, win_w, scale, st); break;
            case 128: launch_sad_band<128>(img_l, img_r, lut, cost, views, h, w, win_h, win_w, scale, st); break;
            default: launch_sad_band<256>(img_l, img_r, lut, cost, views, h, w, win_h, win_w, scale, st); break;
        }
        lc.note("sad_volume_band_kernel");
        return;
    }
    const size_t smem = (size_t)(SAD_TJ + 2 * rw) * d * sizeof(uint16_t);
    if (smem > 48 * 1024 || win_h > 257) {
        lc.err = cudaErrorInvalidValue;
        return;
    }
    dim3 grid((w + SAD_TJ - 1) / SAD_TJ, h, views);
    sad_volume_kernel<<<grid, 256, smem, st>>>(img_l, img_r, cost, h, w, d, win_h, win_w, scale);
    lc.note("sad_volume_kernel");
}

// src/cost.cpp:33-59 / :62-96 as dense volumes of the left view, [h][w][d] floats, as Solver::build_dsi (src/Solver.cpp:96-117)
// would fill them.  kind: 1 = SSD, 2 = CT.
void launch_window_cost_volume(int kind, const uint8_t* img_l, const uint8_t* img_r, float* cost, int h, int w, int d, int win_h, int win_w,
                               int scale, cudaStream_t st, LaunchCounter& lc) {
    const int rh = win_h / 2, rw = win_w / 2;
    if (kind == 1) {
        const size_t smem = (size_t)(SAD_TJ + 2 * rw) * d * sizeof(unsigned);
        if (smem > 200 * 1024 || (long long)win_h * win_w * 65025ll >= (1ll << 24)) {
            lc.err = cudaErrorInvalidValue;
            lc.failed = "ssd_volume_kernel (window too large for exact accumulation / shared memory)";
            return;
        }
        if (smem > 48 * 1024) cudaFuncSetAttribute(ssd_volume_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        ssd_volume_kernel<<<dim3((w + SAD_TJ - 1) / SAD_TJ, h), 256, smem, st>>>(img_l, img_r, cost, h, w, d, win_h, win_w, scale);
        lc.note("ssd_volume_kernel");
    } else if (kind == 2) {
        const size_t smem = (size_t)(2 * rh + 1) * ((CT_VOL_TJ + 2 * rw) + (CT_VOL_TJ + 2 * rw + d - 1));
        if (smem > 200 * 1024 || (2 * rh + 1) * (2 * rw + 1) - 1 > 63) {  // the reference's 64-bit strings hold 63 taps
            lc.err = cudaErrorInvalidValue;
            lc.failed = "ct_volume_kernel (window too large)";
            return;
        }
        if (smem > 48 * 1024) cudaFuncSetAttribute(ct_volume_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        ct_volume_kernel<<<dim3((w + CT_VOL_TJ - 1) / CT_VOL_TJ, h), 256, smem, st>>>(img_l, img_r, cost, h, w, d, win_h, win_w, scale);
        lc.note("ct_volume_kernel");
    } else {
        lc.err = cudaErrorInvalidValue;
    }
}

}  // namespace sgmb200
