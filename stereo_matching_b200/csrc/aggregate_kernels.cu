// aggregate_kernels.cu — SGM path aggregation, path sum, WTA, uniqueness and sub-pixel fit (sm_100a).
//
// Replaces src/SGM.cpp:82-418 (+ :460-797 for the right view) and Solver::compute_subpixel
// (src/Solver.cpp:569-597) of the reference.
//
// Shape: one warp walks one scanline of one path direction; the D disparities of the current pixel
// live in registers spread over the 32 lanes (LaneLayout), d+-1 neighbours come from two lane rotations
// per chunk, the min over disparities is a single Blackwell CREDUX.MIN.F32 (redux.sync.min.f32) and the 3-way
// min of the recurrence is one FMNMX3 (costs can be negative under a sky mask, so the integer DPX forms
// would not be exact; see common.cuh).  Diagonal lines wrap around the image sides and
// re-initialise there exactly where the reference does (src/SGM.cpp:266,282,330,346), so every warp is
// independent of every other one: no inter-block communication.
//
// The float sum of the paths keeps the reference's association order (src/SGM.cpp:386-390):
//     S = (((L1+L2)+L3)+L4) + (((L5+L6)+L7)+L8)
// acc_a carries the first bracket, acc_b the second up to L7; the L8 walker finishes S in registers and
// runs WTA / uniqueness / sub-pixel on it, so S itself never goes to HBM (unless the debug tap asks).
#include "common.cuh"
#include "kernels.h"
#include "sgm_device.cuh"

namespace sgmb200 {

namespace {

enum PathMode {
    PM_WRITE = 0,     // io = L
    PM_ACCUM = 1,     // io = io + L
    PM_FINAL_A = 2,   // S = a + L            (4-path: a = (L1+L2)+L3, L = L4)
    PM_FINAL_AB = 3   // S = a + (b + L)      (8-path: a = ((L1+L2)+L3)+L4, b = (L5+L6)+L7, L = L8)
};

struct PathArgs {
    const float* cost;
    float* io;        // WRITE / ACCUM target
    const float* in_a;
    const float* in_b;
    float* pathsum;   // optional S tap
    uint16_t* disp;
    float* subpixel;
    int h, w;
    int di, dj;       // direction of travel: predecessor of (i, j) is (i - di, j - dj)
    float P1, P2, ratio;
    int min_gap;
    size_t vol_stride;  // floats per view volume
    size_t img_stride;  // pixels per view
};

constexpr int WARPS_PER_BLOCK = 4;

template <int D, int MODE, int PF>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32) sgm_path_kernel(const PathArgs a) {
    using LL = LaneLayout<D>;
    constexpr int E = LL::E;
    const int lane = threadIdx.x & 31;
    const int line = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
    const int view = blockIdx.y;
    const int h = a.h, w = a.w, di = a.di, dj = a.dj;
    const int nlines = di == 0 ? h : w;
    const int len = di == 0 ? w : h;
    if (line >= nlines) return;

    const float* __restrict__ cost = a.cost + view * a.vol_stride;
    float* io = (MODE == PM_WRITE || MODE == PM_ACCUM) ? a.io + view * a.vol_stride : nullptr;
    const float* in_a = (MODE == PM_FINAL_A || MODE == PM_FINAL_AB) ? a.in_a + view * a.vol_stride : nullptr;
    const float* in_b = (MODE == PM_FINAL_AB) ? a.in_b + view * a.vol_stride : nullptr;

    // start pixel of the line
    int ci, cj;
    if (di == 0) {
        ci = line;
        cj = dj > 0 ? 0 : w - 1;
    } else {
        ci = di > 0 ? 0 : h - 1;
        cj = line;
    }
    int li = ci, lj = cj;  // load cursor (runs PF steps ahead)

    float Cb[PF][E];
    float Xb[PF][E];
    float Yb[PF][E];

    auto advance = [&](int& i, int& j) {
        i += di;
        j += dj;
        if (di != 0) {
            if (j >= w)
                j = 0;
            else if (j < 0)
                j = w - 1;
        }
    };
    auto issue = [&](int slot_i, int slot_j, float (&c)[E], float (&x)[E], float (&y)[E]) {
        const size_t off = ((size_t)slot_i * w + slot_j) * D;
        load_pixel_ro<D>(cost + off, lane, c);
        if (MODE == PM_ACCUM) load_pixel<D>(io + off, lane, x);
        if (MODE == PM_FINAL_A || MODE == PM_FINAL_AB) load_pixel<D>(in_a + off, lane, x);
        if (MODE == PM_FINAL_AB) load_pixel<D>(in_b + off, lane, y);
    };

#pragma unroll
    for (int u = 0; u < PF; ++u) {
        if (u < len) {
            issue(li, lj, Cb[u], Xb[u], Yb[u]);
            advance(li, lj);
        }
    }

    float L[E];
    float m = 0.f;
    for (int s0 = 0; s0 < len; s0 += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int s = s0 + u;
            if (s < len) {
                const bool init = (s == 0) || (di != 0 && ((dj > 0 && cj == 0) || (dj < 0 && cj == w - 1)));
                if (init) {
                    float lm = Cb[u][0];
#pragma unroll
                    for (int k = 0; k < E; ++k) {
                        L[k] = Cb[u][k];
                        lm = fminf(lm, L[k]);
                    }
                    m = warp_min_f32(lm);
                } else {
                    path_step<D>(L, Cb[u], m, lane, a.P1, a.P2);
                }
                const size_t pix = (size_t)ci * w + cj;
                const size_t off = pix * D;
                if (MODE == PM_WRITE) {
                    store_pixel<D>(io + off, lane, L);
                } else if (MODE == PM_ACCUM) {
                    float o[E];
#pragma unroll
                    for (int k = 0; k < E; ++k) o[k] = __fadd_rn(Xb[u][k], L[k]);
                    store_pixel<D>(io + off, lane, o);
                } else {
                    float S[E];
#pragma unroll
                    for (int k = 0; k < E; ++k) {
                        if (MODE == PM_FINAL_AB)
                            S[k] = __fadd_rn(Xb[u][k], __fadd_rn(Yb[u][k], L[k]));
                        else
                            S[k] = __fadd_rn(Xb[u][k], L[k]);
                    }
                    if (a.pathsum) store_pixel<D>(a.pathsum + view * a.vol_stride + off, lane, S);
                    int dd;
                    float sub;
                    wta_pixel<D>(S, lane, a.ratio, a.min_gap, true, dd, sub);
                    if (lane == 0) {
                        a.disp[view * a.img_stride + pix] = (uint16_t)dd;
                        a.subpixel[view * a.img_stride + pix] = sub;
                    }
                }
                if (s + PF < len) {
                    issue(li, lj, Cb[u], Xb[u], Yb[u]);
                    advance(li, lj);
                }
                advance(ci, cj);
            }
        }
    }
}

// WTA straight on a cost volume (BM): one warp per pixel, grid-stride.
template <int D>
__global__ void __launch_bounds__(256) wta_volume_kernel(const float* __restrict__ vol, uint16_t* disp, float* subpixel,
                                                         size_t npix_total, float ratio, int min_gap, int parabola) {
    using LL = LaneLayout<D>;
    const int lane = threadIdx.x & 31;
    const size_t warp0 = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const size_t nwarps = (size_t)gridDim.x * (blockDim.x >> 5);
    for (size_t p = warp0; p < npix_total; p += nwarps) {
        float S[LL::E];
        load_pixel_ro<D>(vol + p * D, lane, S);
        int dd;
        float sub;
        wta_pixel<D>(S, lane, ratio, min_gap, parabola != 0, dd, sub);
        if (lane == 0) {
            disp[p] = (uint16_t)dd;
            subpixel[p] = sub;
        }
    }
}

template <int D, int MODE>
void launch_path(const PathArgs& pa, int views, cudaStream_t st, LaunchCounter& lc) {
    constexpr int PF = (LaneLayout<D>::E >= 8) ? 2 : 4;
    const int nlines = pa.di == 0 ? pa.h : pa.w;
    dim3 grid((nlines + WARPS_PER_BLOCK - 1) / WARPS_PER_BLOCK, views);
    sgm_path_kernel<D, MODE, PF><<<grid, WARPS_PER_BLOCK * 32, 0, st>>>(pa);
    lc.note("sgm_path_kernel");
}

template <int D>
void aggregate_impl(const AggregateArgs& a, cudaStream_t st, LaunchCounter& lc) {
    PathArgs pa;
    pa.cost = a.cost;
    pa.in_a = a.acc_a;
    pa.in_b = a.acc_b;
    pa.pathsum = a.pathsum;
    pa.disp = a.disp;
    pa.subpixel = a.subpixel;
    pa.h = a.h;
    pa.w = a.w;
    pa.P1 = a.P1;
    pa.P2 = a.P2;
    pa.ratio = a.unique_ratio;
    pa.min_gap = a.min_gap;
    pa.vol_stride = (size_t)a.h * a.w * a.d;
    pa.img_stride = (size_t)a.h * a.w;
    auto go = [&](int di, int dj, float* io, int mode) {
        pa.di = di;
        pa.dj = dj;
        pa.io = io;
        switch (mode) {
            case PM_WRITE: launch_path<D, PM_WRITE>(pa, a.views, st, lc); break;
            case PM_ACCUM: launch_path<D, PM_ACCUM>(pa, a.views, st, lc); break;
            case PM_FINAL_A: launch_path<D, PM_FINAL_A>(pa, a.views, st, lc); break;
            default: launch_path<D, PM_FINAL_AB>(pa, a.views, st, lc); break;
        }
    };
    if (a.paths == 8) {
        // second bracket first, so the final walker (L8) finds both partial sums complete
        go(+1, +1, a.acc_b, PM_WRITE);   // L5  top-left  -> bottom-right   (src/SGM.cpp:266-280)
        go(+1, -1, a.acc_b, PM_ACCUM);   // L6  top-right -> bottom-left    (:282-296)
        go(-1, +1, a.acc_b, PM_ACCUM);   // L7  bottom-left -> top-right    (:330-344)
    }
    go(0, +1, a.acc_a, PM_WRITE);        // L1  left -> right                (:82-119)
    go(0, -1, a.acc_a, PM_ACCUM);        // L2  right -> left                (:122-159)
    go(+1, 0, a.acc_a, PM_ACCUM);        // L3  top -> down                  (:162-199)
    if (a.paths == 8) {
        go(-1, 0, a.acc_a, PM_ACCUM);    // L4  down -> top                  (:202-239)
        go(-1, -1, nullptr, PM_FINAL_AB);  // L8 bottom-right -> top-left (:346-360) + sum + WTA (:373-418) + sub-pixel
    } else {
        go(-1, 0, nullptr, PM_FINAL_A);  // L4 + sum + WTA
    }
}

}  // namespace

void launch_aggregate(const AggregateArgs& a, cudaStream_t st, LaunchCounter& lc) {
    switch (a.d) {
        case 32: aggregate_impl<32>(a, st, lc); break;
        case 64: aggregate_impl<64>(a, st, lc); break;
        case 128: aggregate_impl<128>(a, st, lc); break;
        case 256: aggregate_impl<256>(a, st, lc); break;
        default: lc.err = cudaErrorInvalidValue; break;
    }
}

void launch_wta(const float* vol, uint16_t* disp, float* subpixel, int views, int h, int w, int d, float unique_ratio,
                int min_gap, int parabola, cudaStream_t st, LaunchCounter& lc) {
    const size_t npix = (size_t)views * h * w;  // view volumes / images are contiguous
    const int blocks = 148 * 8;
    switch (d) {
        case 32: wta_volume_kernel<32><<<blocks, 256, 0, st>>>(vol, disp, subpixel, npix, unique_ratio, min_gap, parabola); break;
        case 64: wta_volume_kernel<64><<<blocks, 256, 0, st>>>(vol, disp, subpixel, npix, unique_ratio, min_gap, parabola); break;
        case 128: wta_volume_kernel<128><<<blocks, 256, 0, st>>>(vol, disp, subpixel, npix, unique_ratio, min_gap, parabola); break;
        case 256: wta_volume_kernel<256><<<blocks, 256, 0, st>>>(vol, disp, subpixel, npix, unique_ratio, min_gap, parabola); break;
        default: lc.err = cudaErrorInvalidValue; return;
    }
    lc.note("wta_volume_kernel");
}

}  // namespace sgmb200
