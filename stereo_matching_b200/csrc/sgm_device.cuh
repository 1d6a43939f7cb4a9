// sgm_device.cuh — device-side building blocks shared by the aggregation kernels: one SGM path step on a
// pixel's disparities held across a warp, and the WTA / uniqueness / sub-pixel selection on a summed pixel.
#pragma once

#include "common.cuh"

namespace sgmb200 {

// L <- min(Lp[d], min(Lp[d-1], Lp[d+1]) + P1, m + P2) + (C[d] - m)       (src/SGM.cpp:105-108)
// min(a + P1, b + P1) == min(a, b) + P1 exactly because float rounding is monotone; at d = 0 / D-1 the
// clamped neighbour is the element itself (src/SGM.cpp:96-97), i.e. no constraint.
template <int D>
__device__ __forceinline__ void path_step(float (&L)[LaneLayout<D>::E], const float (&C)[LaneLayout<D>::E], float& m,
                                          int lane, float P1, float P2) {
    using LL = LaneLayout<D>;
    constexpr int E = LL::E, CW = LL::CW, NCH = LL::NCH;
    float up[NCH], dn[NCH];
#pragma unroll
    for (int q = 0; q < NCH; ++q) {
        up[q] = __shfl_sync(FULL_MASK, L[q * CW + CW - 1], (lane + 31) & 31);  // last value of lane-1 (rotated)
        dn[q] = __shfl_sync(FULL_MASK, L[q * CW], (lane + 1) & 31);            // first value of lane+1 (rotated)
    }
    const float mP2 = __fadd_rn(m, P2);
    float nL[E];
#pragma unroll
    for (int q = 0; q < NCH; ++q) {
#pragma unroll
        for (int e = 0; e < CW; ++e) {
            const int k = q * CW + e;
            float lo, hi;
            if (e > 0)
                lo = L[k - 1];
            else
                lo = lane > 0 ? up[q] : (q > 0 ? up[q > 0 ? q - 1 : 0] : L[k]);
            if (e < CW - 1)
                hi = L[k + 1];
            else
                hi = lane < 31 ? dn[q] : (q < NCH - 1 ? dn[q < NCH - 1 ? q + 1 : q] : L[k]);
            const float nb = __fadd_rn(fminf(lo, hi), P1);
            const float t = min3f(nb, L[k], mP2);
            nL[k] = __fadd_rn(t, __fsub_rn(C[k], m));
        }
    }
    float lm = nL[0];
#pragma unroll
    for (int k = 0; k < E; ++k) {
        L[k] = nL[k];
        lm = fminf(lm, nL[k]);
    }
    m = warp_min_f32(lm);
}

// N independent paths of the SAME pixel advanced together (they share C but not their predecessors): the shuffles of all
// N are issued first, then the arithmetic, then the N warp minima, so the fixed latencies of one path's chain
// (SHFL -> FMNMX -> FADD -> FMNMX3 -> FADD -> CREDUX) are covered by the other paths' instructions.  Same operations
// and rounding as path_step, per path.
template <int D, int N>
__device__ __forceinline__ void path_step_n(float (&L)[N][LaneLayout<D>::E], const float (&C)[LaneLayout<D>::E], float (&m)[N],
                                            int lane, float P1, float P2) {
    using LL = LaneLayout<D>;
    constexpr int E = LL::E, CW = LL::CW, NCH = LL::NCH;
    float up[N][NCH], dn[N][NCH];
#pragma unroll
    for (int n = 0; n < N; ++n)
#pragma unroll
        for (int q = 0; q < NCH; ++q) {
            up[n][q] = __shfl_sync(FULL_MASK, L[n][q * CW + CW - 1], (lane + 31) & 31);
            dn[n][q] = __shfl_sync(FULL_MASK, L[n][q * CW], (lane + 1) & 31);
        }
    float lm[N];
#pragma unroll
    for (int n = 0; n < N; ++n) {
        const float mP2 = __fadd_rn(m[n], P2);
        float nL[E];
#pragma unroll
        for (int q = 0; q < NCH; ++q) {
#pragma unroll
            for (int e = 0; e < CW; ++e) {
                const int k = q * CW + e;
                float lo, hi;
                if (e > 0)
                    lo = L[n][k - 1];
                else
                    lo = lane > 0 ? up[n][q] : (q > 0 ? up[n][q > 0 ? q - 1 : 0] : L[n][k]);
                if (e < CW - 1)
                    hi = L[n][k + 1];
                else
                    hi = lane < 31 ? dn[n][q] : (q < NCH - 1 ? dn[n][q < NCH - 1 ? q + 1 : q] : L[n][k]);
                const float nb = __fadd_rn(fminf(lo, hi), P1);
                const float t = min3f(nb, L[n][k], mP2);
                nL[k] = __fadd_rn(t, __fsub_rn(C[k], m[n]));
            }
        }
        lm[n] = nL[0];
#pragma unroll
        for (int k = 0; k < E; ++k) {
            L[n][k] = nL[k];
            lm[n] = fminf(lm[n], nL[k]);
        }
    }
#pragma unroll
    for (int n = 0; n < N; ++n) m[n] = warp_min_f32(lm[n]);
}

// WTA + uniqueness + sub-pixel on one pixel's summed costs held across the warp, in two halves:
//   wta_reduce  — the warp-wide part: first strict minimum, first occurrence of the smallest value different from it
//                 (src/SGM.cpp:381-405 / src/BM.cpp:59-80) and the winner's two neighbours; the result is warp-uniform;
//   wta_finish  — the scalar part on those six numbers: invalid (D+1) iff min/sec > ratio && |min_d - sec_d| > min_gap
//                 (src/SGM.cpp:406-416), then the parabola through S[d-1], S[d], S[d+1] clamped to D-1
//                 (src/Solver.cpp:577-594).  A caller may run it later, e.g. one lane per pixel for 32 pixels at once.
struct WtaPick {
    float c0, sec;     // minimum; second distinct value, +inf if there is none (costs are finite: the FLT_MAX path)
    int min_d, sec_d;
    float cm, cp;      // S[min_d - 1], S[min_d + 1] (0 where out of range; unused there)
};

template <int D>
__device__ __forceinline__ WtaPick wta_reduce(const float (&S)[LaneLayout<D>::E], int lane) {
    using LL = LaneLayout<D>;
    constexpr int E = LL::E;
    constexpr int NONE = 0x7fffffff;
    WtaPick p;
    float lmin = S[0];
#pragma unroll
    for (int k = 1; k < E; ++k) lmin = fminf(lmin, S[k]);
    const float gmin = warp_min_f32(lmin);
    int lidx = NONE;
    float lsec = INFINITY;
#pragma unroll
    for (int k = 0; k < E; ++k) {
        if (S[k] == gmin)
            lidx = min(lidx, LL::disparity(lane, k));
        else
            lsec = fminf(lsec, S[k]);
    }
    const int min_d = __reduce_min_sync(FULL_MASK, lidx);
    const float gsec = warp_min_f32(lsec);
    int lsidx = NONE;
#pragma unroll
    for (int k = 0; k < E; ++k)
        if (S[k] == gsec) lsidx = min(lsidx, LL::disparity(lane, k));
    const int sec_d = __reduce_min_sync(FULL_MASK, lsidx);
    // neighbours of the winner for the parabola (0 bits where this lane does not own them)
    int bm = 0, bp = 0;
#pragma unroll
    for (int k = 0; k < E; ++k) {
        const int dk = LL::disparity(lane, k);
        if (dk == min_d - 1) bm = __float_as_int(S[k]);
        if (dk == min_d + 1) bp = __float_as_int(S[k]);
    }
    p.c0 = gmin;
    p.sec = gsec;
    p.min_d = min_d;
    p.sec_d = sec_d;
    p.cm = __int_as_float((int)__reduce_or_sync(FULL_MASK, (unsigned)bm));
    p.cp = __int_as_float((int)__reduce_or_sync(FULL_MASK, (unsigned)bp));
    return p;
}

template <int D>
__device__ __forceinline__ void wta_finish(const WtaPick& p, float ratio, int min_gap, bool parabola, int& disp_out, float& sub_out) {
    bool invalid = false;
    if (p.sec != INFINITY)  // a second distinct value exists; otherwise sec_min_cost stays FLT_MAX and the ratio test fails
        invalid = (__fdiv_rn(p.c0, p.sec) > ratio) && (abs(p.min_d - p.sec_d) > min_gap);
    if (invalid) {
        disp_out = D + 1;
        sub_out = (float)(D + 1);
        return;
    }
    disp_out = p.min_d;
    if (!parabola || p.min_d == 0 || p.min_d == D - 1) {
        sub_out = (float)p.min_d;
        return;
    }
    const float num = __fsub_rn(p.cm, p.cp);
    const float den = __fmul_rn(2.0f, __fsub_rn(__fadd_rn(p.cm, p.cp), __fmul_rn(2.0f, p.c0)));
    const float v = __fadd_rn((float)p.min_d, __fdiv_rn(num, den));
    const float cap = (float)(D - 1);
    sub_out = cap < v ? cap : v;  // std::min(v, cap)
}

template <int D>
__device__ __forceinline__ void wta_pixel(const float (&S)[LaneLayout<D>::E], int lane, float ratio, int min_gap,
                                          bool parabola, int& disp_out, float& sub_out) {
    const WtaPick p = wta_reduce<D>(S, lane);
    wta_finish<D>(p, ratio, min_gap, parabola, disp_out, sub_out);
}

}  // namespace sgmb200
