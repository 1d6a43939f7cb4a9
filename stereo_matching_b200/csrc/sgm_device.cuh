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
    // neighbours of the winner for the parabola (0 bits where this lane does not own them).  Fetching them with two uniform
    // shuffles instead (as S16x2Arith::reduce does) was measured neutral to slightly slower here (aggregate 1.221 -> 1.229 ms).
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


// ---------------------------------------------------------------------------------------------
// Arithmetic of the aggregation kernels as a policy, so that the schedule (rings, row barrier, halo exchange) is written
// once.  A pixel's vector is DW 32-bit words spread over the warp as LaneLayout<DW> says; the kernels move the words as
// `float` registers and never look inside them.
//
// F32Arith<D>: the reference's arithmetic.  One word = one float32 cost; parity with src/SGM.cpp is bit for bit.
template <int D>
struct F32Arith {
    static constexpr int DW = D, DISP = D, E = LaneLayout<D>::E;
    __device__ static __forceinline__ float vmin(const float (&v)[E]) {
        float m = v[0];
#pragma unroll
        for (int k = 1; k < E; ++k) m = fminf(m, v[k]);
        return warp_min_f32(m);
    }
    __device__ static __forceinline__ void step(float (&L)[E], const float (&C)[E], float& m, int lane, float P1, float P2) {
        path_step<D>(L, C, m, lane, P1, P2);
    }
    template <int N>
    __device__ static __forceinline__ void step_n(float (&L)[N][E], const float (&C)[E], float (&m)[N], int lane, float P1, float P2) {
        path_step_n<D, N>(L, C, m, lane, P1, P2);
    }
    __device__ static __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    __device__ static __forceinline__ WtaPick reduce(const float (&S)[E], int lane) { return wta_reduce<D>(S, lane); }
};

// S16x2Arith<DW>: the integer mode of SURVEY 8(f) N4 / north_star ("min-plus recurrences use DPX").  With the cost box
// filter switched off (window 1x1) and no sky mask every value on the reference's path is a small non-negative integer
// (Hamming cost <= 62, L <= C + P2, S <= 8 (62 + P2)) that float32 represents exactly, so the SAME recurrence in 16-bit
// integers gives the SAME numbers.  One word = two disparities {d = 2k (low half), d = 2k + 1 (high half)} as s16x2, and
//     L' = min(L[d], min(L[d-1], L[d+1], m + P2 - P1) + P1) - m + C
// is VIMNMX3.S16x2 + VIADDMNMX.S16x2 + IADD3 per word (plus two PRMTs that line the d-1 / d+1 neighbours up); the
// per-half results never leave [0, 2^15), so the 32-bit add / subtract on packed words is exact.  The penalties arrive
// as packed words: P1 = {P1, P1}, P2 = {P2 - P1, P2 - P1} (the engine requires 0 <= P1 <= P2); the running minimum m
// is carried as {m, m}.
template <int DW_>
struct S16x2Arith {
    static constexpr int DW = DW_, DISP = 2 * DW_;
    using LL = LaneLayout<DW_>;
    static constexpr int E = LL::E, CW = LL::CW, NCH = LL::NCH;
    __device__ static __forceinline__ unsigned u(float x) { return __float_as_uint(x); }
    __device__ static __forceinline__ float f(unsigned x) { return __uint_as_float(x); }
    __device__ static __forceinline__ unsigned half_min(unsigned packed) {  // min of the two halves, in the low half
        return __vmins2(packed, packed >> 16) & 0xffffu;
    }
    __device__ static __forceinline__ float vmin(const float (&v)[E]) {
        unsigned m = u(v[0]);
#pragma unroll
        for (int k = 1; k < E; ++k) m = __vmins2(m, u(v[k]));
        return f(__reduce_min_sync(FULL_MASK, half_min(m)) * 0x10001u);
    }
    template <int N>
    __device__ static __forceinline__ void step_n(float (&L)[N][E], const float (&C)[E], float (&m)[N], int lane, float P1f, float P2f) {
        const unsigned P1 = u(P1f), P2mP1 = u(P2f);
        unsigned up[N][NCH], dn[N][NCH];
#pragma unroll
        for (int n = 0; n < N; ++n)
#pragma unroll
            for (int q = 0; q < NCH; ++q) {
                up[n][q] = __shfl_sync(FULL_MASK, u(L[n][q * CW + CW - 1]), (lane + 31) & 31);  // last word of lane-1 (rotated)
                dn[n][q] = __shfl_sync(FULL_MASK, u(L[n][q * CW]), (lane + 1) & 31);            // first word of lane+1 (rotated)
            }
        unsigned lm[N];
#pragma unroll
        for (int n = 0; n < N; ++n) {
            const unsigned mm = u(m[n]), thr = mm + P2mP1;
            unsigned nL[E];
#pragma unroll
            for (int q = 0; q < NCH; ++q) {
#pragma unroll
                for (int e = 0; e < CW; ++e) {
                    const int k = q * CW + e;
                    const unsigned cur = u(L[n][k]);
                    unsigned prev, next;  // the words holding d-1 of the low half / d+1 of the high half
                    if (e > 0)
                        prev = u(L[n][k - 1]);
                    else
                        prev = lane > 0 ? up[n][q] : (q > 0 ? up[n][q > 0 ? q - 1 : 0] : cur << 16);  // d = 0: the element itself (src/SGM.cpp:96)
                    if (e < CW - 1)
                        next = u(L[n][k + 1]);
                    else
                        next = lane < 31 ? dn[n][q] : (q < NCH - 1 ? dn[n][q < NCH - 1 ? q + 1 : q] : cur >> 16);  // d = D-1 (:97)
                    const unsigned dm = __byte_perm(prev, cur, 0x5432);  // {L[d-1] of the low half, of the high half}
                    const unsigned dp = __byte_perm(cur, next, 0x5432);  // {L[d+1] ...}
                    const unsigned x = __vimin3_s16x2(dm, dp, thr);
                    const unsigned t = __viaddmin_s16x2(x, P1, cur);
                    nL[k] = t - mm + u(C[k]);
                }
            }
            lm[n] = nL[0];
#pragma unroll
            for (int k = 0; k < E; ++k) {
                L[n][k] = f(nL[k]);
                lm[n] = __vmins2(lm[n], nL[k]);
            }
            lm[n] = half_min(lm[n]);
        }
#pragma unroll
        for (int n = 0; n < N; ++n) m[n] = f(__reduce_min_sync(FULL_MASK, lm[n]) * 0x10001u);
    }
    __device__ static __forceinline__ void step(float (&L)[E], const float (&C)[E], float& m, int lane, float P1, float P2) {
        float Ln[1][E], mn[1] = {m};
#pragma unroll
        for (int k = 0; k < E; ++k) Ln[0][k] = L[k];
        step_n<1>(Ln, C, mn, lane, P1, P2);
#pragma unroll
        for (int k = 0; k < E; ++k) L[k] = Ln[0][k];
        m = mn[0];
    }
    __device__ static __forceinline__ float add(float a, float b) { return f(u(a) + u(b)); }
    // Same selection as wta_reduce on the integer sums; the six numbers it returns are exact in float32, and wta_finish
    // then does the reference's float division / parabola on them unchanged.  A sum (< 2^15) and its disparity (< 2^8)
    // fit one 32-bit key (S << 8 | d), so "smallest value, first occurrence" is ONE warp-wide integer min (CREDUX.MIN),
    // and the second distinct value is the same min over the keys whose value differs from the first.
    __device__ static __forceinline__ WtaPick reduce(const float (&S)[E], int lane) {
        static_assert(DISP <= 256, "the key holds 8 bits of disparity");
        constexpr unsigned NONE = 0x7fffffffu;
        unsigned key[2 * E];
#pragma unroll
        for (int k = 0; k < E; ++k) {
            const unsigned w = u(S[k]), d0 = 2u * LL::disparity(lane, k);
            key[2 * k] = ((w << 8) & 0xffff00u) | d0;
            key[2 * k + 1] = ((w >> 8) & 0xffff00u) | (d0 + 1u);
        }
        unsigned lk = key[0];
#pragma unroll
        for (int k = 1; k < 2 * E; ++k) lk = min(lk, key[k]);
        const unsigned g1 = __reduce_min_sync(FULL_MASK, lk);
        const unsigned min_d = g1 & 0xffu, above = (g1 | 0xffu) + 1u;  // keys >= above hold a value different from the minimum
        unsigned ls = NONE;
#pragma unroll
        for (int k = 0; k < 2 * E; ++k) ls = min(ls, key[k] >= above ? key[k] : NONE);
        const unsigned g2 = __reduce_min_sync(FULL_MASK, ls);
        // the winner's two neighbours: min_d is warp-uniform, so the word that holds a neighbour is picked by a uniform
        // index and fetched with one shuffle from the lane that owns it (out-of-range neighbours are never used)
        auto fetch = [&](unsigned d) -> unsigned {
            const unsigned widx = (d >> 1) & (unsigned)(DW - 1);             // word of the pixel, wrapped into range
            const unsigned q = widx / (unsigned)LL::CHUNK_SPAN, in = widx % (unsigned)LL::CHUNK_SPAN;
            const unsigned k = q * CW + in % CW;                              // register of the owning lane
            unsigned w = u(S[0]);
#pragma unroll
            for (int e = 1; e < E; ++e) w = k == (unsigned)e ? u(S[e]) : w;
            w = __shfl_sync(FULL_MASK, w, (int)(in / CW));
            return (d & 1u) ? w >> 16 : w & 0xffffu;
        };
        const unsigned cm_i = fetch(min_d - 1u), cp_i = fetch(min_d + 1u);
        WtaPick p;
        p.c0 = (float)(g1 >> 8);
        p.sec = g2 == NONE ? INFINITY : (float)(g2 >> 8);
        p.min_d = (int)min_d;
        p.sec_d = (int)(g2 & 0xffu);
        p.cm = (float)cm_i;
        p.cp = (float)cp_i;
        return p;
    }
};

}  // namespace sgmb200
