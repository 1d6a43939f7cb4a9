// common.cuh — shared device helpers of the sm_100a stereo kernels.
//
// Volume layout everywhere: index = (i * W + j) * D + d (row-major pixels, d fastest), the
// reference's own layout (src/Solver.cpp:164, src/SGM.cpp:91).
//
// Numerics: the reference's hot path is float32 with a fixed operation order;
// every float add / sub / mul / div on the path goes through the round-to-nearest intrinsics
// (__fadd_rn ...), which nvcc never contracts into FMAs, and the library is built with -fmad=false.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace sgmb200 {

constexpr unsigned FULL_MASK = 0xffffffffu;

// Disparities of one pixel spread over the 32 lanes of a warp.  Lane l owns, for chunk q,
// CW consecutive disparities d = q * (32 * CW) + l * CW + e  (e < CW), so that every chunk is one
// fully coalesced 32 * CW * 4-byte warp access (float4 per lane for D >= 128).
template <int D>
struct LaneLayout {
    static_assert(D == 32 || D == 64 || D == 128 || D == 256, "max_disp must be 32, 64, 128 or 256");
    static constexpr int E = D / 32;           // values per lane
    static constexpr int CW = E < 4 ? E : 4;   // contiguous values per lane per chunk
    static constexpr int NCH = E / CW;         // chunks
    static constexpr int CHUNK_SPAN = 32 * CW; // disparities per chunk
    __device__ static __forceinline__ int disparity(int lane, int k) { return (k / CW) * CHUNK_SPAN + lane * CW + (k % CW); }
};

template <int CW>
struct VecIO;
template <>
struct VecIO<1> {
    __device__ static __forceinline__ void load(const float* p, float* v) { v[0] = *p; }
    __device__ static __forceinline__ void load_ro(const float* p, float* v) { v[0] = __ldg(p); }
    __device__ static __forceinline__ void store(float* p, const float* v) { *p = v[0]; }
};
template <>
struct VecIO<2> {
    __device__ static __forceinline__ void load(const float* p, float* v) {
        float2 t = *reinterpret_cast<const float2*>(p);
        v[0] = t.x; v[1] = t.y;
    }
    __device__ static __forceinline__ void load_ro(const float* p, float* v) {
        float2 t = __ldg(reinterpret_cast<const float2*>(p));
        v[0] = t.x; v[1] = t.y;
    }
    __device__ static __forceinline__ void store(float* p, const float* v) { *reinterpret_cast<float2*>(p) = make_float2(v[0], v[1]); }
};
template <>
struct VecIO<4> {
    __device__ static __forceinline__ void load(const float* p, float* v) {
        float4 t = *reinterpret_cast<const float4*>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    __device__ static __forceinline__ void load_ro(const float* p, float* v) {
        float4 t = __ldg(reinterpret_cast<const float4*>(p));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    __device__ static __forceinline__ void store(float* p, const float* v) { *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]); }
};

// One pixel's D values <-> registers of the warp.  `pix` points at d = 0 of the pixel.
template <int D>
__device__ __forceinline__ void load_pixel(const float* pix, int lane, float (&v)[LaneLayout<D>::E]) {
    using LL = LaneLayout<D>;
#pragma unroll
    for (int q = 0; q < LL::NCH; ++q) VecIO<LL::CW>::load(pix + q * LL::CHUNK_SPAN + lane * LL::CW, &v[q * LL::CW]);
}
template <int D>
__device__ __forceinline__ void load_pixel_ro(const float* pix, int lane, float (&v)[LaneLayout<D>::E]) {
    using LL = LaneLayout<D>;
#pragma unroll
    for (int q = 0; q < LL::NCH; ++q) VecIO<LL::CW>::load_ro(pix + q * LL::CHUNK_SPAN + lane * LL::CW, &v[q * LL::CW]);
}
template <int D>
__device__ __forceinline__ void store_pixel(float* pix, int lane, const float (&v)[LaneLayout<D>::E]) {
    using LL = LaneLayout<D>;
#pragma unroll
    for (int q = 0; q < LL::NCH; ++q) VecIO<LL::CW>::store(pix + q * LL::CHUNK_SPAN + lane * LL::CW, &v[q * LL::CW]);
}

// Costs on the path are finite floats but NOT sign-definite: with a sky mask the in-place recursive box
// filter cancels 999999-sized terms and leaves negative costs below the sky boundary (measured: down to -3.7e4),
// so integer (DPX VIMNMX) tricks on the bit patterns would be wrong.  Blackwell has the float forms natively:
// FMNMX3 (3-input float min, what fminf(fminf(a, b), c) compiles to on sm_100a) and CREDUX.MIN.F32
// (redux.sync.min.f32, one instruction for the min over the 32 lanes).
__device__ __forceinline__ float min3f(float a, float b, float c) { return fminf(fminf(a, b), c); }
__device__ __forceinline__ float warp_min_f32(float v) {
    float r;
    asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
    return r;
}

}  // namespace sgmb200
