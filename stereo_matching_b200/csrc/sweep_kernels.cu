// sweep_kernels.cu — the fused SGM aggregation schedule for sm_100a (engine v2).
//
// The line-walker schedule of aggregate_kernels.cu reads the cost volume eight times and round-trips the two
// partial sums after every path: 88 B per disparity-pixel per view.  This file brings that to the 48 B of the
// model in SURVEY.md 8(d) with three launches per frame:
//
//   1. sgm_hpair_kernel   L1 (->) and L2 (<-) of every row, both views, in ONE launch; L1 goes to acc_a, L2 to
//                         acc_b (float add commutes, so L1+L2 can be formed later).        read C 2x4, write 2x4 = 16 B
//   2. sgm_sweep_kernel<DOWN>  row-synchronous sweep top -> bottom, one warp per column holding L3 in registers
//                         and exchanging L5 / L6 with its column neighbours once per row; writes
//                         A = (L1+L2)+L3 over acc_a and B = L5+L6 over acc_b.             read C,L1,L2, write A,B = 20 B
//   3. sgm_sweep_kernel<UP>    the mirror image bottom -> top with L4, L7, L8; finishes
//                         S = (A+L4) + ((B+L7)+L8) in registers (the reference's association order,
//                         src/SGM.cpp:386-390) and runs WTA / uniqueness / sub-pixel on it.  read C,A,B = 12 B
//
// Column neighbours inside a block exchange through shared memory (one barrier per row); the first / last
// column of a block exchange with the neighbouring block through a small L2-resident halo ring in global memory
// whose payload carries its own row tag (HaloIO below: no flags, no fences).  A block only ever waits for its two
// neighbours' PREVIOUS row, so all blocks advance within one row of their neighbours; the kernel is launched
// cooperatively so that every block is resident, and every wait has a time-out that raises an error flag instead of
// hanging.  Launch shapes (columns per block, hardware warps, which warp scheduler the block-edge columns issue on):
// wide_shape() / sweep_column_of_warp() below.
//
// The kernels are written once for both arithmetics: they move a pixel's vector as 32-bit words (`float` registers) and
// leave every operation on the values to a policy class AR (sgm_device.cuh): F32Arith<D> is the reference's float32
// arithmetic (one word = one cost), S16x2Arith<D/2> the integer / DPX mode (one word = two 16-bit costs, SGM_FLAG_INTEGER).
// The template parameter D of a kernel is the number of WORDS per pixel.
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "common.cuh"
#include "kernels.h"
#include "sgm_device.cuh"

namespace sgmb200 {

namespace {

constexpr int HALO_RING = 8;  // rows of halo kept per block edge (>= 4 is required, see HaloIO)

struct SweepArgs {
    const float* cost;
    float* acc_a;
    float* acc_b;
    float* pathsum;
    uint16_t* disp;
    float* subpixel;
    float* halo;     // [views][nblk][2][HALO_RING] tagged vectors (HaloIO)
    int cols;        // columns per block; the LAST block of a row takes what is left (at most the kernel's NW, see fused_impl)
    unsigned tag_base;  // launch sequence number << 12: tags of this launch are tag_base + row + 1
    int* error;      // set to 1 when a halo wait timed out
    int debug;       // perf experiments only (SGM_B200_SWEEP_DEBUG): 1 = skip halo waits, 2 = skip row barrier waits (WRONG results)
    int h, w;
    int view0;       // first view handled by this launch (blockIdx.y is added)
    float P1, P2, ratio;
    int min_gap;
    size_t vol_stride, img_stride;
};

// perf experiments (tuning build only): SweepArgs::debug & 1 skips the halo waits, & 2 the row barrier waits (WRONG results)
__device__ __forceinline__ bool debug_skip(const SweepArgs& a, int bit) {
#ifdef SGM_B200_TUNING
    return (a.debug & bit) != 0;
#else
    (void)a;
    (void)bit;
    return false;
#endif
}

template <int E>
__device__ __forceinline__ float vec_min(const float (&v)[E]) {
    float m = v[0];
#pragma unroll
    for (int k = 1; k < E; ++k) m = fminf(m, v[k]);
    return m;
}

// ---------------------------------------------------------------------------------------------
// Cross-block halo exchange without flags or fences.  A diagonal vector (D floats) that has to reach the neighbouring
// block is written to an L2-resident ring slot as 64-bit words {value, tag}; the tag names the launch and the row, so
// the consumer simply polls the payload itself (L1-bypassing loads) until every word carries the tag it expects.  One
// L2 round trip per hop, nothing else on the critical path.
//
// Ring safety (RING >= 4): a block publishes row r+RING into the slot of row r only after it has consumed its
// neighbour's row r+RING-2 vector, which the neighbour published after finishing row r+RING-3, i.e. after it had
// read this block's row r+RING-4 >= r vector.
template <int D>
struct HaloIO {
    static constexpr int E = LaneLayout<D>::E;
    static constexpr int SLOT_WORDS = E * 32;  // 64-bit {value, tag} words per vector
    // Every 64-bit word carries its own tag and is moved by 64-bit scalar (or v2.u64) accesses: however the
    // hardware or ptxas splits an access, a word is never torn, so a value is valid iff its own tag matches.
    __device__ static __forceinline__ unsigned long long pack(float v, unsigned tag) {
        return (unsigned long long)__float_as_uint(v) | ((unsigned long long)tag << 32);
    }
    __device__ static __forceinline__ void publish(unsigned long long* slot, int lane, const float (&v)[E], unsigned tag) {
        if constexpr (E == 1) {
            asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(slot + lane), "l"(pack(v[0], tag)) : "memory");
        } else {
#pragma unroll
            for (int q = 0; q < E / 2; ++q)
                asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(slot + (q * 32 + lane) * 2), "l"(pack(v[2 * q], tag)),
                             "l"(pack(v[2 * q + 1], tag))
                             : "memory");
        }
    }
    // Early, non-blocking half of consume: the loads are issued (typically a whole step before the data is needed) and
    // accept() later checks the tags; the L2 round trip then overlaps the work in between.
    __device__ static __forceinline__ void peek(const unsigned long long* slot, int lane, unsigned long long (&raw)[E]) {
        if constexpr (E == 1) {
            asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(raw[0]) : "l"(slot + lane) : "memory");
        } else {
#pragma unroll
            for (int q = 0; q < E / 2; ++q)
                asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];"
                             : "=l"(raw[2 * q]), "=l"(raw[2 * q + 1])
                             : "l"(slot + (q * 32 + lane) * 2)
                             : "memory");
        }
    }
    __device__ static __forceinline__ bool accept(const unsigned long long (&raw)[E], float (&v)[E], unsigned tag) {
        bool ok = true;
#pragma unroll
        for (int k = 0; k < E; ++k) {
            ok = ok && (unsigned)(raw[k] >> 32) == tag;
            v[k] = __uint_as_float((unsigned)raw[k]);
        }
        return __all_sync(FULL_MASK, ok);
    }
    // Polls until the whole vector of `tag` has arrived; false when it timed out.
    __device__ static __forceinline__ bool consume(const unsigned long long* slot, int lane, float (&v)[E], unsigned tag) {
#pragma unroll 1
        for (int spin = 0; spin < (1 << 22); ++spin) {
            bool ok = true;
            if constexpr (E == 1) {
                unsigned long long x;
                asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(x) : "l"(slot + lane) : "memory");
                ok = (unsigned)(x >> 32) == tag;
                v[0] = __uint_as_float((unsigned)x);
            } else {
#pragma unroll
                for (int q = 0; q < E / 2; ++q) {
                    unsigned long long x, y;
                    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(x), "=l"(y) : "l"(slot + (q * 32 + lane) * 2) : "memory");
                    ok = ok && (unsigned)(x >> 32) == tag && (unsigned)(y >> 32) == tag;
                    v[2 * q] = __uint_as_float((unsigned)x);
                    v[2 * q + 1] = __uint_as_float((unsigned)y);
                }
            }
            if (__all_sync(FULL_MASK, ok)) return true;
        }
        return false;
    }
};

// ---- shared-memory helpers on 32-bit shared-space addresses (so that every offset folds into the instruction) ----
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
template <int D>
__device__ __forceinline__ void lds_pixel(unsigned addr, float (&v)[LaneLayout<D>::E]) {  // addr already includes lane * CW * 4
    using LL = LaneLayout<D>;
#pragma unroll
    for (int q = 0; q < LL::NCH; ++q) {
        const unsigned p = addr + q * LL::CHUNK_SPAN * 4;
        if constexpr (LL::CW == 4)
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v[q * 4]), "=f"(v[q * 4 + 1]), "=f"(v[q * 4 + 2]), "=f"(v[q * 4 + 3]) : "r"(p));
        else if constexpr (LL::CW == 2)
            asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v[q * 2]), "=f"(v[q * 2 + 1]) : "r"(p));
        else
            asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v[q]) : "r"(p));
    }
}
template <int D>
__device__ __forceinline__ void sts_pixel(unsigned addr, const float (&v)[LaneLayout<D>::E]) {
    using LL = LaneLayout<D>;
#pragma unroll
    for (int q = 0; q < LL::NCH; ++q) {
        const unsigned p = addr + q * LL::CHUNK_SPAN * 4;
        if constexpr (LL::CW == 4)
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(p), "f"(v[q * 4]), "f"(v[q * 4 + 1]), "f"(v[q * 4 + 2]), "f"(v[q * 4 + 3]) : "memory");
        else if constexpr (LL::CW == 2)
            asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(p), "f"(v[q * 2]), "f"(v[q * 2 + 1]) : "memory");
        else
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(p), "f"(v[q]) : "memory");
    }
}
__device__ __forceinline__ float lds_f32(unsigned addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts_f32(unsigned addr, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory"); }

// Row barrier of a block as an mbarrier: a warp ARRIVES as soon as its diagonal vectors of the row are in shared memory
// and only WAITS at the top of the next row, so the row's output work (sums, stores, WTA) overlaps the wait.
__device__ __forceinline__ void mbar_init(unsigned bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(bar) : "memory");
}
constexpr unsigned MBAR_SUSPEND_NS = 2000;  // upper bound of one hardware sleep; a completed phase wakes the warp at once
// Bounded: a phase that never completes (a lost arrival would be a bug, a block that never became resident a launch
// problem) must end as an error code of the frame, not as a hung GPU.  After 2^20 tries of up to 2 us each the warp
// records error 2 and exits; the warps waiting for it follow the same way, so the launch ends within seconds.  The
// counter lives inside the PTX loop: the path of a wait that succeeds at once is one instruction longer than unbounded.
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity, int* err) {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t.reg .u32 n;\n\tmov.u32 n, 0;\n"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"  // %2: suspend-time hint (ns): sleep, do not spin
        "@p bra WAIT_DONE;\n\t"
        "add.u32 n, n, 1;\n\t"
        "setp.lt.u32 q, n, 1048576;\n\t"
        "@q bra WAIT_LOOP;\n\t"
        "st.global.u32 [%3], 2;\n\t"
        "exit;\n"
        "WAIT_DONE:\n\t}" ::"r"(bar),
        "r"(parity), "r"(MBAR_SUSPEND_NS), "l"(err)
        : "memory");
}

// One pixel's D values <-> registers, through a pointer that already includes the lane's offset (lane * CW floats).
template <int D>
__device__ __forceinline__ void ldg_ro_vec(const float* p, float (&v)[LaneLayout<D>::E]) {
    using LL = LaneLayout<D>;
#pragma unroll
    for (int q = 0; q < LL::NCH; ++q) VecIO<LL::CW>::load_ro(p + q * LL::CHUNK_SPAN, &v[q * LL::CW]);
}
template <int D>
__device__ __forceinline__ void ldg_vec(const float* p, float (&v)[LaneLayout<D>::E]) {
    using LL = LaneLayout<D>;
#pragma unroll
    for (int q = 0; q < LL::NCH; ++q) VecIO<LL::CW>::load(p + q * LL::CHUNK_SPAN, &v[q * LL::CW]);
}
template <int D>
__device__ __forceinline__ void stg_vec(float* p, const float (&v)[LaneLayout<D>::E]) {
    using LL = LaneLayout<D>;
#pragma unroll
    for (int q = 0; q < LL::NCH; ++q) VecIO<LL::CW>::store(p + q * LL::CHUNK_SPAN, &v[q * LL::CW]);
}

// ---------------------------------------------------------------------------------------------
// L1 (left -> right, src/SGM.cpp:82-119) and L2 (right -> left, :122-159): one warp per (row, direction, view), the row
// streamed through a per-warp shared-memory ring by cp.async (LDGSTS).  Only 2*H*views independent lines exist, so each
// warp keeps PF pixels in flight.  A deep register prefetch (16 loads in flight per warp) has to share the warp's six load scoreboards, so
// waiting for the oldest load also waits for younger ones and the effective depth collapses (measured: 1100 cycles per
// pixel step, 2.6 TB/s, whatever the image width).  cp.async completion is tracked per commit group, so PF really
// means PF pixels in flight, and the ring costs no registers.  Each lane copies and later reads back its own 16 bytes,
// so no cross-lane synchronisation is needed.
template <int BYTES>
__device__ __forceinline__ void cp_async_lane(unsigned dst, const void* src) {
    if constexpr (BYTES == 16)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
    else if constexpr (BYTES == 8)
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
    else
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}

template <int D, int PF, int WARPS, class AR>
__global__ void __launch_bounds__(WARPS * 32) sgm_hpair_async_kernel(const float* __restrict__ cost, float* __restrict__ out_fwd,
                                                                     float* __restrict__ out_bwd, int h, int w, float P1,
                                                                     float P2, size_t vol_stride) {
    using LL = LaneLayout<D>;
    constexpr int E = LL::E;
    extern __shared__ __align__(16) float hring[];  // [WARPS][PF][D]
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int row = blockIdx.x * WARPS + wid;
    if (row >= h) return;
    const int view = blockIdx.y;
    const bool bwd = blockIdx.z != 0;
    const ptrdiff_t step = bwd ? -(ptrdiff_t)D : (ptrdiff_t)D;
    const size_t first = view * vol_stride + ((size_t)row * w + (bwd ? w - 1 : 0)) * D + lane * LL::CW;
    const float* pc = cost + first;                      // next pixel to fetch
    float* po = (bwd ? out_bwd : out_fwd) + first;       // next pixel to write
    const unsigned ring = smem_u32(hring) + (unsigned)wid * PF * D * 4 + lane * LL::CW * 4;

    auto fetch = [&](int slot) {
#pragma unroll
        for (int q = 0; q < LL::NCH; ++q) cp_async_lane<LL::CW * 4>(ring + slot * D * 4 + q * LL::CHUNK_SPAN * 4, pc + q * LL::CHUNK_SPAN);
        pc += step;
    };
#pragma unroll
    for (int u = 0; u < PF; ++u) {
        if (u < w) fetch(u);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }

    float L[E];
    float m = 0.f;
    for (int s0 = 0; s0 < w; s0 += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int s = s0 + u;
            if (s < w) {
                asm volatile("cp.async.wait_group %0;" ::"n"(PF - 1) : "memory");  // the group of pixel s has landed
                float C[E];
                lds_pixel<D>(ring + u * D * 4, C);
                if (s == 0) {
#pragma unroll
                    for (int k = 0; k < E; ++k) L[k] = C[k];
                    m = AR::vmin(L);
                } else {
                    AR::step(L, C, m, lane, P1, P2);
                }
                stg_vec<D>(po, L);
                po += step;
                if (s + PF < w) fetch(u);  // slot u is free again: the lane has read its own bytes back
                asm volatile("cp.async.commit_group;" ::: "memory");
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Row-synchronous sweep.  UP == false: rows 0 -> h-1 with L3 (src/SGM.cpp:162-199), L5 and L6 (:247-305).
//                         UP == true : rows h-1 -> 0 with L4 (:202-239), L7 and L8 (:311-369) + sum/WTA (:373-418).
// Warp w of block b owns column j = b*NW + w and keeps the vertical path in registers.  "Pl" is the diagonal whose
// predecessor is the LEFT neighbour column (L5 going down, L7 going up), "Pr" the one whose predecessor is the RIGHT
// neighbour column (L6 / L8).  Inside a block the previous row's diagonal vectors travel through shared memory (double
// buffered by row parity, one mbarrier phase per row); across blocks through HaloIO.  Each warp runs its two diagonals
// in the order
//     F = the one a NEIGHBOURING BLOCK is waiting for (if any), published as soon as it is computed,
//     V = the vertical path,
//     G = the one that may have to wait for a neighbouring block,
// so a vector is published about one row and one step before the neighbour needs it, which hides the L2 hop.
// PF (rows of C / A / B kept in flight per warp) is even, so the row parity is a compile-time constant of the
// unrolled body and every shared-memory address is base + immediate.
constexpr int SWEEP_DOWN = 0, SWEEP_UP = 1, SWEEP_UP_KEEP = 2;

// Which column of the block a hardware warp works on (-1: none, the warp exits; NW + e: HELPER of edge column e, below).
// The two block-edge columns carry about a third more work per row than the others (halo accept / publish, two recurrence
// steps one after the other) and every row ends at a block barrier, so they set the pace.  A warp's scheduler is
// (hardware warp id) % 4.  Rule: the edge columns go to hardware warps 3 and 7 (scheduler 3) and, with HELP, their helper
// warps to 11 and 15; the interior columns fill the warps of schedulers 0 - 2 in hardware order and use scheduler 3's
// remaining warps only when those are used up; whatever is left of the HWW hardware warps exits.  With 19 columns on 19
// hardware warps the schedulers host 5 + 5 + 5 interior warps and 2 edge + 2 interior; with 17 columns on 19 hardware warps
// scheduler 3 belongs to the edge warps (and their helpers).  `identity`: tuning-build A/B switch (1242x375 D=128, NW = HWW = 19: 1.249 ms in natural
// order, 1.196 ms with this rule).  Blocks of fewer than 9 columns (the ragged last block of a row) use the natural order.
template <int NW, int HWW, bool HELP>
__device__ __forceinline__ int sweep_column_of_warp(int hw, int nact, bool identity) {
    static_assert(HWW >= NW + (HELP ? 2 : 0), "one hardware warp per column (and helper) at least");
    static_assert(!HELP || NW >= 16, "helper warps sit on hardware warps 11 and 15");
    if (identity || nact < 9 || NW < 13) return hw < nact ? hw : -1;  // (the 8- and 10-column default shapes keep the natural order)
    if (hw == 3) return 0;
    if (hw == 7) return nact - 1;
    if (HELP && hw == 11) return NW;
    if (HELP && hw == 15) return NW + 1;
    constexpr int OUTSIDE = HWW - HWW / 4;      // hardware warps of schedulers 0 - 2
    constexpr int FIRST3 = HELP ? 19 : 11;      // first hardware warp of scheduler 3 that may take an interior column
    const int rank = (hw & 3) != 3 ? hw - ((hw + 1) >> 2) : OUTSIDE + ((hw - FIRST3) >> 2);
    return rank < nact - 2 ? 1 + rank : -1;  // interior columns 1 .. nact-2
}
// Do the edge columns of this block hand their output work to helper warps?  (warp-uniform, the same answer in every warp)
template <bool HELP>
__device__ __forceinline__ bool sweep_block_has_helpers(int nact, bool identity) {
    return HELP && !identity && nact >= 9;
}

// Columns of block blk: a.cols, except that the last block of a row takes the rest (fused_impl folds a short remainder into it
// instead of launching a block of one or two columns that would occupy an SM of its own).
__device__ __forceinline__ int sweep_block_columns(const SweepArgs& a, int blk, int nblk) {
    return blk + 1 < nblk ? a.cols : a.w - blk * a.cols;
}

// EDGE (compile-time copy of a warp-uniform fact): false = an interior warp — both diagonals come from shared memory and
// neither restarts at an image border; such a warp carries none of the halo / border code (in the generic body those
// ~35 instructions per pixel were issued predicated-off by 17 of 19 warps).  true = a block-edge or image-border warp.
// OFFL (edge warps of a block with helpers only): this warp runs the recurrences and the exchanges; its sums, selection and
// stores are done by its helper warp (sweep_helper_body), which gets the vertical path's vector through `vx`.
template <int D, int NW, int MODE, int PATHS, int PF, bool EDGE, class AR, bool OFFL = false>
__device__ __forceinline__ void sweep_body(const SweepArgs& a, float* xch, float* xmin, unsigned bar, float* sring, int wid,
                                           float* vx = nullptr) {
    static_assert(!OFFL || EDGE, "only edge warps have helpers");
    constexpr bool UP = MODE != SWEEP_DOWN, KEEP_S = MODE == SWEEP_UP_KEEP;  // KEEP_S: also write the path sum S (debug tap)
    using LL = LaneLayout<D>;
    using HIO = HaloIO<D>;
    constexpr int E = LL::E;
    constexpr bool DIAG = PATHS == 8;
    constexpr bool TOGETHER = E <= 4;  // see the block-edge branch of the row loop
    static_assert(NW >= 2 && PF % 2 == 0, "a block needs two columns; PF must be even");
    constexpr unsigned VEC_BYTES = D * 4, WARP_BYTES = 2 * VEC_BYTES, PAR_BYTES = NW * WARP_BYTES, MIN_PAR_BYTES = NW * 2 * 4;
    const int lane = threadIdx.x & 31;
    const int blk = blockIdx.x, nblk = gridDim.x;
    const int view = a.view0 + blockIdx.y;
    const int h = a.h, w = a.w;
    const int nact = sweep_block_columns(a, blk, nblk);  // columns (= warps) of this block
    const int j = blk * a.cols + wid;                    // "wid" is the COLUMN of the block this warp works on (sweep_column_of_warp)

    // ---- roles of the two diagonals of this warp ----
    const bool swap = EDGE && DIAG && wid == 0 && blk > 0;  // F = Pr (the left block waits for it), G = Pl (comes from the left block)
    const int f_id = swap ? 1 : 0, g_id = swap ? 0 : 1;
    const bool left_border = EDGE && j == 0, right_border = EDGE && j == w - 1;
    const bool f_init = swap ? right_border : left_border;  // F restarts on every row (image border on its predecessor side)
    const bool g_init = swap ? false : right_border;
    // shared-memory source warp of F; a warp whose F restarts on every row reads its own slot (valid memory, value unused)
    const int f_src_w = f_init ? wid : (swap ? wid + 1 : wid - 1);
    const bool g_global = EDGE && DIAG && (swap || (wid == nact - 1 && blk + 1 < nblk));
    const int g_src_w = (g_global || g_init) ? wid : wid + 1;  // shared-memory source warp of G (own slot when unused)
    const unsigned lane_off = lane * LL::CW * 4;
    const unsigned x0 = smem_u32(xch), m0 = smem_u32(xmin);
    const unsigned f_src = x0 + f_src_w * WARP_BYTES + f_id * VEC_BYTES + lane_off, fm_src = m0 + (f_src_w * 2 + f_id) * 4;
    const unsigned g_src = x0 + g_src_w * WARP_BYTES + g_id * VEC_BYTES + lane_off, gm_src = m0 + (g_src_w * 2 + g_id) * 4;
    const unsigned f_dst = x0 + wid * WARP_BYTES + f_id * VEC_BYTES + lane_off, fm_dst = m0 + (wid * 2 + f_id) * 4;
    const unsigned g_dst = x0 + wid * WARP_BYTES + g_id * VEC_BYTES + lane_off, gm_dst = m0 + (wid * 2 + g_id) * 4;
    // halo ring: [view][blk][2][RING] vectors; which 0 = right-going (Pl of the last column), 1 = left-going (Pr of column 0)
    unsigned long long* halo = reinterpret_cast<unsigned long long*>(a.halo);
    auto ring = [&](int b, int which) { return halo + ((size_t)(view * nblk + b) * 2 + which) * HALO_RING * HIO::SLOT_WORDS; };
    unsigned long long* f_pub = nullptr;
    const unsigned long long* g_sub = nullptr;
    if (EDGE && DIAG) {
        if (swap) f_pub = ring(blk, 1);
        else if (wid == nact - 1 && blk + 1 < nblk) f_pub = ring(blk, 0);
        if (swap) g_sub = ring(blk - 1, 0);
        else if (g_global) g_sub = ring(blk + 1, 1);
    }

    // ---- global pointers: p* run PF rows ahead (prefetch), o* are at the row being finished ----
    const ptrdiff_t rstep = UP ? -(ptrdiff_t)w * D : (ptrdiff_t)w * D;
    const size_t first = ((size_t)(UP ? h - 1 : 0) * w + j) * D + lane * LL::CW;
    const float* pc = a.cost + view * a.vol_stride + first;
    const float* px = a.acc_a + view * a.vol_stride + first;
    const float* py = (DIAG || !UP) ? a.acc_b + view * a.vol_stride + first : nullptr;
    float* oa = a.acc_a + view * a.vol_stride + first;
    float* ob = (DIAG && !UP) ? a.acc_b + view * a.vol_stride + first : nullptr;
    float* os = KEEP_S ? a.pathsum + view * a.vol_stride + first : nullptr;
    const ptrdiff_t pstep = UP ? -(ptrdiff_t)w : (ptrdiff_t)w;
    const size_t opix0 = view * a.img_stride + (size_t)(UP ? h - 1 : 0) * w + j;  // output pixel of sweep row 0
    WtaPick stash = {};  // lane l keeps the selection of sweep row (r & ~31) + l until the 32 rows are finished together
    // Software-pipelined selection (up sweep): the summed vector S of row r-1 stays in registers and its warp-wide selection
    // is issued INSIDE row r's recurrence step, between the barrier wait and the arrival.  Only the step is serial from row
    // to row; sum, selection and stores are not, but a warp issues in order, so leaving them behind the step made every row
    // cost the latency of both chains.  In one basic block ptxas interleaves the two independent chains.
    float Sp[E];
#pragma unroll
    for (int k = 0; k < E; ++k) Sp[k] = 0.f;
    // (tuning build, bit 32: the edge columns of a full block produce NO outputs — WRONG results; bounds what handing their
    // sums / selection / stores to helper warps could gain)
    const bool no_out = OFFL || (EDGE && nact >= 9 && debug_skip(a, 32));
    const unsigned v_dst = OFFL ? smem_u32(vx) + (wid == 0 ? 0u : VEC_BYTES) + lane * LL::CW * 4 : 0u;  // [parity][edge][d]
    auto select_pending = [&](int rp) {  // rp = sweep row of the vector held in Sp
        if (no_out) return;
        const WtaPick pk = AR::reduce(Sp, lane);
        if (lane == (rp & 31)) stash = pk;
    };
    // the scalar half (uniqueness division, parabola, stores) once per 32 rows with one lane per row
    auto flush_rows = [&](int rlast) {
        const int rr = (rlast & ~31) + lane;
        if (rr <= rlast) {
            int dd;
            float sub;
            wta_finish<AR::DISP>(stash, a.ratio, a.min_gap, true, dd, sub);
            const size_t o = opix0 + (ptrdiff_t)rr * pstep;
            a.disp[o] = (uint16_t)dd;
            a.subpixel[o] = sub;
        }
    };

    // C / A / B rows stream through a per-warp shared-memory ring filled by cp.async: PF rows in flight per warp cost no
    // registers and are tracked per commit group (a deep register prefetch would alias the load scoreboards).
    constexpr unsigned SLOT_BYTES = 3 * VEC_BYTES;
    const unsigned rbase = smem_u32(sring) + (unsigned)wid * PF * SLOT_BYTES + lane_off;
    auto fetch = [&](int slot) {
#pragma unroll
        for (int q = 0; q < LL::NCH; ++q) {
            const unsigned dst = rbase + slot * SLOT_BYTES + q * LL::CHUNK_SPAN * 4;
            cp_async_lane<LL::CW * 4>(dst, pc + q * LL::CHUNK_SPAN);
            if (!debug_skip(a, 4) && !no_out) {  // (tuning build: bit 4 = do not fetch the A / B rows: WRONG results, LSU experiment)
                cp_async_lane<LL::CW * 4>(dst + VEC_BYTES, px + q * LL::CHUNK_SPAN);
                if (DIAG || !UP) cp_async_lane<LL::CW * 4>(dst + 2 * VEC_BYTES, py + q * LL::CHUNK_SPAN);
            }
        }
        pc += rstep;
        px += rstep;
        if (DIAG || !UP) py += rstep;
    };
#pragma unroll
    for (int u = 0; u < PF; ++u) {
        if (u < h) fetch(u);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }

    float V[E];
    float mV = 0.f;
    bool dead = false;  // a halo wait timed out: stop waiting, keep the barrier phases going
    // block-edge warp: loads of the neighbour block's vector for the NEXT row, issued right after this row's vectors are
    // published (the neighbour publishes its own early in the row, so it is usually there); checked at the top of the
    // next row, where a miss falls back to a second look and then to polling
    unsigned long long early[E];
    bool early_ok = false;
    int n_early = 0, n_peek = 0, n_poll = 0;  // (tuning build, bit 128: where the neighbour block's vector was found, printed per warp)
    for (int r0 = 0; r0 < h; r0 += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int r = r0 + u;
            if (r < h) {
                const unsigned par = u & 1;  // == r & 1 because PF is even: a constant of the unrolled body
                const unsigned tag = a.tag_base + (unsigned)r + 1u;
                float F[E], G[E], C[E];
                // block-edge warp: ask L2 now for the neighbour block's row r-1 vector (published about a row ago), so the
                // round trip overlaps the barrier wait and the F step; the tags are checked where the vector is needed
                unsigned long long peeked[E];
                float Gin[E];
                const bool peek_ok = EDGE && DIAG && g_global && r > 0 && !dead && !debug_skip(a, 1);
                bool got = false;
                float mGin = 0.f;
                if (peek_ok) {
                    if (early_ok) got = HIO::accept(early, Gin, tag - 1u);
                    n_early += got;
                    if (!got) HIO::peek(g_sub + (size_t)((r - 1) % HALO_RING) * HIO::SLOT_WORDS, lane, peeked);
                    else if (TOGETHER) mGin = AR::vmin(Gin);  // its warp-wide minimum: ahead of the barrier wait, off the row's critical path
                }
                asm volatile("cp.async.wait_group %0;" ::"n"(PF - 1) : "memory");  // this row's C / A / B have landed
                lds_pixel<D>(rbase + u * SLOT_BYTES, C);
                if (DIAG && r > 0 && !debug_skip(a, 2)) mbar_wait(bar, par ^ 1u, a.error);  // every warp's row r-1 vectors are in shared memory
                // ---- the three recurrences of this pixel (P[0] = F, P[1] = V, P[2] = G) ----
                // F, V and G only depend on the previous row, not on each other, so they advance together
                // (path_step_n) unless G has to wait for a neighbouring block: then F goes first and is published,
                // and V and G follow together.
                if (r == 0) {
#pragma unroll
                    for (int k = 0; k < E; ++k) F[k] = V[k] = G[k] = C[k];
                    mV = AR::vmin(V);
                    if (DIAG) {
                        if (f_pub) HIO::publish(f_pub, lane, F, tag);
                        sts_pixel<D>(f_dst + par * PAR_BYTES, F);
                        sts_pixel<D>(g_dst + par * PAR_BYTES, G);
                        if (lane == 0) {
                            sts_f32(fm_dst + par * MIN_PAR_BYTES, mV);
                            sts_f32(gm_dst + par * MIN_PAR_BYTES, mV);
                        }
                    }
                } else if (!DIAG) {
                    AR::step(V, C, mV, lane, a.P1, a.P2);
                    if (UP) select_pending(r - 1);
                } else {
                    float P[3][E], mP[3];
                    // previous-row vectors: F and (unless it comes from another block) G from shared memory
                    // (a warp whose diagonal restarts on this row still steps a vector — its own — and drops the result below)
                    lds_pixel<D>(f_src + (par ^ 1u) * PAR_BYTES, P[0]);
                    mP[0] = lds_f32(fm_src + (par ^ 1u) * MIN_PAR_BYTES);
#pragma unroll
                    for (int k = 0; k < E; ++k) P[1][k] = V[k];
                    mP[1] = mV;
                    if (!g_global) {
                        lds_pixel<D>(g_src + (par ^ 1u) * PAR_BYTES, P[2]);
                        mP[2] = lds_f32(gm_src + (par ^ 1u) * MIN_PAR_BYTES);
                        AR::template step_n<3>(P, C, mP, lane, a.P1, a.P2);
                        if (UP) select_pending(r - 1);
                    } else if (TOGETHER && got && !debug_skip(a, 64)) {
                        // block-edge warp whose neighbour-block vector arrived early (the usual case: blocks run in step and
                        // the neighbour published it about a row ago): all three paths advance together, as in an interior warp
                        // (1242x375: -2.3 % at D = 128, -2.4 % at D = 64, -4.4 % in the integer mode; with 8 values per lane,
                        // i.e. D = 256, the three-path step spills and 1080p gets 8 % slower, hence TOGETHER)
#pragma unroll
                        for (int k = 0; k < E; ++k) P[2][k] = Gin[k];
                        mP[2] = mGin;
                        AR::template step_n<3>(P, C, mP, lane, a.P1, a.P2);
                        if (f_init) {
#pragma unroll
                            for (int k = 0; k < E; ++k) P[0][k] = C[k];
                            mP[0] = AR::vmin(C);
                        }
                        if (f_pub) HIO::publish(f_pub + (size_t)(r % HALO_RING) * HIO::SLOT_WORDS, lane, P[0], tag);
                        if (UP) select_pending(r - 1);
                    } else {
                        // block-edge warp: F first (a neighbouring block is waiting for it), then V and G together
                        float P0[1][E], m0[1] = {mP[0]};
#pragma unroll
                        for (int k = 0; k < E; ++k) P0[0][k] = P[0][k];
                        AR::template step_n<1>(P0, C, m0, lane, a.P1, a.P2);
                        if (f_init) {
#pragma unroll
                            for (int k = 0; k < E; ++k) P0[0][k] = C[k];
                            m0[0] = AR::vmin(C);
                        }
                        if (f_pub) HIO::publish(f_pub + (size_t)(r % HALO_RING) * HIO::SLOT_WORDS, lane, P0[0], tag);
                        if (UP) select_pending(r - 1);  // the neighbour's vector may still be on its way: use the time
                        float P12[2][E], m12[2];
                        if (peek_ok && !got) {
                            const bool second = HIO::accept(peeked, Gin, tag - 1u);
                            n_peek += second;
                            n_poll += !second;
                            if (!second && !HIO::consume(g_sub + (size_t)((r - 1) % HALO_RING) * HIO::SLOT_WORDS, lane, Gin, tag - 1u)) {
                                dead = true;
                                if (lane == 0) atomicExch(a.error, 1);
                            }
                        }
#pragma unroll
                        for (int k = 0; k < E; ++k) P12[1][k] = Gin[k];
#pragma unroll
                        for (int k = 0; k < E; ++k) P12[0][k] = V[k];
                        m12[0] = mV;
                        m12[1] = AR::vmin(P12[1]);
                        AR::template step_n<2>(P12, C, m12, lane, a.P1, a.P2);
#pragma unroll
                        for (int k = 0; k < E; ++k) {
                            P[0][k] = P0[0][k];
                            P[1][k] = P12[0][k];
                            P[2][k] = P12[1][k];
                        }
                        mP[0] = m0[0];
                        mP[1] = m12[0];
                        mP[2] = m12[1];
                    }
                    if (!g_global) {  // image-border columns restart their diagonal on every row (src/SGM.cpp:266,282,330,346)
                        if (f_init) {
#pragma unroll
                            for (int k = 0; k < E; ++k) P[0][k] = C[k];
                            mP[0] = AR::vmin(C);
                        }
                        if (g_init) {
#pragma unroll
                            for (int k = 0; k < E; ++k) P[2][k] = C[k];
                            mP[2] = AR::vmin(C);
                        }
                    }  // (f_pub != nullptr implies g_global, handled above)
#pragma unroll
                    for (int k = 0; k < E; ++k) {
                        F[k] = P[0][k];
                        V[k] = P[1][k];
                        G[k] = P[2][k];
                    }
                    mV = mP[1];
                    sts_pixel<D>(f_dst + par * PAR_BYTES, F);
                    sts_pixel<D>(g_dst + par * PAR_BYTES, G);
                    if (lane == 0) {
                        sts_f32(fm_dst + par * MIN_PAR_BYTES, mP[0]);
                        sts_f32(gm_dst + par * MIN_PAR_BYTES, mP[2]);
                    }
                }
                if (OFFL) sts_pixel<D>(v_dst + par * 2 * VEC_BYTES, V);
                if (DIAG) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar);  // this warp's row r vectors are published
                    early_ok = EDGE && g_global && r + 1 < h && !dead && !debug_skip(a, 1);
                    if (early_ok) HIO::peek(g_sub + (size_t)(r % HALO_RING) * HIO::SLOT_WORDS, lane, early);
                }
                // ---- outputs (overlap the other warps' arrival) ----
                float X[E], Y[E];
                if (no_out) {
                } else
                if (!debug_skip(a, 8)) {  // (tuning build: bit 8 = do not read the A / B rows back from the ring: WRONG results)
                    lds_pixel<D>(rbase + u * SLOT_BYTES + VEC_BYTES, X);
                    if (DIAG || !UP) lds_pixel<D>(rbase + u * SLOT_BYTES + 2 * VEC_BYTES, Y);
                } else {
#pragma unroll
                    for (int k = 0; k < E; ++k) X[k] = Y[k] = C[k];
                }
                if (no_out) {
                } else if (!UP) {
                    float A[E];
#pragma unroll
                    for (int k = 0; k < E; ++k) A[k] = AR::add(AR::add(X[k], Y[k]), V[k]);  // (L1 + L2) + L3
                    stg_vec<D>(oa, A);
                    if (DIAG) {
                        float B[E];
#pragma unroll
                        for (int k = 0; k < E; ++k) B[k] = AR::add(F[k], G[k]);  // L5 + L6 (float add commutes)
                        stg_vec<D>(ob, B);
                        ob += rstep;
                    }
                    oa += rstep;
                } else {
                    float S[E];
                    if (!DIAG) {
#pragma unroll
                        for (int k = 0; k < E; ++k) S[k] = AR::add(X[k], V[k]);
                    } else if (!swap) {
#pragma unroll
                        for (int k = 0; k < E; ++k)  // (A + L4) + ((B + L7) + L8), src/SGM.cpp:386-390
                            S[k] = AR::add(AR::add(X[k], V[k]), AR::add(AR::add(Y[k], F[k]), G[k]));
                    } else {
#pragma unroll
                        for (int k = 0; k < E; ++k)
                            S[k] = AR::add(AR::add(X[k], V[k]), AR::add(AR::add(Y[k], G[k]), F[k]));
                    }
                    if (KEEP_S) {
                        stg_vec<D>(os, S);
                        os += rstep;
                    }
                    // selection: see select_pending (runs inside the next row's step); here only the rare flush of 32 rows
                    if (r > 0 && ((r - 1) & 31) == 31) flush_rows(r - 1);
#pragma unroll
                    for (int k = 0; k < E; ++k) Sp[k] = S[k];
                }
                if (r + PF < h) fetch(u);  // slot u is free: this lane has read its own bytes back
                asm volatile("cp.async.commit_group;" ::: "memory");
            }
        }
    }
    if (EDGE && debug_skip(a, 128) && g_global && lane == 0 && (blk == 1 || blk == nblk / 2) && blockIdx.y == 0)
        printf("sweep %s blk %d col %d: neighbour vector found early %d, at the second look %d, by polling %d\n", UP ? "up" : "down", blk, wid,
               n_early, n_peek, n_poll);
    if (UP && !no_out) {  // drain the pipeline: the last row's selection and the rows still stashed
        select_pending(h - 1);
        flush_rows(h - 1);
    }
}

// Helper warp of a block-edge column (which = 0: the block's first column, 1: its last).  The edge warps are the critical path
// of the block (see sweep_column_of_warp); everything of their pixel that the NEXT row does not wait for is done here
// instead: the row's A / B vectors come through this warp's own cp.async ring, the three path vectors of the row from shared
// memory (F and G are published there for the neighbouring columns anyway, V through `vx`), then the same sums, selection
// and stores as sweep_body, in the same order.  The helper takes part in the row barrier: its arrival for row r + 1 says
// "row r has been read", which is what allows row r + 2 to overwrite the same parity.
template <int D, int NW, int MODE, int PF, class AR>
__device__ __forceinline__ void sweep_helper_body(const SweepArgs& a, const float* xch, const float* vx, unsigned bar, float* sring,
                                                  int which, int nact) {
    constexpr bool UP = MODE != SWEEP_DOWN, KEEP_S = MODE == SWEEP_UP_KEEP;
    using LL = LaneLayout<D>;
    constexpr int E = LL::E;
    static_assert(PF % 2 == 0, "PF must be even");
    constexpr unsigned VEC_BYTES = D * 4, WARP_BYTES = 2 * VEC_BYTES, PAR_BYTES = NW * WARP_BYTES, SLOT_BYTES = 3 * VEC_BYTES;
    const int lane = threadIdx.x & 31;
    const int blk = blockIdx.x;
    const int view = a.view0 + blockIdx.y;
    const int h = a.h, w = a.w;
    const int col = which ? nact - 1 : 0;
    const int j = blk * a.cols + col;
    const bool swap = col == 0 && blk > 0;  // as in sweep_body: slot 0 / 1 of the column hold F / G = Pr / Pl instead of Pl / Pr
    const int f_id = swap ? 1 : 0, g_id = swap ? 0 : 1;
    const unsigned lane_off = lane * LL::CW * 4;
    const unsigned f_src = smem_u32(xch) + col * WARP_BYTES + f_id * VEC_BYTES + lane_off;
    const unsigned g_src = smem_u32(xch) + col * WARP_BYTES + g_id * VEC_BYTES + lane_off;
    const unsigned v_src = smem_u32(vx) + which * VEC_BYTES + lane_off;

    const ptrdiff_t rstep = UP ? -(ptrdiff_t)w * D : (ptrdiff_t)w * D;
    const size_t first = ((size_t)(UP ? h - 1 : 0) * w + j) * D + lane * LL::CW;
    const float* px = a.acc_a + view * a.vol_stride + first;
    const float* py = a.acc_b + view * a.vol_stride + first;
    float* oa = a.acc_a + view * a.vol_stride + first;
    float* ob = !UP ? a.acc_b + view * a.vol_stride + first : nullptr;
    float* os = KEEP_S ? a.pathsum + view * a.vol_stride + first : nullptr;
    const ptrdiff_t pstep = UP ? -(ptrdiff_t)w : (ptrdiff_t)w;
    const size_t opix0 = view * a.img_stride + (size_t)(UP ? h - 1 : 0) * w + j;
    WtaPick stash = {};
    auto flush_rows = [&](int rlast) {  // as in sweep_body
        const int rr = (rlast & ~31) + lane;
        if (rr <= rlast) {
            int dd;
            float sub;
            wta_finish<AR::DISP>(stash, a.ratio, a.min_gap, true, dd, sub);
            const size_t o = opix0 + (ptrdiff_t)rr * pstep;
            a.disp[o] = (uint16_t)dd;
            a.subpixel[o] = sub;
        }
    };
    const unsigned rbase = smem_u32(sring) + (unsigned)(NW + which) * PF * SLOT_BYTES + lane_off;  // the rings after the columns'
    auto fetch = [&](int slot) {
#pragma unroll
        for (int q = 0; q < LL::NCH; ++q) {
            const unsigned dst = rbase + slot * SLOT_BYTES + q * LL::CHUNK_SPAN * 4;
            cp_async_lane<LL::CW * 4>(dst + VEC_BYTES, px + q * LL::CHUNK_SPAN);
            cp_async_lane<LL::CW * 4>(dst + 2 * VEC_BYTES, py + q * LL::CHUNK_SPAN);
        }
        px += rstep;
        py += rstep;
    };
#pragma unroll
    for (int u = 0; u < PF; ++u) {
        if (u < h) fetch(u);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    if (lane == 0) mbar_arrive(bar);  // row 0: nothing read yet
    for (int r0 = 0; r0 < h; r0 += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int r = r0 + u;
            if (r < h) {
                const unsigned par = u & 1;  // == r & 1
                float F[E], G[E], V[E], X[E], Y[E];
                asm volatile("cp.async.wait_group %0;" ::"n"(PF - 1) : "memory");
                mbar_wait(bar, par, a.error);  // every warp's row r vectors are in shared memory
                lds_pixel<D>(f_src + par * PAR_BYTES, F);
                lds_pixel<D>(g_src + par * PAR_BYTES, G);
                lds_pixel<D>(v_src + par * 2 * VEC_BYTES, V);
                lds_pixel<D>(rbase + u * SLOT_BYTES + VEC_BYTES, X);
                lds_pixel<D>(rbase + u * SLOT_BYTES + 2 * VEC_BYTES, Y);
                if (r + 1 < h) {  // arrival for row r + 1: row r has been read (the loads above are complete: F feeds the fence)
                    float keep = F[0] + G[0] + V[0];
                    asm volatile("" ::"f"(keep) : "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar);
                }
                if (!UP) {
                    float A[E], B[E];
#pragma unroll
                    for (int k = 0; k < E; ++k) A[k] = AR::add(AR::add(X[k], Y[k]), V[k]);  // (L1 + L2) + L3
                    stg_vec<D>(oa, A);
#pragma unroll
                    for (int k = 0; k < E; ++k) B[k] = AR::add(F[k], G[k]);  // L5 + L6
                    stg_vec<D>(ob, B);
                    oa += rstep;
                    ob += rstep;
                } else {
                    float S[E];
                    if (!swap) {
#pragma unroll
                        for (int k = 0; k < E; ++k)  // (A + L4) + ((B + L7) + L8), src/SGM.cpp:386-390
                            S[k] = AR::add(AR::add(X[k], V[k]), AR::add(AR::add(Y[k], F[k]), G[k]));
                    } else {
#pragma unroll
                        for (int k = 0; k < E; ++k)
                            S[k] = AR::add(AR::add(X[k], V[k]), AR::add(AR::add(Y[k], G[k]), F[k]));
                    }
                    if (KEEP_S) {
                        stg_vec<D>(os, S);
                        os += rstep;
                    }
                    const WtaPick pk = AR::reduce(S, lane);
                    if (lane == (r & 31)) stash = pk;
                    if ((r & 31) == 31) flush_rows(r);
                }
                if (r + PF < h) fetch(u);
                asm volatile("cp.async.commit_group;" ::: "memory");
            }
        }
    }
    if (UP && ((h - 1) & 31) != 31) flush_rows(h - 1);
}

// floats of the diagonal exchange buffer at the start of the dynamic shared memory (53 KB for 13 columns at D = 256: more than
// static shared memory may hold)
template <int D, int NW, int PATHS>
__host__ __device__ constexpr int sweep_xch_floats() {
    return PATHS == 8 ? 2 * NW * 2 * D : 0;
}

template <int D, int NW, int MODE, int PATHS, int PF, int MINB, class AR, int HWW = NW, bool HELP = false>
__global__ void __launch_bounds__(HWW * 32, MINB) sgm_sweep_kernel(const SweepArgs a) {
    constexpr bool DIAG = PATHS == 8;
    static_assert(!HELP || DIAG, "helper warps exist for the 8-path sweeps only");
    __shared__ float xmin[DIAG ? 2 * NW * 2 : 1];                   // [parity][warp][0 = Pl, 1 = Pr] min of the vector
    __shared__ __align__(16) float vx[HELP ? 2 * 2 * D : 4];        // [parity][edge column][d]: vertical path, for the helper warps
    __shared__ __align__(8) unsigned long long bar_mem;
    extern __shared__ __align__(16) float dyn_smem[];
    float* xch = dyn_smem;                                 // [parity][warp][0 = Pl, 1 = Pr][d] (8 paths only)
    float* sring = dyn_smem + sweep_xch_floats<D, NW, PATHS>();  // [NW (+ 2 helpers)][PF][3][D]
    const int blk = blockIdx.x;
    const int nact = sweep_block_columns(a, blk, gridDim.x);  // columns (= warps) of this block
    const bool helpers = sweep_block_has_helpers<HELP>(nact, debug_skip(a, 16));
    const unsigned bar = smem_u32(&bar_mem);
    if (DIAG) {
        if (threadIdx.x == 0) mbar_init(bar, nact + (helpers ? 2 : 0));  // one arrival per active warp and row
        __syncthreads();
    }
    const int wid = sweep_column_of_warp<NW, HWW, HELP>(threadIdx.x >> 5, nact, debug_skip(a, 16));
    if (wid < 0) return;  // a spare hardware warp, or a column beyond the image
    if constexpr (HELP) {
        if (wid >= NW) {
            sweep_helper_body<D, NW, MODE, PF, AR>(a, xch, vx, bar, sring, wid - NW, nact);
            return;
        }
    }
    const int j = blk * a.cols + wid;
    // block-edge warps (they exchange with a neighbouring block) and image-border columns take the general body
    const bool edge = DIAG && (wid == 0 || wid == nact - 1 || j == 0 || j == a.w - 1);
    if (HELP && edge && helpers)
        sweep_body<D, NW, MODE, PATHS, PF, true, AR, HELP>(a, xch, xmin, bar, sring, wid, vx);
    else if (edge)
        sweep_body<D, NW, MODE, PATHS, PF, true, AR>(a, xch, xmin, bar, sring, wid);
    else
        sweep_body<D, NW, MODE, PATHS, PF, false, AR>(a, xch, xmin, bar, sring, wid);
}

constexpr int SWEEP_NW = 8;  // smallest block width in columns (sizes the halo ring)

// Launch shape of the sweeps: NW columns (warps) per block, PF rows in flight per warp, MINB blocks per SM.
template <int NW_, int PF_, int MINB_, int HWW_ = NW_, bool HELP_ = false>
struct SweepCfg {
    static constexpr int NW = NW_, PF = PF_, MINB = MINB_, HWW = HWW_;  // HWW: hardware warps per block (>= NW: spare warps idle)
    static constexpr bool HELP = HELP_;  // two of the spare warps do the output work of the block-edge columns (8 paths)
};

template <int D, class CFG, int PATHS>
constexpr size_t sweep_smem_bytes() {  // dynamic shared memory of a sweep block: exchange buffer + the warps' rings
    return ((size_t)sweep_xch_floats<D, CFG::NW, PATHS>() + (size_t)(CFG::NW + (CFG::HELP ? 2 : 0)) * CFG::PF * 3 * D) * sizeof(float);
}

template <int D, int MODE, int PATHS, class CFG, class AR>
cudaError_t launch_sweep(const SweepArgs& sa, int nblk, int views, bool cooperative, cudaStream_t st) {
    auto kern = sgm_sweep_kernel<D, CFG::NW, MODE, PATHS, CFG::PF, CFG::MINB, AR, CFG::HWW, CFG::HELP>;
    constexpr size_t ring_bytes = sweep_smem_bytes<D, CFG, PATHS>();
    cudaError_t attr = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ring_bytes);  // per device: cheap, so always
    if (attr != cudaSuccess) return attr;
    if (!cooperative) {
        kern<<<dim3(nblk, views), CFG::HWW * 32, ring_bytes, st>>>(sa);
        return cudaGetLastError();
    }
    void* params[] = {const_cast<SweepArgs*>(&sa)};
    return cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kern), dim3(nblk, views), dim3(CFG::HWW * 32), params, ring_bytes, st);
}

// blocks per image row for `cols` columns per block in a kernel that can take `capacity` columns (see fused_impl)
inline int sweep_blocks(int w, int cols, int capacity) {
    int nblk = (w + cols - 1) / cols;
    if (nblk > 1 && w - (nblk - 1) * cols + cols <= capacity) --nblk;  // the last block absorbs the remainder
    return nblk;
}

template <int D, int PATHS, class CFG, class AR>
int sweep_capacity(int device) {
    static int cached[16] = {};  // per device; the answer only depends on the kernel and the device
    if (device >= 0 && device < 16 && cached[device]) return cached[device] - 1;
    int per_sm[3] = {0, 0, 0}, sms = 0, coop = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device);
    // the ring may exceed the default 48 KB of dynamic shared memory: opt in before asking for the occupancy
    auto ask = [&](auto kern, int& out) {
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sweep_smem_bytes<D, CFG, PATHS>());
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&out, kern, CFG::HWW * 32, sweep_smem_bytes<D, CFG, PATHS>());
    };
    ask(sgm_sweep_kernel<D, CFG::NW, SWEEP_DOWN, PATHS, CFG::PF, CFG::MINB, AR, CFG::HWW, CFG::HELP>, per_sm[0]);
    ask(sgm_sweep_kernel<D, CFG::NW, SWEEP_UP, PATHS, CFG::PF, CFG::MINB, AR, CFG::HWW, CFG::HELP>, per_sm[1]);
    ask(sgm_sweep_kernel<D, CFG::NW, SWEEP_UP_KEEP, PATHS, CFG::PF, CFG::MINB, AR, CFG::HWW, CFG::HELP>, per_sm[2]);
    const int least = per_sm[0] < per_sm[1] ? (per_sm[0] < per_sm[2] ? per_sm[0] : per_sm[2]) : (per_sm[1] < per_sm[2] ? per_sm[1] : per_sm[2]);
    const int cap = coop ? sms * least : 0;
    if (device >= 0 && device < 16) cached[device] = cap + 1;
    return cap;
}

// D = 32-bit words per pixel: the number of disparities for F32Arith, half of it for S16x2Arith (a.d stays the number
// of disparities).
// `cols`: columns per block (default: the kernel's capacity CFG::NW).  With fewer, a remainder that still fits the capacity is
// folded into the last block: 1242 columns = 72 blocks of 17 + one of 18 per view instead of 73 + a block of ONE column, i.e.
// 146 blocks on the 148 SMs of a B200, so a long single-block kernel of another frame in flight (the median) finds an SM
// without holding up the whole cooperative grid.
template <int D, int PATHS, class CFG, class AR = F32Arith<D>>
bool fused_impl(const AggregateArgs& a, cudaStream_t st, LaunchCounter& lc, int cols = CFG::NW) {
    const int nblk = sweep_blocks(a.w, cols, CFG::NW);
    int views_per_launch = a.views;
    if (PATHS == 8) {
        int device = 0;
        cudaGetDevice(&device);
        const int cap = sweep_capacity<D, PATHS, CFG, AR>(device);
        bool split = false;  // one view per launch: taken when both views do not fit a resident grid (1080p); forced in tuning builds
#ifdef SGM_B200_TUNING
        split = std::getenv("SGM_B200_SWEEP_SPLIT_VIEWS") != nullptr;
#endif
        if (nblk * a.views > cap || split) views_per_launch = 1;
        if (nblk > cap) return false;  // image too wide for a resident row-synchronous sweep: caller falls back to line walkers
    }
    const size_t vol = (size_t)a.h * a.w * D;
    {
        constexpr int PFH = (LaneLayout<D>::E >= 8) ? 8 : 16, HW = 4;
        constexpr size_t ring_bytes = (size_t)HW * PFH * D * sizeof(float);  // 32 KB for D >= 128
        sgm_hpair_async_kernel<D, PFH, HW, AR><<<dim3((a.h + HW - 1) / HW, a.views, 2), HW * 32, ring_bytes, st>>>(
                a.cost, a.acc_a, a.acc_b, a.h, a.w, a.P1, a.P2, vol);
        lc.note("sgm_hpair_kernel");
    }

    SweepArgs sa;
    sa.cost = a.cost;
    sa.acc_a = a.acc_a;
    sa.acc_b = a.acc_b;
    sa.pathsum = a.pathsum;
    sa.disp = a.disp;
    sa.subpixel = a.subpixel;
    sa.halo = a.halo;
    sa.error = a.error;
    sa.debug = 0;
#ifdef SGM_B200_TUNING
    {
        static const int dbg = [] { const char* e = std::getenv("SGM_B200_SWEEP_DEBUG"); return e ? std::atoi(e) : 0; }();
        sa.debug = dbg;
    }
#endif
    sa.cols = cols;
    sa.h = a.h;
    sa.w = a.w;
    sa.P1 = a.P1;
    sa.P2 = a.P2;
    sa.ratio = a.unique_ratio;
    sa.min_gap = a.min_gap;
    sa.vol_stride = vol;
    sa.img_stride = (size_t)a.h * a.w;
    bool cooperative = PATHS == 8;
#ifdef SGM_B200_TUNING
    // experiment: plain launches (blocks start as SMs free up instead of all at once).  NOT safe with several frames in flight:
    // two half-resident sweep grids wait for each other until the time-out.
    static const bool plain = std::getenv("SGM_B200_SWEEP_PLAIN_LAUNCH") != nullptr;
    if (plain) cooperative = false;
#endif
    for (int up = 0; up < 2; ++up) {
        // every sweep launch gets its own tag range; after 2^20 launches the ring is wiped and the sequence restarts
        if (*a.launch_seq >= (1u << 20) - 1u) {
            cudaMemsetAsync(a.halo, 0, fused_halo_bytes(a.views, a.w, a.d), st);
            *a.launch_seq = 0;
        }
        sa.tag_base = ++(*a.launch_seq) << 12;
        for (int v0 = 0; v0 < a.views; v0 += views_per_launch) {
            sa.view0 = v0;
            cudaError_t e = !up         ? launch_sweep<D, SWEEP_DOWN, PATHS, CFG, AR>(sa, nblk, views_per_launch, cooperative, st)
                            : a.pathsum ? launch_sweep<D, SWEEP_UP_KEEP, PATHS, CFG, AR>(sa, nblk, views_per_launch, cooperative, st)
                                        : launch_sweep<D, SWEEP_UP, PATHS, CFG, AR>(sa, nblk, views_per_launch, cooperative, st);
            if (e != cudaSuccess && lc.err == cudaSuccess) {
                lc.err = e;
                lc.failed = "sgm_sweep_kernel (launch)";
            }
            lc.note("sgm_sweep_kernel");
        }
    }
    return true;
}

// Default shapes per D; SGM_B200_SWEEP_CFG=<n> selects an alternative for D = 128, 8 paths (tuning knob).
template <int D>
struct DefaultCfg {
    using type = SweepCfg<8, (LaneLayout<D>::E >= 4) ? 2 : 4, (LaneLayout<D>::E >= 8) ? 2 : 3>;
};
template <>
struct DefaultCfg<128> {
    using type = SweepCfg<10, 4, 2>;  // 1242x375: 250 blocks of 10 columns, 2 per SM
};

// Automatic launch shape.  Rows advance in lock-step, so the time of a frame's sweep goes with
//     (launches: 1, or 2 when both views do not fit a resident grid) x (columns on the most crowded SM).
// The default shapes (8 or 10 columns, 2 - 3 blocks per SM) are right for narrow images; when they would stack blocks on
// some SMs, ONE wide block per SM is better:
//   17 columns per block in the 19-column kernel (19 hardware warps): the two block-edge warps have a warp scheduler to
//           themselves (sweep_column_of_warp), and a remainder of one or two columns is folded into the last block of the
//           row.  KITTI width: 2 x 73 blocks on the 148 SMs of a B200.  1242x375 D=128: 1.22 -> 1.12 ms.
//   Wide19: 19 columns per block, for widths whose 17-column grid does not fit the device (the edge warps share their
//           scheduler with two interior warps; 1242x375 D=128: 1.20 ms).
//   Wide13: 13 columns on 15 hardware warps: 1920 columns = 148 blocks, one view per launch.  The only wide shape for 8
//           values per lane (D = 256): up to 16 warps a block may use 128 registers, with 17 - 20 only 96 (five warps on one
//           scheduler's 16 K registers), which that body does not fit.
// All of them keep two rows in flight per warp (four cost more spill traffic in the up sweep than they hide latency); 21
// hardware warps would cap the kernels at 80 registers (1242x375 D=128: 1.25 ms).
using Wide13 = SweepCfg<13, 2, 1, 15>;
using Wide19 = SweepCfg<19, 2, 1, 19>;
template <class Def, bool UP_TO_4_PER_LANE>
int wide_shape(int views, int w) {
    int device = 0, sms = 0;
    cudaGetDevice(&device);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    auto cost = [&](int nw, int per_sm, int capacity) {  // 0: the grid cannot be resident
        const int nblk = sweep_blocks(w, nw, capacity), cap = sms * per_sm;
        if (nblk > cap) return 0;
        const int launches = views * nblk > cap ? views : 1;
        const int per_launch = launches == 1 ? views * nblk : nblk;
        return launches * nw * ((per_launch + sms - 1) / sms);
    };
    int best = 0, best_cost = cost(Def::NW, Def::MINB, Def::NW);
    const int candidates[3] = {17, 19, 13};
    for (int nw : candidates) {
        if (!UP_TO_4_PER_LANE && nw != 13) continue;
        const int c = cost(nw, 1, nw == 13 ? Wide13::NW : Wide19::NW);
        if (c && (!best_cost || c < best_cost)) {
            best = nw;
            best_cost = c;
        }
    }
    return best;
}

template <int D>
bool fused_dispatch(const AggregateArgs& a, cudaStream_t st, LaunchCounter& lc) {
    if (a.paths != 8) return fused_impl<D, 4, typename DefaultCfg<D>::type>(a, st, lc);
#ifdef SGM_B200_TUNING
    if constexpr (D == 256) {
        static const bool def256 = std::getenv("SGM_B200_SWEEP_CFG") && std::atoi(std::getenv("SGM_B200_SWEEP_CFG")) == 13;
        if (def256) return fused_impl<D, 8, typename DefaultCfg<D>::type>(a, st, lc);
    }
    if constexpr (D == 64) {  // a short list for D = 64 (every entry costs three kernels of compile time)
        static const int cfg64 = [] {
            const char* e = std::getenv("SGM_B200_SWEEP_CFG");
            return e ? std::atoi(e) : 0;
        }();
        switch (cfg64) {
            case 16: return fused_impl<D, 8, SweepCfg<19, 2, 1>>(a, st, lc);
            case 30: return fused_impl<D, 8, SweepCfg<17, 2, 1, 19, true>>(a, st, lc);
            default: break;
        }
    }
    if constexpr (D == 128) {
        static const int cfg = [] {
            const char* e = std::getenv("SGM_B200_SWEEP_CFG");
            return e ? std::atoi(e) : 0;
        }();
        switch (cfg) {
            case 12: return fused_impl<D, 8, SweepCfg<19, 4, 1>>(a, st, lc);
            case 13: return fused_impl<D, 8, typename DefaultCfg<D>::type>(a, st, lc);
            case 16: return fused_impl<D, 8, SweepCfg<19, 2, 1>>(a, st, lc);
            case 24: return fused_impl<D, 8, SweepCfg<17, 2, 1, 19>>(a, st, lc);        // 17-column kernel: the remainder gets a block of its own
            case 50: {  // the 19-column kernel with SGM_B200_SWEEP_COLS columns per block
                static const int cols = [] { const char* e = std::getenv("SGM_B200_SWEEP_COLS"); return e ? std::atoi(e) : 17; }();
                return fused_impl<D, 8, Wide19>(a, st, lc, cols < 13 ? 13 : (cols > 19 ? 19 : cols));
            }
            case 21: return fused_impl<D, 8, SweepCfg<17, 4, 1, 19>>(a, st, lc);        // schedulers host 5 5 5 | 2 edge warps
            case 26: return fused_impl<D, 8, SweepCfg<17, 2, 1, 18>>(a, st, lc);        // 5 5 4 | 2 + 1
            case 27: return fused_impl<D, 8, SweepCfg<17, 2, 1, 17>>(a, st, lc);        // 5 4 4 | 2 + 2
            case 30: return fused_impl<D, 8, SweepCfg<17, 2, 1, 19, true>>(a, st, lc);  // helper warps: 5 5 5 | 2 edge + 2 helpers
            case 33: return fused_impl<D, 8, SweepCfg<18, 2, 1, 20, true>>(a, st, lc);  // 5 5 5 | 2 + 2 + 1
            default: break;
        }
    }
#endif
    constexpr bool NARROW = LaneLayout<D>::E <= 4;
    if constexpr (LaneLayout<D>::E <= 8) {
        switch (wide_shape<typename DefaultCfg<D>::type, NARROW>(a.views, a.w)) {
            case 13: return fused_impl<D, 8, Wide13>(a, st, lc);
            case 17:
                if constexpr (NARROW) return fused_impl<D, 8, Wide19>(a, st, lc, 17);
                break;
            case 19:
                if constexpr (NARROW) return fused_impl<D, 8, Wide19>(a, st, lc);
                break;
            default: break;
        }
    }
    return fused_impl<D, 8, typename DefaultCfg<D>::type>(a, st, lc);
}

// Integer mode (S16x2Arith): DW = a.d / 2 words per pixel; the penalties travel as packed s16x2 bit patterns.
template <int DW>
bool fused_dispatch_packed(const AggregateArgs& a0, cudaStream_t st, LaunchCounter& lc) {
    AggregateArgs a = a0;
    const unsigned p1 = (unsigned)a0.P1, p2m = (unsigned)(a0.P2 - a0.P1);
    const unsigned P1w = p1 * 0x10001u, P2w = p2m * 0x10001u;
    static_assert(sizeof(float) == sizeof(unsigned), "carrier");
    memcpy(&a.P1, &P1w, sizeof(float));
    memcpy(&a.P2, &P2w, sizeof(float));
    using AR = S16x2Arith<DW>;
    if (a.paths != 8) return fused_impl<DW, 4, typename DefaultCfg<DW>::type, AR>(a, st, lc);
#ifdef SGM_B200_TUNING
    static const int cfg = [] {
        const char* e = std::getenv("SGM_B200_ISWEEP_CFG");
        return e ? std::atoi(e) : 0;
    }();
    switch (cfg) {
        case 1: return fused_impl<DW, 8, SweepCfg<10, 4, 2>, AR>(a, st, lc);
        case 3: return fused_impl<DW, 8, SweepCfg<19, 4, 1>, AR>(a, st, lc);
        case 11: return fused_impl<DW, 8, SweepCfg<19, 2, 1, 19>, AR>(a, st, lc);
        case 13: return fused_impl<DW, 8, SweepCfg<17, 2, 1, 19, true>, AR>(a, st, lc);
        default: break;
    }
#endif
    using Def = typename DefaultCfg<DW>::type;
    constexpr bool NARROW = LaneLayout<DW>::E <= 4;
    if constexpr (LaneLayout<DW>::E <= 8) {
        switch (wide_shape<Def, NARROW>(a.views, a.w)) {
            case 13: return fused_impl<DW, 8, Wide13, AR>(a, st, lc);
            case 17:
                if constexpr (NARROW) return fused_impl<DW, 8, Wide19, AR>(a, st, lc, 17);
                break;
            case 19:
                if constexpr (NARROW) return fused_impl<DW, 8, Wide19, AR>(a, st, lc);
                break;
            default: break;
        }
    }
    return fused_impl<DW, 8, Def, AR>(a, st, lc);
}

}  // namespace

size_t fused_halo_bytes(int views, int w, int d) {
    const size_t words = (size_t)(d / 32) * 32;  // HaloIO<D>::SLOT_WORDS
    return (size_t)views * ((w + SWEEP_NW - 1) / SWEEP_NW) * 2 * HALO_RING * words * sizeof(unsigned long long);
}

bool launch_aggregate_fused(const AggregateArgs& a, cudaStream_t st, LaunchCounter& lc) {
    if (!a.halo || !a.launch_seq || !a.error || !a.acc_b || a.h > 4000) return false;
    if (a.packed16) {
        switch (a.d) {
            case 64: return fused_dispatch_packed<32>(a, st, lc);
            case 128: return fused_dispatch_packed<64>(a, st, lc);
            case 256: return fused_dispatch_packed<128>(a, st, lc);
            default: return false;
        }
    }
    switch (a.d) {
        case 32: return fused_dispatch<32>(a, st, lc);
        case 64: return fused_dispatch<64>(a, st, lc);
        case 128: return fused_dispatch<128>(a, st, lc);
        case 256: return fused_dispatch<256>(a, st, lc);
        default: return false;
    }
}

}  // namespace sgmb200
