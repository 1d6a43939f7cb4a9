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
// guarded by per-block progress counters (release / acquire).  A block only ever waits for its two neighbours'
// PREVIOUS row, so all blocks advance within one row of their neighbours; the kernel is launched cooperatively so
// that every block is resident, and every spin has a time-out that raises an error flag instead of hanging.
#include "common.cuh"
#include "kernels.h"
#include "sgm_device.cuh"

namespace sgmb200 {

namespace {

constexpr int HALO_RING = 4;                       // rows of halo kept per block edge (2 would suffice)
constexpr long long SPIN_TIMEOUT_CYCLES = 4000000000LL;  // ~2 s at 2 GHz

struct SweepArgs {
    const float* cost;
    float* acc_a;
    float* acc_b;
    float* pathsum;
    uint16_t* disp;
    float* subpixel;
    float* halo;     // [views][nblk][2][HALO_RING][D]
    int* progress;   // [views][nblk] rows published by each block
    int* error;      // set to 1 when a spin timed out
    int h, w;
    int view0;       // first view handled by this launch (blockIdx.y is added)
    float P1, P2, ratio;
    int min_gap;
    size_t vol_stride, img_stride;
};

__device__ __forceinline__ int ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <int D>
__device__ __forceinline__ void load_pixel_cg(const float* pix, int lane, float (&v)[LaneLayout<D>::E]) {
    using LL = LaneLayout<D>;
#pragma unroll
    for (int q = 0; q < LL::NCH; ++q) {
        const float* p = pix + q * LL::CHUNK_SPAN + lane * LL::CW;
        if constexpr (LL::CW == 4) {
            const float4 t = __ldcg(reinterpret_cast<const float4*>(p));
            v[q * 4 + 0] = t.x;
            v[q * 4 + 1] = t.y;
            v[q * 4 + 2] = t.z;
            v[q * 4 + 3] = t.w;
        } else if constexpr (LL::CW == 2) {
            const float2 t = __ldcg(reinterpret_cast<const float2*>(p));
            v[q * 2 + 0] = t.x;
            v[q * 2 + 1] = t.y;
        } else {
            v[q] = __ldcg(p);
        }
    }
}

template <int E>
__device__ __forceinline__ float vec_min(const float (&v)[E]) {
    float m = v[0];
#pragma unroll
    for (int k = 1; k < E; ++k) m = fminf(m, v[k]);
    return m;
}

// ---------------------------------------------------------------------------------------------
// L1 (left -> right, src/SGM.cpp:82-119) and L2 (right -> left, :122-159): one warp per (row, direction, view).
// Only 2*H*views independent lines exist, so each warp keeps PF pixels of the (contiguous) row in flight in
// registers to cover HBM latency.
template <int D, int PF>
__global__ void __launch_bounds__(128) sgm_hpair_kernel(const float* __restrict__ cost, float* __restrict__ out_fwd,
                                                        float* __restrict__ out_bwd, int h, int w, float P1, float P2,
                                                        size_t vol_stride) {
    using LL = LaneLayout<D>;
    constexpr int E = LL::E;
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (row >= h) return;
    const int view = blockIdx.y;
    const bool bwd = blockIdx.z != 0;
    const size_t base = view * vol_stride + (size_t)row * w * D;
    const float* c = cost + base;
    float* o = (bwd ? out_bwd : out_fwd) + base;
    const int j0 = bwd ? w - 1 : 0, dj = bwd ? -1 : 1;

    float Cb[PF][E];
#pragma unroll
    for (int u = 0; u < PF; ++u)
        if (u < w) load_pixel_ro<D>(c + (size_t)(j0 + u * dj) * D, lane, Cb[u]);

    float L[E];
    float m = 0.f;
    for (int s0 = 0; s0 < w; s0 += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int s = s0 + u;
            if (s < w) {
                const int j = j0 + s * dj;
                if (s == 0) {
#pragma unroll
                    for (int k = 0; k < E; ++k) L[k] = Cb[u][k];
                    m = warp_min_f32(vec_min<E>(L));
                } else {
                    path_step<D>(L, Cb[u], m, lane, P1, P2);
                }
                store_pixel<D>(o + (size_t)j * D, lane, L);
                if (s + PF < w) load_pixel_ro<D>(c + (size_t)(j + PF * dj) * D, lane, Cb[u]);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Row-synchronous sweep.  UP == false: rows 0 -> h-1 with L3 (src/SGM.cpp:162-199), L5 and L6 (:247-305).
//                         UP == true : rows h-1 -> 0 with L4 (:202-239), L7 and L8 (:311-369) + sum/WTA (:373-418).
// Warp w of block b owns column j = b*NW + w.  "dl" is the diagonal whose predecessor is the LEFT neighbour column
// (L5 going down, L7 going up), "dr" the one whose predecessor is the RIGHT neighbour column (L6 / L8).  After a row
// every warp publishes its dl ("to the right") and dr ("to the left") vectors; the next row's neighbours pick them up.
template <int D, int NW, bool UP, int PATHS, int PF>
__global__ void __launch_bounds__(NW * 32, (LaneLayout<D>::E >= 8) ? 2 : 3) sgm_sweep_kernel(const SweepArgs a) {
    using LL = LaneLayout<D>;
    constexpr int E = LL::E;
    constexpr bool DIAG = PATHS == 8;
    __shared__ __align__(16) float xch[DIAG ? 2 : 1][DIAG ? NW : 1][2][DIAG ? D : 4];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int blk = blockIdx.x, nblk = gridDim.x;
    const int view = a.view0 + blockIdx.y;
    const int h = a.h, w = a.w;
    const int j = blk * NW + wid;
    const bool active = j < w;
    const float* __restrict__ cost = a.cost + view * a.vol_stride;
    float* acc_a = a.acc_a + view * a.vol_stride;
    float* acc_b = (DIAG || !UP) ? a.acc_b + view * a.vol_stride : nullptr;
    // halo ring slots: [view][blk][which][HALO_RING][D]; which 0 = "to the right" (read by blk+1), 1 = "to the left" (read by blk-1)
    float* halo_mine = DIAG ? a.halo + ((size_t)(view * nblk + blk) * 2) * HALO_RING * D : nullptr;
    const float* halo_left = DIAG && blk > 0 ? a.halo + ((size_t)(view * nblk + blk - 1) * 2 + 0) * HALO_RING * D : nullptr;
    const float* halo_right = DIAG && blk + 1 < nblk ? a.halo + ((size_t)(view * nblk + blk + 1) * 2 + 1) * HALO_RING * D : nullptr;
    int* progress = a.progress + view * nblk;

    const size_t row_stride = (size_t)w * D;
    const size_t col_off = (size_t)(active ? j : 0) * D;
    auto row_of = [&](int r) { return UP ? h - 1 - r : r; };

    float Cb[PF][E], Xb[PF][E], Yb[PF][E];
    auto issue = [&](int r, float (&c)[E], float (&x)[E], float (&y)[E]) {
        const size_t off = (size_t)row_of(r) * row_stride + col_off;
        load_pixel_ro<D>(cost + off, lane, c);
        load_pixel<D>(acc_a + off, lane, x);
        if (DIAG || !UP) load_pixel<D>(acc_b + off, lane, y);
    };
    if (active) {
#pragma unroll
        for (int u = 0; u < PF; ++u)
            if (u < h) issue(u, Cb[u], Xb[u], Yb[u]);
    }

    float V[E];
    float mV = 0.f;
    bool dead = false;  // set when a spin timed out: keep hitting the barriers, stop waiting
    for (int r0 = 0; r0 < h; r0 += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int r = r0 + u;
            if (r < h) {
                const int par = r & 1;
                float Pl[E], Pr[E];
                if (active) {
                    const int i = row_of(r);
                    // ---- predecessors of the two diagonals (previous row of the neighbour columns) ----
                    const bool init_l = !DIAG || r == 0 || j == 0;
                    const bool init_r = !DIAG || r == 0 || j == w - 1;
                    if (DIAG && !init_l) {
                        if (wid > 0) {
                            load_pixel<D>(&xch[par ^ 1][wid - 1][0][0], lane, Pl);
                        } else {
                            if (lane == 0 && !dead) {
                                const long long t0 = clock64();
                                while (ld_acquire(progress + blk - 1) < r) {
                                    if (clock64() - t0 > SPIN_TIMEOUT_CYCLES) {
                                        atomicExch(a.error, 1);
                                        dead = true;
                                        break;
                                    }
                                }
                            }
                            dead = __shfl_sync(FULL_MASK, (int)dead, 0) != 0;
                            load_pixel_cg<D>(halo_left + (size_t)((r - 1) % HALO_RING) * D, lane, Pl);
                        }
                    }
                    if (DIAG && !init_r) {
                        if (wid < NW - 1) {
                            load_pixel<D>(&xch[par ^ 1][wid + 1][1][0], lane, Pr);
                        } else {
                            if (lane == 0 && !dead) {
                                const long long t0 = clock64();
                                while (ld_acquire(progress + blk + 1) < r) {
                                    if (clock64() - t0 > SPIN_TIMEOUT_CYCLES) {
                                        atomicExch(a.error, 1);
                                        dead = true;
                                        break;
                                    }
                                }
                            }
                            dead = __shfl_sync(FULL_MASK, (int)dead, 0) != 0;
                            load_pixel_cg<D>(halo_right + (size_t)((r - 1) % HALO_RING) * D, lane, Pr);
                        }
                    }
                    // ---- the three recurrences ----
                    if (r == 0) {
#pragma unroll
                        for (int k = 0; k < E; ++k) V[k] = Cb[u][k];
                        mV = warp_min_f32(vec_min<E>(V));
                    } else {
                        path_step<D>(V, Cb[u], mV, lane, a.P1, a.P2);
                    }
                    if (DIAG) {
                        if (init_l) {
#pragma unroll
                            for (int k = 0; k < E; ++k) Pl[k] = Cb[u][k];
                        } else {
                            float ml = warp_min_f32(vec_min<E>(Pl));
                            path_step<D>(Pl, Cb[u], ml, lane, a.P1, a.P2);
                        }
                        if (init_r) {
#pragma unroll
                            for (int k = 0; k < E; ++k) Pr[k] = Cb[u][k];
                        } else {
                            float mr = warp_min_f32(vec_min<E>(Pr));
                            path_step<D>(Pr, Cb[u], mr, lane, a.P1, a.P2);
                        }
                        // ---- publish for the next row's neighbours ----
                        store_pixel<D>(&xch[par][wid][0][0], lane, Pl);
                        store_pixel<D>(&xch[par][wid][1][0], lane, Pr);
                        if (wid == NW - 1 && blk + 1 < nblk) store_pixel<D>(halo_mine + (size_t)(0 * HALO_RING + r % HALO_RING) * D, lane, Pl);
                        if (wid == 0 && blk > 0) store_pixel<D>(halo_mine + (size_t)(1 * HALO_RING + r % HALO_RING) * D, lane, Pr);
                    }
                    // ---- outputs ----
                    const size_t pix = (size_t)i * w + j;
                    const size_t off = pix * D;
                    if (!UP) {
                        float A[E];
#pragma unroll
                        for (int k = 0; k < E; ++k) {
                            A[k] = __fadd_rn(__fadd_rn(Xb[u][k], Yb[u][k]), V[k]);  // (L1 + L2) + L3
                        }
                        store_pixel<D>(acc_a + off, lane, A);
                        if (DIAG) {
                            float B[E];
#pragma unroll
                            for (int k = 0; k < E; ++k) B[k] = __fadd_rn(Pl[k], Pr[k]);
                            store_pixel<D>(acc_b + off, lane, B);
                        }
                    } else {
                        float S[E];
#pragma unroll
                        for (int k = 0; k < E; ++k) {
                            const float sa = __fadd_rn(Xb[u][k], V[k]);
                            S[k] = DIAG ? __fadd_rn(sa, __fadd_rn(__fadd_rn(Yb[u][k], Pl[k]), Pr[k])) : sa;
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
                    if (r + PF < h) issue(r + PF, Cb[u], Xb[u], Yb[u]);
                }
                if (DIAG) {
                    __syncthreads();  // this row's xch / halo stores are done block-wide
                    if (threadIdx.x == 0) {
                        __threadfence();
                        st_release(progress + blk, r + 1);
                    }
                }
            }
        }
    }
}

constexpr int SWEEP_NW = 8;

template <int D, bool UP, int PATHS>
cudaError_t launch_sweep(const SweepArgs& sa, int nblk, int views, bool cooperative, cudaStream_t st) {
    constexpr int PF = (LaneLayout<D>::E >= 8) ? 2 : 3;
    auto kern = sgm_sweep_kernel<D, SWEEP_NW, UP, PATHS, PF>;
    if (!cooperative) {
        kern<<<dim3(nblk, views), SWEEP_NW * 32, 0, st>>>(sa);
        return cudaGetLastError();
    }
    void* params[] = {const_cast<SweepArgs*>(&sa)};
    return cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kern), dim3(nblk, views), dim3(SWEEP_NW * 32), params, 0, st);
}

template <int D, int PATHS>
int sweep_capacity(int device) {
    constexpr int PF = (LaneLayout<D>::E >= 8) ? 2 : 3;
    int per_sm_dn = 0, per_sm_up = 0, sms = 0, coop = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_dn, sgm_sweep_kernel<D, SWEEP_NW, false, PATHS, PF>, SWEEP_NW * 32, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_up, sgm_sweep_kernel<D, SWEEP_NW, true, PATHS, PF>, SWEEP_NW * 32, 0);
    if (!coop) return 0;
    return sms * (per_sm_dn < per_sm_up ? per_sm_dn : per_sm_up);
}

template <int D, int PATHS>
bool fused_impl(const AggregateArgs& a, cudaStream_t st, LaunchCounter& lc) {
    const int nblk = (a.w + SWEEP_NW - 1) / SWEEP_NW;
    int views_per_launch = a.views;
    if (PATHS == 8) {
        int device = 0;
        cudaGetDevice(&device);
        const int cap = sweep_capacity<D, PATHS>(device);
        if (nblk * a.views > cap) views_per_launch = 1;
        if (nblk > cap) return false;  // image too wide for a resident row-synchronous sweep: caller falls back to line walkers
    }
    const size_t vol = (size_t)a.h * a.w * a.d;
    constexpr int PFH = (LaneLayout<D>::E >= 8) ? 8 : 16;
    sgm_hpair_kernel<D, PFH><<<dim3((a.h + 3) / 4, a.views, 2), 128, 0, st>>>(a.cost, a.acc_a, a.acc_b, a.h, a.w, a.P1, a.P2, vol);
    lc.note("sgm_hpair_kernel");

    SweepArgs sa;
    sa.cost = a.cost;
    sa.acc_a = a.acc_a;
    sa.acc_b = a.acc_b;
    sa.pathsum = a.pathsum;
    sa.disp = a.disp;
    sa.subpixel = a.subpixel;
    sa.halo = a.halo;
    sa.progress = a.progress;
    sa.error = a.error;
    sa.h = a.h;
    sa.w = a.w;
    sa.P1 = a.P1;
    sa.P2 = a.P2;
    sa.ratio = a.unique_ratio;
    sa.min_gap = a.min_gap;
    sa.vol_stride = vol;
    sa.img_stride = (size_t)a.h * a.w;
    for (int up = 0; up < 2; ++up) {
        if (PATHS == 8) cudaMemsetAsync(a.progress, 0, sizeof(int) * (size_t)a.views * nblk, st);
        for (int v0 = 0; v0 < a.views; v0 += views_per_launch) {
            sa.view0 = v0;
            cudaError_t e = up ? launch_sweep<D, true, PATHS>(sa, nblk, views_per_launch, PATHS == 8, st)
                               : launch_sweep<D, false, PATHS>(sa, nblk, views_per_launch, PATHS == 8, st);
            if (e != cudaSuccess && lc.err == cudaSuccess) {
                lc.err = e;
                lc.failed = "sgm_sweep_kernel (launch)";
            }
            lc.note("sgm_sweep_kernel");
        }
    }
    return true;
}

template <int D>
bool fused_dispatch(const AggregateArgs& a, cudaStream_t st, LaunchCounter& lc) {
    return a.paths == 8 ? fused_impl<D, 8>(a, st, lc) : fused_impl<D, 4>(a, st, lc);
}

}  // namespace

size_t fused_halo_floats(int views, int w, int d) { return (size_t)views * ((w + SWEEP_NW - 1) / SWEEP_NW) * 2 * HALO_RING * d; }
size_t fused_progress_ints(int views, int w) { return (size_t)views * ((w + SWEEP_NW - 1) / SWEEP_NW); }

bool launch_aggregate_fused(const AggregateArgs& a, cudaStream_t st, LaunchCounter& lc) {
    if (!a.halo || !a.progress || !a.error || !a.acc_b) return false;
    switch (a.d) {
        case 32: return fused_dispatch<32>(a, st, lc);
        case 64: return fused_dispatch<64>(a, st, lc);
        case 128: return fused_dispatch<128>(a, st, lc);
        case 256: return fused_dispatch<256>(a, st, lc);
        default: return false;
    }
}

}  // namespace sgmb200
