// post_kernels.cu — LR-consistency check, hole-filling median and speckle filter (sm_100a).
//
// Replaces src/SGM.cpp:803-818 and Solver::post_filter (src/Solver.cpp:600-635, speckle_filter_new
// :514-566) of the reference, with the reference's serial (single-thread) semantics.
#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "kernels.h"

namespace sgmb200 {

namespace {

// Scratch of the allocation-free speckle entry point (sgm_speckle_filter_static): labels [MAX] then areas [MAX], 16 MiB.
// It lives in the module image: every GPU that loads the library holds it until the process exits.  In HOST code this
// name is only a stand-in for the symbol, never a device address: speckle_static_scratch() below asks the runtime for the
// address (cudaGetSymbolAddress, once per device) and the launcher passes that pointer to the kernels.
static __device__ __align__(256) int g_speckle_static_scratch[2 * (size_t)SPECKLE_STATIC_MAX_PIXELS];

// src/SGM.cpp:803-818: for j >= dl: dr = right(i, int(j - dl/scale)); |dl - dr| > thresh -> dl = D+1.
// Writes the result twice: `out_keep` (the LR tap) and `out_work` (the buffer the median updates in place).
__global__ void lr_check_kernel(const float* __restrict__ left, const float* __restrict__ right, float* __restrict__ out_keep,
                                float* __restrict__ out_work, int h, int w, int d, int scale, float thresh) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    if (j >= w) return;
    const size_t p = (size_t)i * w + j;
    float dl = left[p];
    if ((float)j >= dl) {
        const int jr = (int)__fsub_rn((float)j, __fdiv_rn(dl, (float)scale));
        // 0 <= jr <= j holds for every finite dl >= 0; the guard only keeps a non-finite dl (which the reference
        // would index out of bounds with) from faulting
        if (jr >= 0 && jr < w) {
            const float dr = right[(size_t)i * w + jr];
            if (fabsf(__fsub_rn(dl, dr)) > thresh) dl = (float)(d + 1);
        }
    }
    out_keep[p] = dl;
    out_work[p] = dl;
}

// src/Solver.cpp:604-630.  The reference fills holes IN PLACE in raster order, so a pixel sees the fills of
// earlier pixels in its window.  Pixel (i, j) depends on rows i-rh..i-1 up to column j+rw and on row i left
// of j; the wavefront t = j + (rw+1)*i orders all of those strictly before it.  One block walks the
// wavefront, one row per thread (rows strided over the block when h > blockDim).  `src` is the read-only
// pre-filter image (a pixel is only ever written by its own step, so src tells whether it is a hole without
// touching the buffer being updated); holes gather their window from `f` through L2 (__ldcg).
template <int WH, int WW>
__global__ void __launch_bounds__(1024) median_fill_kernel(const float* __restrict__ src, float* f, int h, int w, int d) {
    constexpr int RH = WH / 2, RW = WW / 2, K = RW + 1, NWIN = WH * WW;
    const float vmax = (float)(d - 1);
    const int steps = (w - 1) + K * (h - 1) + 1;
    for (int t = 0; t < steps; ++t) {
        for (int i = threadIdx.x; i < h; i += blockDim.x) {
            const int j = t - K * i;
            if (j < RW || j >= w - RW || i < RH || i >= h - RH) continue;
            if (__ldg(src + (size_t)i * w + j) <= vmax) continue;  // not a hole
            int vals[NWIN];
            int cnt = 0;
#pragma unroll
            for (int a = -RH; a <= RH; ++a)
#pragma unroll
                for (int b = -RW; b <= RW; ++b) {
                    const float v = __ldcg(f + (size_t)(i + a) * w + (j + b));
                    const bool ok = v <= vmax;
                    vals[(a + RH) * WW + (b + RW)] = ok ? (int)v : 0x7fffffff;  // float -> int truncation (:619)
                    cnt += ok ? 1 : 0;
                }
            if (cnt > NWIN / 2) {
                // sorted[cnt / 2] by rank counting (ties broken by position): invalid entries sort last
                const int want = cnt / 2;
                int pick = 0;
#pragma unroll
                for (int x = 0; x < NWIN; ++x) {
                    int rank = 0;
#pragma unroll
                    for (int y = 0; y < NWIN; ++y) rank += (vals[y] < vals[x] || (vals[y] == vals[x] && y < x)) ? 1 : 0;
                    if (rank == want) pick = vals[x];
                }
                f[(size_t)i * w + j] = (float)pick;
            }
        }
        __syncthreads();
    }
}

// Fast form of the same wavefront.  Every row keeps a 32-column ring in shared memory that holds the CURRENT
// value of its pixels around the wavefront (final values behind the row's cursor, original values ahead of it);
// with the skew K = RW+1 the window of any hole lies within +-(K*RH + RW) <= 8 columns of each neighbouring row's
// cursor, so holes gather their 25 values from shared memory instead of L2.  Original values stream into the ring
// with cp.async (LDGSTS, global -> shared without registers) issued LEAD steps before they are needed, one commit
// group per step, so a step costs one barrier and a handful of instructions; the median of a filled hole is found
// by bisection on the value (ints in [0, D-1]).
template <int WH, int WW, int RPT>
__global__ void __launch_bounds__(1024) median_fill_ring_kernel(const float* __restrict__ src, float* __restrict__ f, int h, int w,
                                                                int d) {
    constexpr int RH = WH / 2, RW = WW / 2, K = RW + 1, NWIN = WH * WW, RING = 32, AHEAD = K * RH + RW + 1, LEAD = 8;
    static_assert(K * RH + RW <= 8 && AHEAD + LEAD + K * RH + RW < RING, "window does not fit the ring");
    extern __shared__ float ring[];  // [h][RING]
    const float vmax = (float)(d - 1);
    const int T = blockDim.x;
    // Row i's cursor at step t is column c = t - K*i.  At step t it requests the original of column c + AHEAD + LEAD;
    // the request of step t-LEAD (column c + AHEAD) has landed before this step's barrier.
    const int t_end = (w - 1) + K * (h - 1);
    for (int t = -(AHEAD + LEAD); t <= t_end; ++t) {
#pragma unroll
        for (int rr = 0; rr < RPT; ++rr) {
            const int i = threadIdx.x + rr * T;
            if (i < h) {
                const int cf = t - K * i + AHEAD + LEAD;
                if (cf >= 0 && cf < w) {
                    const unsigned dst = (unsigned)__cvta_generic_to_shared(ring + (size_t)i * RING + (cf & (RING - 1)));
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src + (size_t)i * w + cf));
                }
            }
        }
        asm volatile("cp.async.commit_group;");
        asm volatile("cp.async.wait_group %0;" ::"n"(LEAD - 1));
            __syncthreads();  // ring contents of this step's feeds and of the previous step's fills are visible
#pragma unroll
            for (int rr = 0; rr < RPT; ++rr) {
                const int i = threadIdx.x + rr * T;
                if (i >= RH && i < h - RH) {
                    const int c = t - K * i;
                    if (c >= RW && c < w - RW) {
                        const float own = ring[(size_t)i * RING + (c & (RING - 1))];
                        if (!(own <= vmax)) {  // a hole
                            int vals[NWIN];
                            int cnt = 0;
#pragma unroll
                            for (int a = -RH; a <= RH; ++a)
#pragma unroll
                                for (int b = -RW; b <= RW; ++b) {
                                    const float v = ring[(size_t)(i + a) * RING + ((c + b) & (RING - 1))];
                                    const bool ok = v <= vmax;
                                    vals[(a + RH) * WW + (b + RW)] = ok ? (int)v : 0x7fffffff;  // float -> int truncation (:619)
                                    cnt += ok ? 1 : 0;
                                }
                            if (cnt > NWIN / 2) {
                                // sorted[cnt / 2] = smallest v with #(vals <= v) >= cnt/2 + 1; bisection over [0, D-1]
                                const int need = cnt / 2 + 1;
                                int lo = 0, hi = d - 1;
                                while (lo < hi) {
                                    const int mid = (lo + hi) >> 1;
                                    int le = 0;
#pragma unroll
                                    for (int x = 0; x < NWIN; ++x) le += vals[x] <= mid ? 1 : 0;
                                    if (le >= need)
                                        hi = mid;
                                    else
                                        lo = mid + 1;
                                }
                                const float fill = (float)lo;
                                f[(size_t)i * w + c] = fill;
                                // publish for later pixels: nobody reads this ring slot before the next barrier
                                ring[(size_t)i * RING + (c & (RING - 1))] = fill;
                            }
                        }
                    }
                }
            }
        // the fills written above are ordered before the next step's gathers by the next step's barrier
    }
}

// Batched form of the ring wavefront.  With a skew of K columns per row, a row may advance B = K - RW consecutive columns
// between two barriers: a hole at (i, c + b), b < B, needs row i-1 final up to column c + b + RW <= c + K - 1, which row
// i-1 (cursor c + K) finished in earlier batches, and rows below have not touched what it reads (their cursors trail by
// K, 2K, ...).  Same raster-order result with (W + K*H) / B barriers instead of W + 3*H.
template <int WH, int WW, int K, int RING, int LEADB, int COOP_MAX>
__global__ void __launch_bounds__(1024) median_fill_batch_kernel(const float* __restrict__ src, float* __restrict__ f, int h, int w,
                                                                  int d) {
    constexpr int RH = WH / 2, RW = WW / 2, B = K - RW, NWIN = WH * WW, LOOK = K * RH + RW, AHEAD = LOOK + 1;
    static_assert(B >= 1 && AHEAD + LEADB * B + B + LOOK < RING && (RING & (RING - 1)) == 0, "window does not fit the ring");
    // rows are RS = RING + 1 floats apart: with one row per lane a power-of-two stride would put all 32 lanes of a warp on
    // the same shared-memory bank (a 32-way conflict on every column check, feed and gather)
    constexpr int RS = RING + 1;
    extern __shared__ float ring[];  // [h][RS]
    const float vmax = (float)(d - 1);
    // Values that never change are pinned in registers through opaque moves: left alone, the compiler re-derives them in
    // every batch (S2R tid, the shared window base, the row's limits), which lengthens the loop's dependent chain.
    auto pin = [](unsigned v) {
        unsigned r;
        asm volatile("mov.u32 %0, %1;" : "=r"(r) : "r"(v));
        return r;
    };
    const int i = (int)pin(threadIdx.x);  // one row per thread (h <= blockDim.x)
    const int lane = i & 31;
    static_assert(NWIN <= 32, "the window must fit a warp");
    // COOP_MAX: up to this many holes per warp and column are filled one after the other by the whole warp
    const bool row_ok = i < h;
    const bool interior_row = i >= RH && i < h - RH;
    float* myring = ring + (size_t)(row_ok ? i : 0) * RS;
    const float* mysrc = src + (size_t)(row_ok ? i : 0) * w;
    const unsigned myring_s = pin((unsigned)__cvta_generic_to_shared(myring));  // shared-space address: offsets fold into the cp.async
    // batch n: row i's cursor is c = n*B - K*i.  First batch: the top row needs its first AHEAD + LEADB*B columns requested.
    const int n_first = -((AHEAD + LEADB * B + B - 1) / B);
    const int n_last = ((w - 1) + K * (h - 1)) / B;
    // The batch loop is a chain of dependent instructions executed ~W + K*H times, so everything that does not change is
    // kept out of it: the cursor and the feed address advance by B, range tests are single unsigned compares.
    int c = n_first * B - K * i;
    long long feed = (long long)(size_t)mysrc + 4ll * (c + AHEAD + LEADB * B);  // address of column c + AHEAD + LEADB*B of this row
    const unsigned feed_lim = pin(row_ok ? (unsigned)w : 0u);                  // columns [0, w) are fed
    const unsigned hole_lim = pin(interior_row ? (unsigned)(w - 2 * RW) : 0u);  // columns [RW, w - RW) can be holes
    for (int n = n_first; n <= n_last; ++n, c += B, feed += 4 * B) {
        {
            const int cf = c + AHEAD + LEADB * B;
#pragma unroll
            for (int b = 0; b < B; ++b) {
                const int x = cf + b;
                if ((unsigned)x < feed_lim) {
                    const unsigned dst = myring_s + (unsigned)(x & (RING - 1)) * 4u;
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(feed + 4 * b));
                }
            }
        }
        asm volatile("cp.async.commit_group;");
        asm volatile("cp.async.wait_group %0;" ::"n"(LEADB));
        __syncthreads();  // feeds of batch n - LEADB and the fills of batch n - 1 are visible
        // Holes of this batch are filled warp-cooperatively: the warp votes which of its 32 rows have a hole at column
        // c + b, then for each such hole 25 lanes gather one window element each and the median comes from a bisection
        // on ballots (the values are integers in [0, D-1]).  Columns b = 0 .. B-1 go in order because a fill changes what
        // the next column of the same row sees; different rows of a batch are independent (see above).
        // Which pixels of the batch are holes is known up front (a fill only ever writes the hole it fills, and every
        // position is examined once), so the B shared-memory reads and votes are issued together instead of one
        // dependent read -> compare -> vote chain per column; columns without a hole in this warp then cost nothing.
        unsigned votes[B];
#pragma unroll
        for (int b = 0; b < B; ++b) {
            const int x = c + b;
            float v;  // any slot is readable; the range test decides whether it counts
            asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(myring_s + (unsigned)(x & (RING - 1)) * 4u));
            votes[b] = __ballot_sync(FULL_MASK, (unsigned)(x - RW) < hole_lim && !(v <= vmax));
        }
#pragma unroll 1
        for (int b = 0; b < B; ++b) {
            unsigned todo = votes[0];
#pragma unroll
            for (int q = 1; q < B; ++q) todo = b == q ? votes[q] : todo;
            if (todo == 0) continue;
            const int x = c + b;
            const bool hole = (todo >> lane) & 1u;
            if (__popc(todo) > COOP_MAX) {
                // many rows of this warp have a hole in this column (occlusion bands, the invalid left margin): every such
                // lane fills its own hole, all of them in parallel, with the same bisection on its private window
                if (hole) {
                    int vals[NWIN];
                    int cnt = 0, vlo = 0x7fffffff, vhi = 0;
#pragma unroll
                    for (int a = -RH; a <= RH; ++a)
#pragma unroll
                        for (int e = -RW; e <= RW; ++e) {
                            const float v = ring[(size_t)(i + a) * RS + ((x + e) & (RING - 1))];
                            const bool ok = v <= vmax;
                            const int iv = ok ? (int)v : 0x7fffffff;  // float -> int truncation (:619)
                            vals[(a + RH) * WW + (e + RW)] = iv;
                            cnt += ok ? 1 : 0;
                            vlo = min(vlo, iv);
                            vhi = ok ? max(vhi, iv) : vhi;
                        }
                    if (cnt > NWIN / 2) {
                        // the median lies between the smallest and the largest valid value: neighbouring disparities are
                        // close, so the bisection usually needs one or two rounds instead of log2(D)
                        const int need = cnt / 2 + 1;
                        int lo = vlo, hi = vhi;
                        while (lo < hi) {
                            const int mid = (lo + hi) >> 1;
                            int le = 0;
#pragma unroll
                            for (int q = 0; q < NWIN; ++q) le += vals[q] <= mid ? 1 : 0;
                            if (le >= need)
                                hi = mid;
                            else
                                lo = mid + 1;
                        }
                        const float fill = (float)lo;
                        f[(size_t)i * w + x] = fill;
                        myring[x & (RING - 1)] = fill;
                    }
                }
                __syncwarp();
                todo = 0;
            }
            while (todo) {
                const int src = __ffs(todo) - 1;
                todo &= todo - 1;
                const int i0 = (int)(threadIdx.x & ~31u) + src;
                const int x0 = __shfl_sync(FULL_MASK, x, src);
                bool ok = false;
                int val = 0x7fffffff;
                if (lane < NWIN) {
                    const int a = lane / WW - RH, e = lane % WW - RW;
                    const float v = ring[(size_t)(i0 + a) * RS + ((x0 + e) & (RING - 1))];
                    ok = v <= vmax;
                    if (ok) val = (int)v;  // float -> int truncation (src/Solver.cpp:619)
                }
                const int cnt = __popc(__ballot_sync(FULL_MASK, ok));
                if (cnt > NWIN / 2) {
                    // sorted[cnt / 2] = the smallest value v with at least cnt / 2 + 1 window values <= v: a bisection on
                    // ballots between the smallest and the largest valid value (two warp reductions).  Neighbouring
                    // disparities are close, so it takes zero to three rounds; ranking every value against every other
                    // (25 shuffles) took longer.
                    int lo = __reduce_min_sync(FULL_MASK, val);
                    int hi = __reduce_max_sync(FULL_MASK, ok ? val : -1);
                    const int need = cnt / 2 + 1;
                    while (lo < hi) {
                        const int mid = (lo + hi) >> 1;
                        const int le = __popc(__ballot_sync(FULL_MASK, val <= mid));  // invalid entries are INT_MAX
                        if (le >= need)
                            hi = mid;
                        else
                            lo = mid + 1;
                    }
                    if (lane == 0) {
                        const float fill = (float)lo;
                        f[(size_t)i0 * w + x0] = fill;
                        ring[(size_t)i0 * RS + (x0 & (RING - 1))] = fill;  // later columns of this row, and everyone after the barrier
                    }
                }
                __syncwarp();
            }
        }
    }
}

// ---- speckle filter: exact 4-connected components under |a - b| < max_dis (src/Solver.cpp:514-566) -----
// Union-find with min-index roots.  Labels only ever decrease and always stay inside the component, so a
// stale read merely costs extra hops.
__device__ __forceinline__ int uf_root(const int* label, int x) {
    while (true) {
        const int p = __ldcg(label + x);
        if (p == x) return x;
        x = p;
    }
}
// Root with path halving, for the merge pass: every second node on the way is re-pointed at its grandparent, so the
// chains through the rows of a tall component (one hop per row, each an L2 round trip) shrink as the merge proceeds.
// Safe beside the atomicMin of uf_unite: a parent always has a smaller index than its child (no cycles), only non-roots
// are re-pointed (a node that has a parent never becomes a root again, so a successful hang of a root is never
// overwritten), and the new target is an ancestor, i.e. still inside the component.
__device__ __forceinline__ int uf_root_halving(int* label, int x) {
    while (true) {
        const int p = __ldcg(label + x);
        if (p == x) return x;
        const int g = __ldcg(label + p);
        if (g == p) return p;
        __stcg(label + x, g);
        x = g;
    }
}
__device__ __forceinline__ void uf_unite(int* label, int a, int b) {
    while (true) {
        a = uf_root_halving(label, a);
        b = uf_root_halving(label, b);
        if (a == b) return;
        if (a < b) {
            const int t = a;
            a = b;
            b = t;
        }
        const int old = atomicMin(label + a, b);  // a > b: hang root a under b
        if (old == a) return;
        a = old;  // a was no longer a root: keep merging from where it points now
    }
}

// Pass 1: horizontal runs.  One warp per row; label = index of the first pixel of the maximal run of
// horizontally linked pixels; area cleared.
__global__ void speckle_rows_kernel(const float* __restrict__ f, int* __restrict__ label, int* __restrict__ area, int h, int w,
                                    float max_dis) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= h) return;
    const float* row = f + (size_t)i * w;
    int carry = 0;  // start column of the run that reaches into this chunk
    for (int j0 = 0; j0 < w; j0 += 32) {
        const int j = j0 + lane;
        bool starts = true;  // does a new run start at j?
        if (j < w && j > 0) starts = !(fabsf(__fsub_rn(row[j], row[j - 1])) < max_dis);
        if (j >= w) starts = true;
        const unsigned m = __ballot_sync(FULL_MASK, starts);
        const unsigned upto = m & (0xffffffffu >> (31 - lane));  // starts at lanes <= mine
        const int start = upto ? j0 + (31 - __clz(upto)) : carry;
        if (j < w) {
            label[(size_t)i * w + j] = i * w + start;
            area[(size_t)i * w + j] = 0;
        }
        // run reaching past this chunk starts at the last start flag of the chunk (or carries on)
        carry = m ? j0 + (31 - __clz(m)) : carry;
    }
}

// Pass 2: vertical links.  Redundant unions inside a pair of vertically linked runs are skipped: pixel j only
// unites with the pixel below when its left neighbour does not already carry the same pair of runs down.
__global__ void speckle_merge_kernel(const float* __restrict__ f, int* label, int h, int w, float max_dis) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int i = blockIdx.y;
    if (j >= w || i >= h - 1) return;
    const size_t p = (size_t)i * w + j;
    const float a = f[p], b = f[p + w];
    if (!(fabsf(__fsub_rn(a, b)) < max_dis)) return;
    if (j > 0) {
        const float al = f[p - 1], bl = f[p + w - 1];
        const bool same_pair = fabsf(__fsub_rn(al, bl)) < max_dis && fabsf(__fsub_rn(a, al)) < max_dis &&
                               fabsf(__fsub_rn(b, bl)) < max_dis;
        if (same_pair) return;
    }
    uf_unite(label, (int)p, (int)(p + w));
}

// Pass 3: flatten and count.  Lanes of a warp that share a root add once.
__global__ void speckle_count_kernel(int* label, int* area, int n) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = p < n;
    int root = -1 - (int)(threadIdx.x & 31);  // distinct dummy per dead lane
    if (live) {
        root = uf_root(label, p);
        label[p] = root;
    }
    const unsigned peers = __match_any_sync(FULL_MASK, root);
    if (live && (int)(threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(area + root, __popc(peers));
}

// Pass 4: components of area <= max_size become `value` (:555-565).
__global__ void speckle_apply_kernel(const float* __restrict__ f_in, float* __restrict__ f_out, const int* __restrict__ label,
                                     const int* __restrict__ area, int n, float value, int max_size) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    f_out[p] = area[label[p]] <= max_size ? value : f_in[p];
}

}  // namespace

void launch_lr_check(const float* left, const float* right, float* out_keep, float* out_work, int h, int w, int d, int scale,
                     float thresh, cudaStream_t st, LaunchCounter& lc) {
    dim3 grid((w + 255) / 256, h);
    lr_check_kernel<<<grid, 256, 0, st>>>(left, right, out_keep, out_work, h, w, d, scale, thresh);
    lc.note("lr_check_kernel");
}

void launch_median_fill(const float* src, float* f, int h, int w, int d, int win_h, int win_w, cudaStream_t st,
                        LaunchCounter& lc) {
    int threads = ((h + 31) / 32) * 32;
    if (threads > 1024) threads = 1024;
    const size_t ring_bytes = (size_t)h * 32 * sizeof(float);
    static const bool legacy = std::getenv("SGM_B200_MEDIAN_LEGACY") != nullptr;
    if (!legacy && win_h == 5 && win_w == 5 && h <= 1024 && (size_t)h * 65 * sizeof(float) <= 200 * 1024) {
        // batched wavefront with 64-column rings
        const size_t bytes = (size_t)h * 65 * sizeof(float);  // rows padded to RING + 1 floats
        // Shape: skew 4 (2 columns per barrier), feeds 3 batches ahead, up to 2 cooperative fills per warp and column.  The
        // time goes with the number of column steps W + K*H, not with the number of barriers (measured, K = 3 ... 10:
        // profiles/r01_v13_summary.md), so the smallest skew that still amortises the barrier wins.
        // SGM_B200_MEDIAN_SHAPE=6 selects the earlier skew-6 shape, =40 / =41 the per-lane-only / one-cooperative-fill forms.
        static const int shape = [] { const char* e = std::getenv("SGM_B200_MEDIAN_SHAPE"); return e ? std::atoi(e) : 4; }();
        auto run = [&](auto kern) {
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
            kern<<<1, threads, bytes, st>>>(src, f, h, w, d);
        };
        switch (shape) {
            case 6: run(median_fill_batch_kernel<5, 5, 6, 64, 3, 2>); break;
            case 40: run(median_fill_batch_kernel<5, 5, 4, 64, 3, 0>); break;
            case 41: run(median_fill_batch_kernel<5, 5, 4, 64, 3, 1>); break;
            default: run(median_fill_batch_kernel<5, 5, 4, 64, 3, 2>); break;
        }
    } else if (win_h == 5 && win_w == 5 && ring_bytes <= 200 * 1024 && h <= 2048 && d <= 65536) {
        // shared-memory ring form (the common case); the L2 form below stays as the fallback for very tall images
        if (h <= 1024) {
            cudaFuncSetAttribute(median_fill_ring_kernel<5, 5, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ring_bytes);
            median_fill_ring_kernel<5, 5, 1><<<1, threads, ring_bytes, st>>>(src, f, h, w, d);
        } else {
            cudaFuncSetAttribute(median_fill_ring_kernel<5, 5, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ring_bytes);
            median_fill_ring_kernel<5, 5, 2><<<1, 1024, ring_bytes, st>>>(src, f, h, w, d);
        }
    } else if (win_h == 5 && win_w == 5)
        median_fill_kernel<5, 5><<<1, threads, 0, st>>>(src, f, h, w, d);
    else if (win_h == 3 && win_w == 3)
        median_fill_kernel<3, 3><<<1, threads, 0, st>>>(src, f, h, w, d);
    else {
        lc.err = cudaErrorInvalidValue;
        return;
    }
    lc.note("median_fill_kernel");
}

void launch_speckle(const float* f, float* f_out, int* labels, int* area, int h, int w, float value, int max_size,
                    float max_dis, cudaStream_t st, LaunchCounter& lc) {
    const int n = h * w;
    speckle_rows_kernel<<<(h + 7) / 8, 256, 0, st>>>(f, labels, area, h, w, max_dis);
    lc.note("speckle_rows_kernel");
    dim3 grid((w + 255) / 256, h);
    speckle_merge_kernel<<<grid, 256, 0, st>>>(f, labels, h, w, max_dis);
    lc.note("speckle_merge_kernel");
    speckle_count_kernel<<<(n + 255) / 256, 256, 0, st>>>(labels, area, n);
    lc.note("speckle_count_kernel");
    speckle_apply_kernel<<<(n + 255) / 256, 256, 0, st>>>(f, f_out, labels, area, n, value, max_size);
    lc.note("speckle_apply_kernel");
}

// Device address of g_speckle_static_scratch on the CURRENT device (which must be `device`), looked up once per device.
static cudaError_t speckle_static_scratch(int device, int** out) {
    constexpr int MAX_DEVICES = 64;
    static std::mutex mu;
    static int* cached[MAX_DEVICES] = {};
    if (device < 0 || device >= MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!cached[device]) {
        int cur = -1;
        cudaError_t c = cudaGetDevice(&cur);
        if (c != cudaSuccess) return c;
        if (cur != device) return cudaErrorInvalidDevice;
        void* ptr = nullptr;
        c = cudaGetSymbolAddress(&ptr, g_speckle_static_scratch);
        if (c != cudaSuccess) return c;
        cached[device] = static_cast<int*>(ptr);
    }
    *out = cached[device];
    return cudaSuccess;
}

cudaError_t launch_speckle_static(int device, const float* f, float* f_out, int h, int w, float value, int max_size, float max_dis,
                                  cudaStream_t st, LaunchCounter& lc) {
    if ((size_t)h * w > (size_t)SPECKLE_STATIC_MAX_PIXELS) return cudaErrorInvalidValue;
    int* scratch = nullptr;
    cudaError_t c = speckle_static_scratch(device, &scratch);
    if (c != cudaSuccess) return c;
    // nothing to clear: speckle_rows_kernel writes every label and zeroes every area count of the h * w pixels before the
    // later passes read them, so what the previous call left in the array is never seen
    launch_speckle(f, f_out, scratch, scratch + SPECKLE_STATIC_MAX_PIXELS, h, w, value, max_size, max_dis, st, lc);
    return lc.err;
}

}  // namespace sgmb200
