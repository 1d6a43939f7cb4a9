// post_kernels.cu — LR-consistency check, hole-filling median and speckle filter (sm_100a).
//
// Replaces src/SGM.cpp:803-818 and Solver::post_filter (src/Solver.cpp:600-635, speckle_filter_new
// :514-566) of the reference, with the reference's serial (single-thread) semantics.
#include <cstdio>
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

// ---------------------------------------------------------------------------------------------
// LR check (src/SGM.cpp:803-818) fused with the first half of the hole-filling median (src/Solver.cpp:604-630).
//
// The reference fills holes IN PLACE in raster order: an interior pixel whose value is invalid (> D-1) takes
// sorted[cnt / 2] of the valid values of its WH x WW window (truncated to int) when more than half of the window is valid,
// and later pixels see that fill.  Only a hole can change, and only another hole earlier in raster order inside its
// window ("causal" cells: the rows above and the cells to the left) can change what it sees.  So:
//   * a hole whose valid neighbours plus causal holes cannot exceed half the window can NEVER be filled, whatever happens
//     before it; it needs no ordering at all.  On real frames that leaves a few hundred CANDIDATES out of ~10^4 holes
//     (the interior of occlusion bands and of the invalid left margin drops out at once);
//   * the candidates form a shallow dependency DAG (depth 20 - 70 on SGM frames, a few hundred on block matching).
// lr_classify_kernel does the LR check for a tile plus halo in shared memory, writes the checked image (twice: the LR tap
// and the image the median updates) and one bit per pixel: "candidate".  median_solve_kernel (one block) then compacts
// the bits in raster order and resolves the candidates in shared memory.  Replaces a 2700-step wavefront over every pixel.
template <int WH, int WW>
struct MedianWin {
    static constexpr int RH = WH / 2, RW = WW / 2, NWIN = WH * WW, NB = NWIN - 1, NC = RH * WW + RW;  // neighbours; causal ones first
    __device__ static __forceinline__ void offset(int t, int& a, int& b) {  // t-th neighbour in row-major order, centre skipped
        const int u = t < NC ? t : t + 1;
        a = u / WW - RH;
        b = u % WW - RW;
    }
};

constexpr int LRC_TX = 32, LRC_TY = 8;

template <int WH, int WW>
__global__ void __launch_bounds__(LRC_TX* LRC_TY) lr_classify_kernel(const float* __restrict__ left, const float* __restrict__ right,
                                                                     float* __restrict__ out_keep, float* __restrict__ out_work,
                                                                     unsigned* __restrict__ bitmap, int h, int w, int d, int scale,
                                                                     float thresh) {
    using MW = MedianWin<WH, WW>;
    constexpr int RH = MW::RH, RW = MW::RW, TH = LRC_TY + 2 * RH, TW = LRC_TX + 2 * RW;
    __shared__ float tile[TH][TW + 1];
    const int i0 = blockIdx.y * LRC_TY, j0 = blockIdx.x * LRC_TX;
    const float vmax = (float)(d - 1);
    for (int t = threadIdx.y * LRC_TX + threadIdx.x; t < TH * TW; t += LRC_TX * LRC_TY) {
        const int ti = t / TW, tj = t % TW, i = i0 - RH + ti, j = j0 - RW + tj;
        float dl = INFINITY;  // outside the image: never read by an interior pixel's window
        if (i >= 0 && i < h && j >= 0 && j < w) {
            const size_t p = (size_t)i * w + j;
            dl = left[p];
            if (right && (float)j >= dl) {  // src/SGM.cpp:803-818: for j >= dl: dr = right(i, int(j - dl/scale)); |dl - dr| > thresh -> D+1
                const int jr = (int)__fsub_rn((float)j, __fdiv_rn(dl, (float)scale));
                // 0 <= jr <= j holds for every finite dl >= 0; the guard only keeps a non-finite dl (which the reference
                // would index out of bounds with) from faulting
                if (jr >= 0 && jr < w) {
                    const float dr = right[(size_t)i * w + jr];
                    if (fabsf(__fsub_rn(dl, dr)) > thresh) dl = (float)(d + 1);
                }
            }
            if (ti >= RH && ti < RH + LRC_TY && tj >= RW && tj < RW + LRC_TX) {
                out_keep[p] = dl;
                out_work[p] = dl;
            }
        }
        tile[ti][tj] = dl;
    }
    __syncthreads();
    const int i = i0 + threadIdx.y, j = j0 + threadIdx.x;
    auto interior = [&](int y, int x) { return y >= RH && y < h - RH && x >= RW && x < w - RW; };
    bool cand = false;
    if (i < h && j < w && interior(i, j) && !(tile[threadIdx.y + RH][threadIdx.x + RW] <= vmax)) {
        int reach = 0;  // valid neighbours + causal holes that might still be filled
#pragma unroll
        for (int t = 0; t < MW::NB; ++t) {
            int a, b;
            MW::offset(t, a, b);
            const bool ok = tile[threadIdx.y + RH + a][threadIdx.x + RW + b] <= vmax;
            reach += (ok || (t < MW::NC && interior(i + a, j + b))) ? 1 : 0;
        }
        cand = reach > MW::NWIN / 2;
    }
    const unsigned word = __ballot_sync(FULL_MASK, cand);  // one warp = one tile row = one bitmap word
    if (threadIdx.x == 0 && i < h) bitmap[(size_t)i * gridDim.x + blockIdx.x] = word;
}

// Second half: ONE block resolves the candidates.
//   1. ordered compaction of the candidate bits: per-row counts (warp per row, lane per bitmap word), block scan over the
//      rows, then every set bit gets its raster-order index; `wordoff` keeps the index of the first candidate of every
//      bitmap word, so "which candidate is pixel q" is one look-up plus a popcount;
//   2. in chunks of CHUNK candidates (raster order): every candidate's window goes to shared memory as 16-bit entries —
//      a literal value (already final: a non-hole, a hole that can never be filled, or a fill of an earlier chunk) or a
//      REFERENCE to an earlier candidate of the same chunk;
//   3. warps resolve groups of 32 consecutive candidates (no block barrier): a candidate whose references are all resolved
//      takes its median (the warp evaluates it together), one that can no longer reach half a window of valid values is
//      dropped early.  The first unresolved candidate in raster order has no unresolved reference, so there is always
//      progress.
// Window entries: 0 .. 0x7ffe literal (int)value (the reference truncates to int, src/Solver.cpp:619), 0x7fff invalid,
// 0x8000 | k reference to candidate k of the chunk.  val[k]: 0xffff unresolved, else like a literal.
constexpr unsigned short MED_INVALID = 0x7fff, MED_UNRESOLVED = 0xffff;

template <int WH, int WW, int CHUNK>
__global__ void __launch_bounds__(512) median_solve_kernel(float* f, const unsigned* __restrict__ bitmap, int* wordoff, int* rowoff, int* list,
                                                            int h, int w, int d, unsigned nap_ns) {
    using MW = MedianWin<WH, WW>;
    constexpr int NB = MW::NB, NC = MW::NC, HALF = MW::NWIN / 2;
    static_assert(NB <= 32, "one lane per window entry");
    extern __shared__ __align__(16) unsigned char msm[];
    constexpr int NBP = NB + 2;  // row pitch: NB window entries + the reference mask + pad = an odd number of 32-bit words (no bank collisions between rows)
    unsigned short* win = reinterpret_cast<unsigned short*>(msm);            // [CHUNK][NBP]
    unsigned short* val = win + (size_t)CHUNK * NBP;                           // [CHUNK]
    unsigned char* litv = reinterpret_cast<unsigned char*>(val + CHUNK);       // [CHUNK] literal valid entries of the window
    __shared__ int wsum[32];
    __shared__ int total_s;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nwarp = blockDim.x >> 5;
    const int nw = (w + 31) / 32;
    const float vmax = (float)(d - 1);
#ifdef SGM_B200_TUNING
    long long tstamp[8];
    int nstamp = 0;
#define MED_STAMP() do { __syncthreads(); if (nstamp < 8) tstamp[nstamp++] = clock64(); } while (0)
#else
#define MED_STAMP() do { } while (0)
#endif
    MED_STAMP();

    // ---- 1a. per-row counts and row-relative word offsets ----
    // (one block does all of this, so global latency is hidden by keeping RB rows in flight per warp)
    constexpr int RB = 4;
    for (int i0 = wid * RB; i0 < h; i0 += nwarp * RB) {
        int total[RB];
#pragma unroll
        for (int r = 0; r < RB; ++r) total[r] = 0;
        for (int jw0 = 0; jw0 < nw; jw0 += 32) {
            const int jw = jw0 + lane;
            int c[RB];
#pragma unroll
            for (int r = 0; r < RB; ++r) c[r] = (jw < nw && i0 + r < h) ? __popc(bitmap[(size_t)(i0 + r) * nw + jw]) : 0;
#pragma unroll
            for (int r = 0; r < RB; ++r) {
                int inc = c[r];
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int x = __shfl_up_sync(FULL_MASK, inc, o);
                    if (lane >= o) inc += x;
                }
                if (jw < nw && i0 + r < h) wordoff[(size_t)(i0 + r) * nw + jw] = total[r] + inc - c[r];
                total[r] += __shfl_sync(FULL_MASK, inc, 31);
            }
        }
#pragma unroll
        for (int r = 0; r < RB; ++r)
            if (lane == 0 && i0 + r < h) rowoff[i0 + r] = total[r];
    }
    __syncthreads();
    MED_STAMP();
    // ---- 1b. exclusive scan of the row counts (contiguous rows per thread, then the classic two-level scan) ----
    {
        const int per = (h + (int)blockDim.x - 1) / (int)blockDim.x, r0 = tid * per, r1 = min(h, r0 + per);
        int mine = 0;
        for (int r = r0; r < r1; ++r) mine += rowoff[r];
        int inc = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int x = __shfl_up_sync(FULL_MASK, inc, o);
            if (lane >= o) inc += x;
        }
        if (lane == 31) wsum[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            const int v = lane < nwarp ? wsum[lane] : 0;
            int winc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int x = __shfl_up_sync(FULL_MASK, winc, o);
                if (lane >= o) winc += x;
            }
            wsum[lane] = winc - v;
            if (lane == 31) total_s = winc;
        }
        __syncthreads();
        int run = wsum[wid] + inc - mine;
        for (int r = r0; r < r1; ++r) {
            const int c = rowoff[r];
            rowoff[r] = run;
            run += c;
        }
    }
    __syncthreads();
    const int ncand = total_s;
    MED_STAMP();
    if (ncand == 0) return;
    // ---- 1c. absolute word offsets and the candidate list ----
    for (int i0 = wid * RB; i0 < h; i0 += nwarp * RB) {
        int base[RB];
#pragma unroll
        for (int r = 0; r < RB; ++r) base[r] = i0 + r < h ? rowoff[i0 + r] : 0;
        for (int jw = lane; jw < nw; jw += 32) {
            int k[RB];
            unsigned word[RB];
#pragma unroll
            for (int r = 0; r < RB; ++r) {
                const bool in = i0 + r < h;
                const size_t wi = (size_t)(in ? i0 + r : 0) * nw + jw;
                k[r] = in ? wordoff[wi] : 0;
                word[r] = in ? bitmap[wi] : 0u;
            }
#pragma unroll
            for (int r = 0; r < RB; ++r) {
                if (i0 + r >= h) continue;
                int kk = k[r] + base[r];
                wordoff[(size_t)(i0 + r) * nw + jw] = kk;
                for (unsigned wd = word[r]; wd; wd &= wd - 1) list[kk++] = (i0 + r) * w + jw * 32 + (__ffs(wd) - 1);
            }
        }
    }
    __syncthreads();
    MED_STAMP();

    // ---- 2 / 3. chunks of candidates ----
    for (int base = 0; base < ncand; base += CHUNK) {
        const int cn = min(CHUNK, ncand - base);
        for (int c = tid; c < cn; c += blockDim.x) {
            const int p = list[base + c], i = p / w, j = p - i * w;
            // all loads of the window first (independent, so they overlap), decisions afterwards
            float fv[NB];
            unsigned bw[NC];
#pragma unroll
            for (int t = 0; t < NB; ++t) {
                int a, b;
                MW::offset(t, a, b);
                fv[t] = f[(size_t)(i + a) * w + (j + b)];
                if (t < NC) bw[t] = bitmap[(size_t)(i + a) * nw + ((j + b) >> 5)];
            }
            int nlit = 0;
            unsigned refmask = 0;
#pragma unroll
            for (int t = 0; t < NB; ++t) {
                int a, b;
                MW::offset(t, a, b);
                const int y = i + a, x = j + b;
                unsigned short e = 0;
                bool is_ref = false;
                if (t < NC) {  // a causal cell may be an earlier candidate of this chunk
                    if ((bw[t] >> (x & 31)) & 1u) {
                        const int k = wordoff[(size_t)y * nw + (x >> 5)] + __popc(bw[t] & ((1u << (x & 31)) - 1u));
                        if (k >= base) {
                            e = (unsigned short)(0x8000u | (unsigned)(k - base));
                            is_ref = true;
                            refmask |= 1u << t;
                        }
                    }
                }
                if (!is_ref) {  // final already (earlier chunks wrote their fills back before this chunk was built)
                    const bool ok = fv[t] <= vmax;
                    e = ok ? (unsigned short)(int)fv[t] : MED_INVALID;
                    nlit += ok ? 1 : 0;
                }
                win[(size_t)c * NBP + t] = e;
            }
            win[(size_t)c * NBP + NB] = (unsigned short)refmask;  // the row's spare entry: which causal entries are references
            litv[c] = (unsigned char)nlit;
            val[c] = MED_UNRESOLVED;
        }
        __syncthreads();
        MED_STAMP();
        // Resolution without block barriers.  A warp takes one GROUP of 32 consecutive candidates at a time (groups wid,
        // wid + nwarp, ...) and iterates on it until all 32 are resolved: a candidate whose references are all resolved takes
        // its median, one that can no longer reach half a window of valid values is dropped early.  Dependencies mostly run
        // between neighbouring candidates (the hole to the left, the holes one and two rows up), i.e. inside a group or into
        // the groups just before it.  A reference always points to a LOWER index: the lowest unfinished group never waits
        // for anybody, every warp of the block is resident, so this terminates.  val[] entries go unresolved -> final
        // exactly once (one 16-bit store): whoever sees a final value sees the right one.
        //
        // Measured alternatives on the B200 (cycles of this phase, 1242x375: SGM frame with 1842 candidates / SAD block
        // matching with 3065, see profiles/r02_median_solver.md): this form 177 k / 1.35 M; block-barrier rounds with one
        // thread per candidate ~180 k / 3.4 M; per-lane instead of cooperative medians 294 k / 1.67 M; one candidate per
        // warp, round-robin 158 k / 1.10 M; the same after peeling the never-fillable candidates to a fixpoint in parallel
        // rounds (75 k / 116 k for the peeling) 145 k / 1.17 M; ONE warp walking the survivors in order 481 k / 1.76 M.
        // They all end up near (number of filled holes) x (400 - 1000 cycles): consecutive holes are neighbours, so the
        // medians form one long chain.  Later: the median by rank counting instead of bisection (below) 166 k / 1.29 M; with it,
        // looking only at the LATEST open reference per round plus forwarding a fill to the next lane in registers 497 k / 1.30 M
        // (the never-fillable candidates of an SGM frame are then found in chain order instead of at once); all open references
        // looked at with independent, predicated loads instead of a loop over the set bits (shipped) 112 k / 1.02 M: on a
        // block-matching frame the chain runs DOWN the invalid left band, 3 - 5 dependent fills per image row, one round each.
        volatile unsigned short* vval = val;
        for (int g = wid; g * 32 < cn; g += nwarp) {
            const int c = g * 32 + lane;
            bool unresolved = c < cn;
            const unsigned short* wc = win + (size_t)(unresolved ? c : 0) * NBP;
            unsigned refmask = wc[NB];             // references still unresolved
            int known = litv[unresolved ? c : 0];  // valid entries seen so far: literals + references resolved to a value
            // the candidate's causal references, fetched once (entries that are literals are never used as an index)
            int ref[NC];
#pragma unroll
            for (int t = 0; t < NC; ++t) ref[t] = wc[t] & 0x7fffu;
            for (;;) {
                bool ready = false, progress = false;
                if (unresolved) {
                    // look at every open reference: NC independent, predicated look-ups issued together.  (A loop over the set
                    // bits of refmask — find the bit, load the entry, load the value, branch — costs its full latency per
                    // open reference, ~1000 cycles per round for a hole in the middle of a band of holes.)
                    constexpr unsigned OPEN = 0xffffu;  // == MED_UNRESOLVED
                    unsigned xs[NC];
#pragma unroll
                    for (int t = 0; t < NC; ++t) {
                        xs[t] = OPEN;
                        if ((refmask >> t) & 1u) xs[t] = vval[ref[t]];
                    }
#pragma unroll
                    for (int t = 0; t < NC; ++t) {
                        if (xs[t] != OPEN) {  // (only open references were looked at) a resolved reference is consumed once
                            refmask &= ~(1u << t);
                            known += xs[t] != 0x7fffu ? 1 : 0;  // MED_INVALID
                            progress = true;
                        }
                    }
                    if (known + __popc(refmask) <= HALF) {  // can no longer reach half a window (or all known, too few): invalid
                        vval[c] = MED_INVALID;
                        unresolved = false;
                        progress = true;
                    } else {
                        ready = refmask == 0;                // known > HALF: takes the median of its window
                    }
                }
                // ready candidates are evaluated by the whole warp, one after the other: lane t holds window entry t, the
                // median is the entry at position cnt / 2 of the sorted valid entries
                for (unsigned todo = __ballot_sync(FULL_MASK, ready); todo; todo &= todo - 1) {
                    const int src = __ffs(todo) - 1, cs = g * 32 + src;
                    int x = MED_INVALID;
                    if (lane < NB) {
                        x = win[(size_t)cs * NBP + lane];
                        if (x & 0x8000) x = vval[x & 0x7fff];  // resolved: cs is ready
                    }
                    const bool ok = x != MED_INVALID;
                    const int cnt = __popc(__ballot_sync(FULL_MASK, ok));
                    // rank by counting: (value, lane) keys are distinct, so the valid entries get the ranks 0 .. cnt-1 of the
                    // sorted order; NB independent shuffles instead of a bisection of ~8 dependent ballot / popcount rounds
                    // (166 k / 1.29 M cycles against 177 k / 1.35 M: the evaluation itself is not what the chain costs)
                    const int key = ok ? ((x << 5) | lane) : 0x7fffffff;
                    int rank = 0;
#pragma unroll
                    for (int t = 0; t < NB; ++t) rank += (__shfl_sync(FULL_MASK, key, t) < key) ? 1 : 0;
                    const unsigned hit = __ballot_sync(FULL_MASK, ok && rank == cnt / 2);  // exactly one lane: cnt > HALF >= 1
                    const int med = __shfl_sync(FULL_MASK, x, __ffs(hit) - 1);
                    if (lane == src) {
                        vval[cs] = (unsigned short)med;
                        unresolved = false;
                    }
                    progress = true;
                }
                __syncwarp();
                if (__ballot_sync(FULL_MASK, unresolved) == 0) break;
                // nothing moved: the group waits for an earlier group.  A nap here (nap_ns = 100, tuning build knob
                // SGM_B200_MEDIAN_NAP) costs this kernel 2 - 7 % when it runs alone and changes nothing for the frames in flight
                // beside it (bench.py with 0 / 100 on one box: 669 / 668 frames/s, block matching 2087 / 2056), so the product naps 0.
                if (nap_ns && __ballot_sync(FULL_MASK, progress) == 0) __nanosleep(nap_ns);
            }
        }
        __syncthreads();
        for (int c = tid; c < cn; c += blockDim.x) {
            const unsigned short v = val[c];
            if (v != MED_INVALID) f[list[base + c]] = (float)v;
        }
        __syncthreads();  // the fills are visible to the next chunk's window build
        MED_STAMP();
    }
#ifdef SGM_B200_TUNING
    if (tid == 0) {
        printf("median_solve: ncand %d, cycles since start:", ncand);
        for (int k = 1; k < nstamp; ++k) printf(" %lld", tstamp[k] - tstamp[0]);
        printf("  (1a rows, 1b scan, 1c list, build, resolve+writeback, ...)\n");
    }
#endif
#undef MED_STAMP
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

template <int WH, int WW>
static void launch_lr_median_impl(const float* left, const float* right, float* out_lr, float* out_med, int* scratch_a, int* scratch_b, int h,
                                  int w, int d, int scale, float thresh, cudaStream_t st, LaunchCounter& lc) {
    constexpr int CHUNK = 3072;
    const int nw = (w + 31) / 32;
    unsigned* bitmap = reinterpret_cast<unsigned*>(scratch_a);  // [h][nw]
    int* wordoff = scratch_a + (size_t)h * nw;                  // [h][nw]
    int* rowoff = wordoff + (size_t)h * nw;                     // [h]
    dim3 grid(nw, (h + LRC_TY - 1) / LRC_TY);
    lr_classify_kernel<WH, WW><<<grid, dim3(LRC_TX, LRC_TY), 0, st>>>(left, right, out_lr, out_med, bitmap, h, w, d, scale, thresh);
    lc.note("lr_classify_kernel");
    constexpr size_t smem = (size_t)CHUNK * ((MedianWin<WH, WW>::NB + 2) * 2 + 2 + 1);
    auto kern = median_solve_kernel<WH, WW, CHUNK>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    unsigned nap_ns = 0;
#ifdef SGM_B200_TUNING
    {
        static const int nap = [] { const char* e = std::getenv("SGM_B200_MEDIAN_NAP"); return e ? std::atoi(e) : 0; }();
        nap_ns = (unsigned)nap;
    }
#endif
#ifdef SGM_B200_TUNING
    {
        // experiment: a block small enough (threads x registers, shared memory) to sit on an SM beside a sweep block of another
        // frame in flight, so that a 148-block cooperative sweep never waits for the SM the median is on
        static const int small = [] { const char* e = std::getenv("SGM_B200_MEDIAN_SMALL"); return e ? std::atoi(e) : 0; }();
        if (small > 0) {
            constexpr int SCHUNK = 1024;
            constexpr size_t ssmem = (size_t)SCHUNK * ((MedianWin<WH, WW>::NB + 2) * 2 + 2 + 1);
            auto skern = median_solve_kernel<WH, WW, SCHUNK>;
            cudaFuncSetAttribute(skern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssmem);
            skern<<<1, small, ssmem, st>>>(out_med, bitmap, wordoff, rowoff, scratch_b, h, w, d, nap_ns);
            lc.note("median_solve_kernel");
            return;
        }
    }
#endif
    kern<<<1, 512, smem, st>>>(out_med, bitmap, wordoff, rowoff, scratch_b, h, w, d, nap_ns);
    lc.note("median_solve_kernel");
}

// scratch_a: >= 2 * h * ceil(w / 32) + h ints, scratch_b: >= h * w ints (the speckle filter's label / area buffers serve).
void launch_lr_median(const float* left, const float* right, float* out_lr, float* out_med, int* scratch_a, int* scratch_b, int h, int w,
                      int d, int scale, float thresh, int win_h, int win_w, cudaStream_t st, LaunchCounter& lc) {
    if ((size_t)2 * h * ((w + 31) / 32) + h > (size_t)h * w) {
        lc.err = cudaErrorInvalidValue;
        lc.failed = "launch_lr_median (image too narrow for the scratch layout)";
        return;
    }
    if (win_h == 5 && win_w == 5)
        launch_lr_median_impl<5, 5>(left, right, out_lr, out_med, scratch_a, scratch_b, h, w, d, scale, thresh, st, lc);
    else if (win_h == 3 && win_w == 3)
        launch_lr_median_impl<3, 3>(left, right, out_lr, out_med, scratch_a, scratch_b, h, w, d, scale, thresh, st, lc);
    else {
        lc.err = cudaErrorInvalidValue;
        lc.failed = "launch_lr_median (window)";
    }
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
