// engine.cu — the C ABI of include/sgm_b200.h: owns device volumes, pinned staging, the stream and the
// per-frame kernel schedule.  Host side of the drop-in boundary for the reference's Solver / SGM / BM
// classes (inc/Solver.h:23-70, src/SGM.cpp:32-834, src/BM.cpp:9-105).  No CPU compute path exists here:
// every stage is a kernel from cost_kernels.cu / aggregate_kernels.cu / post_kernels.cu.
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <vector>

#include "engine_internal.h"
#include "kernels.h"

using namespace sgmb200;

namespace {

thread_local std::string g_create_error;

int fail(sgm_engine* e, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (e)
        e->err = buf;
    else
        g_create_error = buf;
    return code;
}

#define CU_TRY(e, call)                                                                                   \
    do {                                                                                                  \
        cudaError_t _c = (call);                                                                          \
        if (_c != cudaSuccess) return fail((e), SGM_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(_c));   \
    } while (0)

// src/cost.cpp:29 returns sum / win_h / win_w: two float divisions of an integer sum <= 255 * win_h * win_w.  The SAD
// kernel looks the result up instead of dividing (host IEEE division == __fdiv_rn).
std::vector<float> sad_division_table(int win_h, int win_w) {
    std::vector<float> t((size_t)255 * win_h * win_w + 1);
    for (size_t s = 0; s < t.size(); ++s) {
        volatile float q = (float)s / (float)win_h;
        t[s] = q / (float)win_w;
    }
    return t;
}

template <typename T>
cudaError_t dev_alloc(sgm_engine* e, T** p, size_t count) {
    cudaError_t c = cudaMalloc(reinterpret_cast<void**>(p), count * sizeof(T));
    if (c == cudaSuccess) e->device_bytes += count * sizeof(T);
    return c;
}

void release(sgm_engine* e) {
    if (!e) return;
    cudaSetDevice(e->p.device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    for (int v = 0; v < 2; ++v) {
        cudaFree(e->d_in[v]);
        cudaFree(e->d_sky[v]);
        cudaFreeHost(e->h_in[v]);
        cudaFreeHost(e->h_sky[v]);
    }
    cudaFree(e->d_img);
    cudaFree(e->d_blur);
    cudaFree(e->d_census);
    cudaFree(e->d_cost);
    cudaFree(e->d_acc_a);
    cudaFree(e->d_acc_b);
    cudaFree(e->d_pathsum);
    cudaFree(e->d_disp);
    cudaFree(e->d_sub);
    cudaFree(e->d_lr);
    cudaFree(e->d_med);
    cudaFree(e->d_final);
    cudaFree(e->d_label);
    cudaFree(e->d_area);
    cudaFree(e->d_halo);
    cudaFree(e->d_sad_lut);
    cudaFree(e->d_error);
    cudaFree(e->d_view);
    cudaFree(e->d_pc_counts);
    cudaFree(e->d_pc_offsets);
    cudaFree(e->d_pc_xyz);
    cudaFree(e->d_pc_int);
    cudaFreeHost(e->h_error);
    cudaFreeHost(e->h_final);
    cudaFreeHost(e->h_disp8);
    cudaFreeHost(e->h_sub_r);
    cudaFree(e->d_disp8);
    if (e->ev_done) cudaEventDestroy(e->ev_done);
    for (auto& ev : e->ev)
        if (ev) cudaEventDestroy(ev);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

}  // namespace

extern "C" {

int sgm_abi_version(void) { return SGM_B200_ABI_VERSION; }

void sgm_default_params(sgm_params* p, int height, int width, int scale, int max_disp, int algo) {
    if (!p) return;
    std::memset(p, 0, sizeof(*p));
    p->height = height;
    p->width = width;
    p->scale = scale;
    p->max_disp = max_disp;
    p->algo = algo;
    p->paths = 8;            // inc/SGM.h:7
    p->P1 = 10;              // src/SGM.cpp:27
    p->P2 = 100;             // src/SGM.cpp:28
    p->census_h = 7;         // inc/Solver.h:10
    p->census_w = 9;         // inc/Solver.h:11
    p->box_h = 3;            // inc/Solver.h:12
    p->box_w = 5;            // inc/Solver.h:13
    p->unique_ratio = 0.7f;  // inc/Solver.h:14
    p->lr_thresh = 1.0f;     // inc/Solver.h:16
    p->median_h = 5;         // inc/Solver.h:17
    p->median_w = 5;         // inc/Solver.h:18
    p->speckle_size = 1000;  // inc/Solver.h:19
    p->speckle_dis = 2;      // inc/Solver.h:20
    p->sad_h = 9;
    p->sad_w = 9;
    p->device = 0;
    p->flags = 0;
}

const char* sgm_last_error(const sgm_engine* e) { return e ? e->err.c_str() : g_create_error.c_str(); }

int sgm_create(sgm_engine** out, const sgm_params* params) {
    if (!out || !params) return fail(nullptr, SGM_ERR_INVALID, "sgm_create: null argument");
    *out = nullptr;
    const sgm_params& p = *params;
    // src/Solver.cpp:6-10 asserts; here they are error codes
    if (p.height <= 0 || p.width <= 0 || p.scale <= 0 || p.max_disp <= 0)
        return fail(nullptr, SGM_ERR_INVALID, "h, w, scale, max_disp must be positive");
    if (p.scale != 1 && p.scale != 2) return fail(nullptr, SGM_ERR_INVALID, "scale must be 1 or 2");
    if (p.max_disp != 32 && p.max_disp != 64 && p.max_disp != 128 && p.max_disp != 256)
        return fail(nullptr, SGM_ERR_UNSUPPORTED, "max_disp must be 32, 64, 128 or 256");
    if (p.algo != SGM_ALGO_SGM && p.algo != SGM_ALGO_BM && p.algo != SGM_ALGO_BM_SAD)
        return fail(nullptr, SGM_ERR_INVALID, "unknown algo %d", p.algo);
    if (p.paths != 4 && p.paths != 8) return fail(nullptr, SGM_ERR_INVALID, "paths must be 4 or 8");
    const int ih = p.height / p.scale, iw = p.width / p.scale;
    const int bw = p.box_w / p.scale, bh = p.box_h / p.scale;
    if (ih < 16 || iw < 16) return fail(nullptr, SGM_ERR_INVALID, "working image %dx%d is smaller than 16x16", iw, ih);
    if (bw < 1 || bw > 5 || bh < 1 || bh > 5) return fail(nullptr, SGM_ERR_UNSUPPORTED, "box window / scale must be in 1..5");
    if (!((p.median_h == 5 && p.median_w == 5) || (p.median_h == 3 && p.median_w == 3)))
        return fail(nullptr, SGM_ERR_UNSUPPORTED, "median window must be 5x5 or 3x3");
    if (p.algo != SGM_ALGO_BM_SAD) {
        const int ch = p.census_h / p.scale, cw = p.census_w / p.scale;
        if (ch / 2 > 3 || cw / 2 > 4 || ch < 1 || cw < 1)
            return fail(nullptr, SGM_ERR_UNSUPPORTED, "census window / scale must fit 7x9");
        if (!hamming_rows_fit(iw, p.max_disp)) return fail(nullptr, SGM_ERR_UNSUPPORTED, "working image too wide (%d) for the Hamming kernels' row staging", iw);
    } else {
        const int sh = p.sad_h / p.scale, sw = p.sad_w / p.scale;
        if (sh < 1 || sw < 1 || sh > 31 || sw > 31) return fail(nullptr, SGM_ERR_UNSUPPORTED, "SAD window / scale must be in 1..31");
        if ((size_t)(32 + 2 * (sw / 2)) * p.max_disp * 2 > 48 * 1024)
            return fail(nullptr, SGM_ERR_UNSUPPORTED, "SAD window too wide for this max_disp");
    }
    if (p.flags & SGM_FLAG_INTEGER) {
        // every value on the path must be an integer that fits 15 bits: raw Hamming costs (no cost filter), no sky fill
        if (p.algo != SGM_ALGO_SGM) return fail(nullptr, SGM_ERR_UNSUPPORTED, "SGM_FLAG_INTEGER needs SGM_ALGO_SGM");
        if (bw != 1 || bh != 1) return fail(nullptr, SGM_ERR_UNSUPPORTED, "SGM_FLAG_INTEGER needs box_w / scale == box_h / scale == 1 (no cost filter)");
        if (p.max_disp < 64) return fail(nullptr, SGM_ERR_UNSUPPORTED, "SGM_FLAG_INTEGER needs max_disp 64, 128 or 256");
        if (p.P1 < 0 || p.P2 < p.P1 || 8 * (62 + p.P2) > 32767) return fail(nullptr, SGM_ERR_UNSUPPORTED, "SGM_FLAG_INTEGER needs 0 <= P1 <= P2 <= 4033");
        if (p.flags & SGM_FLAG_LINE_WALKERS) return fail(nullptr, SGM_ERR_UNSUPPORTED, "SGM_FLAG_INTEGER has no line-walker schedule");
    }
    int ndev = 0;
    cudaError_t c = cudaGetDeviceCount(&ndev);
    if (c != cudaSuccess || ndev <= 0)
        return fail(nullptr, SGM_ERR_CUDA, "no CUDA device (%s): this engine has no CPU path",
                    c != cudaSuccess ? cudaGetErrorString(c) : "device count 0");
    if (p.device < 0 || p.device >= ndev) return fail(nullptr, SGM_ERR_INVALID, "device %d out of range (0..%d)", p.device, ndev - 1);

    sgm_engine* e = new (std::nothrow) sgm_engine();
    if (!e) return fail(nullptr, SGM_ERR_NOMEM, "out of host memory");
    e->p = p;
    e->ih = ih;
    e->iw = iw;
    e->D = p.max_disp;
    e->views = p.algo == SGM_ALGO_BM ? 1 : 2;
    e->n = (size_t)ih * iw;
    e->N = e->n * e->D;
    e->timing = (p.flags & SGM_FLAG_STAGE_TIMING) != 0;
    e->integer = (p.flags & SGM_FLAG_INTEGER) != 0;
    e->vol_words = e->integer ? e->N / 2 : e->N;
    const size_t full = (size_t)p.height * p.width;
    bool ok = cudaSetDevice(p.device) == cudaSuccess;
    ok = ok && cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) == cudaSuccess;
    for (int k = 0; ok && k <= SGM_STAGE_COUNT; ++k) ok = cudaEventCreate(&e->ev[k]) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&e->ev_done, cudaEventBlockingSync | cudaEventDisableTiming) == cudaSuccess;
    for (int v = 0; ok && v < 2; ++v) {
        ok = ok && dev_alloc(e, &e->d_in[v], full) == cudaSuccess;
        ok = ok && dev_alloc(e, &e->d_sky[v], e->n) == cudaSuccess;
        ok = ok && cudaMallocHost(reinterpret_cast<void**>(&e->h_in[v]), full) == cudaSuccess;
        ok = ok && cudaMallocHost(reinterpret_cast<void**>(&e->h_sky[v]), e->n) == cudaSuccess;
    }
    ok = ok && dev_alloc(e, &e->d_img, 2 * e->n) == cudaSuccess;
    ok = ok && dev_alloc(e, &e->d_blur, 2 * e->n) == cudaSuccess;
    ok = ok && dev_alloc(e, &e->d_census, 2 * e->n) == cudaSuccess;
    ok = ok && dev_alloc(e, &e->d_cost, e->views * e->vol_words) == cudaSuccess;
    if (p.algo == SGM_ALGO_SGM) {
        ok = ok && dev_alloc(e, &e->d_acc_a, e->views * e->vol_words) == cudaSuccess;
        ok = ok && dev_alloc(e, &e->d_acc_b, e->views * e->vol_words) == cudaSuccess;
        const size_t halo_bytes = fused_halo_bytes(e->views, iw, e->D);
        ok = ok && dev_alloc(e, &e->d_halo, halo_bytes / sizeof(float)) == cudaSuccess;
        ok = ok && cudaMemset(e->d_halo, 0, halo_bytes) == cudaSuccess;
#ifdef SGM_B200_TUNING
        // test hook (tuning build only): start the launch sequence of the halo tags near its wrap-around (normally reached after ~5e5 frames)
        if (const char* s0 = std::getenv("SGM_B200_HALO_SEQ_START")) e->halo_seq = (unsigned)std::strtoul(s0, nullptr, 0);
#endif
        ok = ok && dev_alloc(e, &e->d_error, 1) == cudaSuccess;
        ok = ok && cudaMemset(e->d_error, 0, sizeof(int)) == cudaSuccess;
        ok = ok && cudaMallocHost(reinterpret_cast<void**>(&e->h_error), sizeof(int)) == cudaSuccess;
        if (ok) *e->h_error = 0;
        if (p.flags & SGM_FLAG_KEEP_VOLUMES) ok = ok && dev_alloc(e, &e->d_pathsum, e->views * e->vol_words) == cudaSuccess;
    }
    if (p.algo == SGM_ALGO_BM_SAD) {
        const int sh = p.sad_h / p.scale, sw = p.sad_w / p.scale;
        std::vector<float> lut = sad_division_table(sh, sw);
        ok = ok && dev_alloc(e, &e->d_sad_lut, lut.size()) == cudaSuccess;
        ok = ok && cudaMemcpy(e->d_sad_lut, lut.data(), lut.size() * sizeof(float), cudaMemcpyHostToDevice) == cudaSuccess;
    }
    ok = ok && dev_alloc(e, &e->d_disp, 2 * e->n) == cudaSuccess;
    ok = ok && dev_alloc(e, &e->d_sub, 2 * e->n) == cudaSuccess;
    ok = ok && dev_alloc(e, &e->d_lr, e->n) == cudaSuccess;
    ok = ok && dev_alloc(e, &e->d_med, e->n) == cudaSuccess;
    ok = ok && dev_alloc(e, &e->d_final, e->n) == cudaSuccess;
    ok = ok && dev_alloc(e, &e->d_label, e->n) == cudaSuccess;
    ok = ok && dev_alloc(e, &e->d_area, e->n) == cudaSuccess;
    ok = ok && cudaMallocHost(reinterpret_cast<void**>(&e->h_final), e->n * sizeof(float)) == cudaSuccess;
    e->host_taps = (p.flags & SGM_FLAG_HOST_TAPS) != 0;
    if (e->host_taps) {
        ok = ok && dev_alloc(e, &e->d_disp8, 2 * e->n) == cudaSuccess;
        ok = ok && cudaMallocHost(reinterpret_cast<void**>(&e->h_disp8), 2 * e->n) == cudaSuccess;
        ok = ok && cudaMallocHost(reinterpret_cast<void**>(&e->h_sub_r), e->n * sizeof(float)) == cudaSuccess;
    }
    if (!ok) {
        cudaError_t last = cudaGetLastError();
        fail(nullptr, last == cudaErrorMemoryAllocation ? SGM_ERR_NOMEM : SGM_ERR_CUDA, "sgm_create: allocation failed (%s)",
             cudaGetErrorString(last));
        release(e);
        return last == cudaErrorMemoryAllocation ? SGM_ERR_NOMEM : SGM_ERR_CUDA;
    }
    *out = e;
    return SGM_OK;
}

void sgm_destroy(sgm_engine* e) { release(e); }

}  // extern "C"

namespace {

// Queues every kernel of one frame on the engine's stream.  left / right are device pointers to
// full-resolution 8-bit images; sky masks (nullable) are dense device buffers at working resolution.
int run_frame(sgm_engine* e, const uint8_t* left, size_t lstride, const uint8_t* right, size_t rstride, const uint8_t* sky_l,
              const uint8_t* sky_r, LaunchCounter& lc) {
    const sgm_params& p = e->p;
    const int h = e->ih, w = e->iw, D = e->D, s = p.scale;
    cudaStream_t st = e->stream;
    const bool is_bm = p.algo != SGM_ALGO_SGM;

    // ---- cost stage ----
    if (e->timing) cudaEventRecord(e->ev[SGM_STAGE_COST], st);
    const int row_step = is_bm ? 1 : s;  // src/BM.cpp:23-24 reads source row i, src/SGM.cpp:46-47 row i*scale
    launch_subsample(left, lstride, s, row_step, e->d_img, h, w, st, lc);
    launch_subsample(right, rstride, s, row_step, e->d_img + e->n, h, w, st, lc);
    if (p.algo == SGM_ALGO_BM_SAD) {
        launch_sad_volume(e->d_img, e->d_img + e->n, e->d_sad_lut, e->d_cost, e->views, h, w, D, p.sad_h / s, p.sad_w / s, s, st, lc);
    } else {
        launch_blur3(e->d_img, e->d_blur, h, w, st, lc);                                   // src/Solver.cpp:124-125
        launch_census(e->d_blur, e->d_census, h, w, p.census_h / s, p.census_w / s, st, lc);  // :127-139
        if (e->integer)
            launch_hamming_s16x2(e->d_census, e->d_census + e->n, sky_l, sky_r, 2u * (unsigned)p.P2, e->d_cost, e->views, h, w, D, s, st,
                                 lc);  // :143-248, raw; masked pixels: 2 * P2 stands in for 999999
        else {
        launch_hamming_hfilter(e->d_census, e->d_census + e->n, sky_l, sky_r, e->d_cost, e->views, h, w, D, s, p.box_w / s, st,
                               lc);                                                          // :143-248, :296-330
        launch_vfilter(e->d_cost, e->views, h, w, D, p.box_h / s, st, lc);                  // :333-368
        }
    }

    // ---- aggregation / selection ----
    if (e->timing) cudaEventRecord(e->ev[SGM_STAGE_AGGREGATE], st);
    if (p.algo == SGM_ALGO_SGM) {
        AggregateArgs a;
        a.cost = e->d_cost;
        a.acc_a = e->d_acc_a;
        a.acc_b = e->d_acc_b;
        a.pathsum = e->d_pathsum;
        a.disp = e->d_disp;
        a.subpixel = e->d_sub;
        a.views = e->views;
        a.h = h;
        a.w = w;
        a.d = D;
        a.paths = p.paths;
        a.P1 = (float)p.P1;
        a.P2 = (float)p.P2;
        a.unique_ratio = p.unique_ratio;
        a.min_gap = 1;
        a.halo = e->d_halo;
        a.launch_seq = &e->halo_seq;
        a.error = e->d_error;
        a.packed16 = e->integer ? 1 : 0;
        // fused three-launch schedule (sweep_kernels.cu) unless disabled or the image is too wide for a resident sweep
        const bool want_fused = !(p.flags & SGM_FLAG_LINE_WALKERS);
        e->fused = want_fused && launch_aggregate_fused(a, st, lc);
        if (!e->fused && e->integer)
            return fail(e, SGM_ERR_UNSUPPORTED, "integer mode: the image is too wide for a resident sweep and there is no other integer schedule");
        if (!e->fused) launch_aggregate(a, st, lc);
        if (e->fused) cudaMemcpyAsync(e->h_error, e->d_error, sizeof(int), cudaMemcpyDeviceToHost, st);
    } else {
        // src/BM.cpp:52-91; census-BM gets compute_subpixel on the filtered cost, SAD-BM keeps integers
        launch_wta(e->d_cost, e->d_disp, e->d_sub, e->views, h, w, D, p.unique_ratio, 2, p.algo == SGM_ALGO_BM ? 1 : 0, st, lc);
    }

    // ---- post stage ----
    if (e->timing) cudaEventRecord(e->ev[SGM_STAGE_POST], st);
    // src/SGM.cpp:803-818 (LR check; BM has one view and none) + src/Solver.cpp:604-630 (hole-filling median)
    launch_lr_median(e->d_sub, e->views == 2 ? e->d_sub + e->n : nullptr, e->d_lr, e->d_med, e->d_label, e->d_area, h, w, D, s, p.lr_thresh,
                     p.median_h, p.median_w, st, lc);
    launch_speckle(e->d_med, e->d_final, e->d_label, e->d_area, h, w, (float)(D + 1), p.speckle_size / s, (float)p.speckle_dis, st,
                   lc);                                                              // src/Solver.cpp:634
    if (e->timing) cudaEventRecord(e->ev[SGM_STAGE_DOWNLOAD], st);
    if (lc.err != cudaSuccess)
        return fail(e, SGM_ERR_CUDA, "%s failed: %s", lc.failed ? lc.failed : "kernel launch", cudaGetErrorString(lc.err));
    return SGM_OK;
}

int check_images(sgm_engine* e, const void* left, size_t ls, const void* right, size_t rs, const void* sl, size_t sls, const void* sr,
                 size_t srs) {
    if (!e) return SGM_ERR_INVALID;
    if (!left || !right) return fail(e, SGM_ERR_INVALID, "left / right image is null");
    if (ls < (size_t)e->p.width || rs < (size_t)e->p.width) return fail(e, SGM_ERR_INVALID, "image stride smaller than width");
    if ((sl && sls < (size_t)e->iw) || (sr && srs < (size_t)e->iw)) return fail(e, SGM_ERR_INVALID, "sky mask stride smaller than width");
    // integer mode: a masked pixel's fill is 2 * P2 (stand-in of 999999), and 8 * (2 * P2 + P2) must fit 15 bits
    if (e->integer && (sl || sr) && 24 * e->p.P2 > 32767)
        return fail(e, SGM_ERR_UNSUPPORTED, "integer mode with sky masks needs P2 <= 1365 (the masked pixels' stand-in cost 2 * P2 must fit 16-bit sums)");
    if (e->busy) return fail(e, SGM_ERR_STATE, "a frame is already in flight on this handle: call sgm_wait first");
    return SGM_OK;
}

void pack_rows(uint8_t* dst, const uint8_t* src, size_t stride, int rows, int cols) {
    if (stride == (size_t)cols) {
        std::memcpy(dst, src, (size_t)rows * cols);
        return;
    }
    for (int r = 0; r < rows; ++r) std::memcpy(dst + (size_t)r * cols, src + (size_t)r * stride, cols);
}


bool debug_sync() {
#ifdef SGM_B200_TUNING
    static const bool on = std::getenv("SGM_B200_DEBUG_SYNC") != nullptr;  // synchronise after every launch to localise a fault
    return on;
#else
    return false;
#endif
}

// A frame failed after work may already have been queued: nothing of it may be read as a result, and the pinned staging
// must not be overwritten by the next sgm_enqueue while an earlier copy is still in flight.
int abandon_frame(sgm_engine* e, int rc) {
    cudaStreamSynchronize(e->stream);
    (void)cudaGetLastError();
    e->busy = false;
    e->have_result = false;
    return rc;
}

// Common tail of sgm_wait / engine_wait_blocking once the stream has drained (c = what the synchronisation returned).
int finish_frame(sgm_engine* e, cudaError_t c) {
    e->busy = false;
    e->have_result = false;  // set again below, only when the frame is good
    if (c != cudaSuccess) return fail(e, SGM_ERR_CUDA, "frame failed: %s", cudaGetErrorString(c));
    if (e->h_error && *e->h_error) {
        const int code = *e->h_error;
        *e->h_error = 0;
        cudaMemset(e->d_error, 0, sizeof(int));
        return fail(e, SGM_ERR_CUDA, "a device-side wait timed out (code %d: blocks of a cooperative kernel not co-resident?)", code);
    }
    e->have_result = true;
    if (e->timing) {
        for (int k = 0; k < SGM_STAGE_COUNT; ++k) {
            float ms = 0.f;
            const int nxt = k + 1 == SGM_STAGE_COUNT ? SGM_STAGE_COUNT : k + 1;
            if (cudaEventElapsedTime(&ms, e->ev[k], e->ev[nxt]) != cudaSuccess) ms = -1.f;
            e->stage_ms[k] = ms;
        }
        (void)cudaGetLastError();
    }
    return SGM_OK;
}

}  // namespace

namespace sgmb200 {

int engine_stage(sgm_engine* e, const uint8_t* left, size_t ls, const uint8_t* right, size_t rs, const uint8_t* sky_l, size_t sls,
                 const uint8_t* sky_r, size_t srs) {
    int rc = check_images(e, left, ls, right, rs, sky_l, sls, sky_r, srs);
    if (rc != SGM_OK) return rc;
    const int H = e->p.height, W = e->p.width;
    pack_rows(e->h_in[0], left, ls, H, W);
    pack_rows(e->h_in[1], right, rs, H, W);
    if (sky_l) pack_rows(e->h_sky[0], sky_l, sls, e->ih, e->iw);
    if (sky_r) pack_rows(e->h_sky[1], sky_r, srs, e->ih, e->iw);
    e->staged_sky[0] = sky_l != nullptr;
    e->staged_sky[1] = sky_r != nullptr;
    return SGM_OK;
}

int engine_launch_staged(sgm_engine* e) {
    if (!e) return SGM_ERR_INVALID;
    if (e->busy) return fail(e, SGM_ERR_STATE, "a frame is already in flight on this handle: call sgm_wait first");
    CU_TRY(e, cudaSetDevice(e->p.device));
    const size_t HW = (size_t)e->p.height * e->p.width;
    cudaStream_t st = e->stream;
    cudaEventRecord(e->ev[SGM_STAGE_UPLOAD], st);
    CU_TRY(e, cudaMemcpyAsync(e->d_in[0], e->h_in[0], HW, cudaMemcpyHostToDevice, st));
    CU_TRY(e, cudaMemcpyAsync(e->d_in[1], e->h_in[1], HW, cudaMemcpyHostToDevice, st));
    for (int v = 0; v < 2; ++v)
        if (e->staged_sky[v]) CU_TRY(e, cudaMemcpyAsync(e->d_sky[v], e->h_sky[v], e->n, cudaMemcpyHostToDevice, st));
    LaunchCounter lc;
    lc.stream = st;
    lc.sync_each = debug_sync();
    int rc = run_frame(e, e->d_in[0], e->p.width, e->d_in[1], e->p.width, e->staged_sky[0] ? e->d_sky[0] : nullptr,
                       e->staged_sky[1] ? e->d_sky[1] : nullptr, lc);
    e->launches = lc.launches;
    if (rc != SGM_OK) return abandon_frame(e, rc);
    cudaError_t c = cudaMemcpyAsync(e->h_final, e->d_final, e->n * sizeof(float), cudaMemcpyDeviceToHost, st);
    if (c == cudaSuccess && e->host_taps) {
        // the integer and right-view results the reference keeps in its protected Mats (inc/Solver.h:43-46), narrowed on the device
        launch_narrow_disp(e->d_disp, e->d_disp8, (int)(e->views * e->n), st, lc);
        c = lc.err;
        if (c == cudaSuccess) c = cudaMemcpyAsync(e->h_disp8, e->d_disp8, e->views * e->n, cudaMemcpyDeviceToHost, st);
        if (c == cudaSuccess && e->views == 2)
            c = cudaMemcpyAsync(e->h_sub_r, e->d_sub + e->n, e->n * sizeof(float), cudaMemcpyDeviceToHost, st);
        e->launches = lc.launches;
    }
    if (c != cudaSuccess) return abandon_frame(e, fail(e, SGM_ERR_CUDA, "result download: %s", cudaGetErrorString(c)));
    cudaEventRecord(e->ev[SGM_STAGE_COUNT], st);
    cudaEventRecord(e->ev_done, st);
    e->busy = true;
    e->result_on_host = true;
    return SGM_OK;
}

int engine_wait_blocking(sgm_engine* e) {
    if (!e) return SGM_ERR_INVALID;
    if (!e->busy) return SGM_OK;
    CU_TRY(e, cudaSetDevice(e->p.device));
    cudaError_t c = cudaEventSynchronize(e->ev_done);          // sleeps instead of spinning
    if (c == cudaSuccess) c = cudaStreamSynchronize(e->stream);  // already drained: returns at once
    return finish_frame(e, c);
}

const float* engine_host_result(const sgm_engine* e) { return (e && e->have_result && e->result_on_host) ? e->h_final : nullptr; }

}  // namespace sgmb200

extern "C" {

int sgm_enqueue(sgm_engine* e, const uint8_t* left, size_t ls, const uint8_t* right, size_t rs, const uint8_t* sky_l, size_t sls,
                const uint8_t* sky_r, size_t srs) {
    int rc = engine_stage(e, left, ls, right, rs, sky_l, sls, sky_r, srs);
    if (rc != SGM_OK) return rc;
    return engine_launch_staged(e);
}

int sgm_enqueue_device(sgm_engine* e, const void* left, size_t ls, const void* right, size_t rs, const void* sky_l, size_t sls,
                       const void* sky_r, size_t srs) {
    int rc = check_images(e, left, ls, right, rs, sky_l, sls, sky_r, srs);
    if (rc != SGM_OK) return rc;
    CU_TRY(e, cudaSetDevice(e->p.device));
    cudaStream_t st = e->stream;
    const uint8_t* sl = static_cast<const uint8_t*>(sky_l);
    const uint8_t* sr = static_cast<const uint8_t*>(sky_r);
    cudaEventRecord(e->ev[SGM_STAGE_UPLOAD], st);
    if (sl && sls != (size_t)e->iw) {  // densify a pitched mask
        CU_TRY(e, cudaMemcpy2DAsync(e->d_sky[0], e->iw, sl, sls, e->iw, e->ih, cudaMemcpyDeviceToDevice, st));
        sl = e->d_sky[0];
    }
    if (sr && srs != (size_t)e->iw) {
        CU_TRY(e, cudaMemcpy2DAsync(e->d_sky[1], e->iw, sr, srs, e->iw, e->ih, cudaMemcpyDeviceToDevice, st));
        sr = e->d_sky[1];
    }
    // The left image is read again after the frame (sgm_point_cloud takes its grey values): keep an engine-owned copy, so
    // the caller may release or overwrite its buffer as soon as sgm_wait has returned.
    const size_t W = (size_t)e->p.width;
    CU_TRY(e, cudaMemcpy2DAsync(e->d_in[0], W, left, ls, W, (size_t)e->p.height, cudaMemcpyDeviceToDevice, st));
    LaunchCounter lc;
    lc.stream = st;
    lc.sync_each = debug_sync();
    rc = run_frame(e, e->d_in[0], W, static_cast<const uint8_t*>(right), rs, sl, sr, lc);
    e->launches = lc.launches;
    if (rc != SGM_OK) return abandon_frame(e, rc);
    cudaEventRecord(e->ev[SGM_STAGE_COUNT], st);
    cudaEventRecord(e->ev_done, st);
    e->busy = true;
    e->result_on_host = false;
    return SGM_OK;
}

int sgm_wait(sgm_engine* e) {
    if (!e) return SGM_ERR_INVALID;
    if (!e->busy) return SGM_OK;
    CU_TRY(e, cudaSetDevice(e->p.device));
    cudaError_t c = cudaStreamSynchronize(e->stream);
    return finish_frame(e, c);
}

int sgm_process(sgm_engine* e, const uint8_t* left, size_t ls, const uint8_t* right, size_t rs, const uint8_t* sky_l, size_t sls,
                const uint8_t* sky_r, size_t srs) {
    int rc = sgm_enqueue(e, left, ls, right, rs, sky_l, sls, sky_r, srs);
    if (rc != SGM_OK) return rc;
    return sgm_wait(e);
}

static int ready(sgm_engine* e) {
    if (!e) return SGM_ERR_INVALID;
    if (e->busy) {
        int rc = sgm_wait(e);
        if (rc != SGM_OK) return rc;
    }
    if (!e->have_result) return fail(e, SGM_ERR_STATE, "no frame has been processed yet");
    if (cudaSetDevice(e->p.device) != cudaSuccess) return fail(e, SGM_ERR_CUDA, "cudaSetDevice failed");
    return SGM_OK;
}

int sgm_get_disparity_f32(sgm_engine* e, float* dst, size_t dst_stride) {
    int rc = ready(e);
    if (rc != SGM_OK) return rc;
    if (!dst || dst_stride < e->iw * sizeof(float)) return fail(e, SGM_ERR_INVALID, "bad destination");
    if (e->result_on_host) {
        for (int r = 0; r < e->ih; ++r)
            std::memcpy(reinterpret_cast<char*>(dst) + (size_t)r * dst_stride, e->h_final + (size_t)r * e->iw, e->iw * sizeof(float));
        return SGM_OK;
    }
    CU_TRY(e, cudaMemcpy2D(dst, dst_stride, e->d_final, e->iw * sizeof(float), e->iw * sizeof(float), e->ih, cudaMemcpyDeviceToHost));
    return SGM_OK;
}

int sgm_get_disparity_u16(sgm_engine* e, int view, uint16_t* dst, size_t dst_stride) {
    int rc = ready(e);
    if (rc != SGM_OK) return rc;
    if (!dst || dst_stride < e->iw * sizeof(uint16_t) || view < 0 || view >= e->views) return fail(e, SGM_ERR_INVALID, "bad destination / view");
    CU_TRY(e, cudaMemcpy2D(dst, dst_stride, e->d_disp + (size_t)view * e->n, e->iw * sizeof(uint16_t), e->iw * sizeof(uint16_t), e->ih,
                           cudaMemcpyDeviceToHost));
    return SGM_OK;
}

int sgm_get_disparity_u8(sgm_engine* e, int view, uint8_t* dst, size_t dst_stride) {
    int rc = ready(e);
    if (rc != SGM_OK) return rc;
    if (!dst || dst_stride < (size_t)e->iw || view < 0 || view >= e->views) return fail(e, SGM_ERR_INVALID, "bad destination / view");
    if (e->host_taps && e->result_on_host) {  // already on the host: narrowed on the device, copied with the result
        // D = 256: the byte saturates at 255 (the reference's uchar Mats cannot hold D+1 = 257 either, src/Solver.cpp:16-19)
        for (int r = 0; r < e->ih; ++r) std::memcpy(dst + (size_t)r * dst_stride, e->h_disp8 + (size_t)view * e->n + (size_t)r * e->iw, e->iw);
        return SGM_OK;
    }
    if (e->D > 128) return fail(e, SGM_ERR_UNSUPPORTED, "8-bit disparity needs max_disp <= 128 (invalid = D+1 must fit a byte) or SGM_FLAG_HOST_TAPS (saturating)");
    std::vector<uint16_t> tmp(e->n);
    CU_TRY(e, cudaMemcpy(tmp.data(), e->d_disp + (size_t)view * e->n, e->n * sizeof(uint16_t), cudaMemcpyDeviceToHost));
    for (int r = 0; r < e->ih; ++r)
        for (int k = 0; k < e->iw; ++k) dst[(size_t)r * dst_stride + k] = (uint8_t)tmp[(size_t)r * e->iw + k];  // uchar Mat, src/Solver.cpp:18-19
    return SGM_OK;
}

int sgm_get_subpixel_beta(sgm_engine* e, float* dst, size_t dst_stride) {
    int rc = ready(e);
    if (rc != SGM_OK) return rc;
    if (!dst || dst_stride < e->iw * sizeof(float)) return fail(e, SGM_ERR_INVALID, "bad destination");
    if (e->views < 2) return fail(e, SGM_ERR_STATE, "this mode has no right view");
    if (e->host_taps && e->result_on_host) {
        for (int r = 0; r < e->ih; ++r)
            std::memcpy(reinterpret_cast<char*>(dst) + (size_t)r * dst_stride, e->h_sub_r + (size_t)r * e->iw, e->iw * sizeof(float));
        return SGM_OK;
    }
    CU_TRY(e, cudaMemcpy2D(dst, dst_stride, e->d_sub + e->n, e->iw * sizeof(float), e->iw * sizeof(float), e->ih, cudaMemcpyDeviceToHost));
    return SGM_OK;
}

const float* sgm_disparity_device(sgm_engine* e) { return e ? e->d_final : nullptr; }

size_t sgm_tap_bytes(const sgm_engine* e, int which) {
    if (!e) return 0;
    switch (which) {
        case SGM_TAP_CENSUS: return e->n * sizeof(uint64_t);
        case SGM_TAP_COST:
        case SGM_TAP_PATHSUM: return e->N * sizeof(float);
        case SGM_TAP_DISP: return e->n * sizeof(uint16_t);
        case SGM_TAP_SUBPIXEL:
        case SGM_TAP_LR:
        case SGM_TAP_MEDIAN:
        case SGM_TAP_FINAL: return e->n * sizeof(float);
        case SGM_TAP_BLUR: return e->n;
        default: return 0;
    }
}

int sgm_get_tap(sgm_engine* e, int which, int view, void* dst, size_t bytes) {
    int rc = ready(e);
    if (rc != SGM_OK) return rc;
    if (!dst || bytes != sgm_tap_bytes(e, which) || bytes == 0) return fail(e, SGM_ERR_INVALID, "tap %d: bad destination size", which);
    if (view < 0 || view > 1) return fail(e, SGM_ERR_INVALID, "view must be 0 or 1");
    const void* src = nullptr;
    switch (which) {
        case SGM_TAP_CENSUS: src = e->p.algo == SGM_ALGO_BM_SAD ? nullptr : e->d_census + (size_t)view * e->n; break;
        case SGM_TAP_BLUR: src = e->p.algo == SGM_ALGO_BM_SAD ? nullptr : e->d_blur + (size_t)view * e->n; break;
        case SGM_TAP_COST:
            // SGM mode: acc/cost buffers are scratch but the filtered cost volume itself is never overwritten
            src = view < e->views ? e->d_cost + (size_t)view * e->vol_words : nullptr;
            break;
        case SGM_TAP_PATHSUM: src = (e->d_pathsum && view < e->views) ? e->d_pathsum + (size_t)view * e->vol_words : nullptr; break;
        case SGM_TAP_DISP: src = view < e->views ? e->d_disp + (size_t)view * e->n : nullptr; break;
        case SGM_TAP_SUBPIXEL: src = view < e->views ? e->d_sub + (size_t)view * e->n : nullptr; break;
        case SGM_TAP_LR: src = view == 0 ? e->d_lr : nullptr; break;
        case SGM_TAP_MEDIAN: src = view == 0 ? e->d_med : nullptr; break;
        case SGM_TAP_FINAL: src = view == 0 ? e->d_final : nullptr; break;
        default: break;
    }
    if (!src) return fail(e, SGM_ERR_STATE, "tap %d / view %d is not available in this mode", which, view);
    if (e->integer && (which == SGM_TAP_COST || which == SGM_TAP_PATHSUM)) {
        // integer mode: the volume holds 16-bit values; the tap keeps its float layout (every value is exact in float32)
        CU_TRY(e, cudaMemcpy(dst, src, bytes / 2, cudaMemcpyDeviceToHost));
        const uint16_t* in = static_cast<const uint16_t*>(dst);
        float* out = static_cast<float*>(dst);
        for (size_t t = e->N; t-- > 0;) out[t] = (float)in[t];
        return SGM_OK;
    }
    CU_TRY(e, cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
    return SGM_OK;
}

int sgm_get_size(const sgm_engine* e, int* out_h, int* out_w, int* out_d) {
    if (!e) return SGM_ERR_INVALID;
    if (out_h) *out_h = e->ih;
    if (out_w) *out_w = e->iw;
    if (out_d) *out_d = e->D;
    return SGM_OK;
}
int sgm_get_launch_count(const sgm_engine* e) { return e ? e->launches : -1; }
int sgm_get_stage_ms(sgm_engine* e, float* ms, int count) {
    int rc = ready(e);
    if (rc != SGM_OK) return rc;
    if (!e->timing) return fail(e, SGM_ERR_STATE, "engine was created without SGM_FLAG_STAGE_TIMING");
    for (int k = 0; k < count && k < SGM_STAGE_COUNT; ++k) ms[k] = e->stage_ms[k];
    return SGM_OK;
}
void* sgm_get_stream(sgm_engine* e) { return e ? (void*)e->stream : nullptr; }
size_t sgm_device_bytes(const sgm_engine* e) { return e ? e->device_bytes : 0; }

int sgm_render_debug_view(sgm_engine* e, uint8_t* dst, size_t dst_stride) {
    int rc = ready(e);
    if (rc != SGM_OK) return rc;
    const size_t row_bytes = (size_t)e->iw * 3;
    if (!dst || dst_stride < row_bytes) return fail(e, SGM_ERR_INVALID, "bad destination");
    if (!e->d_view) CU_TRY(e, dev_alloc(e, &e->d_view, 2 * e->n * 3));
    LaunchCounter lc;
    lc.stream = e->stream;
    launch_debug_view(e->d_img, e->d_final, e->d_view, e->ih, e->iw, e->D, e->stream, lc);
    if (lc.err != cudaSuccess) return fail(e, SGM_ERR_CUDA, "debug view: %s", cudaGetErrorString(lc.err));
    CU_TRY(e, cudaMemcpy2DAsync(dst, dst_stride, e->d_view, row_bytes, row_bytes, 2 * (size_t)e->ih, cudaMemcpyDeviceToHost, e->stream));
    CU_TRY(e, cudaStreamSynchronize(e->stream));
    return SGM_OK;
}

int sgm_point_cloud(sgm_engine* e, const sgm_camera* cam, double* xyz, uint8_t* intensity, size_t capacity, size_t* count) {
    int rc = ready(e);
    if (rc != SGM_OK) return rc;
    if (!cam || !count) return fail(e, SGM_ERR_INVALID, "camera / count is null");
    const int nb = cloud_blocks((int)e->n);
    if (!e->d_pc_counts) {
        CU_TRY(e, dev_alloc(e, &e->d_pc_counts, (size_t)nb));
        CU_TRY(e, dev_alloc(e, &e->d_pc_offsets, (size_t)nb + 1));
        CU_TRY(e, dev_alloc(e, &e->d_pc_xyz, e->n * 3));
        CU_TRY(e, dev_alloc(e, &e->d_pc_int, e->n));
    }
    LaunchCounter lc;
    lc.stream = e->stream;
    launch_point_cloud(e->d_final, e->d_in[0], (size_t)e->p.width, e->ih, e->iw, cam->fx, cam->fy, cam->cx, cam->cy, cam->baseline,
                       cam->max_range, (float)(e->D + 1), e->p.scale, e->d_pc_counts, e->d_pc_offsets, e->d_pc_xyz, e->d_pc_int, (int)e->n,
                       e->stream, lc);
    if (lc.err != cudaSuccess) return fail(e, SGM_ERR_CUDA, "point cloud: %s", cudaGetErrorString(lc.err));
    int total = 0;
    CU_TRY(e, cudaMemcpyAsync(&total, e->d_pc_offsets + nb, sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    CU_TRY(e, cudaStreamSynchronize(e->stream));
    *count = (size_t)total;
    const size_t take = (size_t)total < capacity ? (size_t)total : capacity;
    if (xyz && take) CU_TRY(e, cudaMemcpy(xyz, e->d_pc_xyz, take * 3 * sizeof(double), cudaMemcpyDeviceToHost));
    if (intensity && take) CU_TRY(e, cudaMemcpy(intensity, e->d_pc_int, take, cudaMemcpyDeviceToHost));
    return SGM_OK;
}

int sgm_speckle_filter_static(int device, const float* in_device, float* out_device, int height, int width, float invalid_value,
                              int max_size, float max_dis, void* stream) {
    static_assert(SPECKLE_STATIC_MAX_PIXELS == SGM_SPECKLE_STATIC_MAX_PIXELS, "header and kernels disagree on the static scratch size");
    if (!in_device || !out_device || height <= 0 || width <= 0 || max_size < 0)
        return fail(nullptr, SGM_ERR_INVALID, "sgm_speckle_filter_static: bad argument");
    if ((size_t)height * (size_t)width > (size_t)SGM_SPECKLE_STATIC_MAX_PIXELS)
        return fail(nullptr, SGM_ERR_UNSUPPORTED, "sgm_speckle_filter_static: %d x %d = %zu pixels exceed the static scratch (%d pixels); use sgm_process",
                    width, height, (size_t)height * (size_t)width, (int)SGM_SPECKLE_STATIC_MAX_PIXELS);
    if (cudaSetDevice(device) != cudaSuccess) return fail(nullptr, SGM_ERR_CUDA, "no CUDA device %d: this engine has no CPU path", device);
    LaunchCounter lc;
    lc.stream = static_cast<cudaStream_t>(stream);
    cudaError_t c = launch_speckle_static(device, in_device, out_device, height, width, invalid_value, max_size, max_dis, lc.stream, lc);
    if (c != cudaSuccess)
        return fail(nullptr, SGM_ERR_CUDA, "sgm_speckle_filter_static: %s failed: %s", lc.failed ? lc.failed : "set-up", cudaGetErrorString(c));
    return SGM_OK;
}

int sgm_cost_volume_sad(int device, const uint8_t* left, const uint8_t* right, int height, int width, int max_disp, int win_h, int win_w,
                        int scale, float* out) {
    if (!left || !right || !out || height <= 0 || width <= 0 || max_disp <= 0 || scale <= 0 || win_h <= 0 || win_w <= 0)
        return fail(nullptr, SGM_ERR_INVALID, "sgm_cost_volume_sad: bad argument");
    if ((size_t)(32 + 2 * (win_w / 2)) * max_disp * 2 > 48 * 1024 || win_h > 257)
        return fail(nullptr, SGM_ERR_UNSUPPORTED, "sgm_cost_volume_sad: window too large for this max_disp");
    if (cudaSetDevice(device) != cudaSuccess) return fail(nullptr, SGM_ERR_CUDA, "no CUDA device %d: this engine has no CPU path", device);
    const size_t n = (size_t)height * width, N = n * max_disp;
    uint8_t* d_img = nullptr;
    float* d_cost = nullptr;
    cudaError_t c = cudaMalloc(reinterpret_cast<void**>(&d_img), 2 * n);
    if (c == cudaSuccess) c = cudaMalloc(reinterpret_cast<void**>(&d_cost), N * sizeof(float));
    if (c == cudaSuccess) c = cudaMemcpy(d_img, left, n, cudaMemcpyHostToDevice);
    if (c == cudaSuccess) c = cudaMemcpy(d_img + n, right, n, cudaMemcpyHostToDevice);
    float* d_lut = nullptr;
    const std::vector<float> lut = sad_division_table(win_h, win_w);
    if (c == cudaSuccess) c = cudaMalloc(reinterpret_cast<void**>(&d_lut), lut.size() * sizeof(float));
    if (c == cudaSuccess) c = cudaMemcpy(d_lut, lut.data(), lut.size() * sizeof(float), cudaMemcpyHostToDevice);
    LaunchCounter lc;
    if (c == cudaSuccess) {
        launch_sad_volume(d_img, d_img + n, d_lut, d_cost, 1, height, width, max_disp, win_h, win_w, scale, nullptr, lc);
        c = lc.err;
    }
    if (c == cudaSuccess) c = cudaMemcpy(out, d_cost, N * sizeof(float), cudaMemcpyDeviceToHost);
    cudaFree(d_img);
    cudaFree(d_cost);
    cudaFree(d_lut);
    if (c != cudaSuccess) return fail(nullptr, SGM_ERR_CUDA, "sgm_cost_volume_sad: %s", cudaGetErrorString(c));
    return SGM_OK;
}

namespace {
int window_cost_volume(int kind, const char* name, int device, const uint8_t* left, const uint8_t* right, int height, int width, int max_disp,
                       int win_h, int win_w, int scale, float* out) {
    if (!left || !right || !out || height <= 0 || width <= 0 || max_disp <= 0 || scale <= 0 || win_h <= 0 || win_w <= 0)
        return fail(nullptr, SGM_ERR_INVALID, "%s: bad argument", name);
    if (cudaSetDevice(device) != cudaSuccess) return fail(nullptr, SGM_ERR_CUDA, "no CUDA device %d: this engine has no CPU path", device);
    const size_t n = (size_t)height * width, N = n * max_disp;
    uint8_t* d_img = nullptr;
    float* d_cost = nullptr;
    cudaError_t c = cudaMalloc(reinterpret_cast<void**>(&d_img), 2 * n);
    if (c == cudaSuccess) c = cudaMalloc(reinterpret_cast<void**>(&d_cost), N * sizeof(float));
    if (c == cudaSuccess) c = cudaMemcpy(d_img, left, n, cudaMemcpyHostToDevice);
    if (c == cudaSuccess) c = cudaMemcpy(d_img + n, right, n, cudaMemcpyHostToDevice);
    LaunchCounter lc;
    if (c == cudaSuccess) {
        launch_window_cost_volume(kind, d_img, d_img + n, d_cost, height, width, max_disp, win_h, win_w, scale, nullptr, lc);
        c = lc.err;
    }
    const bool rejected = c == cudaErrorInvalidValue && lc.failed;
    if (c == cudaSuccess) c = cudaMemcpy(out, d_cost, N * sizeof(float), cudaMemcpyDeviceToHost);
    cudaFree(d_img);
    cudaFree(d_cost);
    if (rejected) return fail(nullptr, SGM_ERR_UNSUPPORTED, "%s: %s", name, lc.failed);
    if (c != cudaSuccess) return fail(nullptr, SGM_ERR_CUDA, "%s: %s", name, cudaGetErrorString(c));
    return SGM_OK;
}
}  // namespace

int sgm_cost_volume_ssd(int device, const uint8_t* left, const uint8_t* right, int height, int width, int max_disp, int win_h, int win_w,
                        int scale, float* out) {
    return window_cost_volume(1, "sgm_cost_volume_ssd", device, left, right, height, width, max_disp, win_h, win_w, scale, out);
}

int sgm_cost_volume_ct(int device, const uint8_t* left, const uint8_t* right, int height, int width, int max_disp, int win_h, int win_w,
                       int scale, float* out) {
    return window_cost_volume(2, "sgm_cost_volume_ct", device, left, right, height, width, max_disp, win_h, win_w, scale, out);
}

}  // extern "C"
