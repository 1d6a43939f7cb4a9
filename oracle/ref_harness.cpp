// oracle/_ref harness — TEST INFRASTRUCTURE, not product code.
//
// Wraps the UNMODIFIED reference CPU sources (/root/reference/src/{Solver,SGM,BM,cost}.cpp,
// compiled where they lie by oracle/Makefile against the stand-in headers in oracle/shim/)
// behind a flat C interface so that tests/ and bench.py's cpu_baseline leg can run the real
// reference through ctypes.  Nothing from the reference is copied: this file only derives from
// its classes to reach protected members and re-run protected stages for intermediate taps.
//
// Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) may load
// the resulting oracle/_ref/libsgm_ref.so.
#include <fcntl.h>
#include <omp.h>
#include <unistd.h>

#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>

#include "inc/BM.h"
#include "inc/SGM.h"

// ---- symbols the reference expects from translation units we do not build -------------------
// src/utils.cpp (needs highgui/VideoCapture) -> only get_cur_ms() is used on the path.
double get_cur_ms() {
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

// LKRefine/ (needs Eigen): Solver's ctor calls LKSubPixel::create (src/Solver.cpp:51) but the
// refinement itself is never invoked (src/SGM.cpp:823-824 is commented out).
namespace {
class NoopLK : public LKSubPixel {
public:
    NoopLK(int h, int w, int s, int d) : LKSubPixel(h, w, s, d) {}
    void LKRefine(const cv::Mat&, const cv::Mat&, cv::Mat&) override {}
};
}  // namespace
LKSubPixel::LKSubPixel(int, int, int, int) {}
LKSubPixel::~LKSubPixel() {}
std::shared_ptr<LKSubPixel> LKSubPixel::create(int h, int w, int s, int d) { return std::make_shared<NoopLK>(h, w, s, d); }

namespace {

// Reaches the protected state and stages of Solver through a derived class.
template <class Base>
struct Probe : public Base {
    Probe(int h, int w, int s, int d) : Base(h, w, s, d) {}
    using Solver::build_cost_table;
    using Solver::build_dsi_from_table;
    using Solver::build_dsi_from_table_beta;
    using Solver::cost;
    using Solver::cost_horizontal_filter;
    using Solver::cost_table_l;
    using Solver::cost_table_r;
    using Solver::cost_vertical_filter;
    using Solver::disp;
    using Solver::disp_beta;
    using Solver::filtered_disp;
    using Solver::filtered_disp_beta;
    using Solver::img_h;
    using Solver::img_l;
    using Solver::img_r;
    using Solver::img_w;
    using Solver::max_disp;
    using Solver::scale;
    using Solver::sky_mask;
    using Solver::sky_mask_beta;
};

bool g_quiet = true;

// The reference printf()s per-stage timings; keep them off the caller's stdout when quiet.
struct StdoutMute {
    int saved = -1;
    StdoutMute() {
        if (!g_quiet) return;
        fflush(stdout);
        saved = dup(1);
        int nul = open("/dev/null", O_WRONLY);
        if (nul >= 0) {
            dup2(nul, 1);
            close(nul);
        }
    }
    ~StdoutMute() {
        if (saved < 0) return;
        fflush(stdout);
        dup2(saved, 1);
        close(saved);
    }
};

void copy_mat_u8(const cv::Mat& m, uint8_t* dst) {
    if (!dst) return;
    for (int r = 0; r < m.rows; ++r) std::memcpy(dst + (size_t)r * m.cols, m.ptr<uchar>(r), m.cols);
}
void copy_mat_f32(const cv::Mat& m, float* dst) {
    if (!dst) return;
    for (int r = 0; r < m.rows; ++r) std::memcpy(dst + (size_t)r * m.cols, m.ptr<float>(r), sizeof(float) * m.cols);
}

template <class P>
void subsample_into(P& sv, cv::Mat& L, cv::Mat& R, bool bm_rows) {
    // Mirrors what process() does to its inputs so the stage taps see the same images:
    // SGM reads source row i*scale (src/SGM.cpp:46-47), BM reads source row i (src/BM.cpp:23-24).
    if (sv.scale > 1) {
        sv.img_l.create(sv.img_h, sv.img_w, CV_8UC1);
        sv.img_r.create(sv.img_h, sv.img_w, CV_8UC1);
        for (int i = 0; i < sv.img_h; ++i) {
            const uchar* pl = L.ptr<uchar>(bm_rows ? i : i * sv.scale);
            const uchar* pr = R.ptr<uchar>(bm_rows ? i : i * sv.scale);
            for (int j = 0; j < sv.img_w; ++j) {
                sv.img_l.template ptr<uchar>(i)[j] = pl[j * sv.scale];
                sv.img_r.template ptr<uchar>(i)[j] = pr[j * sv.scale];
            }
        }
    } else {
        sv.img_l = L;
        sv.img_r = R;
    }
}

}  // namespace

extern "C" {

void ref_set_quiet(int q) { g_quiet = q != 0; }
void ref_set_threads(int n) {
    if (n > 0) omp_set_num_threads(n);
}
int ref_max_threads() { return omp_get_max_threads(); }

// Full reference run.  algo: 0 = SGM::process, 1 = BM::process.
// left/right: h*w bytes (full resolution, contiguous).  sky_l/sky_r: (h/scale)*(w/scale) bytes or NULL.
// Outputs (each nullable) are (h/scale)*(w/scale) elements; cost_last receives the volume left in
// Solver::cost when process() returns (SGM: summed path volume S of the beta pass; BM: filtered cost).
// Returns elapsed milliseconds of process() (>= 0), or -1 on bad arguments.
double ref_process(int algo, int h, int w, int scale, int d, const uint8_t* left, const uint8_t* right,
                   const uint8_t* sky_l, const uint8_t* sky_r, uint8_t* disp, uint8_t* disp_beta, float* fdisp,
                   float* fdisp_beta, float* cost_last) {
    if (!left || !right || h <= 0 || w <= 0) return -1.0;
    StdoutMute mute;
    cv::Mat L(h, w, CV_8UC1), R(h, w, CV_8UC1);
    std::memcpy(L.data, left, (size_t)h * w);
    std::memcpy(R.data, right, (size_t)h * w);
    const int ih = h / scale, iw = w / scale;
    cv::Mat SL, SR;
    if (sky_l) {
        SL.create(ih, iw, CV_8UC1);
        std::memcpy(SL.data, sky_l, (size_t)ih * iw);
    }
    if (sky_r) {
        SR.create(ih, iw, CV_8UC1);
        std::memcpy(SR.data, sky_r, (size_t)ih * iw);
    }
    double ms = 0;
    auto run = [&](auto& sv) {
        auto t0 = std::chrono::steady_clock::now();
        sv.process(L, R, SL, SR);
        ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        copy_mat_u8(sv.disp, disp);
        copy_mat_u8(sv.disp_beta, disp_beta);
        copy_mat_f32(sv.get_disp(), fdisp);
        copy_mat_f32(sv.filtered_disp_beta, fdisp_beta);
        if (cost_last) std::memcpy(cost_last, sv.cost, sizeof(float) * (size_t)ih * iw * d);
    };
    if (algo == 0) {
        Probe<SGM> sv(h, w, scale, d);
        run(sv);
    } else if (algo == 1) {
        Probe<BM> sv(h, w, scale, d);
        run(sv);
    } else {
        return -1.0;
    }
    return ms;
}

// Solver::show_disp after one process(): debug_view_out is (2*h/scale) x (w/scale) x 3 BGR (pins oracle/restate.py's
// colormap / debug_view and the device rendering).  Returns 0, or -1 on bad arguments.
int ref_show_disp(int algo, int h, int w, int scale, int d, const uint8_t* left, const uint8_t* right, uint8_t* debug_view_out,
                  float* fdisp) {
    if (!left || !right || !debug_view_out || h <= 0 || w <= 0) return -1;
    StdoutMute mute;
    cv::Mat L(h, w, CV_8UC1), R(h, w, CV_8UC1);
    std::memcpy(L.data, left, (size_t)h * w);
    std::memcpy(R.data, right, (size_t)h * w);
    const int ih = h / scale, iw = w / scale;
    auto run = [&](auto& sv) {
        sv.process(L, R);
        cv::Mat view;
        sv.show_disp(view);
        for (int i = 0; i < 2 * ih; ++i) std::memcpy(debug_view_out + (size_t)i * iw * 3, view.template ptr<uint8_t>(i), (size_t)iw * 3);
        if (fdisp) copy_mat_f32(sv.get_disp(), fdisp);
    };
    if (algo == 0) {
        Probe<SGM> sv(h, w, scale, d);
        run(sv);
    } else if (algo == 1) {
        Probe<BM> sv(h, w, scale, d);
        run(sv);
    } else {
        return -1;
    }
    return 0;
}

// Times `iters` back-to-back process() calls on one Solver instance (as node.cpp would reuse it);
// ms_out[k] = wall time of call k.  Used by bench.py's cpu_baseline / --impl reference legs.
int ref_time_process(int algo, int h, int w, int scale, int d, const uint8_t* left, const uint8_t* right, int iters,
                     double* ms_out) {
    if (!left || !right || iters <= 0 || !ms_out) return -1;
    StdoutMute mute;
    cv::Mat L(h, w, CV_8UC1), R(h, w, CV_8UC1);
    std::memcpy(L.data, left, (size_t)h * w);
    std::memcpy(R.data, right, (size_t)h * w);
    auto loop = [&](auto& sv) {
        for (int k = 0; k < iters; ++k) {
            cv::Mat l = L.clone(), r = R.clone();  // show_disp()/process may alias the inputs
            auto t0 = std::chrono::steady_clock::now();
            sv.process(l, r);
            ms_out[k] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        }
    };
    if (algo == 0) {
        Probe<SGM> sv(h, w, scale, d);
        loop(sv);
    } else if (algo == 1) {
        Probe<BM> sv(h, w, scale, d);
        loop(sv);
    } else {
        return -1;
    }
    return 0;
}

// Stage taps: re-runs the protected stages of Solver in the order process() uses them
// (src/SGM.cpp:66-72 and :449-454) and copies the intermediate buffers out.
// beta = 0: left-view volume, 1: right-view ("beta") volume.  All outputs nullable.
// A solver kept across frames, the way node.cpp:49,93 uses it (one make_shared<SGM>, process() per frame): bench.py's reference
// arm times frame after frame of a stream of DIFFERENT pairs without paying the 2 GB allocation (and its first-touch page
// faults) per frame.  ref_open -> handle; ref_run -> wall milliseconds of one process() (optionally copies get_disp() out).
struct RefSession {
    int algo, h, w, scale, d;
    Probe<SGM>* sgm = nullptr;
    Probe<BM>* bm = nullptr;
};
void* ref_open(int algo, int h, int w, int scale, int d) {
    if (algo != 0 && algo != 1) return nullptr;
    StdoutMute mute;
    RefSession* s = new RefSession{algo, h, w, scale, d};
    if (algo == 0)
        s->sgm = new Probe<SGM>(h, w, scale, d);
    else
        s->bm = new Probe<BM>(h, w, scale, d);
    return s;
}
double ref_run(void* session, const uint8_t* left, const uint8_t* right, float* fdisp_out) {
    RefSession* s = static_cast<RefSession*>(session);
    if (!s || !left || !right) return -1.0;
    StdoutMute mute;
    cv::Mat L(s->h, s->w, CV_8UC1), R(s->h, s->w, CV_8UC1);
    std::memcpy(L.data, left, (size_t)s->h * s->w);
    std::memcpy(R.data, right, (size_t)s->h * s->w);
    auto t0 = std::chrono::steady_clock::now();
    if (s->sgm)
        s->sgm->process(L, R);
    else
        s->bm->process(L, R);
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (fdisp_out) copy_mat_f32(s->sgm ? s->sgm->get_disp() : s->bm->get_disp(), fdisp_out);
    return ms;
}
void ref_close(void* session) {
    RefSession* s = static_cast<RefSession*>(session);
    if (!s) return;
    delete s->sgm;
    delete s->bm;
    delete s;
}

int ref_stages(int algo, int h, int w, int scale, int d, const uint8_t* left, const uint8_t* right, const uint8_t* sky,
               int beta, uint64_t* ct_l, uint64_t* ct_r, float* cost_raw, float* cost_h, float* cost_hv) {
    if (!left || !right) return -1;
    StdoutMute mute;
    cv::Mat L(h, w, CV_8UC1), R(h, w, CV_8UC1);
    std::memcpy(L.data, left, (size_t)h * w);
    std::memcpy(R.data, right, (size_t)h * w);
    const int ih = h / scale, iw = w / scale;
    const size_t n = (size_t)ih * iw, nd = n * d;
    auto run = [&](auto& sv, bool bm_rows) {
        subsample_into(sv, L, R, bm_rows);
        if (sky) {
            cv::Mat S(ih, iw, CV_8UC1);
            std::memcpy(S.data, sky, n);
            if (beta)
                sv.sky_mask_beta = S;
            else
                sv.sky_mask = S;
        }
        sv.build_cost_table();
        if (ct_l) std::memcpy(ct_l, sv.cost_table_l, n * sizeof(uint64_t));
        if (ct_r) std::memcpy(ct_r, sv.cost_table_r, n * sizeof(uint64_t));
        if (beta)
            sv.build_dsi_from_table_beta();
        else
            sv.build_dsi_from_table();
        if (cost_raw) std::memcpy(cost_raw, sv.cost, nd * sizeof(float));
        sv.cost_horizontal_filter(COST_WIN_W / scale);
        if (cost_h) std::memcpy(cost_h, sv.cost, nd * sizeof(float));
        sv.cost_vertical_filter(COST_WIN_H / scale);
        if (cost_hv) std::memcpy(cost_hv, sv.cost, nd * sizeof(float));
    };
    if (algo == 0) {
        Probe<SGM> sv(h, w, scale, d);
        run(sv, false);
    } else if (algo == 1) {
        Probe<BM> sv(h, w, scale, d);
        run(sv, true);
    } else {
        return -1;
    }
    return 0;
}

// The reference's free cost functions (inc/cost.h:7-14) on whole images, for the SAD/SSD/CT rows.
// kind: 0 = SAD, 1 = SSD, 2 = CT.  out: h*w*d floats laid out [i][j][d].
int ref_cost_volume(int kind, int h, int w, int d, int win_h, int win_w, int scale, const uint8_t* left,
                    const uint8_t* right, float* out) {
    if (!left || !right || !out) return -1;
    cv::Mat L(h, w, CV_8UC1), R(h, w, CV_8UC1);
    std::memcpy(L.data, left, (size_t)h * w);
    std::memcpy(R.data, right, (size_t)h * w);
#pragma omp parallel for
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j)
            for (int k = 0; k < d; ++k) {
                float v;
                if (kind == 0)
                    v = SAD(L, R, cv::Point(j, i), k, win_h, win_w, nullptr, scale);
                else if (kind == 1)
                    v = SSD(L, R, cv::Point(j, i), k, win_h, win_w, nullptr, scale);
                else
                    v = (float)CT(L, R, cv::Point(j, i), k, win_h, win_w, nullptr, scale);
                out[((size_t)i * w + j) * d + k] = v;
            }
    return 0;
}

int ref_hamming_cost(uint64_t a, uint64_t b) { return hamming_cost(a, b); }

// The shim's GaussianBlur on its own, so tests can pin it against cv2.
int ref_gaussian_blur(int h, int w, const uint8_t* src, uint8_t* dst, double sx, double sy) {
    cv::Mat S(h, w, CV_8UC1), D;
    std::memcpy(S.data, src, (size_t)h * w);
    cv::GaussianBlur(S, D, cv::Size(3, 3), sx, sy);
    std::memcpy(dst, D.data, (size_t)h * w);
    return 0;
}

}  // extern "C"
