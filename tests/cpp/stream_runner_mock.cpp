// stream_runner_mock.cpp — CPU test of the frame-stream runner's host logic (stereo_matching_b200/csrc/stream_runner.cu compiled
// as plain C++) against a MOCK engine: no GPU, no kernels.  The mock "device" is a thread-free fake: engine_launch_staged marks the
// handle busy and remembers the staged frame's first bytes, engine_wait_blocking sleeps a little and writes a result derived
// from them.  What is tested is the runner: ordering, back-pressure, views held / released, producer and consumer on two
// threads, one thread alone, several device lanes, failing frames.  TEST INFRASTRUCTURE (tests/test_stream_cpu.py builds and runs it).
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "../../stereo_matching_b200/csrc/engine_internal.h"

// ---- the few CUDA runtime symbols stream_runner.cu touches ----
extern "C" cudaError_t cudaSetDevice(int) { return cudaSuccess; }

// ---- mock engine ----
static std::atomic<int> g_live_engines{0};
static std::atomic<int> g_fail_frame{-1};   // payload value whose frame fails in engine_wait_blocking
extern "C" {
int sgm_create(sgm_engine** out, const sgm_params* p) {
    if (p->max_disp == 48) return SGM_ERR_UNSUPPORTED;
    sgm_engine* e = new sgm_engine();
    e->p = *p;
    e->ih = p->height / p->scale;
    e->iw = p->width / p->scale;
    e->n = (size_t)e->ih * e->iw;
    e->h_in[0] = new uint8_t[(size_t)p->height * p->width];
    e->h_final = new float[e->n];
    ++g_live_engines;
    *out = e;
    return SGM_OK;
}
void sgm_destroy(sgm_engine* e) {
    if (!e) return;
    delete[] e->h_in[0];
    delete[] e->h_final;
    delete e;
    --g_live_engines;
}
int sgm_wait(sgm_engine* e) { e->busy = false; return SGM_OK; }
const char* sgm_last_error(const sgm_engine* e) { return e ? e->err.c_str() : "mock create error"; }
int sgm_get_launch_count(const sgm_engine*) { return 20; }
int sgm_get_size(const sgm_engine* e, int* h, int* w, int* d) {
    if (h) *h = e->ih;
    if (w) *w = e->iw;
    if (d) *d = e->p.max_disp;
    return SGM_OK;
}
}
namespace sgmb200 {
int engine_stage(sgm_engine* e, const uint8_t* left, size_t, const uint8_t*, size_t, const uint8_t*, size_t, const uint8_t*, size_t) {
    if (e->busy) { e->err = "busy"; return SGM_ERR_STATE; }
    if (!left) { e->err = "left / right image is null"; return SGM_ERR_INVALID; }
    std::memcpy(e->h_in[0], left, 4);  // the test encodes the frame's payload in the first 4 bytes
    return SGM_OK;
}
int engine_launch_staged(sgm_engine* e) {
    if (e->busy) { e->err = "busy at launch"; return SGM_ERR_STATE; }
    e->busy = true;
    e->launches = 20;
    return SGM_OK;
}
int engine_wait_blocking(sgm_engine* e) {
    std::this_thread::sleep_for(std::chrono::microseconds(50 + std::rand() % 200));
    int payload;
    std::memcpy(&payload, e->h_in[0], 4);
    e->busy = false;
    if (payload == g_fail_frame.load()) {
        e->have_result = false;
        e->err = "mock device fault";
        return SGM_ERR_CUDA;
    }
    for (size_t k = 0; k < e->n; ++k) e->h_final[k] = (float)payload + (float)(k % 7);
    e->have_result = true;
    e->result_on_host = true;
    return SGM_OK;
}
const float* engine_host_result(const sgm_engine* e) { return (e->have_result && e->result_on_host) ? e->h_final : nullptr; }
}  // namespace sgmb200

// ---- the tests ----
#define CHECK(cond)                                                                  \
    do {                                                                             \
        if (!(cond)) {                                                               \
            std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);   \
            std::exit(1);                                                            \
        }                                                                            \
    } while (0)

static const int H = 8, W = 16;
static void frame_image(int payload, std::vector<uint8_t>& img) {
    img.assign((size_t)H * W, 0);
    std::memcpy(img.data(), &payload, 4);
}
static bool result_ok(const float* r, int payload) {
    for (int k = 0; k < H * W; ++k)
        if (r[k] != (float)payload + (float)(k % 7)) return false;
    return true;
}

static void two_threads(int n_dev, int inflight, int frames, bool views) {
    sgm_params p;
    std::memset(&p, 0, sizeof(p));
    p.height = H, p.width = W, p.scale = 1, p.max_disp = 64;
    std::vector<int> devs((size_t)n_dev, 0);
    sgm_stream* st = nullptr;
    CHECK(sgm_stream_create(&st, &p, devs.data(), n_dev, inflight) == SGM_OK);
    std::atomic<bool> ok{true};
    std::thread producer([&] {
        std::vector<uint8_t> img;
        for (int k = 0; k < frames && ok; ++k) {
            frame_image(1000 + k, img);
            uint64_t idx = 0;
            if (sgm_stream_submit(st, img.data(), W, img.data(), W, nullptr, 0, nullptr, 0, &idx) != SGM_OK || idx != (uint64_t)k) ok = false;
        }
    });
    std::vector<float> out((size_t)H * W);
    for (int k = 0; k < frames; ++k) {
        while (sgm_stream_pending(st) == 0 && ok) std::this_thread::yield();
        uint64_t idx = 0;
        if (views) {
            const float* v = nullptr;
            CHECK(sgm_stream_next(st, nullptr, 0, &v, &idx) == SGM_OK);
            CHECK(idx == (uint64_t)k && v && result_ok(v, 1000 + k));
            CHECK(sgm_stream_release(st) == SGM_OK);  // with two threads the consumer gives the slot back itself (the one-thread test covers the implicit release)
        } else {
            CHECK(sgm_stream_next(st, out.data(), W * sizeof(float), nullptr, &idx) == SGM_OK);
            CHECK(idx == (uint64_t)k && result_ok(out.data(), 1000 + k));
        }
    }
    producer.join();
    CHECK(ok);
    CHECK(sgm_stream_launches(st) == (uint64_t)frames * 20);
    sgm_stream_destroy(st);
    CHECK(g_live_engines == 0);
}

static void one_thread(int n_dev, int inflight, int frames, bool views) {
    sgm_params p;
    std::memset(&p, 0, sizeof(p));
    p.height = H, p.width = W, p.scale = 1, p.max_disp = 64;
    sgm_stream* st = nullptr;
    CHECK(sgm_stream_create(&st, &p, nullptr, n_dev, inflight) == SGM_OK);
    const int cap = n_dev * inflight;
    std::vector<uint8_t> img;
    std::vector<float> out((size_t)H * W);
    int got = 0;
    auto collect = [&] {
        uint64_t idx = 0;
        const float* v = nullptr;
        CHECK(sgm_stream_next(st, views ? nullptr : out.data(), views ? 0 : W * sizeof(float), views ? &v : nullptr, &idx) == SGM_OK);
        CHECK(idx == (uint64_t)got && result_ok(views ? v : out.data(), 1000 + got));
        ++got;
    };
    for (int k = 0; k < frames; ++k) {
        // a held view occupies its slot and counts as pending; a caller on one thread gives it back (sgm_stream_release) when
        // it is done with it, before a submit that would otherwise wait for itself
        if (sgm_stream_pending(st) == cap) {
            collect();
            if (views) {
                CHECK(sgm_stream_pending(st) == cap);  // the view still holds the slot
                CHECK(sgm_stream_release(st) == SGM_OK);
            }
            CHECK(sgm_stream_pending(st) == cap - 1);
        }
        frame_image(1000 + k, img);
        CHECK(sgm_stream_submit(st, img.data(), W, img.data(), W, nullptr, 0, nullptr, 0, nullptr) == SGM_OK);
    }
    while (got < frames) collect();  // consecutive sgm_stream_next calls give the previous view back themselves
    CHECK(sgm_stream_release(st) == SGM_OK && sgm_stream_pending(st) == 0);
    CHECK(sgm_stream_next(st, out.data(), W * sizeof(float), nullptr, nullptr) == SGM_ERR_STATE);  // nothing pending: no wait
    sgm_stream_destroy(st);
    CHECK(g_live_engines == 0);
}

static void failures() {
    sgm_params p;
    std::memset(&p, 0, sizeof(p));
    p.height = H, p.width = W, p.scale = 1, p.max_disp = 64;
    sgm_stream* st = nullptr;
    CHECK(sgm_stream_create(&st, &p, nullptr, 1, 0) == SGM_ERR_INVALID && st == nullptr);
    p.max_disp = 48;
    CHECK(sgm_stream_create(&st, &p, nullptr, 2, 2) == SGM_ERR_UNSUPPORTED && st == nullptr && g_live_engines == 0);
    p.max_disp = 64;
    CHECK(sgm_stream_create(&st, &p, nullptr, 2, 2) == SGM_OK);
    std::vector<uint8_t> img;
    std::vector<float> out((size_t)H * W);
    CHECK(sgm_stream_submit(st, nullptr, W, nullptr, W, nullptr, 0, nullptr, 0, nullptr) == SGM_ERR_INVALID);  // rejected: not a frame
    CHECK(sgm_stream_pending(st) == 0);
    g_fail_frame = 1001;
    for (int k = 0; k < 3; ++k) {
        frame_image(1000 + k, img);
        CHECK(sgm_stream_submit(st, img.data(), W, img.data(), W, nullptr, 0, nullptr, 0, nullptr) == SGM_OK);
    }
    uint64_t idx = 9;
    CHECK(sgm_stream_next(st, out.data(), W * sizeof(float), nullptr, &idx) == SGM_OK && idx == 0);
    CHECK(sgm_stream_next(st, out.data(), W * sizeof(float), nullptr, &idx) == SGM_ERR_CUDA && idx == 1);   // the failing frame
    CHECK(std::strstr(sgm_stream_last_error(st), "mock device fault") != nullptr);
    CHECK(sgm_stream_next(st, out.data(), W * sizeof(float), nullptr, &idx) == SGM_OK && idx == 2 && result_ok(out.data(), 1002));
    g_fail_frame = -1;
    frame_image(1003, img);  // the slot of the failed frame is usable again
    CHECK(sgm_stream_submit(st, img.data(), W, img.data(), W, nullptr, 0, nullptr, 0, &idx) == SGM_OK && idx == 3);
    sgm_stream_destroy(st);  // with a frame still pending
    CHECK(g_live_engines == 0);
}

int main() {
    for (int views = 0; views < 2; ++views) {
        two_threads(1, 1, 200, views);
        two_threads(1, 5, 500, views);
        two_threads(3, 2, 500, views);
        two_threads(8, 5, 2000, views);
        one_thread(1, 1, 50, views);
        one_thread(2, 3, 200, views);
    }
    failures();
    std::printf("stream runner mock: all checks passed\n");
    return 0;
}
