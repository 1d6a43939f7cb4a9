// stream_runner.cu — sgm_stream: a stream of independent stereo pairs over 1..N GPUs of one box behind one handle.
//
// The reference's caller (node.cpp:49,93,104) owns one Solver and calls process() once per frame; Solver objects keep no
// state between frames, so the only multi-GPU structure on this path is "frame k goes to device k mod N" (SURVEY 8(e)):
// no collective, no peer traffic.  Per device: `inflight` engine handles (each with its own stream, device volumes and
// pinned staging for inputs and result) and ONE host thread that issues the launches, so the producer thread only packs
// images into pinned memory and the consumer thread only waits (sleeping, cudaEventBlockingSync) and reads.
//
// Slot life cycle (one slot = one engine handle), every transition under the stream's mutex:
//     FREE --submit: engine_stage (producer thread)--> STAGED --engine_launch_staged (device thread)--> LAUNCHED
//          --sgm_stream_next: engine_wait_blocking (consumer thread)--> HELD (result pointer handed out)
//          --sgm_stream_release / the following sgm_stream_next--> FREE
// Frame k lives in slot (k / N) mod inflight of device k mod N, so both the producer and the consumer know where a frame
// is without a queue, and results come out in submission order by construction.
#include <cuda_runtime.h>

#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <deque>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "engine_internal.h"

using namespace sgmb200;

namespace {

enum SlotState { SLOT_FREE = 0, SLOT_STAGED, SLOT_LAUNCHED, SLOT_HELD };

struct Slot {
    sgm_engine* engine = nullptr;
    SlotState state = SLOT_FREE;
    int launch_rc = SGM_OK;  // what engine_launch_staged returned (reported by sgm_stream_next)
};

struct DeviceLane {
    int device = 0;
    std::vector<Slot> slots;
    std::deque<int> to_launch;  // slot indices in submission order
    std::thread worker;
};

thread_local std::string g_stream_create_error;

}  // namespace

struct sgm_stream {
    sgm_params params;
    int n_devices = 0, inflight = 0, ih = 0, iw = 0;
    std::vector<DeviceLane> lanes;
    std::mutex mu;
    std::condition_variable cv;
    bool stopping = false;
    uint64_t submitted = 0, delivered = 0;  // frame counters: next index to submit / to hand out
    uint64_t launches = 0;
    bool holding = false;                   // the slot of frame delivered - 1 is still HELD
    std::string err;
};

namespace {

int sfail(sgm_stream* s, int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (s)
        s->err = buf;
    else
        g_stream_create_error = buf;
    return code;
}

Slot& slot_of(sgm_stream* s, uint64_t k, DeviceLane** lane_out = nullptr) {
    DeviceLane& lane = s->lanes[(size_t)(k % (uint64_t)s->n_devices)];
    if (lane_out) *lane_out = &lane;
    return lane.slots[(size_t)((k / (uint64_t)s->n_devices) % (uint64_t)s->inflight)];
}

// One per device: issues the launches of staged frames in submission order.  Launching from a dedicated thread keeps the
// ~20 launches of a frame off the producer's thread and gives every GPU its own issuing thread.
void device_worker(sgm_stream* s, DeviceLane* lane) {
    cudaSetDevice(lane->device);
    std::unique_lock<std::mutex> lock(s->mu);
    for (;;) {
        s->cv.wait(lock, [&] { return s->stopping || !lane->to_launch.empty(); });
        if (lane->to_launch.empty()) return;  // stopping and drained
        const int idx = lane->to_launch.front();
        lane->to_launch.pop_front();
        Slot& sl = lane->slots[(size_t)idx];
        lock.unlock();
        const int rc = engine_launch_staged(sl.engine);  // H2D + kernels + D2H queued on the handle's stream
        const int n = sgm_get_launch_count(sl.engine);
        lock.lock();
        sl.launch_rc = rc;
        sl.state = SLOT_LAUNCHED;
        if (rc == SGM_OK && n > 0) s->launches += (uint64_t)n;
        s->cv.notify_all();
    }
}

void release_held(sgm_stream* s) {  // mu held
    if (!s->holding) return;
    slot_of(s, s->delivered - 1).state = SLOT_FREE;
    s->holding = false;
    s->cv.notify_all();
}

}  // namespace

extern "C" {

const char* sgm_stream_last_error(const sgm_stream* s) { return s ? s->err.c_str() : g_stream_create_error.c_str(); }

void sgm_stream_destroy(sgm_stream* s) {
    if (!s) return;
    {
        std::lock_guard<std::mutex> lock(s->mu);
        s->stopping = true;
        s->cv.notify_all();
    }
    for (auto& lane : s->lanes)
        if (lane.worker.joinable()) lane.worker.join();  // drains what was staged
    for (auto& lane : s->lanes)
        for (auto& sl : lane.slots)
            if (sl.engine) {
                sgm_wait(sl.engine);
                sgm_destroy(sl.engine);
            }
    delete s;
}

int sgm_stream_create(sgm_stream** out, const sgm_params* params, const int* devices, int n_devices, int inflight) {
    if (!out || !params) return sfail(nullptr, SGM_ERR_INVALID, "sgm_stream_create: null argument");
    *out = nullptr;
    if (n_devices < 1 || n_devices > 64 || inflight < 1 || inflight > 64)
        return sfail(nullptr, SGM_ERR_INVALID, "sgm_stream_create: need 1..64 devices and 1..64 frames in flight per device");
    sgm_stream* s = new (std::nothrow) sgm_stream();
    if (!s) return sfail(nullptr, SGM_ERR_NOMEM, "out of host memory");
    s->params = *params;
    s->n_devices = n_devices;
    s->inflight = inflight;
    s->lanes.resize((size_t)n_devices);
    for (int g = 0; g < n_devices; ++g) {
        DeviceLane& lane = s->lanes[(size_t)g];
        lane.device = devices ? devices[g] : g;  // NULL = devices 0 .. n_devices-1
        lane.slots.resize((size_t)inflight);
        for (int k = 0; k < inflight; ++k) {
            sgm_params p = *params;
            p.device = lane.device;
            const int rc = sgm_create(&lane.slots[(size_t)k].engine, &p);
            if (rc != SGM_OK) {
                sfail(nullptr, rc, "sgm_stream_create: device %d, handle %d: %s", lane.device, k, sgm_last_error(nullptr));
                sgm_stream_destroy(s);
                return rc;
            }
        }
    }
    sgm_get_size(s->lanes[0].slots[0].engine, &s->ih, &s->iw, nullptr);
    for (auto& lane : s->lanes) lane.worker = std::thread(device_worker, s, &lane);
    *out = s;
    return SGM_OK;
}

int sgm_stream_submit(sgm_stream* s, const uint8_t* left, size_t ls, const uint8_t* right, size_t rs, const uint8_t* sky_l, size_t sls,
                      const uint8_t* sky_r, size_t srs, uint64_t* frame_index) {
    if (!s) return SGM_ERR_INVALID;
    std::unique_lock<std::mutex> lock(s->mu);
    const uint64_t k = s->submitted;
    DeviceLane* lane = nullptr;
    Slot& sl = slot_of(s, k, &lane);
    // back-pressure: the slot still carries frame k - n_devices * inflight until the consumer has taken it AND given a
    // view of it back (sgm_stream_release, or its next sgm_stream_next): a result that is being read is never overwritten
    s->cv.wait(lock, [&] { return sl.state == SLOT_FREE; });
    lock.unlock();
    // the copy into pinned staging runs outside the lock: device threads and the consumer keep going
    const int rc = engine_stage(sl.engine, left, ls, right, rs, sky_l, sls, sky_r, srs);
    lock.lock();
    if (rc != SGM_OK) return sfail(s, rc, "frame %llu: %s", (unsigned long long)k, sgm_last_error(sl.engine));  // nothing was queued
    sl.state = SLOT_STAGED;
    lane->to_launch.push_back((int)(&sl - lane->slots.data()));
    s->submitted = k + 1;
    if (frame_index) *frame_index = k;
    s->cv.notify_all();
    return SGM_OK;
}

int sgm_stream_next(sgm_stream* s, float* dst, size_t dst_stride, const float** view, uint64_t* frame_index) {
    if (!s) return SGM_ERR_INVALID;
    if (dst && dst_stride < (size_t)s->iw * sizeof(float)) return sfail(s, SGM_ERR_INVALID, "sgm_stream_next: destination stride too small");
    std::unique_lock<std::mutex> lock(s->mu);
    release_held(s);  // the pointer handed out by the previous call dies here
    if (s->delivered == s->submitted) return sfail(s, SGM_ERR_STATE, "sgm_stream_next: no frame is pending");
    const uint64_t k = s->delivered;
    Slot& sl = slot_of(s, k);
    s->cv.wait(lock, [&] { return sl.state == SLOT_LAUNCHED; });
    const int launch_rc = sl.launch_rc;
    lock.unlock();
    int rc = launch_rc;
    if (rc == SGM_OK) rc = engine_wait_blocking(sl.engine);  // sleeps until the frame's D2H has landed
    const float* res = rc == SGM_OK ? engine_host_result(sl.engine) : nullptr;
    if (rc == SGM_OK && !res) rc = SGM_ERR_STATE;
    if (rc == SGM_OK && dst)
        for (int r = 0; r < s->ih; ++r)
            std::memcpy(reinterpret_cast<char*>(dst) + (size_t)r * dst_stride, res + (size_t)r * s->iw, (size_t)s->iw * sizeof(float));
    lock.lock();
    s->delivered = k + 1;
    if (frame_index) *frame_index = k;
    if (rc != SGM_OK) {
        sfail(s, rc, "frame %llu: %s", (unsigned long long)k, sgm_last_error(sl.engine));
        sl.state = SLOT_FREE;  // a failed frame holds nothing
        if (view) *view = nullptr;
        s->cv.notify_all();
        return rc;
    }
    if (view) {
        *view = res;  // stays valid (slot HELD) until the next sgm_stream_next / destroy
        sl.state = SLOT_HELD;
        s->holding = true;
    } else {
        sl.state = SLOT_FREE;
        s->cv.notify_all();
    }
    return SGM_OK;
}

int sgm_stream_release(sgm_stream* s) {
    if (!s) return SGM_ERR_INVALID;
    std::lock_guard<std::mutex> lock(s->mu);
    release_held(s);
    return SGM_OK;
}

int sgm_stream_pending(const sgm_stream* s) {
    if (!s) return -1;
    std::lock_guard<std::mutex> lock(const_cast<sgm_stream*>(s)->mu);
    return (int)(s->submitted - s->delivered) + (s->holding ? 1 : 0);  // a held view still occupies its slot
}

uint64_t sgm_stream_launches(const sgm_stream* s) {
    if (!s) return 0;
    std::lock_guard<std::mutex> lock(const_cast<sgm_stream*>(s)->mu);
    return s->launches;
}

}  // extern "C"
