// engine_internal.h — the engine handle behind include/sgm_b200.h and the few entry points the frame-stream runner
// (stream_runner.cu) needs beyond the public C ABI.  Internal to libsgm_b200.so.
#pragma once

#include <cuda_runtime.h>

#include <string>

#include "../../include/sgm_b200.h"

struct sgm_engine {
    sgm_params p;
    int ih = 0, iw = 0, D = 0, views = 1;
    size_t n = 0, N = 0;  // pixels, pixels * D (working resolution)
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[SGM_STAGE_COUNT + 1] = {};
    cudaEvent_t ev_done = nullptr;  // end of the frame, cudaEventBlockingSync: lets a waiting host thread sleep (stream runner)
    bool timing = false, have_result = false, result_on_host = false, busy = false;
    bool staged_sky[2] = {false, false};  // masks of the frame packed by engine_stage
    bool host_taps = false;               // SGM_FLAG_HOST_TAPS: disp / disp_beta (8 bit) and the right view's sub-pixel image follow the result to the host
    bool integer = false;  // SGM_FLAG_INTEGER: volumes hold s16x2 words (two disparities per 32-bit word)
    size_t vol_words = 0;  // 32-bit words per view of a volume: N (float mode) or N / 2 (integer mode)
    int launches = 0;
    float stage_ms[SGM_STAGE_COUNT] = {};
    size_t device_bytes = 0;
    std::string err;

    // device
    uint8_t* d_in[2] = {nullptr, nullptr};   // full-resolution staging of the host inputs (dense)
    uint8_t* d_sky[2] = {nullptr, nullptr};  // [n] each
    uint8_t* d_img = nullptr;                // [2][n] working-resolution images
    uint8_t* d_blur = nullptr;               // [2][n]
    uint64_t* d_census = nullptr;            // [2][n]
    float* d_cost = nullptr;                 // [views][N]
    float* d_acc_a = nullptr;                // [views][N]
    float* d_acc_b = nullptr;                // [views][N]
    float* d_pathsum = nullptr;              // [views][N] (SGM_FLAG_KEEP_VOLUMES)
    uint16_t* d_disp = nullptr;              // [2][n]
    float* d_sub = nullptr;                  // [2][n]
    float* d_lr = nullptr;                   // [n]
    float* d_med = nullptr;                  // [n]
    float* d_final = nullptr;                // [n]
    int* d_label = nullptr;                  // [n]
    int* d_area = nullptr;                   // [n]
    float* d_sad_lut = nullptr;              // SAD block matching: (sum / win_h) / win_w for every integer sum
    float* d_halo = nullptr;                 // fused aggregation: cross-block halo ring (tagged payload)
    unsigned halo_seq = 0;                   // fused aggregation: sweep launches so far (tag source)
    int* d_error = nullptr;                  // fused aggregation: time-out flag
    int* h_error = nullptr;                  // pinned copy of d_error
    bool fused = false;                      // last frame used the fused schedule
    // consumers of the result (allocated on first use)
    uint8_t* d_view = nullptr;               // [2*ih][iw][3] debug view
    int* d_pc_counts = nullptr;              // point cloud: per-block counts / offsets
    int* d_pc_offsets = nullptr;
    double* d_pc_xyz = nullptr;              // [n][3]
    uint8_t* d_pc_int = nullptr;             // [n]
    uint8_t* d_disp8 = nullptr;              // [2][n] saturated 8-bit copy of d_disp (SGM_FLAG_HOST_TAPS)
    // pinned host
    uint8_t* h_in[2] = {nullptr, nullptr};
    uint8_t* h_sky[2] = {nullptr, nullptr};
    float* h_final = nullptr;
    uint8_t* h_disp8 = nullptr;              // [2][n] (SGM_FLAG_HOST_TAPS)
    float* h_sub_r = nullptr;                // [n]    (SGM_FLAG_HOST_TAPS)
};


namespace sgmb200 {

// sgm_enqueue in two halves, so that the copy into pinned staging and the launches can run on different host threads:
// engine_stage validates and packs the caller's images (and masks) into the handle's pinned staging (no CUDA call);
// engine_launch_staged queues H2D + all kernels + D2H of the staged frame on the handle's stream.
int engine_stage(sgm_engine* e, const uint8_t* left, size_t ls, const uint8_t* right, size_t rs, const uint8_t* sky_l, size_t sls,
                 const uint8_t* sky_r, size_t srs);
int engine_launch_staged(sgm_engine* e);
// Blocks (yielding the CPU: cudaEventBlockingSync) until the frame in flight has finished, then runs sgm_wait's checks.
int engine_wait_blocking(sgm_engine* e);
// Pinned host copy of the final disparity of the last host-path frame (dense, iw floats per row).
const float* engine_host_result(const sgm_engine* e);

}  // namespace sgmb200
