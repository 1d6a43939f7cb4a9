// kernels.h — host-side launchers of the sm_100a stereo kernels (internal to libsgm_b200.so).
// Each launcher cites the reference stage it replaces; every launch goes through `count` so the
// engine can report how many kernels a frame took.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace sgmb200 {

struct LaunchCounter {
    int launches = 0;
    cudaError_t err = cudaSuccess;
    const char* failed = nullptr;   // name of the first launch that failed
    bool sync_each = false;         // SGM_B200_DEBUG_SYNC=1: synchronise after every launch to localise a fault
    cudaStream_t stream = nullptr;
    void note(const char* name = "kernel") {
        ++launches;
        cudaError_t e = cudaGetLastError();
        if (e == cudaSuccess && sync_each) e = cudaStreamSynchronize(stream);
        if (err == cudaSuccess && e != cudaSuccess) {
            err = e;
            failed = name;
        }
    }
};

// ---- cost stage (cost_kernels.cu) ------------------------------------------------------------
// src/SGM.cpp:40-56 / src/BM.cpp:17-33: nearest-neighbour decimation (row_step = scale for SGM, 1 for BM).
void launch_subsample(const uint8_t* src, size_t src_stride, int col_step, int row_step, uint8_t* dst, int ih, int iw,
                      cudaStream_t st, LaunchCounter& lc);
// src/Solver.cpp:124-125: 3x3 Gaussian (sigma 2, 1) in OpenCV's 8-bit fixed point; both images in one launch.
void launch_blur3(const uint8_t* src /*[2][h*w]*/, uint8_t* dst /*[2][h*w]*/, int h, int w, cudaStream_t st,
                  LaunchCounter& lc);
// src/cost.cpp:99-129 (CT_pts) for every pixel of both images in one launch.
void launch_census(const uint8_t* img /*[2][h*w]*/, uint64_t* table /*[2][h*w]*/, int h, int w, int win_h, int win_w,
                   cudaStream_t st, LaunchCounter& lc);
// src/Solver.cpp:143-248 fused with cost_horizontal_filter (:296-330): view v of `cost` (stride n*d floats) gets
// the Hamming volume of view v, recursively box-filtered along j.  sky[v] nullable.
void launch_hamming_hfilter(const uint64_t* ct_l, const uint64_t* ct_r, const uint8_t* sky_l, const uint8_t* sky_r,
                            float* cost, int views, int h, int w, int d, int scale, int win, cudaStream_t st,
                            LaunchCounter& lc);
// Both census rows of an image row (plus max_disp entries of padding) are staged in shared memory by the Hamming kernels:
// false for images too wide for that (about 12,000 pixels at D = 256).
bool hamming_rows_fit(int w, int d);
// Integer mode: the raw Hamming volume (src/Solver.cpp:143-248 without cost filter) as packed s16x2 words,
// [views][h*w*d/2] 32-bit words.  Pixels masked by sky_l / sky_r (nullable; 255 = sky) cost 0 at d = 0 and sky_big elsewhere
// (the 16-bit stand-in of the reference's 999999, see the kernel).
void launch_hamming_s16x2(const uint64_t* ct_l, const uint64_t* ct_r, const uint8_t* sky_l, const uint8_t* sky_r, unsigned sky_big, void* cost,
                          int views, int h, int w, int d, int scale, cudaStream_t st, LaunchCounter& lc);
// src/Solver.cpp:333-368 in place on `views` volumes.
void launch_vfilter(float* cost, int views, int h, int w, int d, int win, cudaStream_t st, LaunchCounter& lc);
// src/cost.cpp:4-30 (SAD) for every (pixel, d) of `views` views (view 1 = right view by symmetry).
void launch_sad_volume(const uint8_t* img_l, const uint8_t* img_r, const float* lut, float* cost, int views, int h, int w, int d,
                       int win_h, int win_w, int scale, cudaStream_t st, LaunchCounter& lc);

// src/cost.cpp:33-59 (SSD, kind 1) / :62-96 (CT, kind 2) for every (pixel, d) of the left view.
void launch_window_cost_volume(int kind, const uint8_t* img_l, const uint8_t* img_r, float* cost, int h, int w, int d, int win_h, int win_w,
                               int scale, cudaStream_t st, LaunchCounter& lc);

// ---- aggregation (aggregate_kernels.cu) ------------------------------------------------------
struct AggregateArgs {
    const float* cost;  // [views][n*d]   filtered cost volume C
    float* acc_a;       // [views][n*d]   scratch: ((L1+L2)+L3)+L4
    float* acc_b;       // [views][n*d]   scratch: ((L5+L6)+L7)
    float* pathsum;     // [views][n*d]   optional: S written out (debug tap), else nullptr
    uint16_t* disp;     // [views][n]
    float* subpixel;    // [views][n]
    int views, h, w, d, paths;
    float P1, P2, unique_ratio;
    int min_gap;        // 1 for SGM (src/SGM.cpp:409), 2 for BM (src/BM.cpp:82)
    int packed16 = 0;   // integer mode: cost / acc_a / acc_b / pathsum hold s16x2 words (two disparities per 32-bit word), fused schedule only
    // scratch of the fused schedule (sweep_kernels.cu); nullptr -> line walkers only
    float* halo = nullptr;          // fused_halo_bytes() bytes, zeroed once at allocation
    unsigned* launch_seq = nullptr; // HOST counter of sweep launches that used `halo` (tags of the halo payload)
    int* error = nullptr;           // 1 int, set when a cross-block wait timed out
};
// src/SGM.cpp:82-418 + compute_subpixel (src/Solver.cpp:569-597): the path recurrences, their sum in the
// reference's association order, WTA, uniqueness and the sub-pixel parabola.
void launch_aggregate(const AggregateArgs& a, cudaStream_t st, LaunchCounter& lc);
// The same result with the fused three-launch schedule of sweep_kernels.cu (48 B instead of 88 B per disparity-pixel
// and view).  Returns false without launching anything when the image is too wide for a resident sweep.
bool launch_aggregate_fused(const AggregateArgs& a, cudaStream_t st, LaunchCounter& lc);
size_t fused_halo_bytes(int views, int w, int d);
// src/BM.cpp:52-91 (+ compute_subpixel when parabola != 0): WTA / uniqueness straight on a cost volume.
void launch_wta(const float* vol, uint16_t* disp, float* subpixel, int views, int h, int w, int d, float unique_ratio,
                int min_gap, int parabola, cudaStream_t st, LaunchCounter& lc);

// ---- post stage (post_kernels.cu) --------------------------------------------------------------
// src/SGM.cpp:803-818 (LR check of the left sub-pixel image against the right one; right == nullptr: none, BM) fused with the
// hole-filling median of src/Solver.cpp:604-630 (exact raster-order semantics).  out_lr = the checked image (tap), out_med =
// the checked image with the holes filled.  scratch_a / scratch_b: int buffers of h * w elements each (contents destroyed).
void launch_lr_median(const float* left, const float* right, float* out_lr, float* out_med, int* scratch_a, int* scratch_b, int h, int w,
                      int d, int scale, float thresh, int win_h, int win_w, cudaStream_t st, LaunchCounter& lc);
// src/Solver.cpp:514-566 with serial semantics: f_out = f_in with small components set to `value`;
// labels / area are n-int scratch buffers.
void launch_speckle(const float* f_in, float* f_out, int* labels, int* area, int h, int w, float value, int max_size,
                    float max_dis, cudaStream_t st, LaunchCounter& lc);

// The same filter with labels / area in a file-scope static __device__ array of post_kernels.cu instead of caller scratch
// (sgm_speckle_filter_static: no allocation anywhere).  `device` must be the current device; h * w <= SPECKLE_STATIC_MAX_PIXELS.
constexpr int SPECKLE_STATIC_MAX_PIXELS = 1 << 21;  // == SGM_SPECKLE_STATIC_MAX_PIXELS of include/sgm_b200.h (1920x1080 fits)
cudaError_t launch_speckle_static(int device, const float* f_in, float* f_out, int h, int w, float value, int max_size, float max_dis,
                                  cudaStream_t st, LaunchCounter& lc);

// ---- consumers of the result (view_kernels.cu) ---------------------------------------------------
// Solver::show_disp + colormap (src/Solver.cpp:55-93, :652-707): out is a dense (2*h) x w x 3 BGR image.
void launch_debug_view(const uint8_t* gray, const float* disp, uint8_t* out, int h, int w, int max_disp, cudaStream_t st,
                       LaunchCounter& lc);
// uint16 disparities -> the reference's CV_8UC1 Mats (inc/Solver.h:43), saturating at 255.
void launch_narrow_disp(const uint16_t* disp, uint8_t* out, int n, cudaStream_t st, LaunchCounter& lc);
// node.cpp:119-143: ordered list of camera-frame points (xyz doubles) + grey values; counts / offsets are scratch of
// cloud_blocks(h*w) and cloud_blocks(h*w) + 1 ints, offsets[cloud_blocks] receives the number of points.
int cloud_blocks(int n);
void launch_point_cloud(const float* disp, const uint8_t* img, size_t img_stride, int h, int w, float fx, float fy, float cx, float cy,
                        double baseline, float max_range, float invalid, int scale, int* counts, int* offsets, double* xyz,
                        uint8_t* intensity, int capacity, cudaStream_t st, LaunchCounter& lc);

}  // namespace sgmb200
