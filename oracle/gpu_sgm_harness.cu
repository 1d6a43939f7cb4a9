// oracle/_ref harness for the reference's CUDA class GPU_SGM (gpu_sgm/, compiled UNMODIFIED from /root/reference by
// oracle/Makefile `make gpu_sgm`).  TEST / BASELINE INFRASTRUCTURE: GPU_SGM is a different, approximate algorithm that
// the reference's node does not call (node.cpp:10,50,98-102 are commented out); it is only TIMED as "the reference's
// GPU path" next to our engine (bench.py --impl gpu_sgm_reference, tests/perf_configs.py), never compared bit-wise.
#include <chrono>
#include <cstdint>
#include <cstring>

#include "gpu_sgm/inc/SGM.cuh"

double get_cur_ms() {
    using namespace std::chrono;
    return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

extern "C" {

// Runs `iters` frames through one GPU_SGM object (constructed once, like node.cpp would) and returns the wall time of
// each process() call including a device synchronise (the class leaves its D2H copy unsynchronised, SGM.cu:230).
// out_disp (nullable) receives the last frame's float disparity (h/s x w/s).  Returns 0, or a cudaError_t.
int gpu_sgm_time(int h, int w, int s, int d, const uint8_t* left, const uint8_t* right, int iters, double* ms, float* out_disp) {
    cv::Mat l(h, w, CV_8UC1, const_cast<uint8_t*>(left)), r(h, w, CV_8UC1, const_cast<uint8_t*>(right));
    GPU_SGM sv(h, w, s, d);
    for (int k = 0; k < iters; ++k) {
        const double t0 = get_cur_ms();
        sv.process(l, r);
        const cudaError_t e = cudaDeviceSynchronize();
        ms[k] = get_cur_ms() - t0;
        if (e != cudaSuccess) return (int)e;
    }
    if (out_disp) {
        const cv::Mat& f = sv.get_disp();
        for (int i = 0; i < f.rows; ++i) std::memcpy(out_disp + (size_t)i * f.cols, f.ptr<float>(i), f.cols * sizeof(float));
    }
    return (int)cudaGetLastError();
}

}  // extern "C"
