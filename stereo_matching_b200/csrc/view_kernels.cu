// view_kernels.cu — consumers of the final disparity that the reference runs on the CPU right after process():
//   * the debug view of Solver::show_disp (src/Solver.cpp:55-93) with Solver::colormap (:652-707);
//   * the disparity -> camera-frame point cloud loop of node.cpp:119-143.
// Both are per-pixel and embarrassingly parallel; the point cloud additionally needs an ORDER-PRESERVING compaction
// (the reference push_back()s in raster order), done with per-block counts, a scan of the counts and a scatter.
#include "common.cuh"
#include "kernels.h"

namespace sgmb200 {

namespace {

// src/Solver.cpp:652-707: black for invalid, else a blue -> cyan -> green -> yellow -> red ramp of v * (256 / D) with the
// reference's float -> uchar truncations (the green channel of the last two segments goes through double).
__device__ __forceinline__ uchar3 colormap_bgr(float v, int max_disp) {
    if (v > (float)(max_disp - 1)) return make_uchar3(0, 0, 0);
    v = __fmul_rn(v, (float)(256 / max_disp));
    if (v <= 51.f) return make_uchar3(255, (unsigned char)(int)__fmul_rn(v, 5.f), 0);
    if (v <= 102.f) {
        v = __fsub_rn(v, 51.f);
        return make_uchar3((unsigned char)(int)__fsub_rn(255.f, __fmul_rn(v, 5.f)), 255, 0);
    }
    if (v <= 153.f) {
        v = __fsub_rn(v, 102.f);
        return make_uchar3(0, 255, (unsigned char)(int)__fmul_rn(v, 5.f));
    }
    if (v <= 204.f) {
        v = __fsub_rn(v, 153.f);
        const unsigned char t = (unsigned char)(int)__dadd_rn(__ddiv_rn(__dmul_rn(128.0, (double)v), 51.0), 0.5);
        return make_uchar3(0, (unsigned char)(255 - t), 255);
    }
    v = __fsub_rn(v, 204.f);
    const unsigned char t = (unsigned char)(int)__dadd_rn(__ddiv_rn(__dmul_rn(127.0, (double)v), 51.0), 0.5);
    return make_uchar3(0, (unsigned char)(127 - t), 255);
}

// out: (2*h) x w x 3, dense.  Rows 0..h-1 = the working-resolution left image as grey BGR; the colour map is then
// copied over rows h-1 .. 2h-2 (the reference's Rect(0, img_h - 1, ...), src/Solver.cpp:91); row 2h-1 stays zero.
__global__ void debug_view_kernel(const uint8_t* __restrict__ gray, const float* __restrict__ disp, uint8_t* __restrict__ out, int h, int w,
                                  int max_disp) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const int row = blockIdx.y;  // 0 .. 2h-1
    if (j >= w) return;
    uchar3 px = make_uchar3(0, 0, 0);
    if (row >= h - 1 && row < 2 * h - 1) {
        px = colormap_bgr(disp[(size_t)(row - (h - 1)) * w + j], max_disp);
    } else if (row < h - 1) {
        const unsigned char g = gray[(size_t)row * w + j];
        px = make_uchar3(g, g, g);
    }
    uint8_t* o = out + ((size_t)row * w + j) * 3;
    o[0] = px.x;
    o[1] = px.y;
    o[2] = px.z;
}

// ---- point cloud (node.cpp:119-143) -----------------------------------------------------------------------------
// keep(i, j): disparity != invalid and Zc = (fx+fy)/2.0 * baseline / (d + 1e-6) <= max_range, in the reference's mixed
// float / double arithmetic ((fx + fy) is a float add, everything after it double).
struct CamArgs {
    float fx, fy, cx, cy;
    double baseline;
    float max_range;
    float invalid;
    int scale;
};

__device__ __forceinline__ bool cloud_point(const CamArgs& c, float d, int i, int j, double& X, double& Y, double& Z) {
    if (d == c.invalid) return false;
    const double f2 = __ddiv_rn((double)__fadd_rn(c.fx, c.fy), 2.0);
    Z = __ddiv_rn(__dmul_rn(f2, c.baseline), __dadd_rn((double)d, 1e-6));
    if (Z > (double)c.max_range) return false;
    X = __ddiv_rn(__dmul_rn((double)__fsub_rn((float)(j * c.scale), c.cx), Z), (double)c.fx);
    Y = __ddiv_rn(__dmul_rn((double)__fsub_rn((float)(i * c.scale), c.cy), Z), (double)c.fy);
    return true;
}

constexpr int PC_BLOCK = 256;

__global__ void cloud_count_kernel(const float* __restrict__ disp, int n, int w, CamArgs c, int* __restrict__ block_counts) {
    const int p = blockIdx.x * PC_BLOCK + threadIdx.x;
    double X, Y, Z;
    const bool keep = p < n && cloud_point(c, disp[p], p / w, p % w, X, Y, Z);
    const int cnt = __syncthreads_count(keep);
    if (threadIdx.x == 0) block_counts[blockIdx.x] = cnt;
}

// exclusive scan of the block counts in one block (nblocks is a few thousand); total goes to offsets[nblocks]
__global__ void cloud_scan_kernel(const int* __restrict__ counts, int* __restrict__ offsets, int nblocks) {
    __shared__ int warp_sums[32];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nblocks; base += 1024) {
        const int idx = base + threadIdx.x;
        const int v = idx < nblocks ? counts[idx] : 0;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(FULL_MASK, incl, o);
            if ((threadIdx.x & 31) >= o) incl += t;
        }
        if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            int s = warp_sums[threadIdx.x];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(FULL_MASK, s, o);
                if (threadIdx.x >= o) s += t;
            }
            warp_sums[threadIdx.x] = s;  // inclusive over warps
        }
        __syncthreads();
        const int warp_excl = (threadIdx.x >> 5) ? warp_sums[(threadIdx.x >> 5) - 1] : 0;
        if (idx < nblocks) offsets[idx] = carry + warp_excl + incl - v;
        __syncthreads();
        if (threadIdx.x == 0) carry += warp_sums[31];
        __syncthreads();
    }
    if (threadIdx.x == 0) offsets[nblocks] = carry;
}

__global__ void cloud_scatter_kernel(const float* __restrict__ disp, const uint8_t* __restrict__ img, size_t img_stride, int n, int w,
                                     CamArgs c, const int* __restrict__ offsets, double* __restrict__ xyz, uint8_t* __restrict__ intensity,
                                     int capacity) {
    __shared__ int warp_base[PC_BLOCK / 32];
    const int p = blockIdx.x * PC_BLOCK + threadIdx.x;
    const int i = p / w, j = p % w;
    double X = 0, Y = 0, Z = 0;
    const bool keep = p < n && cloud_point(c, disp[p], i, j, X, Y, Z);
    const unsigned m = __ballot_sync(FULL_MASK, keep);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) warp_base[wid] = __popc(m);
    __syncthreads();
    int base = offsets[blockIdx.x];
    for (int k = 0; k < wid; ++k) base += warp_base[k];
    const int slot = base + __popc(m & ((1u << lane) - 1u));
    if (keep && slot < capacity) {
        xyz[(size_t)slot * 3 + 0] = X;
        xyz[(size_t)slot * 3 + 1] = Y;
        xyz[(size_t)slot * 3 + 2] = Z;
        intensity[slot] = img[(size_t)i * img_stride + j];  // node.cpp:138 img_l.at<uchar>(i, j)
    }
}

// Solver::disp / disp_beta are CV_8UC1 (inc/Solver.h:43): the engine's uint16 disparities narrowed for the host shells.
__global__ void narrow_disp_kernel(const uint16_t* __restrict__ disp, uint8_t* __restrict__ out, int n) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < n) out[p] = (uint8_t)min((int)disp[p], 255);
}

}  // namespace

void launch_narrow_disp(const uint16_t* disp, uint8_t* out, int n, cudaStream_t st, LaunchCounter& lc) {
    narrow_disp_kernel<<<(n + 255) / 256, 256, 0, st>>>(disp, out, n);
    lc.note("narrow_disp_kernel");
}

void launch_debug_view(const uint8_t* gray, const float* disp, uint8_t* out, int h, int w, int max_disp, cudaStream_t st, LaunchCounter& lc) {
    dim3 grid((w + 255) / 256, 2 * h);
    debug_view_kernel<<<grid, 256, 0, st>>>(gray, disp, out, h, w, max_disp);
    lc.note("debug_view_kernel");
}

int cloud_blocks(int n) { return (n + PC_BLOCK - 1) / PC_BLOCK; }

void launch_point_cloud(const float* disp, const uint8_t* img, size_t img_stride, int h, int w, float fx, float fy, float cx, float cy,
                        double baseline, float max_range, float invalid, int scale, int* counts, int* offsets, double* xyz,
                        uint8_t* intensity, int capacity, cudaStream_t st, LaunchCounter& lc) {
    const int n = h * w, nb = cloud_blocks(n);
    CamArgs c{fx, fy, cx, cy, baseline, max_range, invalid, scale};
    cloud_count_kernel<<<nb, PC_BLOCK, 0, st>>>(disp, n, w, c, counts);
    lc.note("cloud_count_kernel");
    cloud_scan_kernel<<<1, 1024, 0, st>>>(counts, offsets, nb);
    lc.note("cloud_scan_kernel");
    cloud_scatter_kernel<<<nb, PC_BLOCK, 0, st>>>(disp, img, img_stride, n, w, c, offsets, xyz, intensity, capacity);
    lc.note("cloud_scatter_kernel");
}

}  // namespace sgmb200
