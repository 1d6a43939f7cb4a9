// Microbenchmark (dev helper): read bandwidth of the sweeps' access pattern without any arithmetic.
// One warp per image column walks down the rows; per row it reads NA arrays x 512 B (D = 128 floats) through a cp.async
// ring of PF rows (like sgm_sweep_kernel) or with plain LDG.128 + register prefetch.  Also: a plain linear streaming read.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

constexpr int D = 128;

template <int NA, int PF, int NW, bool WRITE>
__global__ void __launch_bounds__(NW * 32) colstream_cpasync(const float* __restrict__ base, float* __restrict__ wbase, size_t arr_stride, int h, int w, float* sink) {
    extern __shared__ __align__(16) float ring[];  // [NW][PF][NA][D]
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int j = blockIdx.x * NW + wid, view = blockIdx.y;
    if (j >= w) return;
    const size_t vol = (size_t)h * w * D;
    const float* p = base + view * vol + (size_t)j * D + lane * 4;
    float* o = wbase + view * vol + (size_t)j * D + lane * 4;
    const unsigned rb = (unsigned)__cvta_generic_to_shared(ring) + ((wid * PF) * NA * D + lane * 4) * 4;
    const size_t rstep = (size_t)w * D;
    auto fetch = [&](int slot) {
#pragma unroll
        for (int a = 0; a < NA; ++a)
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(rb + (slot * NA + a) * D * 4), "l"(p + a * arr_stride) : "memory");
        p += rstep;
    };
#pragma unroll
    for (int u = 0; u < PF; ++u) {
        if (u < h) fetch(u);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    float acc = 0.f;
    for (int r0 = 0; r0 < h; r0 += PF) {
#pragma unroll
        for (int u = 0; u < PF; ++u) {
            const int r = r0 + u;
            if (r < h) {
                asm volatile("cp.async.wait_group %0;" ::"n"(PF - 1) : "memory");
                float4 s = make_float4(0, 0, 0, 0);
#pragma unroll
                for (int a = 0; a < NA; ++a) {
                    float4 v;
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(rb + (u * NA + a) * D * 4));
                    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
                }
                acc += s.x + s.y + s.z + s.w;
                if (WRITE) {
                    *reinterpret_cast<float4*>(o) = s;
                    *reinterpret_cast<float4*>(o + arr_stride) = s;
                    o += rstep;
                }
                if (r + PF < h) fetch(u);
                asm volatile("cp.async.commit_group;" ::: "memory");
            }
        }
    }
    if (acc == 12345.678f) sink[0] = acc;
}

__global__ void linear_read(const float4* __restrict__ p, size_t n, float* sink) {
    float acc = 0.f;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i + 3 * stride < n; i += 4 * stride) {
        float4 a = __ldcs(p + i), b = __ldcs(p + i + stride), c = __ldcs(p + i + 2 * stride), d = __ldcs(p + i + 3 * stride);
        acc += a.x + b.y + c.z + d.w;
    }
    if (acc == 12345.678f) sink[0] = acc;
}

template <class F>
float time_ms(F f, int reps = 5) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    f();
    cudaEventRecord(a);
    for (int i = 0; i < reps; ++i) f();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    return ms / reps;
}

template <int NA, int PF, int NW, bool WRITE>
void run(const float* buf, float* wbuf, size_t arr_stride, int h, int w, float* sink, const char* name) {
    auto k = colstream_cpasync<NA, PF, NW, WRITE>;
    const size_t smem = (size_t)NW * PF * NA * D * 4;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    dim3 grid((w + NW - 1) / NW, 2);
    const float ms = time_ms([&] { k<<<grid, NW * 32, smem>>>(buf, wbuf, arr_stride, h, w, sink); });
    const double rd = 2.0 * NA * h * w * D * 4, wr = WRITE ? 2.0 * 2 * h * w * D * 4 : 0;
    printf("%-40s NA=%d PF=%d NW=%2d blocks=%3d: %.3f ms  read %.2f TB/s  total %.2f TB/s  (%s)\n", name, NA, PF, NW, grid.x * 2, ms, rd / ms / 1e9,
           (rd + wr) / ms / 1e9, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    const int h = 375, w = 1242;
    const size_t vol = (size_t)h * w * D, arr_stride = 2 * vol;  // [array][view][h][w][D]
    float *buf, *wbuf, *sink;
    cudaMalloc(&buf, 3 * arr_stride * 4);
    cudaMalloc(&wbuf, 2 * arr_stride * 4);
    cudaMalloc(&sink, 16);
    cudaMemset(buf, 0, 3 * arr_stride * 4);
    {
        const size_t n = 3 * arr_stride / 4;
        const float ms = time_ms([&] { linear_read<<<148 * 8, 256>>>(reinterpret_cast<const float4*>(buf), n, sink); });
        printf("linear streaming read of %.2f GB: %.3f ms = %.2f TB/s\n", n * 16 / 1e9, ms, n * 16 / ms / 1e9);
    }
    run<3, 4, 19, false>(buf, wbuf, arr_stride, h, w, sink, "up-sweep pattern (3 reads)");
    run<3, 8, 19, false>(buf, wbuf, arr_stride, h, w, sink, "up-sweep pattern (3 reads)");
    run<3, 4, 10, false>(buf, wbuf, arr_stride, h, w, sink, "up-sweep pattern (3 reads)");
    run<3, 8, 10, false>(buf, wbuf, arr_stride, h, w, sink, "up-sweep pattern (3 reads)");
    run<3, 2, 19, false>(buf, wbuf, arr_stride, h, w, sink, "up-sweep pattern (3 reads)");
    run<1, 8, 19, false>(buf, wbuf, arr_stride, h, w, sink, "one array");
    run<3, 4, 19, true>(buf, wbuf, arr_stride, h, w, sink, "down-sweep pattern (3 reads, 2 writes)");
    run<3, 8, 19, true>(buf, wbuf, arr_stride, h, w, sink, "down-sweep pattern (3 reads, 2 writes)");
    run<3, 4, 10, true>(buf, wbuf, arr_stride, h, w, sink, "down-sweep pattern (3 reads, 2 writes)");
    return 0;
}
