// Microbenchmark (dev helper): issue throughput and latency of the warp-wide reductions and shuffles used by the sweeps.
#include <cstdio>
#include <cuda_runtime.h>

template <int OP, bool DEP>
__global__ void bench(float* out, long long* cyc, int iters) {
    float v[8];
    int iv[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        v[k] = threadIdx.x * 1.5f + k;
        iv[k] = threadIdx.x * 3 + k;
    }
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int s = DEP ? (k + 7) & 7 : k;  // DEP: chain through the previous result
            if (OP == 0) {
                float r;
                asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v[s]));
                v[k] = r + 1.0f;
            } else if (OP == 1) {
                iv[k] = __reduce_min_sync(0xffffffffu, iv[s]) + 1;
            } else if (OP == 2) {
                iv[k] = (int)__reduce_or_sync(0xffffffffu, (unsigned)iv[s]) + 1;
            } else if (OP == 3) {
                v[k] = __shfl_sync(0xffffffffu, v[s], (threadIdx.x + 1) & 31) + 1.0f;
            } else {
                v[k] = fminf(v[s], v[(s + 1) & 7]) + 1.0f;
            }
        }
    }
    const long long t1 = clock64();
    __syncthreads();
    float acc = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) acc += v[k] + iv[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0 && blockIdx.x == 0) cyc[0] = t1 - t0;
}

int main() {
    float* out;
    long long* cyc;
    cudaMalloc(&out, 148 * 1024 * 4);
    cudaMalloc(&cyc, 8);
    const char* names[] = {"redux.sync.min.f32 (CREDUX)", "__reduce_min_sync s32 (REDUX)", "__reduce_or_sync (REDUX)", "shfl.sync.idx", "fminf + fadd"};
    const int iters = 2000;
    for (int op = 0; op < 5; ++op)
        for (int dep = 0; dep < 2; ++dep)
            for (int warps : {1, 4, 8, 16, 32}) {
                auto launch = [&](auto k) { k<<<148, warps * 32>>>(out, cyc, iters); };
                if (op == 0) dep ? launch(bench<0, true>) : launch(bench<0, false>);
                if (op == 1) dep ? launch(bench<1, true>) : launch(bench<1, false>);
                if (op == 2) dep ? launch(bench<2, true>) : launch(bench<2, false>);
                if (op == 3) dep ? launch(bench<3, true>) : launch(bench<3, false>);
                if (op == 4) dep ? launch(bench<4, true>) : launch(bench<4, false>);
                long long c = 0;
                cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
                const double per_op_warp = (double)c / (iters * 8.0);
                printf("%-32s %s warps/SM=%2d: %.1f cycles per op per warp, %.2f cycles per op per SM\n", names[op], dep ? "dependent  " : "independent", warps,
                       per_op_warp, per_op_warp / warps);
            }
    return 0;
}
