// Microbenchmark (dev helper): one-way latency of the tagged-payload exchange between two CTAs on different SMs.
// CTA 0 writes a 512-byte vector of {value, tag} words, CTA 1 polls it and answers; time per one-way hop in cycles.
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void pingpong(unsigned long long* buf, int iters, long long* out, int other) {
    if (blockIdx.x != 0 && blockIdx.x != other) return;
    const int me = blockIdx.x == 0 ? 0 : 1;
    const int lane = threadIdx.x;
    unsigned long long* mine = buf + me * 128;        // I write here (64 words of 16 B = v2.u64 per lane x 2)
    unsigned long long* theirs = buf + (1 - me) * 128;
    long long t0 = clock64();
    for (int it = 1; it <= iters; ++it) {
        if (me == 0 || it > 1 || true) {
            if (me == 1) {  // wait first
                bool ok;
                do {
                    unsigned long long x, y, z, w;
                    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(x), "=l"(y) : "l"(theirs + lane * 2) : "memory");
                    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(z), "=l"(w) : "l"(theirs + 64 + lane * 2) : "memory");
                    ok = (x >> 32) == it && (y >> 32) == it && (z >> 32) == it && (w >> 32) == it;
                } while (!__all_sync(0xffffffffu, ok));
            }
            const unsigned long long v = ((unsigned long long)it << 32) | lane;
            if (MODE == 0) {
                asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(mine + lane * 2), "l"(v), "l"(v) : "memory");
                asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(mine + 64 + lane * 2), "l"(v), "l"(v) : "memory");
            } else if (MODE == 1) {
                asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(mine + lane * 2), "l"(v), "l"(v) : "memory");
                asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(mine + 64 + lane * 2), "l"(v), "l"(v) : "memory");
                __threadfence();
            } else {
                asm volatile("st.global.cg.v2.u64 [%0], {%1, %2};" ::"l"(mine + lane * 2), "l"(v), "l"(v) : "memory");
                asm volatile("st.global.cg.v2.u64 [%0], {%1, %2};" ::"l"(mine + 64 + lane * 2), "l"(v), "l"(v) : "memory");
            }
            if (me == 0) {
                bool ok;
                do {
                    unsigned long long x, y, z, w;
                    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(x), "=l"(y) : "l"(theirs + lane * 2) : "memory");
                    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(z), "=l"(w) : "l"(theirs + 64 + lane * 2) : "memory");
                    ok = (x >> 32) == it && (y >> 32) == it && (z >> 32) == it && (w >> 32) == it;
                } while (!__all_sync(0xffffffffu, ok));
            }
        }
    }
    long long t1 = clock64();
    if (lane == 0 && me == 0) out[0] = t1 - t0;
}

// plain load latency of an L2-resident line (dependent chain)
__global__ void l2_latency(unsigned long long* buf, int iters, long long* out) {
    unsigned long long idx = 0;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        unsigned long long x;
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(x) : "l"(buf + idx) : "memory");
        idx = x & 1;
    }
    long long t1 = clock64();
    out[0] = t1 - t0 + idx;
}

int main() {
    unsigned long long* buf;
    long long* out;
    cudaMalloc(&buf, 4096);
    cudaMalloc(&out, 64);
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int iters = 2000;
    for (int mode = 0; mode < 3; ++mode)
        for (int other : {1, 2, sms / 2, sms - 1}) {
            cudaMemset(buf, 0, 4096);
            if (mode == 0) pingpong<0><<<sms, 32>>>(buf, iters, out, other);
            if (mode == 1) pingpong<1><<<sms, 32>>>(buf, iters, out, other);
            if (mode == 2) pingpong<2><<<sms, 32>>>(buf, iters, out, other);
            long long c = 0;
            cudaError_t e = cudaMemcpy(&c, out, 8, cudaMemcpyDeviceToHost);
            printf("mode %d (0 st.relaxed.gpu, 1 + threadfence, 2 st.cg) peer block %3d: %.0f cycles per one-way hop (%s)\n", mode, other,
                   (double)c / iters / 2, cudaGetErrorString(e));
        }
    cudaMemset(buf, 0, 4096);
    l2_latency<<<1, 1>>>(buf, 10000, out);
    long long c = 0;
    cudaMemcpy(&c, out, 8, cudaMemcpyDeviceToHost);
    printf("dependent ld.relaxed.gpu of an L2 line: %.0f cycles\n", (double)c / 10000);
    return 0;
}
