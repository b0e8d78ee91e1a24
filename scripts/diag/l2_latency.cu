// Diagnostic: CTA<->CTA signalling latency through L2 on one GPU (two co-resident CTAs ping-pong on a flag word).
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__global__ void pingpong(volatile u64* f, int rounds, int mode, long long* out) {
    if (threadIdx.x != 0) return;
    const int me = blockIdx.x;          // 0 or 1
    volatile u64* mine = f + 32 * me; volatile u64* theirs = f + 32 * (1 - me);
    long long t0 = clock64();
    for (int i = 1; i <= rounds; ++i) {
        if ((i & 1) == me) {
            if (mode == 1) __threadfence();
            *theirs = (u64)i;
        } else {
            long long s = clock64();
            while (*mine < (u64)i) { if (clock64() - s > 2000000000LL) { out[me] = -1; return; } }
            if (mode == 1) __threadfence();
        }
    }
    out[me] = clock64() - t0;
}
int main() {
    u64* f; long long* out; cudaMalloc(&f, 1024); cudaMalloc(&out, 16);
    for (int mode = 0; mode < 2; ++mode) {
        cudaMemset(f, 0, 1024);
        pingpong<<<2, 32>>>(f, 4000, mode, out);
        long long c[2]; cudaMemcpy(c, out, 16, cudaMemcpyDeviceToHost);
        printf("mode %d (%s): %.0f cycles per one-way hop\n", mode, mode ? "fence + flag" : "flag only", (double)c[0] / 4000);
    }
    return 0;
}
