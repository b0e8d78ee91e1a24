// walk_pattern.cu -- does the ORDER in which resident warps walk a [T+1][B][128] float array bound the Burgers pair-loss kernel?
// Every warp copies the 33 frames of one row (512 B per frame, frame stride B * 512 B) from s to g, one frame in flight (as the r1 walk)
// or four.  phase = 0: all warps start at frame 0 (neighbouring warps touch neighbouring 512-byte blocks at about the same time);
// phase = 5: warp w starts at frame (5 w) mod 33 (the drift of the equal-range partition: 36.9 pairs per warp over rows of 32).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/walk_pattern scripts/diag/walk_pattern.cu ; run: /tmp/walk_pattern
#include <cstdio>
#include <cuda_runtime.h>

template <int DEPTH>
__global__ void __launch_bounds__(384, 2) walk(const float4* __restrict__ s, float4* __restrict__ g, int B, int T1, int phase, int dep) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= B) return;
    const size_t f4 = (size_t)B * 32;
    int f = (warp * phase) % T1;
    float4 q[DEPTH];
    int fl = f;
#pragma unroll
    for (int i = 0; i < DEPTH; ++i) { q[i] = __ldcs(s + fl * f4 + (size_t)warp * 32 + lane); fl = fl + 1 == T1 ? 0 : fl + 1; }
    float acc = 0.0f;
    for (int k = 0; k < T1; k += DEPTH) {
#pragma unroll
        for (int i = 0; i < DEPTH; ++i) {
            if (k + i >= T1) break;
            float4 v = q[i];
            if (k + i + DEPTH < T1) { q[i] = __ldcs(s + fl * f4 + (size_t)warp * 32 + lane); fl = fl + 1 == T1 ? 0 : fl + 1; }
            // dependent arithmetic between load and store, `dep` rounds (0 = pure copy)
            for (int r = 0; r < dep; ++r) { v.x = fmaf(v.x, 1.0001f, acc); acc = v.x * 0.5f; }
            __stcs(g + f * f4 + (size_t)warp * 32 + lane, v);
            f = f + 1 == T1 ? 0 : f + 1;
        }
    }
}

int main() {
    const int B = 148 * 2 * 12, T1 = 33;
    const size_t bytes = (size_t)B * T1 * 512;
    float4 *s, *g;
    cudaMalloc(&s, bytes * 2); cudaMalloc(&g, bytes * 2);
    cudaMemset(s, 0, bytes * 2); cudaMemset(g, 0, bytes * 2);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int depth = 1; depth <= 4; depth *= 4)
        for (int phase : {0, 5, 7})
            for (int dep : {0, 64}) {
                float best = 1e9f;
                for (int rep = 0; rep < 5; ++rep) {
                    cudaEventRecord(a);
                    for (int it = 0; it < 20; ++it) {
                        const float4* src = s + (it & 1) * (bytes / 16);
                        float4* dst = g + (it & 1) * (bytes / 16);
                        if (depth == 1) walk<1><<<148 * 2, 384>>>(src, dst, B, T1, phase, dep); else walk<4><<<148 * 2, 384>>>(src, dst, B, T1, phase, dep);
                    }
                    cudaEventRecord(b); cudaEventSynchronize(b);
                    float ms; cudaEventElapsedTime(&ms, a, b);
                    best = ms / 20 < best ? ms / 20 : best;
                }
                printf("depth %d phase %d dep %2d: %7.2f us  %7.1f GB/s\n", depth, phase, dep, best * 1e3, 2.0 * bytes / best / 1e6);
            }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
