// f32x2_throughput.cu -- does packed fp32 (add/mul/fma .f32x2, sm_100) buy issue slots on B200?
// Each kernel runs ITERS x 8 independent dependency chains per thread; reports warp-instructions per cycle per SM and
// "fp32 lane-ops" per cycle per SM for: scalar FADD (rn), packed FADD2, scalar + ALU mix, packed + ALU mix.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o f32x2_throughput f32x2_throughput.cu
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 4096;

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float seed, int alu) {
    float a[8];
    int q[4];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 0.001f + i;
#pragma unroll
    for (int i = 0; i < 4; ++i) q[i] = threadIdx.x + i + alu;
    const float c = seed * 1.0001f;
    for (int it = 0; it < ITERS; ++it) {
        if (MODE == 0 || MODE == 2) {       // 8 scalar adds
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = __fadd_rn(a[i], c);
        } else if (MODE == 1 || MODE == 3) {                 // 4 packed adds = the same 8 lane-ops
#pragma unroll
            for (int i = 0; i < 8; i += 2) {
                float2 r = __fadd2_rn(make_float2(a[i], a[i + 1]), make_float2(c, c));
                a[i] = r.x; a[i + 1] = r.y;
            }
        } else if (MODE == 4) {             // 8 scalar fmul
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = __fmul_rn(a[i], c);
        } else if (MODE == 5) {
#pragma unroll
            for (int i = 0; i < 8; i += 2) {
                float2 r = __fmul2_rn(make_float2(a[i], a[i + 1]), make_float2(c, c));
                a[i] = r.x; a[i + 1] = r.y;
            }
        } else if (MODE == 6) {             // scalar FFMA
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = __fmaf_rn(a[i], c, c);
        } else if (MODE == 7) {
#pragma unroll
            for (int i = 0; i < 8; i += 2) {
                float2 r = __ffma2_rn(make_float2(a[i], a[i + 1]), make_float2(c, c), make_float2(c, c));
                a[i] = r.x; a[i + 1] = r.y;
            }
        }
        if (MODE == 2 || MODE == 3) {       // + 4 ALU-pipe instructions (LOP3 / SHF chains)
#pragma unroll
            for (int i = 0; i < 4; ++i) q[i] = (q[i] ^ (q[i] >> 3)) & 0x7fffffff;
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + (float)(q[0] + q[1] + q[2] + q[3]);
}

template <int MODE>
void run(const char* name, int fp_inst, int alu_inst) {
    int dev = 0, sms = 0, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    const int ctas = sms * 8;
    float* out;
    cudaMalloc(&out, sizeof(float) * ctas * 256);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<ctas, 256>>>(out, 1.0f, 0);
    cudaEventRecord(e0);
    for (int r = 0; r < 5; ++r) k<MODE><<<ctas, 256>>>(out, 1.0f, r);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double cycles = ms * 1e-3 / 5 * khz * 1e3;
    const double warps_per_sm = 8.0 * 8;
    const double inst = warps_per_sm * ITERS * (fp_inst + alu_inst);
    printf("%-28s %8.3f ms/launch  warp-inst/cycle/SM %.2f  fp32 lane-ops/cycle/SM %.1f  (max clock %d MHz assumed)\n", name, ms / 5,
           inst / cycles, warps_per_sm * ITERS * 8 * 32 / cycles, khz / 1000);
    cudaFree(out);
}

int main() {
    run<0>("FADD x8", 8, 0);
    run<1>("FADD2 x4", 4, 0);
    run<4>("FMUL x8", 8, 0);
    run<5>("FMUL2 x4", 4, 0);
    run<6>("FFMA x8", 8, 0);
    run<7>("FFMA2 x4", 4, 0);
    run<2>("FADD x8 + 8 ALU", 8, 8);
    run<3>("FADD2 x4 + 8 ALU", 4, 8);
    return 0;
}
