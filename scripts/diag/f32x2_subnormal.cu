// do the packed fp32 instructions keep subnormals (like scalar FADD/FMUL/FFMA without .ftz)?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(float* o, float a, float b) {
    float2 x = make_float2(a, a), y = make_float2(b, b);
    float2 r1 = __fadd2_rn(x, y), r2 = __fmul2_rn(x, make_float2(0.5f, 0.5f)), r3 = __ffma2_rn(x, make_float2(-1.f, -1.f), y);
    o[0] = r1.x; o[1] = r2.x; o[2] = r3.x;
    o[3] = __fadd_rn(a, b); o[4] = __fmul_rn(a, 0.5f); o[5] = __fmaf_rn(a, -1.f, b);
}
int main() {
    float* o; cudaMallocManaged(&o, 64);
    k<<<1, 1>>>(o, 3e-40f, 1e-40f);
    cudaDeviceSynchronize();
    printf("packed: add %g mul %g fma %g | scalar: add %g mul %g fma %g\n", o[0], o[1], o[2], o[3], o[4], o[5]);
    return 0;
}
