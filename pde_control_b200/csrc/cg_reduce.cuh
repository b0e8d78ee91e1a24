// cg_reduce.cuh -- block reductions shared by the conjugate-gradient kernels (cg.cu, cg_tiled.cu).
#pragma once
#include "pdec_common.cuh"

namespace pdec {

// math.divide_no_nan (PF/math/blas.py:44,54)
__device__ __forceinline__ float div_no_nan(float a, float b) { return b == 0.0f ? 0.0f : a / b; }
// the same with the hardware reciprocal (<= 2 ulp): for the CG coefficients of the on-chip kernels, where every thread of the CTA
// waits on this one division twice per iteration and the IEEE sequence is ~150 cycles of dependent instructions.  x and r are
// updated with the SAME rounded coefficient, so the recurrence stays consistent; iterates are not the parity oracle.
__device__ __forceinline__ float div_no_nan_fast(float a, float b) { return b == 0.0f ? 0.0f : __fdividef(a, b); }

// Four sums and one maximum in ONE block reduction (one barrier).  Stage 1 is a transposing butterfly: at xor 16 the half-warps
// trade two of the four sums, at xor 8 one of the remaining two, so lanes 0/8/16/24 end up with the warp totals of s1/s2/s4/s5
// after 6 shuffles (+5 for the maximum).  Stage 2: lane (h, w') sums the partials of warp w' for (s1|s2, s4|s5, max), a 16-lane
// butterfly, then the half-warps swap.  Every lane of every warp gets bit-identical totals (fixed order, commutative pairs).
template <int WARPS>
__device__ __forceinline__ void block_reduce5(float& s1, float& s2, float& s4, float& s5, float& mx, float* red /* [5][16] */) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const bool h16 = (lane & 16) != 0, h8 = (lane & 8) != 0;
    float k0 = h16 ? s4 : s1, k1 = h16 ? s5 : s2;
    k0 += __shfl_xor_sync(0xffffffffu, h16 ? s1 : s4, 16);
    k1 += __shfl_xor_sync(0xffffffffu, h16 ? s2 : s5, 16);
    float keep = h8 ? k1 : k0;
    keep += __shfl_xor_sync(0xffffffffu, h8 ? k0 : k1, 8);
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) keep += __shfl_xor_sync(0xffffffffu, keep, o);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((lane & 7) == 0) red[(lane >> 3) * 16 + w] = keep;
    if (lane == 1) red[64 + w] = mx;
    __syncthreads();
    const int wq = lane & 15, h = lane >> 4;
    float A = 0.0f, B = 0.0f, C = 0.0f;
    if (wq < WARPS) { A = red[h * 16 + wq]; B = red[(2 + h) * 16 + wq]; C = red[64 + wq]; }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
        A += __shfl_xor_sync(0xffffffffu, A, o);
        B += __shfl_xor_sync(0xffffffffu, B, o);
        C = fmaxf(C, __shfl_xor_sync(0xffffffffu, C, o));
    }
    const float A2 = __shfl_xor_sync(0xffffffffu, A, 16), B2 = __shfl_xor_sync(0xffffffffu, B, 16);
    s1 = h ? A2 : A; s2 = h ? A : A2;
    s4 = h ? B2 : B; s5 = h ? B : B2;
    mx = C;
}

}  // namespace pdec
