// pdec_common.cuh -- shared device helpers for the sm_100a physics kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/pdec.h"

#define PDEC_FULL_MASK 0xffffffffu

// `closed` argument of the 2-D entry points -> 4-bit side mask (bit 0: y-lower side solid, 1: y-upper, 2: x-lower, 3: x-upper).
// 0 = OPEN everywhere, 1 (or any value without bit 4) = CLOSED everywhere, PDEC_CLOSED_SIDES(...) = per side (include/pdec.h).
__host__ __device__ __forceinline__ unsigned pdec_closed_sides(int closed) {
    return (closed & 0x10) ? (unsigned)(closed & 0xF) : (closed ? 0xFu : 0u);
}

namespace pdec {

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(PDEC_FULL_MASK, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(PDEC_FULL_MASK, v, o));
    return v;
}

// cp.async (LDGSTS) global -> shared, 16 or 4 bytes
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async4(void* smem, const void* gmem) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// Warp-cooperative async copy of one row of n floats, global -> shared (vectorised when 16B-aligned).
__device__ __forceinline__ void warp_row_load_async(float* srow, const float* grow, int n, bool vec4) {
    const int lane = lane_id();
    if (vec4) {
        for (int i = lane * 4; i < n; i += 128) cp_async16(srow + i, grow + i);
    } else {
        for (int i = lane; i < n; i += 32) cp_async4(srow + i, grow + i);
    }
}
// Warp-cooperative store of one row shared -> global (float4 when aligned).
__device__ __forceinline__ void warp_row_store(float* grow, const float* srow, int n, bool vec4) {
    const int lane = lane_id();
    if (vec4) {
        for (int i = lane * 4; i < n; i += 128)
            *reinterpret_cast<float4*>(grow + i) = *reinterpret_cast<const float4*>(srow + i);
    } else {
        for (int i = lane; i < n; i += 32) grow[i] = srow[i];
    }
}

// Deterministic scatter-add into a warp-private shared-memory row: the transpose of a gather, accumulated in
// a fixed order (instruction order, then ascending lane id inside one instruction) without float atomics.
// All 32 lanes must call; inactive lanes pass active=false.
__device__ __forceinline__ void warp_ordered_scatter_add(float* acc, int key, float val, bool active) {
    const int lane = lane_id();
    const int k = active ? key : (-1 - lane);            // unique key for inactive lanes: no peers
    const unsigned peers = __match_any_sync(PDEC_FULL_MASK, k);
    const int rank = __popc(peers & ((1u << lane) - 1u));
    int r = 0;
    while (true) {
        const unsigned pending = __ballot_sync(PDEC_FULL_MASK, active && rank >= r);
        if (!pending) break;
        if (active && rank == r) acc[key] += val;
        __syncwarp();
        ++r;
    }
}

// numpy.linspace(0.5/n, 1-0.5/n, n)[i] in float64, cast to float32
// (reference: CenteredGrid.getpoints, PF/physics/field/grid.py:176-181).
__device__ __forceinline__ float cell_center_local(int i, int n) {
    const double start = 0.5 / (double)n;
    const double stop = 1.0 - 0.5 / (double)n;
    if (n == 1) return (float)start;
    const double step = (stop - start) / (double)(n - 1);
    const double v = (i == n - 1) ? stop : __dadd_rn(__dmul_rn((double)i, step), start);
    return (float)v;
}

// clamp a float index into [0, n-1] (NaN -> 0) and convert
__device__ __forceinline__ int clamp_index(float f, int n) {
    f = fminf(fmaxf(f, 0.0f), (float)(n - 1));
    return (int)f;
}
__device__ __forceinline__ int wrap_index(float f, int n) {
    // tf.mod(tf.mod(x, n) + n, n) on a float holding an integer value (PF/tf/tf_backend.py:521-522)
    const float nf = (float)n;
    float m = fmodf(f, nf);
    if (m < 0.0f) m += nf;
    int i = (int)m;
    if (i >= n) i -= n;
    if (i < 0) i = 0;
    return i;
}

// ---- interpolation set-up: PF/tf/tf_backend.py:462-476 --------------------------------------------
template <int MODE>
__device__ __forceinline__ void interp_setup(float xi, int n, int& fi, int& ci, float& w0, float& w1) {
    const float base = floorf(xi);
    if (MODE == PDEC_MODE_CONSTANT) {
        const float hi = (float)(n - 1);
        const float fc = fminf(fmaxf(base, 0.0f), hi);
        const float cc = fminf(fmaxf(base + 1.0f, 0.0f), hi);
        w0 = xi - fc;  // weight of the ceil sample
        w1 = cc - xi;  // weight of the floor sample
        fi = (int)fc;
        ci = (int)cc;
    } else {
        w0 = xi - base;
        w1 = 1.0f - w0;
        if (MODE == PDEC_MODE_REPLICATE) {
            fi = clamp_index(base, n);
            ci = clamp_index(base + 1.0f, n);
        } else {
            fi = wrap_index(base, n);
            ci = wrap_index(base + 1.0f, n);
        }
    }
}

// ---- order-independent accumulation: 64-bit fixed point (resample.cu, the advection adjoints of fluid.cu) -------------------------
// header[0] = bits of max|g| (non-negative floats order like their bit patterns; written with an integer atomicMax).
// scale = 2^(bits - exponent of max|g|): every |contribution| <= max|g| maps below 2^bits; integer adds commute, so the sums do not
// depend on the order the atomics land in: bit-reproducible without float atomics.
__device__ __forceinline__ int fixed_shift(const unsigned* header, int bits, bool& valid) {
    const unsigned u = header[0];                                   // bits of max|g|: a non-negative finite float
    valid = u != 0u && u < 0x7f800000u;
    const int e = (int)(u >> 23) - 126;                             // max|g| = f * 2^e, f in [0.5, 1) (as frexpf for normal numbers)
    return min(max(bits - e, -120), 120);
}
__device__ __forceinline__ float fixed_scale(const unsigned* header, int bits) {
    bool valid;
    const int k = fixed_shift(header, bits, valid);
    return valid ? __int_as_float((unsigned)(k + 127) << 23) : 0.0f;     // 2^k, exact
}
// 1 / scale = 2^-k (exact); 0 when there is nothing to convert
__device__ __forceinline__ float fixed_inv_scale(const unsigned* header, int bits) {
    bool valid;
    const int k = fixed_shift(header, bits, valid);
    return valid ? __int_as_float((unsigned)(127 - k) << 23) : 0.0f;
}
// fixed-point sum -> float: ONE rounding (integer -> float), then an exact power-of-two scaling
__device__ __forceinline__ float fixed_to_float(unsigned long long acc, float inv_scale) { return __ll2float_rn((long long)acc) * inv_scale; }
__device__ __forceinline__ void fixed_add(unsigned long long* cell, float v, float scale) {     // two's-complement add
    const float sv = v * scale;
    if (sv == sv && fabsf(sv) < 9.0e18f) atomicAdd(cell, (unsigned long long)__float2ll_rn(sv));
}

}  // namespace pdec

#define PDEC_CHECK_LAUNCH()                                   \
    do {                                                      \
        cudaError_t e__ = cudaPeekAtLastError();              \
        if (e__ != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;  \
    } while (0)
