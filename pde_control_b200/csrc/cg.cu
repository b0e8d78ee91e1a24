// cg.cu -- conjugate-gradient Poisson pressure solve for sm_100a (F7 adjoint reuses it).
//
// Operator (PF/physics/pressuresolver/sparse.py:114-221, no obstacles): (A p)[i,j] = sum of in-domain neighbours
// + diag * p[i,j], diag = -4 (open) or -(number of in-domain neighbours) clamped <= -1 (closed).
// Recurrence (PF/math/blas.py:37-55):  t = m.Am;  a = (m.r)/t;  x += a m;  r -= a Am;  m = r - ((r.Am) m)/t;  Am = A m,
// with divide_no_nan, stop when max|r| < accuracy.
//
// Two kernels:
//  * cg_smem_kernel: ONE CTA PER BATCH ELEMENT, persistent over all iterations (grids up to 16384 cells, e.g. 128x128).
//    x, r, Am live in registers (16 cells per thread), the search direction m lives in shared memory (the only vector the
//    5-point stencil reads across threads).  Per iteration: matvec + three dots + three axpys with NO global-memory traffic,
//    two warp-shuffle block reductions ((m.Am, m.r) and (r.Am, max|r|)) and three __syncthreads.  HBM is touched once
//    (rhs in, p out).  The iteration count is decided on the device -> CUDA-graph capturable.
//  * cg_global_kernel: cooperative persistent kernel for large grids (vectors in global memory / L2), two passes per
//    iteration (matvec fused with the dots of pass 1; axpys fused with the dot of pass 2; the m-update is folded into the
//    next matvec), grid-wide barriers between passes, deterministic two-stage reductions.
#include "pdec_common.cuh"
#include "cg_reduce.cuh"
#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace pdec {


// neighbour-availability nibble: bit0 up (i-1), bit1 down (i+1), bit2 left (j-1), bit3 right (j+1)
__device__ __forceinline__ unsigned cell_nibble(int i, int j, int ny, int nx) {
    return (i > 0 ? 1u : 0u) | (i < ny - 1 ? 2u : 0u) | (j > 0 ? 4u : 0u) | (j < nx - 1 ? 8u : 0u);
}
// diagonal = -(accessible neighbours) clamped <= -1 (sparse.py:186-213): an in-domain neighbour always counts, the outside counts on
// OPEN sides only.  `sides` = 4-bit mask of SOLID sides in the nibble's bit order (y-lower, y-upper, x-lower, x-upper)
__device__ __forceinline__ float cell_diag(unsigned nib, unsigned sides) {
    const float d = -(float)(4 - __popc(~nib & 0xFu & sides));
    return fminf(d, -1.0f);
}
// sum in the column order of the sorted sparse row (sparse.py:173-176): i-1, j-1, centre, j+1, i+1
__device__ __forceinline__ float stencil_apply(const float* m, int idx, int nx, unsigned nib, float diag) {
    float s = 0.0f;
    if (nib & 1u) s += m[idx - nx];
    if (nib & 4u) s += m[idx - 1];
    s += diag * m[idx];
    if (nib & 8u) s += m[idx + 1];
    if (nib & 2u) s += m[idx + nx];
    return s;
}

template <int THREADS>
__device__ __forceinline__ void block_reduce2(float& a, float& b, float (*red)[32], bool second_is_max) {
    constexpr int NW = THREADS / 32;
    a = warp_sum(a);
    b = second_is_max ? warp_max(b) : warp_sum(b);
    if (NW == 1) return;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { red[0][w] = a; red[1][w] = b; }
    __syncthreads();
    a = lane < NW ? red[0][lane] : 0.0f;
    b = lane < NW ? red[1][lane] : 0.0f;
    a = warp_sum(a);
    b = second_is_max ? warp_max(b) : warp_sum(b);
}

template <int THREADS, int CPT>
__global__ void __launch_bounds__(THREADS, 1) cg_smem_kernel(const float* __restrict__ rhs, const float* __restrict__ guess, float* __restrict__ out,
                                                             int* __restrict__ iters, float* __restrict__ resid, int ny, int nx, int closed,
                                                             float accuracy, int max_it) {
    extern __shared__ __align__(16) float sm_m[];
    __shared__ float redA[2][32], redB[2][32];
    const int N = ny * nx, tid = threadIdx.x;
    const size_t base = (size_t)blockIdx.x * N;
    float x[CPT], r[CPT], am[CPT];
    unsigned long long nibs[(CPT + 15) / 16];
#pragma unroll
    for (int k = 0; k < (CPT + 15) / 16; ++k) nibs[k] = 0ull;
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        const int idx = tid + k * THREADS;
        x[k] = 0.0f; r[k] = 0.0f; am[k] = 0.0f;
        if (idx < N) {
            const int i = idx / nx, j = idx - i * nx;
            nibs[k >> 4] |= (unsigned long long)cell_nibble(i, j, ny, nx) << (4 * (k & 15));
            r[k] = rhs[base + idx];
            if (guess != nullptr) { x[k] = guess[base + idx]; sm_m[idx] = x[k]; }
        }
    }
    if (guess != nullptr) {                 // momentum = k - A x0   (blas.py:24-26)
        __syncthreads();
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            const int idx = tid + k * THREADS;
            if (idx < N) {
                const unsigned nib = (unsigned)(nibs[k >> 4] >> (4 * (k & 15))) & 15u;
                r[k] -= stencil_apply(sm_m, idx, nx, nib, cell_diag(nib, closed));
            }
        }
        __syncthreads();
    }
    float rmax = 0.0f, dummy = 0.0f;
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        const int idx = tid + k * THREADS;
        if (idx < N) { sm_m[idx] = r[k]; rmax = fmaxf(rmax, fabsf(r[k])); }
    }
    block_reduce2<THREADS>(dummy, rmax, redB, true);   // contains a __syncthreads when THREADS > 32
    if (THREADS == 32) __syncwarp();
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        const int idx = tid + k * THREADS;
        if (idx < N) {
            const unsigned nib = (unsigned)(nibs[k >> 4] >> (4 * (k & 15))) & 15u;
            am[k] = stencil_apply(sm_m, idx, nx, nib, cell_diag(nib, closed));
        }
    }
    int it = 0;
    while (it < max_it && (accuracy <= 0.0f || rmax >= accuracy)) {
        float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            const int idx = tid + k * THREADS;
            if (idx < N) { const float m = sm_m[idx]; s1 += m * am[k]; s2 += m * r[k]; }
        }
        block_reduce2<THREADS>(s1, s2, redA, false);
        const float tmp = s1, a = div_no_nan(s2, tmp);
        float s3 = 0.0f, mx = 0.0f;
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            const int idx = tid + k * THREADS;
            if (idx < N) {
                const float m = sm_m[idx];
                x[k] += a * m;
                r[k] -= a * am[k];
                s3 += r[k] * am[k];
                mx = fmaxf(mx, fabsf(r[k]));
            }
        }
        block_reduce2<THREADS>(s3, mx, redB, true);
        // every thread has read its own m above; cross-thread (stencil) reads of the old m finished before reduction A
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            const int idx = tid + k * THREADS;
            if (idx < N) sm_m[idx] = r[k] - div_no_nan(s3 * sm_m[idx], tmp);
        }
        if (THREADS == 32) __syncwarp(); else __syncthreads();
#pragma unroll
        for (int k = 0; k < CPT; ++k) {
            const int idx = tid + k * THREADS;
            if (idx < N) {
                const unsigned nib = (unsigned)(nibs[k >> 4] >> (4 * (k & 15))) & 15u;
                am[k] = stencil_apply(sm_m, idx, nx, nib, cell_diag(nib, closed));
            }
        }
        rmax = mx;
        ++it;
    }
#pragma unroll
    for (int k = 0; k < CPT; ++k) {
        const int idx = tid + k * THREADS;
        if (idx < N) out[base + idx] = x[k];
    }
    if (tid == 0) {
        if (iters) iters[blockIdx.x] = it;
        if (resid) resid[blockIdx.x] = rmax;
    }
}

// -------------------------------------------------------------------------------------------------------------
// cg_block_kernel: production one-CTA-per-element kernel for grids up to 128 x 128 (second generation).
// An earlier tile kernel (4 x 4 cells per thread, vectors in shared memory) was bound by shared-memory instruction issue (88 LDS/STS per
// thread and iteration + 40 reduction shuffles on
// 32 warps).  Here a thread owns a BLOCK of 8 rows x 4 consecutive columns (cells i = 8*warp + ky, j = 4*lane + kx):
//   * r, Am, m of the 32 cells stay in registers (96 registers; 16 warps x 128 registers fill the register file);
//   * left/right neighbours come from registers or ONE shuffle per row and side; up/down neighbours from registers except
//     the block's first/last row, which read the neighbouring warp's edge row from shared memory (2 LDS.128 per thread);
//     only those two edge rows are ever published (2 STS.128);
//   * x is the one vector that does not fit the register file: it lives in shared memory as float4 (8 LDS.128 + 8 STS.128);
//   * out-of-domain neighbours contribute 0 and the diagonal carries the boundary rule (-4 open; -(in-domain neighbours)
//     closed, sparse.py:151-221), so there are no ghost cells to maintain;
//   * both block reductions carry two values in ONE butterfly (the half-warps reduce different values: 5 shuffles), and the
//     second stage reads the 2 x WARPS partials with one LDS per lane.
// Per iteration: 3 barriers, ~24 shuffles, 20 128-bit shared-memory accesses and ~11 FP32 instructions per cell.
// -------------------------------------------------------------------------------------------------------------
// The reduction comes in two halves so that independent work can sit between them (in the shadow of the shuffle latency and the
// barrier): _begin = warp butterfly + one partial per warp and value into shared memory; _end = barrier + second stage.
template <int WARPS>
__device__ __forceinline__ void block_reduce_pair_begin(float a, float b, float* red, const bool second_is_max) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const bool hi = lane >= 16;
    const bool mx = hi && second_is_max;
    float keep = hi ? b : a;
    float recv = __shfl_xor_sync(0xffffffffu, hi ? a : b, 16);
    keep = mx ? fmaxf(keep, recv) : keep + recv;
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
        recv = __shfl_xor_sync(0xffffffffu, keep, o);
        keep = mx ? fmaxf(keep, recv) : keep + recv;
    }
    if ((lane & 15) == 0) red[(lane >> 4) * 16 + w] = keep;
}
template <int WARPS>
__device__ __forceinline__ void block_reduce_pair_end(float& a, float& b, float* red, const bool second_is_max) {
    const int lane = threadIdx.x & 31;
    const bool hi = lane >= 16;
    const bool mx = hi && second_is_max;
    float recv;
    __syncthreads();
    if (WARPS >= 16) {      // throughput-bound iteration: one LDS + five shuffles per lane
        float v = red[lane];                                  // lane = value * 16 + warp
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
            recv = __shfl_xor_sync(0xffffffffu, v, o);
            v = mx ? fmaxf(v, recv) : v + recv;
        }
        recv = __shfl_xor_sync(0xffffffffu, v, 16);
        a = hi ? recv : v;
        b = hi ? v : recv;
        return;
    }
    // latency-bound iteration (<= 8 warps): every lane reads all 2 x WARPS partials (broadcast LDS.128) and adds them in a fixed
    // tree -- ~45 cycles instead of five dependent shuffles; bit-identical totals in every lane of every warp
    const float4* r4 = reinterpret_cast<const float4*>(red);
    float ta[WARPS / 4], tb[WARPS / 4];
#pragma unroll
    for (int q = 0; q < WARPS / 4; ++q) {
        const float4 va = r4[q], vb = r4[4 + q];
        ta[q] = (va.x + va.y) + (va.z + va.w);
        tb[q] = second_is_max ? fmaxf(fmaxf(vb.x, vb.y), fmaxf(vb.z, vb.w)) : (vb.x + vb.y) + (vb.z + vb.w);
    }
#pragma unroll
    for (int q = WARPS / 4; q > 1; q >>= 1)
#pragma unroll
        for (int k = 0; k < q / 2; ++k) {
            ta[k] = ta[2 * k] + ta[2 * k + 1];
            tb[k] = second_is_max ? fmaxf(tb[2 * k], tb[2 * k + 1]) : tb[2 * k] + tb[2 * k + 1];
        }
    a = ta[0];
    b = tb[0];
}
template <int WARPS>
__device__ __forceinline__ void block_reduce_pair(float& a, float& b, float* red, const bool second_is_max) {
    block_reduce_pair_begin<WARPS>(a, b, red, second_is_max);
    block_reduce_pair_end<WARPS>(a, b, red, second_is_max);
}

// PROJ != 0: the whole pressure projection in this ONE launch (divergence_free, PF/physics/fluid.py:214-243).  The right-hand side
// is formed from the staggered field in the prologue and the corrected field is written in the epilogue, so neither the divergence nor
// the pressure makes a round trip through HBM and the two stencil launches around the solve disappear:
//   PROJ_FWD: rhs = div(mask v)                      out = mask v - mask grad(x dx0) / dx       (F4 -> F5 -> F6)
//   PROJ_BWD: rhs = grad^T(g) [+ add_rhs]            out = mask g + mask div^T(x [+ add_x])     (F6^T -> F5 -> F4^T; A symmetric)
// Every expression keeps the operation order (and roundings: explicit _rn intrinsics, no contraction) of the stand-alone kernels
// in fluid.cu, so the fused launch is bit-identical to the three-launch composition.
constexpr int PROJ_NONE = 0, PROJ_FWD = 1, PROJ_BWD = 2;
struct ProjArgs {
    const float* in_y;      // [B, ny+1, nx]  v_y (fwd) / dL/dv'_y (bwd)
    const float* in_x;      // [B, ny, nx+1]
    float* out_y;           // [B, ny+1, nx]  v'_y (fwd) / dL/dv_y (bwd)
    float* out_x;           // [B, ny, nx+1]
    const float* add_rhs;   // bwd: upstream dL/dp or nullptr
    const float* add_x;     // bwd: upstream dL/ddiv or nullptr
    float* store_rhs;       // fwd: divergence output or nullptr
    float dx0, dxy, dxx;    // p *= dx0 (fluid.py:240), gradient / dxy, / dxx (staggered_grid.py:219-231)
};

template <int WARPS, bool FULL, bool FUSED, int PROJ = PROJ_NONE>
__global__ void __launch_bounds__(32 * WARPS, 1) cg_block_kernel(const float* __restrict__ rhs, const float* __restrict__ guess, float* __restrict__ out,
                                                                 int* __restrict__ iters, float* __restrict__ resid, int ny, int nx, int closed,
                                                                 float accuracy, int max_it, const ProjArgs pa = ProjArgs()) {
    constexpr int ROWS = 8 * WARPS;
    extern __shared__ __align__(16) float smem_dyn[];
    __shared__ __align__(16) float redA[32], redB[32], red5[2][80];
    float4* sm_x = reinterpret_cast<float4*>(smem_dyn);          // [ROWS][32] float4 (row pitch 128 floats)
    float4* sm_e = sm_x + ROWS * 32;                            // [2*WARPS + 2][32]: 0 = zeros, 1+2w / 2+2w = first / last row of warp w
    const int tx = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int i0 = 8 * w, j0 = 4 * tx;
    const size_t base = (size_t)blockIdx.x * ny * nx;
    float r[32], am[32], m[32];
    unsigned valid = 0u;
    // diagonal = dcol[kx] + drow[ky]
    float dcol[4], drow[8];
    // `closed` = 4-bit mask of solid sides (pdec_closed_sides): a cell on a solid side has one accessible neighbour fewer
#pragma unroll
    for (int kx = 0; kx < 4; ++kx) dcol[kx] = -4.0f + ((j0 + kx == 0 && (closed & 4)) ? 1.0f : 0.0f) + ((j0 + kx == nx - 1 && (closed & 8)) ? 1.0f : 0.0f);
#pragma unroll
    for (int ky = 0; ky < 8; ++ky) drow[ky] = ((i0 + ky == 0 && (closed & 1)) ? 1.0f : 0.0f) + ((i0 + ky == ny - 1 && (closed & 2)) ? 1.0f : 0.0f);
    const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    for (int q = threadIdx.x; q < (2 * WARPS + 2) * 32; q += 32 * WARPS) sm_e[q] = zero4;
    // staggered component bases of this batch element and the boundary-face mask (closed: solid boundary faces are 0)
    // (recomputed in the epilogue instead of being kept live across the iteration loop, whose register budget is full)
    auto fmask_y = [&](int idx) -> float { return ((idx == 0 && (closed & 1)) || (idx == ny && (closed & 2))) ? 0.0f : 1.0f; };
    auto fmask_x = [&](int idx) -> float { return ((idx == 0 && (closed & 4)) || (idx == nx && (closed & 8))) ? 0.0f : 1.0f; };
#pragma unroll
    for (int ky = 0; ky < 8; ++ky) {
        float4 g4 = zero4, r4 = zero4;
        if (PROJ != PROJ_NONE) {
            float rv[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            const int i = i0 + ky;
#pragma unroll
            for (int kx = 0; kx < 4; ++kx) {
                const int j = j0 + kx;
                if (i < ny && j < nx) {
                    if (!FULL) valid |= 1u << (ky * 4 + kx);
                    const float* y = pa.in_y + (size_t)blockIdx.x * (ny + 1) * nx;
                    const float* x = pa.in_x + (size_t)blockIdx.x * ny * (nx + 1);
                    float v;
                    if (PROJ == PROJ_FWD) {          // mask_divergence_fwd_kernel
                        const float dy = __fsub_rn(__fmul_rn(y[(size_t)(i + 1) * nx + j], fmask_y(i + 1)), __fmul_rn(y[(size_t)i * nx + j], fmask_y(i)));
                        const float dx = __fsub_rn(__fmul_rn(x[(size_t)i * (nx + 1) + j + 1], fmask_x(j + 1)), __fmul_rn(x[(size_t)i * (nx + 1) + j], fmask_x(j)));
                        v = __fadd_rn(dy, dx);
                        if (pa.store_rhs != nullptr) pa.store_rhs[base + (size_t)i * nx + j] = v;
                    } else {                         // subtract_gradient_bwd_kernel, the g_p output (interior faces only)
                        float acc = 0.0f;
                        if (i >= 1) acc = __fsub_rn(acc, __fdiv_rn(y[(size_t)i * nx + j], pa.dxy));
                        if (i + 1 < ny) acc = __fadd_rn(acc, __fdiv_rn(y[(size_t)(i + 1) * nx + j], pa.dxy));
                        if (j >= 1) acc = __fsub_rn(acc, __fdiv_rn(x[(size_t)i * (nx + 1) + j], pa.dxx));
                        if (j + 1 < nx) acc = __fadd_rn(acc, __fdiv_rn(x[(size_t)i * (nx + 1) + j + 1], pa.dxx));
                        v = __fmul_rn(acc, pa.dx0);
                        if (pa.add_rhs != nullptr) v = __fadd_rn(v, pa.add_rhs[base + (size_t)i * nx + j]);
                    }
                    rv[kx] = v;
                }
            }
            r4 = make_float4(rv[0], rv[1], rv[2], rv[3]);
        } else if (FULL) {
            r4 = *reinterpret_cast<const float4*>(rhs + base + (size_t)(i0 + ky) * nx + j0);
            if (guess != nullptr) g4 = *reinterpret_cast<const float4*>(guess + base + (size_t)(i0 + ky) * nx + j0);
        } else {
            float rv[4] = {0.0f, 0.0f, 0.0f, 0.0f}, gv[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
            for (int kx = 0; kx < 4; ++kx)
                if (i0 + ky < ny && j0 + kx < nx) {
                    valid |= 1u << (ky * 4 + kx);
                    rv[kx] = rhs[base + (size_t)(i0 + ky) * nx + j0 + kx];
                    if (guess != nullptr) gv[kx] = guess[base + (size_t)(i0 + ky) * nx + j0 + kx];
                }
            r4 = make_float4(rv[0], rv[1], rv[2], rv[3]);
            g4 = make_float4(gv[0], gv[1], gv[2], gv[3]);
        }
        r[ky * 4 + 0] = r4.x; r[ky * 4 + 1] = r4.y; r[ky * 4 + 2] = r4.z; r[ky * 4 + 3] = r4.w;
        m[ky * 4 + 0] = g4.x; m[ky * 4 + 1] = g4.y; m[ky * 4 + 2] = g4.z; m[ky * 4 + 3] = g4.w;
        sm_x[(i0 + ky) * 32 + tx] = g4;
    }
    __syncthreads();                         // zero-fill of the edge rows completes before anyone publishes
    auto publish = [&]() {                   // the block's first and last row, read by the warps above / below
        sm_e[(1 + 2 * w) * 32 + tx] = make_float4(m[0], m[1], m[2], m[3]);
        sm_e[(2 + 2 * w) * 32 + tx] = make_float4(m[28], m[29], m[30], m[31]);
    };
    auto matvec = [&]() {
        const float4 u4 = sm_e[(2 * w) * 32 + tx], d4 = sm_e[(2 * w + 3) * 32 + tx];
        const float upv[4] = {u4.x, u4.y, u4.z, u4.w}, dnv[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
        for (int ky = 0; ky < 8; ++ky) {
            float left = __shfl_up_sync(0xffffffffu, m[ky * 4 + 3], 1);
            float right = __shfl_down_sync(0xffffffffu, m[ky * 4 + 0], 1);
            if (tx == 0) left = 0.0f;
            if (tx == 31) right = 0.0f;
#pragma unroll
            for (int kx = 0; kx < 4; ++kx) {
                const int c = ky * 4 + kx;
                const float u = ky > 0 ? m[c >= 4 ? c - 4 : 0] : upv[kx];
                const float d = ky < 7 ? m[c + 4 <= 31 ? c + 4 : 31] : dnv[kx];
                const float l = kx > 0 ? m[c >= 1 ? c - 1 : 0] : left;
                const float rr = kx < 3 ? m[c + 1 <= 31 ? c + 1 : 31] : right;
                float dg;
                if (FULL && kx > 0 && kx < 3 && ky > 0 && ky < 7) dg = -4.0f;
                else if (FULL && kx > 0 && kx < 3) dg = -4.0f + drow[ky];
                else if (FULL && ky > 0 && ky < 7) dg = dcol[kx];
                else dg = dcol[kx] + drow[ky];
                const float val = (u + l) + (rr + d) + dg * m[c];
                am[c] = (FULL || (valid & (1u << c))) ? val : 0.0f;
            }
        }
    };
    if (guess != nullptr) {                  // momentum = k - A x0   (blas.py:24-26); m holds x0 here
        publish();
        __syncthreads();
        matvec();
#pragma unroll
        for (int c = 0; c < 32; ++c) r[c] -= am[c];
        __syncthreads();                     // x0 edge rows read before the first real publish
    }
    float rmax = 0.0f;
    int it = 0;
    if (FUSED) {
        // One reduction per iteration.  With q = A m:  a = (m.r)/(m.q) and, since r' = r - a q,
        //   beta = (r'.q)/(m.q) = ((r.q) - a (q.q))/(m.q)
        // so all four dot products are taken BEFORE the update and travel with max|r| (the stop test of blas.py:33-37, applied
        // to the current residual exactly as the two-reduction loop does).  The coefficients differ from the reference's by
        // rounding only; CG iterates are not the parity oracle, the solution is.
#pragma unroll
        for (int c = 0; c < 32; ++c) m[c] = r[c];
        publish();
        __syncthreads();
        matvec();
        while (true) {
            float p1[4] = {0.0f, 0.0f, 0.0f, 0.0f}, p2[4] = {0.0f, 0.0f, 0.0f, 0.0f}, p4[4] = {0.0f, 0.0f, 0.0f, 0.0f},
                  p5[4] = {0.0f, 0.0f, 0.0f, 0.0f}, pm[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
            for (int c = 0; c < 32; ++c) {
                p1[c & 3] += m[c] * am[c];
                p2[c & 3] += m[c] * r[c];
                p4[c & 3] += r[c] * am[c];
                p5[c & 3] += am[c] * am[c];
                pm[c & 3] = fmaxf(pm[c & 3], fabsf(r[c]));
            }
            float s1 = (p1[0] + p1[1]) + (p1[2] + p1[3]), s2 = (p2[0] + p2[1]) + (p2[2] + p2[3]);
            float s4 = (p4[0] + p4[1]) + (p4[2] + p4[3]), s5 = (p5[0] + p5[1]) + (p5[2] + p5[3]);
            float mx = fmaxf(fmaxf(pm[0], pm[1]), fmaxf(pm[2], pm[3]));
            block_reduce5<WARPS>(s1, s2, s4, s5, mx, red5[it & 1]);
            rmax = mx;
            if (!(it < max_it && (accuracy <= 0.0f || rmax >= accuracy))) break;
            const float a = div_no_nan_fast(s2, s1);
            const float beta = div_no_nan_fast(s4 - a * s5, s1);
#pragma unroll
            for (int ky = 0; ky < 8; ++ky) {
                float4 x4 = sm_x[(i0 + ky) * 32 + tx];
                x4.x += a * m[ky * 4 + 0]; x4.y += a * m[ky * 4 + 1]; x4.z += a * m[ky * 4 + 2]; x4.w += a * m[ky * 4 + 3];
                sm_x[(i0 + ky) * 32 + tx] = x4;
#pragma unroll
                for (int kx = 0; kx < 4; ++kx) {
                    const int c = ky * 4 + kx;
                    r[c] -= a * am[c];
                    m[c] = r[c] - beta * m[c];
                }
            }
            publish();                       // edge-row reads of the old m finished before the reduction's barrier
            __syncthreads();
            matvec();
            ++it;
        }
    } else {
        float dummy = 0.0f;
#pragma unroll
        for (int c = 0; c < 32; ++c) { m[c] = r[c]; rmax = fmaxf(rmax, fabsf(r[c])); }
        publish();
        block_reduce_pair<WARPS>(dummy, rmax, redB, true);      // its barrier also orders publish before the matvec
        matvec();
        while (it < max_it && (accuracy <= 0.0f || rmax >= accuracy)) {
            float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
            for (int c = 0; c < 32; ++c) { s1 += m[c] * am[c]; s2 += m[c] * r[c]; }
            block_reduce_pair<WARPS>(s1, s2, redA, false);
            const float tmp = s1, a = div_no_nan_fast(s2, tmp);
            float s3 = 0.0f, mx = 0.0f;
#pragma unroll
            for (int ky = 0; ky < 8; ++ky) {
                float4 x4 = sm_x[(i0 + ky) * 32 + tx];
                x4.x += a * m[ky * 4 + 0]; x4.y += a * m[ky * 4 + 1]; x4.z += a * m[ky * 4 + 2]; x4.w += a * m[ky * 4 + 3];
                sm_x[(i0 + ky) * 32 + tx] = x4;
#pragma unroll
                for (int kx = 0; kx < 4; ++kx) {
                    const int c = ky * 4 + kx;
                    r[c] -= a * am[c];
                    s3 += r[c] * am[c];
                    mx = fmaxf(mx, fabsf(r[c]));
                }
            }
            block_reduce_pair<WARPS>(s3, mx, redB, true);
            const float beta = div_no_nan_fast(s3, tmp);
#pragma unroll
            for (int c = 0; c < 32; ++c) m[c] = r[c] - beta * m[c];
            publish();                       // every edge-row read of the old m happened before reduction A's barrier
            __syncthreads();
            matvec();
            rmax = mx;
            ++it;
        }
    }
    if (PROJ == PROJ_NONE || out != nullptr) {
#pragma unroll
        for (int ky = 0; ky < 8; ++ky) {
            const float4 x4 = sm_x[(i0 + ky) * 32 + tx];
            if (FULL) {
                *reinterpret_cast<float4*>(out + base + (size_t)(i0 + ky) * nx + j0) = x4;
            } else {
                const float xv[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
                for (int kx = 0; kx < 4; ++kx)
                    if (valid & (1u << (ky * 4 + kx))) out[base + (size_t)(i0 + ky) * nx + j0 + kx] = xv[kx];
            }
        }
    }
    if (threadIdx.x == 0) {
        if (iters) iters[blockIdx.x] = it;
        if (resid) resid[blockIdx.x] = rmax;
    }
    if (PROJ != PROJ_NONE) {
        // epilogue: the corrected staggered field from the solution, which is complete in shared memory after this barrier.
        // A thread writes the faces on the low side of its cells; the owners of the last row / column also write face ny / nx.
        __syncthreads();
        const float* xs = reinterpret_cast<const float*>(sm_x);          // x(i, j) = xs[i * 128 + j]
        const size_t base_y = (size_t)blockIdx.x * (ny + 1) * nx, base_x = (size_t)blockIdx.x * ny * (nx + 1);
        const float* y = pa.in_y + base_y;
        const float* x = pa.in_x + base_x;
        auto sol = [&](int i, int j) -> float {
            float v = xs[i * 128 + j];
            if (PROJ == PROJ_BWD && pa.add_x != nullptr) v = __fadd_rn(v, pa.add_x[base + (size_t)i * nx + j]);
            return v;
        };
        auto face_y = [&](int i, int j) {
            const size_t o = (size_t)i * nx + j;
            const float mk = fmask_y(i);
            float res;
            if (PROJ == PROJ_FWD) {              // subtract_gradient_fwd_kernel: replicate pad -> boundary-face gradient 0
                float gr = 0.0f;
                if (i >= 1 && i < ny) gr = __fdiv_rn(__fsub_rn(__fmul_rn(sol(i, j), pa.dx0), __fmul_rn(sol(i - 1, j), pa.dx0)), pa.dxy);
                res = __fsub_rn(__fmul_rn(y[o], mk), __fmul_rn(gr, mk));
            } else {                             // subtract_gradient_bwd (g * mask) + mask_divergence_bwd ((lo - hi) * mask)
                const float lo = i >= 1 ? sol(i - 1, j) : 0.0f, hi = i < ny ? sol(i, j) : 0.0f;
                res = __fadd_rn(__fmul_rn(y[o], mk), __fmul_rn(__fsub_rn(lo, hi), mk));
            }
            pa.out_y[base_y + o] = res;
        };
        auto face_x = [&](int i, int j) {
            const size_t o = (size_t)i * (nx + 1) + j;
            const float mk = fmask_x(j);
            float res;
            if (PROJ == PROJ_FWD) {
                float gr = 0.0f;
                if (j >= 1 && j < nx) gr = __fdiv_rn(__fsub_rn(__fmul_rn(sol(i, j), pa.dx0), __fmul_rn(sol(i, j - 1), pa.dx0)), pa.dxx);
                res = __fsub_rn(__fmul_rn(x[o], mk), __fmul_rn(gr, mk));
            } else {
                const float lo = j >= 1 ? sol(i, j - 1) : 0.0f, hi = j < nx ? sol(i, j) : 0.0f;
                res = __fadd_rn(__fmul_rn(x[o], mk), __fmul_rn(__fsub_rn(lo, hi), mk));
            }
            pa.out_x[base_x + o] = res;
        };
#pragma unroll
        for (int ky = 0; ky < 8; ++ky) {
            const int i = i0 + ky;
            if (i >= ny) break;
#pragma unroll
            for (int kx = 0; kx < 4; ++kx) {
                const int j = j0 + kx;
                if (j >= nx) break;
                face_y(i, j);
                face_x(i, j);
                if (i == ny - 1) face_y(ny, j);
                if (j == nx - 1) face_x(i, nx);
            }
        }
    }
}

// -------------------------------------------------------------------------------------------------------------
// cg_block_masked_kernel: the on-chip block kernel for a domain with obstacles (any ny <= 8 * WARPS, nx <= 128, zero guess).
// The masked matrix (sparse.py:157-221) couples cell c to neighbour k with active_c * active_k, so with
//   a_c  = 1 if c has at least one coupled neighbour, else 0        (low nibble of the stencil code != 0)
//   dg_c = -(accessible neighbours of c), at most -1               (high nibble)
// rows with a_c = 0 (obstacle cells, fluid cells walled in on all sides, cells outside the grid) are 1x1 blocks: x_c = rhs_c / dg_c
// is their exact solution and they take no part in the iteration (r_c = m_c = 0 throughout).  For the others
//   (A m)_c = a_c * (m_up + m_left + m_right + m_down) + dg_c * m_c
// because every neighbour that is not coupled has m = 0.  a and dg live in shared memory (one LDS.128 per four cells each).
// The reference iterates on the decoupled rows too; they are zero in the fluid step (the masked divergence and the incoming pressure
// gradient vanish inside obstacles), where the iterates are therefore the same.
// -------------------------------------------------------------------------------------------------------------
template <int WARPS>
__global__ void __launch_bounds__(32 * WARPS, 1) cg_block_masked_kernel(const float* __restrict__ rhs, float* __restrict__ out, int* __restrict__ iters,
                                                                        float* __restrict__ resid, const unsigned char* __restrict__ code,
                                                                        int ny, int nx, float accuracy, int max_it) {
    constexpr int ROWS = 8 * WARPS;
    extern __shared__ __align__(16) float smem_dyn[];
    __shared__ __align__(16) float redA[32], redB[32];
    float4* sm_x = reinterpret_cast<float4*>(smem_dyn);          // [ROWS][32] float4 (row pitch 128 floats)
    float4* sm_e = sm_x + ROWS * 32;                            // [2*WARPS + 2][32]: 0 = zeros, 1+2w / 2+2w = first / last row of warp w
    float4* sm_a = sm_e + (2 * WARPS + 2) * 32;                 // [ROWS][32]: a
    float4* sm_d = sm_a + ROWS * 32;                            // [ROWS][32]: dg
    const int tx = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int i0 = 8 * w, j0 = 4 * tx;
    const size_t base = (size_t)blockIdx.x * ny * nx;
    float r[32], am[32], m[32];
    const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    for (int q = threadIdx.x; q < (2 * WARPS + 2) * 32; q += 32 * WARPS) sm_e[q] = zero4;
#pragma unroll
    for (int ky = 0; ky < 8; ++ky) {
        float av[4], dv[4], xv[4];
#pragma unroll
        for (int kx = 0; kx < 4; ++kx) {
            const int c = ky * 4 + kx;
            float rv = 0.0f;
            unsigned cd = 0u;
            if (i0 + ky < ny && j0 + kx < nx) {
                rv = rhs[base + (size_t)(i0 + ky) * nx + j0 + kx];
                cd = code[(size_t)(i0 + ky) * nx + j0 + kx];
            }
            const bool coupled = (cd & 15u) != 0u;
            av[kx] = coupled ? 1.0f : 0.0f;
            dv[kx] = coupled ? -(float)(cd >> 4) : 0.0f;
            xv[kx] = (!coupled && cd != 0u) ? rv / -(float)(cd >> 4) : 0.0f;       // decoupled row: solved here
            r[c] = coupled ? rv : 0.0f;
        }
        sm_a[(i0 + ky) * 32 + tx] = make_float4(av[0], av[1], av[2], av[3]);
        sm_d[(i0 + ky) * 32 + tx] = make_float4(dv[0], dv[1], dv[2], dv[3]);
        sm_x[(i0 + ky) * 32 + tx] = make_float4(xv[0], xv[1], xv[2], xv[3]);
    }
    __syncthreads();                         // zero-fill of the edge rows completes before anyone publishes
    auto publish = [&]() {
        sm_e[(1 + 2 * w) * 32 + tx] = make_float4(m[0], m[1], m[2], m[3]);
        sm_e[(2 + 2 * w) * 32 + tx] = make_float4(m[28], m[29], m[30], m[31]);
    };
    auto matvec = [&]() {
        const float4 u4 = sm_e[(2 * w) * 32 + tx], d4 = sm_e[(2 * w + 3) * 32 + tx];
        const float upv[4] = {u4.x, u4.y, u4.z, u4.w}, dnv[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
        for (int ky = 0; ky < 8; ++ky) {
            float left = __shfl_up_sync(0xffffffffu, m[ky * 4 + 3], 1);
            float right = __shfl_down_sync(0xffffffffu, m[ky * 4 + 0], 1);
            if (tx == 0) left = 0.0f;
            if (tx == 31) right = 0.0f;
            const float4 a4 = sm_a[(i0 + ky) * 32 + tx], g4 = sm_d[(i0 + ky) * 32 + tx];
            const float av[4] = {a4.x, a4.y, a4.z, a4.w}, gv[4] = {g4.x, g4.y, g4.z, g4.w};
#pragma unroll
            for (int kx = 0; kx < 4; ++kx) {
                const int c = ky * 4 + kx;
                const float u = ky > 0 ? m[c >= 4 ? c - 4 : 0] : upv[kx];
                const float d = ky < 7 ? m[c + 4 <= 31 ? c + 4 : 31] : dnv[kx];
                const float l = kx > 0 ? m[c >= 1 ? c - 1 : 0] : left;
                const float rr = kx < 3 ? m[c + 1 <= 31 ? c + 1 : 31] : right;
                am[c] = av[kx] * ((u + l) + (rr + d)) + gv[kx] * m[c];
            }
        }
    };
    float rmax = 0.0f, dummy = 0.0f;
    int it = 0;
#pragma unroll
    for (int c = 0; c < 32; ++c) { m[c] = r[c]; rmax = fmaxf(rmax, fabsf(r[c])); }
    publish();
    block_reduce_pair<WARPS>(dummy, rmax, redB, true);      // its barrier also orders publish before the matvec
    matvec();
    while (it < max_it && (accuracy <= 0.0f || rmax >= accuracy)) {
        float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
        for (int c = 0; c < 32; ++c) { s1 += m[c] * am[c]; s2 += m[c] * r[c]; }
        block_reduce_pair<WARPS>(s1, s2, redA, false);
        const float tmp = s1, a = div_no_nan_fast(s2, tmp);
        float s3 = 0.0f, mx = 0.0f;
#pragma unroll
        for (int ky = 0; ky < 8; ++ky) {
            float4 x4 = sm_x[(i0 + ky) * 32 + tx];
            x4.x += a * m[ky * 4 + 0]; x4.y += a * m[ky * 4 + 1]; x4.z += a * m[ky * 4 + 2]; x4.w += a * m[ky * 4 + 3];
            sm_x[(i0 + ky) * 32 + tx] = x4;
#pragma unroll
            for (int kx = 0; kx < 4; ++kx) {
                const int c = ky * 4 + kx;
                r[c] -= a * am[c];
                s3 += r[c] * am[c];
                mx = fmaxf(mx, fabsf(r[c]));
            }
        }
        block_reduce_pair<WARPS>(s3, mx, redB, true);
        const float beta = div_no_nan_fast(s3, tmp);
#pragma unroll
        for (int c = 0; c < 32; ++c) m[c] = r[c] - beta * m[c];
        publish();                           // every edge-row read of the old m happened before reduction A's barrier
        __syncthreads();
        matvec();
        rmax = mx;
        ++it;
    }
#pragma unroll
    for (int ky = 0; ky < 8; ++ky) {
        const float4 x4 = sm_x[(i0 + ky) * 32 + tx];
        const float xv[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
        for (int kx = 0; kx < 4; ++kx)
            if (i0 + ky < ny && j0 + kx < nx) out[base + (size_t)(i0 + ky) * nx + j0 + kx] = xv[kx];
    }
    if (threadIdx.x == 0) {
        if (iters) iters[blockIdx.x] = it;
        if (resid) resid[blockIdx.x] = rmax;
    }
}

template <int WARPS>
static int launch_block_masked(const float* rhs, float* p, int* iterations, float* residual, const unsigned char* code, int B, int ny, int nx,
                               float accuracy, int max_iterations, cudaStream_t st) {
    const size_t smem = (size_t)(3 * 8 * WARPS * 32 + (2 * WARPS + 2) * 32) * sizeof(float4);
    auto kernel = cg_block_masked_kernel<WARPS>;
    if (smem > 48 * 1024 && cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
        return PDEC_ERR_CUDA_LAUNCH;
    kernel<<<B, 32 * WARPS, smem, st>>>(rhs, p, iterations, residual, code, ny, nx, accuracy, max_iterations);
    return PDEC_OK;
}

// -------------------------------------------------------------------------------------------------------------
// Measured and dropped (round 1, profiles/r1n_cg_packed.md): a packed-fp32 twin of cg_block_kernel (vectors as float2 pairs of rows
// q and q + 4, add/fma.rn.f32x2, lane halo through shared memory or shuffles) has 17-20 % fewer instructions per iteration (388-399
// against 482) and is SLOWER: 1.99-2.62 us per iteration at 128x128 against 1.78.  The kernel is bound by the barrier / reduction
// latency chain at 4 warps per scheduler and by the fp32 pipe's lane throughput (which packing does not change), not by issue slots.
// -------------------------------------------------------------------------------------------------------------
// -------------------------------------------------------------------------------------------------------------
// Large grids: cooperative persistent kernel, vectors in global memory (L2 resident up to ~20 MB per vector).
// Unit of work = (batch element b, chunk c of consecutive cells); blocks take units grid-stride.
// Per iteration TWO passes separated by grid-wide barriers:
//   pass 1: m_new = r - (s3 m_old)/t_prev (folded m-update, recomputed for the 4 neighbours; m ping-pongs between two
//           buffers), Am = A m_new, partials of m.Am and m.r
//   pass 2: x += a m, r -= a Am, partials of r.Am and max|r|
// Reductions are two-stage and deterministic: one partial per unit (fixed intra-block tree), then every block re-reduces
// the partials of its element in a fixed order.  All partial slots ping-pong on the iteration parity, so no block can
// overwrite a value another block still reads.  An element that has converged stops updating; its max|r| is carried.
// -------------------------------------------------------------------------------------------------------------
struct CgGlobal {
    const float* rhs; const float* guess; float* x; float* r; float* m[2]; float* am; float* part;
    const unsigned char* code;      // optional [ny * nx] per-cell stencil code (obstacles): low nibble = coupled neighbours, high = -diagonal
    int* iters; float* resid; int* count;
    int B, ny, nx, closed, chunks, max_it, whole_batch;
    float accuracy;
};

enum { S_MAM = 0, S_MR = 1, S_RAM = 2, S_MAX = 3, S_SLOTS = 4 };

__device__ __forceinline__ float* part_ptr(const CgGlobal& P, int parity, int slot, int b) {
    return P.part + (((size_t)parity * S_SLOTS + slot) * P.B + b) * P.chunks;
}
__device__ __forceinline__ float block_sum(float v, float* sh) {
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = (threadIdx.x & 31) < (blockDim.x >> 5) ? sh[threadIdx.x & 31] : 0.0f;
    t = warp_sum(t);
    __syncthreads();
    return t;
}
__device__ __forceinline__ float block_max(float v, float* sh) {
    v = warp_max(v);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = (threadIdx.x & 31) < (blockDim.x >> 5) ? sh[threadIdx.x & 31] : 0.0f;
    t = warp_max(t);
    __syncthreads();
    return t;
}
// fixed-order sum / max over the `chunks` partials of one element (every block gets the same bits)
__device__ __forceinline__ float reduce_chunks_sum(const float* p, int chunks, float* sh) {
    float s = 0.0f;
    for (int c = threadIdx.x; c < chunks; c += blockDim.x) s += __ldcg(p + c);
    return block_sum(s, sh);
}
__device__ __forceinline__ float reduce_chunks_max(const float* p, int chunks, float* sh) {
    float s = 0.0f;
    for (int c = threadIdx.x; c < chunks; c += blockDim.x) s = fmaxf(s, __ldcg(p + c));
    return block_max(s, sh);
}

__global__ void __launch_bounds__(256) cg_global_kernel(CgGlobal P) {
    cg::grid_group grid = cg::this_grid();
    __shared__ float sh[32];
    const int N = P.ny * P.nx, nx = P.nx, ny = P.ny;
    const int units = P.B * P.chunks;
    const int per = (N + P.chunks - 1) / P.chunks;
    // ---- init: x = guess | 0, r = rhs - A x; max|r| -> parity 1 (read by iteration 0) ----
    for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const int b = u / P.chunks, c = u - b * P.chunks;
        const size_t base = (size_t)b * N;
        const int lo = c * per, hi = min(N, lo + per);
        float mx = 0.0f;
        for (int idx = lo + threadIdx.x; idx < hi; idx += blockDim.x) {
            float rr = P.rhs[base + idx], xx = 0.0f;
            if (P.guess) {
                xx = P.guess[base + idx];
                const int i = idx / nx, j = idx - i * nx;
                const unsigned cd = P.code ? P.code[idx] : 0u;
                const unsigned nib = P.code ? (cd & 15u) : cell_nibble(i, j, ny, nx);
                rr -= stencil_apply(P.guess + base, idx, nx, nib, P.code ? -(float)(cd >> 4) : cell_diag(nib, P.closed));
            }
            P.x[base + idx] = xx;
            P.r[base + idx] = rr;
            P.m[1][base + idx] = 0.0f;        // "old" direction of iteration 0 (beta = 0)
            mx = fmaxf(mx, fabsf(rr));
        }
        mx = block_max(mx, sh);
        if (threadIdx.x == 0) {
            part_ptr(P, 1, S_MAX, b)[c] = mx;
            part_ptr(P, 1, S_RAM, b)[c] = 0.0f;
            part_ptr(P, 1, S_MAM, b)[c] = 0.0f;
            if (c == 0) P.count[b] = 0;
        }
    }
    grid.sync();
    for (int it = 0; it < P.max_it; ++it) {
        const int cur = it & 1, prev = cur ^ 1;
        // uniform exit test: is any element still above the accuracy?  (max >= acc  <=>  some chunk partial >= acc)
        if (P.accuracy > 0.0f) {
            int above = 0;
            const float* mxp = part_ptr(P, prev, S_MAX, 0);
            for (int k = threadIdx.x; k < units; k += blockDim.x) above |= (__ldcg(mxp + k) >= P.accuracy) ? 1 : 0;
            if (!__syncthreads_or(above)) break;
        }
        const float* m_old = P.m[prev];
        float* m_new = P.m[cur];
        // ---- pass 1 ----
        for (int u = blockIdx.x; u < units; u += gridDim.x) {
            const int b = u / P.chunks, c = u - b * P.chunks;
            const float rmax = reduce_chunks_max(part_ptr(P, prev, S_MAX, b), P.chunks, sh);
            if (!P.whole_batch && P.accuracy > 0.0f && rmax < P.accuracy) continue;
            const float t_prev = reduce_chunks_sum(part_ptr(P, prev, S_MAM, b), P.chunks, sh);
            const float s3 = reduce_chunks_sum(part_ptr(P, prev, S_RAM, b), P.chunks, sh);
            const size_t base = (size_t)b * N;
            const float* rb = P.r + base;
            const float* mo = m_old + base;
            auto mval = [&](int q) -> float { return __ldcg(rb + q) - div_no_nan(s3 * __ldcg(mo + q), t_prev); };
            const int lo = c * per, hi = min(N, lo + per);
            float s1 = 0.0f, s2 = 0.0f;
            for (int idx = lo + threadIdx.x; idx < hi; idx += blockDim.x) {
                const int i = idx / nx, j = idx - i * nx;
                const unsigned cd = P.code ? P.code[idx] : 0u;
                const unsigned nib = P.code ? (cd & 15u) : cell_nibble(i, j, ny, nx);
                const float mm = mval(idx);
                float a_m = 0.0f;
                if (nib & 1u) a_m += mval(idx - nx);
                if (nib & 4u) a_m += mval(idx - 1);
                a_m += (P.code ? -(float)(cd >> 4) : cell_diag(nib, P.closed)) * mm;
                if (nib & 8u) a_m += mval(idx + 1);
                if (nib & 2u) a_m += mval(idx + nx);
                m_new[base + idx] = mm;
                P.am[base + idx] = a_m;
                s1 += mm * a_m;
                s2 += mm * __ldcg(rb + idx);
            }
            s1 = block_sum(s1, sh);
            s2 = block_sum(s2, sh);
            if (threadIdx.x == 0) {
                part_ptr(P, cur, S_MAM, b)[c] = s1;
                part_ptr(P, cur, S_MR, b)[c] = s2;
            }
        }
        grid.sync();
        // ---- pass 2 ----
        for (int u = blockIdx.x; u < units; u += gridDim.x) {
            const int b = u / P.chunks, c = u - b * P.chunks;
            const float rmax = reduce_chunks_max(part_ptr(P, prev, S_MAX, b), P.chunks, sh);
            if (!P.whole_batch && P.accuracy > 0.0f && rmax < P.accuracy) {       // converged earlier: carry its state forward
                if (threadIdx.x == 0) {
                    part_ptr(P, cur, S_MAX, b)[c] = part_ptr(P, prev, S_MAX, b)[c];
                    part_ptr(P, cur, S_MAM, b)[c] = 0.0f;
                    part_ptr(P, cur, S_RAM, b)[c] = 0.0f;
                }
                continue;
            }
            const float tmp = reduce_chunks_sum(part_ptr(P, cur, S_MAM, b), P.chunks, sh);
            const float mr = reduce_chunks_sum(part_ptr(P, cur, S_MR, b), P.chunks, sh);
            const float a = div_no_nan(mr, tmp);
            const size_t base = (size_t)b * N;
            const int lo = c * per, hi = min(N, lo + per);
            float s3 = 0.0f, mx = 0.0f;
            for (int idx = lo + threadIdx.x; idx < hi; idx += blockDim.x) {
                const float mm = m_new[base + idx], a_m = P.am[base + idx];
                P.x[base + idx] += a * mm;
                const float rr = P.r[base + idx] - a * a_m;
                P.r[base + idx] = rr;
                s3 += rr * a_m;
                mx = fmaxf(mx, fabsf(rr));
            }
            s3 = block_sum(s3, sh);
            mx = block_max(mx, sh);
            if (threadIdx.x == 0) {
                part_ptr(P, cur, S_RAM, b)[c] = s3;
                part_ptr(P, cur, S_MAX, b)[c] = mx;
                if (c == 0) P.count[b] = it + 1;
            }
        }
        grid.sync();
    }
    // the last written parity holds the final max|r|: it is the one NOT equal to the parity of the next iteration.
    // Both parities agree for converged elements; for elements stopped by max_it use the larger iteration's slot.
    for (int b = blockIdx.x; b < P.B; b += gridDim.x) {
        const int n_it = P.count[b];
        const int par = (n_it == 0) ? 1 : ((n_it - 1) & 1);
        const float rmax = reduce_chunks_max(part_ptr(P, par, S_MAX, b), P.chunks, sh);
        if (threadIdx.x == 0) {
            if (P.iters) P.iters[b] = n_it;
            if (P.resid) P.resid[b] = rmax;
        }
    }
}

}  // namespace pdec

using namespace pdec;

#define PDEC_CG_SMEM_MAX_CELLS 16384

static int global_chunks(int N) {
    int c = (N + 8191) / 8192;
    return c < 1 ? 1 : (c > 1024 ? 1024 : c);
}

extern "C" size_t pdec_cg_poisson_2d_workspace_bytes(int B, int ny, int nx) {
    const size_t N = (size_t)ny * nx, b = B > 0 ? B : 1;
    // r, m[2], Am + 2 parities x 4 partial slots + iteration counters; or the tiled kernel's scratch
    const size_t streaming = 4 * b * N * sizeof(float) + 8 * b * global_chunks((int)N) * sizeof(float) + b * sizeof(int) + 1024;
    const size_t tiled = pdec_cg_slab_workspace_bytes(B, ny, nx);
    return streaming > tiled ? streaming : tiled;
}

template <int WARPS>
static int launch_block(const float* rhs, const float* guess, float* p, int* iterations, float* residual, int B, int ny, int nx,
                        int closed, float accuracy, int max_iterations, cudaStream_t st, bool fused) {
    const size_t smem = (size_t)(8 * WARPS * 32 + (2 * WARPS + 2) * 32) * sizeof(float4);
    const bool full = ny == 8 * WARPS && nx == 128 && (reinterpret_cast<uintptr_t>(rhs) & 15) == 0 && (reinterpret_cast<uintptr_t>(p) & 15) == 0 &&
                      (guess == nullptr || (reinterpret_cast<uintptr_t>(guess) & 15) == 0);
    auto kernel = fused ? (full ? cg_block_kernel<WARPS, true, true> : cg_block_kernel<WARPS, false, true>)
                        : (full ? cg_block_kernel<WARPS, true, false> : cg_block_kernel<WARPS, false, false>);
    if (smem > 48 * 1024 && cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
        return PDEC_ERR_CUDA_LAUNCH;
    kernel<<<B, 32 * WARPS, smem, st>>>(rhs, guess, p, iterations, residual, ny, nx, closed, accuracy, max_iterations, ProjArgs());
    return PDEC_OK;
}

// the fused projection (PROJ_FWD / PROJ_BWD) on the two-reduction loop, one CTA per batch element
template <int WARPS, int PROJ>
static int launch_block_proj(float* p, int* iterations, float* residual, int B, int ny, int nx, int closed, float accuracy, int max_iterations,
                             const ProjArgs& pa, cudaStream_t st) {
    const size_t smem = (size_t)(8 * WARPS * 32 + (2 * WARPS + 2) * 32) * sizeof(float4);
    const bool full = ny == 8 * WARPS && nx == 128 && (p == nullptr || (reinterpret_cast<uintptr_t>(p) & 15) == 0);
    auto kernel = full ? cg_block_kernel<WARPS, true, false, PROJ> : cg_block_kernel<WARPS, false, false, PROJ>;
    if (smem > 48 * 1024 && cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
        return PDEC_ERR_CUDA_LAUNCH;
    kernel<<<B, 32 * WARPS, smem, st>>>(nullptr, nullptr, p, iterations, residual, ny, nx, closed, accuracy, max_iterations, pa);
    return PDEC_OK;
}

template <int PROJ>
static int project_2d(float* p, int* iterations, float* residual, int B, int ny, int nx, int closed, float size_y, float size_x, float accuracy,
                      int max_iterations, ProjArgs pa, void* stream) {
    if (B < 0 || ny < 2 || nx < 2 || max_iterations < 0 || !(size_y > 0.0f) || !(size_x > 0.0f)) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!pa.in_y || !pa.in_x || !pa.out_y || !pa.out_x) return PDEC_ERR_INVALID_ARGUMENT;
    if (ny > 128 || nx > 128) return PDEC_ERR_UNSUPPORTED_SIZE;          // larger grids: compose the stand-alone entry points
    pa.dxy = (float)((double)size_y / ny);
    pa.dxx = (float)((double)size_x / nx);
    pa.dx0 = pa.dxy;
    cudaStream_t st = (cudaStream_t)stream;
    closed = (int)pdec_closed_sides(closed);
    int rc;
    if (ny > 64) rc = launch_block_proj<16, PROJ>(p, iterations, residual, B, ny, nx, closed, accuracy, max_iterations, pa, st);
    else if (ny > 32) rc = launch_block_proj<8, PROJ>(p, iterations, residual, B, ny, nx, closed, accuracy, max_iterations, pa, st);
    else rc = launch_block_proj<4, PROJ>(p, iterations, residual, B, ny, nx, closed, accuracy, max_iterations, pa, st);
    if (rc != PDEC_OK) return rc;
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" int pdec_project_2d_supported(int ny, int nx) { return ny >= 2 && nx >= 2 && ny <= 128 && nx <= 128; }

extern "C" int pdec_project_2d_fwd(const float* vy, const float* vx, float* vy_out, float* vx_out, float* p, float* div, int* iterations,
                                   float* residual, int B, int ny, int nx, int closed, float size_y, float size_x, float accuracy,
                                   int max_iterations, void* stream) {
    ProjArgs pa = {};
    pa.in_y = vy; pa.in_x = vx; pa.out_y = vy_out; pa.out_x = vx_out; pa.store_rhs = div;
    return project_2d<PROJ_FWD>(p, iterations, residual, B, ny, nx, closed, size_y, size_x, accuracy, max_iterations, pa, stream);
}

extern "C" int pdec_project_2d_bwd(const float* g_vy_out, const float* g_vx_out, const float* g_p, const float* g_div, float* g_vy, float* g_vx,
                                   int* iterations, float* residual, int B, int ny, int nx, int closed, float size_y, float size_x,
                                   float accuracy, int max_iterations, void* stream) {
    ProjArgs pa = {};
    pa.in_y = g_vy_out; pa.in_x = g_vx_out; pa.out_y = g_vy; pa.out_x = g_vx; pa.add_rhs = g_p; pa.add_x = g_div;
    return project_2d<PROJ_BWD>(nullptr, iterations, residual, B, ny, nx, closed, size_y, size_x, accuracy, max_iterations, pa, stream);
}

template <int THREADS, int CPT = 16>
static int launch_smem(const float* rhs, const float* guess, float* p, int* iterations, float* residual, int B, int ny, int nx,
                       int closed, float accuracy, int max_iterations, cudaStream_t st) {
    const size_t smem = (size_t)ny * nx * sizeof(float);
    auto kernel = cg_smem_kernel<THREADS, CPT>;
    if (smem > 48 * 1024 && cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
        return PDEC_ERR_CUDA_LAUNCH;
    kernel<<<B, THREADS, smem, st>>>(rhs, guess, p, iterations, residual, ny, nx, closed, accuracy, max_iterations);
    return PDEC_OK;
}

static int launch_global(const float* rhs, const float* guess, float* p, int* iterations, float* residual, const unsigned char* code, int B, int ny,
                         int nx, int closed, float accuracy, int max_iterations, bool whole_batch, void* workspace, cudaStream_t st);

extern "C" int pdec_cg_poisson_2d(const float* rhs, const float* guess, float* p, int* iterations, float* residual, int B, int ny, int nx,
                                  int closed, float accuracy, int max_iterations, int stop_mode, void* workspace, size_t workspace_bytes,
                                  void* stream) {
    if (B < 0 || ny < 2 || nx < 2 || max_iterations < 0) return PDEC_ERR_INVALID_ARGUMENT;
    const int variant = (stop_mode >> 8) & 0xff;      // experimental kernel-shape selector (benchmarks only)
    stop_mode &= 0xff;
    if (stop_mode != PDEC_CG_STOP_PER_ELEMENT && stop_mode != PDEC_CG_STOP_WHOLE_BATCH) return PDEC_ERR_INVALID_ARGUMENT;
    // the reference's rule (all elements iterate until the batch-wide max|r| < accuracy, blas.py:33-37) needs a batch-wide
    // exchange every iteration: it runs on the cooperative kernel, whatever the grid size
    const bool whole_batch = stop_mode == PDEC_CG_STOP_WHOLE_BATCH && accuracy > 0.0f && B > 1;
    if (B == 0) return PDEC_OK;
    if (!rhs || !p) return PDEC_ERR_INVALID_ARGUMENT;
    const bool on_chip = !whole_batch && (long long)ny * nx <= PDEC_CG_SMEM_MAX_CELLS;
    if (!on_chip && workspace_bytes < pdec_cg_poisson_2d_workspace_bytes(B, ny, nx)) return PDEC_ERR_WORKSPACE_TOO_SMALL;
    cudaStream_t st = (cudaStream_t)stream;
    const int closed_arg = closed;                    // forwarded unchanged to the tiled entry, which normalises it itself
    closed = (int)pdec_closed_sides(closed);
    const long long N = (long long)ny * nx;
    int rc = PDEC_OK;
    if (on_chip && ny <= 128 && nx <= 128 && (variant == 0 || variant == 4 || variant == 5 || variant >= 6)) {
        // the one-reduction loop (variant 5) saves a barrier but costs a fifth dot product: measured slower (profiles/r1e_*)
        const bool fused = variant == 5;
        if (ny > 64) rc = launch_block<16>(rhs, guess, p, iterations, residual, B, ny, nx, closed, accuracy, max_iterations, st, fused);
        else if (ny > 32) rc = launch_block<8>(rhs, guess, p, iterations, residual, B, ny, nx, closed, accuracy, max_iterations, st, fused);
        else rc = launch_block<4>(rhs, guess, p, iterations, residual, B, ny, nx, closed, accuracy, max_iterations, st, fused);
        if (rc != PDEC_OK) return rc;
        PDEC_CHECK_LAUNCH();
        return PDEC_OK;
    }
    if (on_chip) {
        if (N > 8192 && variant == 1) rc = launch_smem<512, 32>(rhs, guess, p, iterations, residual, B, ny, nx, closed, accuracy, max_iterations, st);
        else if (N > 8192) rc = launch_smem<1024>(rhs, guess, p, iterations, residual, B, ny, nx, closed, accuracy, max_iterations, st);
        else if (N > 4096) rc = launch_smem<512>(rhs, guess, p, iterations, residual, B, ny, nx, closed, accuracy, max_iterations, st);
        else if (N > 2048) rc = launch_smem<256>(rhs, guess, p, iterations, residual, B, ny, nx, closed, accuracy, max_iterations, st);
        else if (N > 1024) rc = launch_smem<128>(rhs, guess, p, iterations, residual, B, ny, nx, closed, accuracy, max_iterations, st);
        else if (N > 512) rc = launch_smem<64>(rhs, guess, p, iterations, residual, B, ny, nx, closed, accuracy, max_iterations, st);
        else rc = launch_smem<32>(rhs, guess, p, iterations, residual, B, ny, nx, closed, accuracy, max_iterations, st);
        if (rc != PDEC_OK) return rc;
        PDEC_CHECK_LAUNCH();
        return PDEC_OK;
    }
    if (N > 0x7fffffffLL || !workspace) return PDEC_ERR_UNSUPPORTED_SIZE;
    // medium grids: a grid of on-chip tiles with one grid-wide barrier per iteration (cg_tiled.cu); it declines what it cannot
    // keep resident (or an initial guess / the whole-batch rule), which then streams through L2 with the cooperative kernel below
    if (!whole_batch && guess == nullptr && variant == 0) {
        rc = pdec_cg_poisson_2d_slab(rhs, p, iterations, residual, B, ny, nx, ny, 0, closed_arg, accuracy, max_iterations, nullptr, workspace,
                                     workspace_bytes, stream);
        if (rc != PDEC_ERR_UNSUPPORTED_SIZE) return rc;
    }
    return launch_global(rhs, guess, p, iterations, residual, nullptr, B, ny, nx, closed, accuracy, max_iterations, whole_batch, workspace, st);
}

static int launch_global(const float* rhs, const float* guess, float* p, int* iterations, float* residual, const unsigned char* code, int B, int ny,
                         int nx, int closed, float accuracy, int max_iterations, bool whole_batch, void* workspace, cudaStream_t st) {
    const long long N = (long long)ny * nx;
    CgGlobal P;
    const size_t n = (size_t)B * N;
    float* w = reinterpret_cast<float*>(workspace);
    P.code = code;
    P.rhs = rhs; P.guess = guess; P.x = p; P.r = w; P.m[0] = w + n; P.m[1] = w + 2 * n; P.am = w + 3 * n; P.part = w + 4 * n;
    P.chunks = global_chunks((int)N);
    P.count = reinterpret_cast<int*>(P.part + (size_t)8 * B * P.chunks);
    P.iters = iterations; P.resid = residual;
    P.B = B; P.ny = ny; P.nx = nx; P.closed = closed; P.max_it = max_iterations; P.accuracy = accuracy;
    P.whole_batch = whole_batch ? 1 : 0;
    int dev = 0, sms = 148, per_sm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, cg_global_kernel, 256, 0) != cudaSuccess || per_sm < 1) return PDEC_ERR_CUDA_LAUNCH;
    long long grid = (long long)sms * per_sm;
    const long long units = (long long)B * P.chunks;
    if (grid > units) grid = units;
    void* args[] = {&P};
    if (cudaLaunchCooperativeKernel((void*)cg_global_kernel, dim3((unsigned)grid), dim3(256), args, 0, st) != cudaSuccess) {
        cudaGetLastError();
        return PDEC_ERR_CUDA_LAUNCH;
    }
    return PDEC_OK;
}

// Masked operator (obstacles): the stencil of every cell comes from `code` [ny * nx] (shared by the batch): bits 0..3 = coupled
// neighbour above / below / left / right (active_self * active_neighbour, sparse.py:186-205), bits 4..7 = minus the diagonal
// (number of accessible neighbours, at least 1, sparse.py:207-213).  Grids up to 128 x 128 with a zero guess and the per-element
// stop rule run on chip (cg_block_masked_kernel), everything else on the streaming kernel (stop_mode bit 8 forces it: benchmarks).
extern "C" int pdec_cg_poisson_2d_masked(const float* rhs, const float* guess, float* p, int* iterations, float* residual,
                                         const unsigned char* code, int B, int ny, int nx, float accuracy, int max_iterations, int stop_mode,
                                         void* workspace, size_t workspace_bytes, void* stream) {
    if (!rhs || !p || !code || !workspace || B < 1 || ny < 1 || nx < 1 || max_iterations < 0) return PDEC_ERR_INVALID_ARGUMENT;
    if ((long long)ny * nx > 0x7fffffffLL) return PDEC_ERR_UNSUPPORTED_SIZE;
    const bool force_streaming = (stop_mode >> 8) & 1;
    stop_mode &= 0xff;
    const bool whole_batch = stop_mode == PDEC_CG_STOP_WHOLE_BATCH && accuracy > 0.0f && B > 1;
    if (!force_streaming && !whole_batch && guess == nullptr && ny <= 128 && nx <= 128) {
        cudaStream_t st = (cudaStream_t)stream;
        int rc;
        if (ny > 64) rc = launch_block_masked<16>(rhs, p, iterations, residual, code, B, ny, nx, accuracy, max_iterations, st);
        else if (ny > 32) rc = launch_block_masked<8>(rhs, p, iterations, residual, code, B, ny, nx, accuracy, max_iterations, st);
        else rc = launch_block_masked<4>(rhs, p, iterations, residual, code, B, ny, nx, accuracy, max_iterations, st);
        if (rc != PDEC_OK) return rc;
        PDEC_CHECK_LAUNCH();
        return PDEC_OK;
    }
    if (workspace_bytes < pdec_cg_poisson_2d_workspace_bytes(B, ny, nx)) return PDEC_ERR_WORKSPACE_TOO_SMALL;
    return launch_global(rhs, guess, p, iterations, residual, code, B, ny, nx, 0, accuracy, max_iterations,
                         stop_mode == PDEC_CG_STOP_WHOLE_BATCH && B > 1, workspace, (cudaStream_t)stream);
}
