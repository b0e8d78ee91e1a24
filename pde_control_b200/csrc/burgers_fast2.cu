// burgers_fast2.cu -- second-generation fused force-loss kernel for the headline case of BASELINE config 2: rows of 128 cells on
// the unit box, 'constant' extrapolation, one diffusion substep (SURVEY rows B1-B6 in one launch).  Same work split as
// burgers_fast.cu (one warp per (row, chunk of frame pairs), lane l owns cells 4l..4l+3, one 16-byte load / store per frame) but
// built around what the B200 issue model charges for (scripts/diag/f32x2_throughput.cu: the integer/ALU pipe runs at half rate,
// a packed fp32 instruction does two lane-operations in one issue slot):
//   * the fp32 chains run on PAIRS of cells with add/mul/fma.rn.f32x2 -- explicit round-to-nearest, so the forward step still
//     rounds exactly like the reference's unfused TF ops (same operation order as burgers_fast.cu / burgers.cu);
//   * shared memory holds the gather source as float2 (u[j], u[j+1]) per cell, so ONE 8-byte load brings both taps of the linear
//     interpolation and one address computation serves both; the transpose accumulator is float2 as well ((floor, ceil) parts of
//     the same back-traced sample), so the scatter is one read-modify-write per sample instead of two;
//   * out-of-domain samples ('constant' rule: exactly +0 and zero gradient, tf_backend.py:466-497) read a dedicated zero slot
//     and scatter into a dummy slot instead of being predicated; their index needs no clamping (one unsigned compare);
//   * index arithmetic is folded: xi = (i + 0.5) - u * (128 dt) - 0.5 is bit-identical to the reference's
//     ((x0 - u dt) - lower) / size * N - 0.5 on the unit box (x0 = (2i+1)/256 exactly; scaling by 128 commutes with rounding);
//     a * (raw / dx^2) = (a / dx^2) * raw because 1/dx^2 = 2^14.
// Slots are skewed by one float2 per 16 (pskew) so that the 16 lanes of a shared-memory wavefront, which touch every 4th cell,
// fall into 16 different bank pairs.
// The adjoint scatter is deterministic without atomics as in burgers_fast.cu: if floor(xi) is strictly increasing over the lanes
// of one warp instruction the targets are distinct (plain read-modify-write); otherwise lanes take turns in ascending order.
#include "pdec_common.cuh"

namespace pdec {

struct Unit2P {
    float dt128;     // float32(dt) * 128 = -dxi/du
    float c;         // float32(dt * viscosity) / dx^2: increment = c * (up + lo - 2 u)
    float gl2, gl1;  // dL/d(l2), dL/d(l1), already divided by B
};

#define F2_ROW 144                    // 128 + 7 skew slots, rounded up; slot 143 = zero (source row) / dummy (accumulator row)
#define F2_DUMMY 143
constexpr int kUnroll =
#ifdef F2_UNROLL
    F2_UNROLL;
#else
    2;
#endif
#ifndef F2_MINB
#define F2_MINB 4
#endif
#ifndef F2_UNROLL
#define F2_UNROLL 2
#endif

__device__ __forceinline__ float2 f2(float x, float y) { return make_float2(x, y); }
__device__ __forceinline__ float2 sub2(float2 a, float2 b) { return __ffma2_rn(b, f2(-1.0f, -1.0f), a); }   // rn(a - b), one rounding

struct Step2 {                        // what the forward step leaves for its adjoint
    int bi[4];                        // floor(xi)
    int slot[4];                      // skewed slot of the sample (dummy when outside)
    float2 w0[2], w1[2];              // weight of the ceil / floor tap, cells (0,1) and (2,3)
    float du[4];                      // u[ceil] - u[floor]
};

// forward step of the row (u01, u23) -> (v01, v23); su2 = this warp's source row, mine = su2 + pskew(4 lane)
__device__ __forceinline__ void step2_forward(const Unit2P& p, float2 X01, float2 X23, float2 u01, float2 u23, float2* su2, float2* mine,
                                              int lane, float2& v01, float2& v23, Step2& S) {
    const float unext = __shfl_down_sync(PDEC_FULL_MASK, u01.x, 1);    // u[4l+4]; lane 31: unused tap of cell 127
    mine[0] = u01;
    mine[1] = f2(u01.y, u23.x);
    mine[2] = u23;
    mine[3] = f2(u23.y, unext);
    __syncwarp();
    const float2 dt2 = f2(p.dt128, p.dt128), half2v = f2(-0.5f, -0.5f), one2 = f2(1.0f, 1.0f);
    // advect.py:35, _box.py:50-52, grid.py:104-105 folded (see the header): xi = ((i + 0.5) - u * 128dt) - 0.5
    const float2 xi01 = __fadd2_rn(sub2(X01, __fmul2_rn(u01, dt2)), half2v);
    const float2 xi23 = __fadd2_rn(sub2(X23, __fmul2_rn(u23, dt2)), half2v);
    const float xi[4] = {xi01.x, xi01.y, xi23.x, xi23.y};
    float bf[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        bf[k] = floorf(xi[k]);
        S.bi[k] = __float2int_rd(xi[k]);                               // saturating: far-away back-traces stay outside
    }
    // TF 'constant' rule (tf_backend.py:466-497): inside 0 <= floor <= N-2 the clamps do nothing: w0 = xi - floor,
    // w1 = (floor + 1) - xi; outside both clamped taps coincide and the sample is exactly +0
    const float2 bf01 = f2(bf[0], bf[1]), bf23 = f2(bf[2], bf[3]);
    S.w0[0] = sub2(xi01, bf01);
    S.w0[1] = sub2(xi23, bf23);
    S.w1[0] = sub2(__fadd2_rn(bf01, one2), xi01);
    S.w1[1] = sub2(__fadd2_rn(bf23, one2), xi23);
    const float w0[4] = {S.w0[0].x, S.w0[0].y, S.w0[1].x, S.w0[1].y};
    const float w1[4] = {S.w1[0].x, S.w1[0].y, S.w1[1].x, S.w1[1].y};
    float adv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int b = S.bi[k];
        S.slot[k] = (unsigned)b < 127u ? b + (b >> 4) : F2_DUMMY;
        const float2 tap = su2[S.slot[k]];                             // (u[floor], u[floor + 1]) or (0, 0)
        adv[k] = __fadd_rn(__fmul_rn(tap.x, w1[k]), __fmul_rn(tap.y, w0[k]));   // tf_backend.py:491-497
        S.du[k] = tap.y - tap.x;
    }
    float left = __shfl_up_sync(PDEC_FULL_MASK, adv[3], 1), right = __shfl_down_sync(PDEC_FULL_MASK, adv[0], 1);
    if (lane == 0) left = 0.0f;                                        // zero padding of the 'constant' field (nd.py:271-283)
    if (lane == 31) right = 0.0f;
    const float2 a01 = f2(adv[0], adv[1]), a23 = f2(adv[2], adv[3]), mid = f2(adv[1], adv[2]);
    const float2 m2 = f2(-2.0f, -2.0f), c2v = f2(p.c, p.c);
    // nd.py:281-282: (up + lo) - 2 u (2 u is exact, so one fused rounding equals the reference's two); grid.py:190 and
    // util.py:41-43 folded into one multiply (power-of-two dx^2); then u + increment
    const float2 raw01 = __ffma2_rn(a01, m2, __fadd2_rn(mid, f2(left, adv[0])));
    const float2 raw23 = __ffma2_rn(a23, m2, __fadd2_rn(f2(adv[3], right), mid));
    v01 = __fadd2_rn(a01, __fmul2_rn(raw01, c2v));
    v23 = __fadd2_rn(a23, __fmul2_rn(raw23, c2v));
}

// adjoint of one step.  h = -(dL/d step output) on entry (the caller keeps the sign of the residual gradient); returns dL/d(step
// input) in (gin01, gin23).  The accumulator row is all-zero on entry (own slots) and again on exit.
__device__ __forceinline__ void step2_adjoint(const Unit2P& p, const Step2& S, float2 h01, float2 h23, float2* sacc2, float2* mine_acc,
                                              int lane, float2& gin01, float2& gin23) {
    float hl = __shfl_up_sync(PDEC_FULL_MASK, h23.y, 1), hr = __shfl_down_sync(PDEC_FULL_MASK, h01.x, 1);
    if (lane == 0) hl = 0.0f;
    if (lane == 31) hr = 0.0f;
    // transpose of the (symmetric, zero-padded) diffusion stencil
    const float2 mid = f2(h01.y, h23.x), m2 = f2(-2.0f, -2.0f), c2v = f2(p.c, p.c);
    const float2 q01 = __ffma2_rn(c2v, __ffma2_rn(h01, m2, __fadd2_rn(mid, f2(hl, h01.x))), h01);
    const float2 q23 = __ffma2_rn(c2v, __ffma2_rn(h23, m2, __fadd2_rn(f2(h23.y, hr), mid)), h23);
    const float q[4] = {q01.x, q01.y, q23.x, q23.y};
    const float w0[4] = {S.w0[0].x, S.w0[0].y, S.w0[1].x, S.w0[1].y};
    const float w1[4] = {S.w1[0].x, S.w1[0].y, S.w1[1].x, S.w1[1].y};
    bool ordered = true;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int prev = __shfl_up_sync(PDEC_FULL_MASK, S.bi[k], 1);
        ordered = ordered && (lane == 0 || S.bi[k] > prev);
    }
    if (__all_sync(PDEC_FULL_MASK, ordered)) {
        // strictly increasing floor(xi) over the lanes: the in-domain targets of one warp instruction are distinct
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float2 a = sacc2[S.slot[k]];
            a.x = fmaf(-q[k], w1[k], a.x);
            a.y = fmaf(-q[k], w0[k], a.y);
            sacc2[S.slot[k]] = a;
            __syncwarp();
        }
    } else {
        // folded back-trace (a shock steeper than one cell per cell): lanes take turns, ascending -- same sums, fixed order
#pragma unroll
        for (int k = 0; k < 4; ++k) {
#pragma unroll 1
            for (int turn = 0; turn < 32; ++turn) {
                if (lane == turn && S.slot[k] != F2_DUMMY) {
                    float2 a = sacc2[S.slot[k]];
                    a.x = fmaf(-q[k], w1[k], a.x);
                    a.y = fmaf(-q[k], w0[k], a.y);
                    sacc2[S.slot[k]] = a;
                }
                __syncwarp();
            }
        }
    }
    // cell j receives the floor part of slot j and the ceil part of slot j - 1, plus the derivative through xi(u)
    const float2 o0 = mine_acc[0], o1 = mine_acc[1], o2 = mine_acc[2], o3 = mine_acc[3];
    float cprev = __shfl_up_sync(PDEC_FULL_MASK, o3.y, 1);
    if (lane == 0) cprev = 0.0f;
    const float2 z = f2(0.0f, 0.0f);
    mine_acc[0] = z; mine_acc[1] = z; mine_acc[2] = z; mine_acc[3] = z;
    const float2 dt2 = f2(p.dt128, p.dt128);
    const float2 e01 = __fmul2_rn(f2(S.du[0], S.du[1]), dt2), e23 = __fmul2_rn(f2(S.du[2], S.du[3]), dt2);   // -du * dxi/du
    gin01 = __ffma2_rn(q01, e01, __fadd2_rn(f2(o0.x, o1.x), f2(cprev, o0.y)));
    gin23 = __ffma2_rn(q23, e23, __fadd2_rn(f2(o2.x, o3.x), f2(o1.y, o2.y)));
    __syncwarp();
}

__device__ __forceinline__ float sgn(float d) { return (d > 0.0f) ? 1.0f : ((d < 0.0f) ? -1.0f : 0.0f); }

template <bool L1TERM>
__global__ void __launch_bounds__(256, F2_MINB) burgers_pair_loss_unit_kernel(Unit2P p, const float* __restrict__ s, float* __restrict__ grad_s,
                                                                      float2* partials, int T, int B, int C, int nchunks,
                                                                      unsigned* done_counter, float* __restrict__ losses, float inv_B) {
    __shared__ float2 smem[8][2][F2_ROW];
    __shared__ float red2[8], red1[8];
    __shared__ unsigned s_last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float2* su2 = smem[warp][0];
    float2* sacc2 = smem[warp][1];
    for (int q = lane; q < F2_ROW; q += 32) { su2[q] = f2(0.0f, 0.0f); sacc2[q] = f2(0.0f, 0.0f); }
    __syncwarp();
    const int own = 4 * lane + (lane >> 2);                           // pskew(4 lane); cells 4l..4l+3 never straddle a skew step
    float2* mine = su2 + own;
    float2* mine_acc = sacc2 + own;
    const float2 X01 = f2((float)(4 * lane) + 0.5f, (float)(4 * lane) + 1.5f), X23 = f2((float)(4 * lane) + 2.5f, (float)(4 * lane) + 3.5f);
    const size_t f4 = (size_t)B * 32;                                 // float4 stride between frames
    const long long units = (long long)B * nchunks;
    const float2 gl2v = f2(p.gl2, p.gl2);
    for (long long unit = (long long)blockIdx.x * 8 + warp; unit < units; unit += (long long)gridDim.x * 8) {
        const int chunk = (int)(unit / B), b = (int)(unit - (long long)chunk * B);      // consecutive warps -> consecutive rows
        const int t0 = chunk * C, t1 = min(T, t0 + C);
        const float4* row = reinterpret_cast<const float4*>(s + (size_t)b * 128) + lane;
        float4* grow = grad_s ? reinterpret_cast<float4*>(grad_s + (size_t)b * 128) + lane : nullptr;
        float2 u01, u23, carry01 = f2(0.0f, 0.0f), carry23 = f2(0.0f, 0.0f), v01, v23;
        Step2 S;
        {
            const float4 q = row[(size_t)t0 * f4];
            u01 = f2(q.x, q.y); u23 = f2(q.z, q.w);
        }
        if (t0 > 0 && grad_s != nullptr) {          // residual gradient of the pair (t0-1, t0): frame t0 is its target
            const float4 q = row[(size_t)(t0 - 1) * f4];
            step2_forward(p, X01, X23, f2(q.x, q.y), f2(q.z, q.w), su2, mine, lane, v01, v23, S);
            __syncwarp();
            const float2 d01 = sub2(u01, v01), d23 = sub2(u23, v23);
            carry01 = __fmul2_rn(d01, gl2v);
            carry23 = __fmul2_rn(d23, gl2v);
            if (L1TERM) {
                carry01.x += p.gl1 * sgn(d01.x); carry01.y += p.gl1 * sgn(d01.y);
                carry23.x += p.gl1 * sgn(d23.x); carry23.y += p.gl1 * sgn(d23.y);
            }
        }
        float2 acc2 = f2(0.0f, 0.0f);
        float acc1 = 0.0f;
        float4 pre = __ldcs(row + (size_t)(t0 + 1) * f4);                // streamed: every frame is read once per chunk
        const float4* pre_ptr = row + (size_t)(t0 + 2) * f4;             // running pointers: no 64-bit multiplies in the loop
        float4* g_ptr = grow + (size_t)t0 * f4;
#pragma unroll kUnroll
        for (int t = t0; t < t1; ++t, pre_ptr += f4, g_ptr += f4) {
            const float2 n01 = f2(pre.x, pre.y), n23 = f2(pre.z, pre.w);
            if (t + 2 <= T) pre = __ldcs(pre_ptr);                       // prefetch: in flight during this pair's arithmetic
            step2_forward(p, X01, X23, u01, u23, su2, mine, lane, v01, v23, S);
            const float2 d01 = sub2(n01, v01), d23 = sub2(n23, v23);     // burgers.py:29
            acc2 = __ffma2_rn(d01, d01, acc2);
            acc2 = __ffma2_rn(d23, d23, acc2);
            acc1 += (fabsf(d01.x) + fabsf(d01.y)) + (fabsf(d23.x) + fabsf(d23.y));
            float2 gd01 = __fmul2_rn(d01, gl2v), gd23 = __fmul2_rn(d23, gl2v);
            if (L1TERM) {
                gd01.x += p.gl1 * sgn(d01.x); gd01.y += p.gl1 * sgn(d01.y);
                gd23.x += p.gl1 * sgn(d23.x); gd23.y += p.gl1 * sgn(d23.y);
            }
            if (grad_s != nullptr) {
                float2 gin01, gin23;
                step2_adjoint(p, S, gd01, gd23, sacc2, mine_acc, lane, gin01, gin23);     // dL/dv = -gd
                gin01 = __fadd2_rn(gin01, carry01);
                gin23 = __fadd2_rn(gin23, carry23);
                __stcs(g_ptr, make_float4(gin01.x, gin01.y, gin23.x, gin23.y));          // written once, never re-read here
            } else {
                __syncwarp();
            }
            carry01 = gd01; carry23 = gd23;
            u01 = n01; u23 = n23;
        }
        if (grad_s != nullptr && t1 == T) __stcs(grow + (size_t)T * f4, make_float4(carry01.x, carry01.y, carry23.x, carry23.y));
        const float a2 = warp_sum(acc2.x + acc2.y);
        acc1 = warp_sum(acc1);
        if (lane == 0) partials[(size_t)b * nchunks + chunk] = make_float2(a2, acc1);
    }
    // second stage fused into the same launch: the CTA that finishes last adds all partials in a FIXED order
    // losses[0] = sum(d^2)/2/B (l2_loss, PF/math/nd.py:115-121), losses[1] = sum|d|/B (l1_loss, nd.py:96-108)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(done_counter, 1u) == gridDim.x - 1u ? 1u : 0u;
    }
    __syncthreads();
    if (s_last) {
        __threadfence();
        float a2 = 0.0f, a1 = 0.0f;
        const int count = B * nchunks;
        for (int i = threadIdx.x; i < count; i += 256) {
            const float2 v = __ldcg(partials + i);
            a2 += v.x;
            a1 += v.y;
        }
        a2 = warp_sum(a2);
        a1 = warp_sum(a1);
        if (lane == 0) { red2[warp] = a2; red1[warp] = a1; }
        __syncthreads();
        if (threadIdx.x == 0) {
            float t2 = 0.0f, t1 = 0.0f;
#pragma unroll
            for (int w = 0; w < 8; ++w) { t2 += red2[w]; t1 += red1[w]; }
            losses[0] = t2 * 0.5f * inv_B;
            losses[1] = t1 * inv_B;
        }
    }
}

}  // namespace pdec

using namespace pdec;

// returns PDEC_OK if it handled the call, 1 if the case is not this kernel's (caller falls back to burgers_fast.cu)
int pdec_burgers_pair_loss_unit(const float* s, float* grad_s, float2* partials, int T, int B, int C, int nchunks, double dt, double viscosity,
                                float gl2, float gl1, unsigned* done_counter, float* losses, cudaStream_t st) {
    Unit2P p;
    const float dtf = (float)dt;
    p.dt128 = dtf * 128.0f;                                   // exact (power-of-two scale)
    p.c = (float)(dt * viscosity) * 16384.0f;                 // a / dx^2, dx^2 = 2^-14: exact scale of float32(dt * nu)
    if (!(p.dt128 < 3.0e38f) || !(p.c < 3.0e38f)) return 1;
    const float invB = 1.0f / (float)B;
    p.gl2 = gl2 * invB; p.gl1 = gl1 * invB;
    const long long units = (long long)B * nchunks;
    long long grid = (units + 7) / 8;
    if (grid > 148LL * 64) grid = 148LL * 64;
    if (gl1 != 0.0f) burgers_pair_loss_unit_kernel<true><<<(unsigned)grid, 256, 0, st>>>(p, s, grad_s, partials, T, B, C, nchunks, done_counter, losses, invB);
    else burgers_pair_loss_unit_kernel<false><<<(unsigned)grid, 256, 0, st>>>(p, s, grad_s, partials, T, B, C, nchunks, done_counter, losses, invB);
    return cudaPeekAtLastError() == cudaSuccess ? PDEC_OK : PDEC_ERR_CUDA_LAUNCH;
}
