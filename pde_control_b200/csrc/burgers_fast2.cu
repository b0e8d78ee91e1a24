// burgers_fast2.cu -- second-generation fused force-loss kernel for the headline case of BASELINE config 2: rows of 128 cells on
// the unit box, 'constant' extrapolation, one diffusion substep (SURVEY rows B1-B6 in one launch).  One warp per row segment,
// lane l owns cells 4l..4l+3 (one 16-byte load / store per frame), as in burgers_fast.cu, but shaped by three measurements
// (profiles/r1k_burgers_summary.md):
//   (1) the integer/ALU pipe runs at half rate and a packed fp32 instruction does two lane-operations in one issue slot
//       (scripts/diag/f32x2_throughput.cu): the fp32 chains run on PAIRS of cells with add/mul/fma.rn.f32x2 -- explicit
//       round-to-nearest, so the forward step still rounds exactly like the reference's unfused TF ops;
//   (2) with the instruction count halved the kernel is bound by shared-memory wavefronts, a third of them bank conflicts of the
//       data-dependent gathers/scatters: the row lives in shared memory TRANSPOSED, word(j) = (j & 3) * 32 + (j >> 2), i.e. bank =
//       j >> 2.  Four ADJACENT cells share a bank, so two lanes of one instruction (whose back-traced indices increase with the
//       lane) can only collide where the flow compresses 4 cells into fewer than 4 (shocks) -- never in smooth or expanding
//       regions -- and a lane's own four cells sit in its own bank (own-row stores / loads are conflict-free);
//   (3) 1024 CTAs over 148 x 3 resident slots left the SMs idle 22 % of the time (tail): the (row, pair) space is cut into
//       equal contiguous ranges, one per resident warp, crossing row boundaries; a range that starts inside a row recomputes
//       the residual of the pair before it (one extra forward step per ~37 pairs).
// Index arithmetic is folded without changing a bit: xi = ((i + 0.5) - u * (128 dt)) - 0.5 equals the reference's
// ((x0 - u dt) - lower) / size * N - 0.5 on the unit box (x0 = (2i+1)/256 exactly; scaling by 128 commutes with rounding) and
// a * (raw / dx^2) = (a / dx^2) * raw because 1/dx^2 = 2^14.  Out-of-domain samples ('constant' rule: exactly +0 and zero
// gradient, tf_backend.py:466-497) are redirected to zero / dummy words instead of being predicated.
// The adjoint scatter is deterministic without atomics as in burgers_fast.cu: if floor(xi) is strictly increasing over the lanes
// of one warp instruction the targets are distinct (plain read-modify-write); otherwise lanes take turns in ascending order.
#include "pdec_common.cuh"

namespace pdec {

struct Unit2P {
    float dt128;     // float32(dt) * 128 = -dxi/du
    float c;         // float32(dt * viscosity) / dx^2: increment = c * (up + lo - 2 u)
    float gl2, gl1;  // dL/d(l2), dL/d(l1), already divided by B
};

// shared-memory row: words 0..127 = cells (transposed), words 128 and 160 = the taps of an out-of-domain sample (zero in the
// source row, scratch in the accumulator row); byte offset of cell j = j + (j & 3) * 127
#define F2_ROW 164
#define F2_OUT 512                    // "floor index" given to out-of-domain samples: taps at bytes 512 and 640
#ifndef F2_MINB
#define F2_MINB 1                     // resident CTAs per SM: one (see F2_MINB_LEAN).  A/B on one box, us per step, round 1 walk: 8x3 54.5, 7x4 53.5, 12x2 52.8,
                                      // 4x6 55.3, 8x4 60.9 and 6x5 63.6 spill; round 2 row-segment walk, 72 registers: 14x2 51.2, 12x2 53.3, 7x4 53.3, 9x3 55.3,
                                      // 16x2 57.8 spills; periodic rows 12x2 55.3, 24x1 53.7; replicate rows 12x2 58.7, 24x1 57.0
#endif
#ifndef F2_WARPS
#define F2_WARPS 24                   // warps per CTA (80 registers)
#endif
#ifndef F2_WARPS_LEAN
#define F2_WARPS_LEAN 28              // warps per CTA of the variant that fits 72 registers without spills (no L1 term, dense gradient, 'constant')
#endif
#ifndef F2_MINB_LEAN
#define F2_MINB_LEAN 1                // resident CTAs per SM of that variant.  ONE CTA of 28 warps, not two of 14: the warp schedulers favour the older
                                      // CTA, which then finishes 5 us before the younger one and leaves the SM half empty for that long (per-warp timers,
                                      // scripts/diag/burgers_tail.py: CTAs 0..147 37 us, CTAs 148..295 42 us; one CTA: all warps 40 us).  14x2 51.2, 28x1 49.2, 24x1 51.4 us
#endif
// Measured and dropped (profiles/r1k_burgers_summary.md, r2ab_burgers_walk.md): several shorter ranges per warp, more CTAs than resident
// slots, floor / ceil parts in two accumulator rows, a second register-staged prefetch, a cp.async ring of 2-5 frames per warp.

__device__ __forceinline__ float2 f2(float x, float y) { return make_float2(x, y); }
__device__ __forceinline__ float2 sub2(float2 a, float2 b) { return __ffma2_rn(b, f2(-1.0f, -1.0f), a); }   // rn(a - b), one rounding

struct Step2 {                        // what the forward step leaves for its adjoint
    int bi[4];                        // floor(xi)
    unsigned offF[4], offC[4];        // byte offsets of the floor / ceil tap inside a row
    float2 w0[2], w1[2];              // weight of the ceil / floor tap, cells (0,1) and (2,3)
    float du[4];                      // u[ceil] - u[floor]
};

__device__ __forceinline__ float lds_off(const float* row, unsigned off) { return *reinterpret_cast<const float*>(reinterpret_cast<const char*>(row) + off); }
__device__ __forceinline__ float* ptr_off(float* row, unsigned off) { return reinterpret_cast<float*>(reinterpret_cast<char*>(row) + off); }

// forward step of the row (u01, u23) -> (v01, v23); su = this warp's source row, mine = su + lane (own cells at mine[32 k]).
// PAD = PDEC_MODE_*.  'circular' rows (tf_backend.py:476-480, 532-538): taps wrap modulo 128, w1 = 1 - w0, the stencil wraps over the
// lanes.  'replicate' rows (:505-518): taps clamp to [0, 127], w1 = 1 - w0 from the UNCLAMPED floor, the stencil repeats the edge cell.
// `after_u()` runs once the frame's registers are dead (row stored, back-trace coordinates taken): the caller reloads them there.
template <int PAD, class AfterU>
__device__ __forceinline__ void step2_forward(const Unit2P& p, float2 X01, float2 X23, float2 u01, float2 u23, float* su, float* mine, int lane,
                                              float2& v01, float2& v23, Step2& S, AfterU&& after_u) {
    mine[0] = u01.x; mine[32] = u01.y; mine[64] = u23.x; mine[96] = u23.y;
    __syncwarp();
    const float2 dt2 = f2(p.dt128, p.dt128), half2v = f2(-0.5f, -0.5f), one2 = f2(1.0f, 1.0f);
    // advect.py:35, _box.py:50-52, grid.py:104-105 folded (see the header): xi = ((i + 0.5) - u * 128dt) - 0.5
    const float2 xi01 = __fadd2_rn(sub2(X01, __fmul2_rn(u01, dt2)), half2v);
    const float2 xi23 = __fadd2_rn(sub2(X23, __fmul2_rn(u23, dt2)), half2v);
    after_u();
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
    if (PAD != PDEC_MODE_CONSTANT) {
        S.w1[0] = sub2(one2, S.w0[0]);
        S.w1[1] = sub2(one2, S.w0[1]);
    } else {
        S.w1[0] = sub2(__fadd2_rn(bf01, one2), xi01);
        S.w1[1] = sub2(__fadd2_rn(bf23, one2), xi23);
    }
    const float w0[4] = {S.w0[0].x, S.w0[0].y, S.w0[1].x, S.w0[1].y};
    const float w1[4] = {S.w1[0].x, S.w1[0].y, S.w1[1].x, S.w1[1].y};
    float adv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if (PAD == PDEC_MODE_CIRCULAR) {
            const unsigned bF = (unsigned)S.bi[k] & 127u, bC = (unsigned)(S.bi[k] + 1) & 127u;
            S.offF[k] = bF + (bF & 3u) * 127u;
            S.offC[k] = bC + (bC & 3u) * 127u;
        } else if (PAD == PDEC_MODE_REPLICATE) {
            const unsigned bF = (unsigned)min(max(S.bi[k], 0), 127), bC = (unsigned)min(max(S.bi[k], -1), 126) + 1u;
            S.offF[k] = bF + (bF & 3u) * 127u;
            S.offC[k] = bC + (bC & 3u) * 127u;
        } else {
            const unsigned b = (unsigned)S.bi[k] < 127u ? (unsigned)S.bi[k] : (unsigned)F2_OUT;
            S.offF[k] = b + (b & 3u) * 127u;
            S.offC[k] = S.offF[k] + ((b & 3u) == 3u ? 4u - 384u : 128u);      // next cell: the next 32-word group, or word + 1 of group 0
        }
        const float tf = lds_off(su, S.offF[k]), tc = lds_off(su, S.offC[k]);     // (u[floor], u[floor + 1]) or (0, 0)
        adv[k] = __fadd_rn(__fmul_rn(tf, w1[k]), __fmul_rn(tc, w0[k]));           // tf_backend.py:491-497
        S.du[k] = tc - tf;
    }
    float left, right;
    if (PAD == PDEC_MODE_CIRCULAR) {                                   // circular padding (nd.py:271-283)
        left = __shfl_sync(PDEC_FULL_MASK, adv[3], (lane + 31) & 31);
        right = __shfl_sync(PDEC_FULL_MASK, adv[0], (lane + 1) & 31);
    } else {
        left = __shfl_up_sync(PDEC_FULL_MASK, adv[3], 1);
        right = __shfl_down_sync(PDEC_FULL_MASK, adv[0], 1);
        if (lane == 0) left = PAD == PDEC_MODE_REPLICATE ? adv[0] : 0.0f;      // edge cell repeated / zero padding
        if (lane == 31) right = PAD == PDEC_MODE_REPLICATE ? adv[3] : 0.0f;
    }
    const float2 a01 = f2(adv[0], adv[1]), a23 = f2(adv[2], adv[3]), mid = f2(adv[1], adv[2]);
    const float2 m2 = f2(-2.0f, -2.0f), c2v = f2(p.c, p.c);
    // nd.py:281-282: (up + lo) - 2 u (2 u is exact, so one fused rounding equals the reference's two); grid.py:190 and
    // util.py:41-43 folded into one multiply (power-of-two dx^2); then u + increment
    const float2 raw01 = __ffma2_rn(a01, m2, __fadd2_rn(mid, f2(left, adv[0])));
    const float2 raw23 = __ffma2_rn(a23, m2, __fadd2_rn(f2(adv[3], right), mid));
    // ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 even under --fmad=false (it never touches the scalar .rn forms; measured:
    // scripts/diag/dump_forward.py): the final sum u + increment is done with scalar adds
    const float2 inc01 = __fmul2_rn(raw01, c2v), inc23 = __fmul2_rn(raw23, c2v);
    v01 = f2(__fadd_rn(adv[0], inc01.x), __fadd_rn(adv[1], inc01.y));
    v23 = f2(__fadd_rn(adv[2], inc23.x), __fadd_rn(adv[3], inc23.y));
}

// adjoint of one step.  h = -(dL/d step output) on entry (the caller keeps the sign of the residual gradient); returns dL/d(step
// input) in (gin01, gin23).  The accumulator row's cell words are all-zero on entry and again on exit.
template <int PAD>
__device__ __forceinline__ void step2_adjoint(const Unit2P& p, const Step2& S, float2 h01, float2 h23, float* sacc, float* mine_acc, int lane,
                                              float2& gin01, float2& gin23) {
    float hl, hr;
    if (PAD == PDEC_MODE_CIRCULAR) {
        hl = __shfl_sync(PDEC_FULL_MASK, h23.y, (lane + 31) & 31);
        hr = __shfl_sync(PDEC_FULL_MASK, h01.x, (lane + 1) & 31);
    } else {                                                           // the stencil matrix is symmetric for every padding
        hl = __shfl_up_sync(PDEC_FULL_MASK, h23.y, 1);
        hr = __shfl_down_sync(PDEC_FULL_MASK, h01.x, 1);
        if (lane == 0) hl = PAD == PDEC_MODE_REPLICATE ? h01.x : 0.0f;
        if (lane == 31) hr = PAD == PDEC_MODE_REPLICATE ? h23.y : 0.0f;
    }
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
        if (PAD == PDEC_MODE_CIRCULAR) {      // wrapped targets are distinct when the unwrapped floors also span less than one period
            const int first = __shfl_sync(PDEC_FULL_MASK, S.bi[k], 0);
            ordered = ordered && ((long long)S.bi[k] - first < 128);
        }
        if (PAD == PDEC_MODE_REPLICATE) {     // the CLAMPED floor taps and the clamped ceil taps must each increase strictly (one lane
                                              // outside per instruction is harmless: its two taps coincide, but in different instructions)
            const int cf = min(max(S.bi[k], 0), 127), cc = min(max(S.bi[k], -1), 126);
            const int pf = __shfl_up_sync(PDEC_FULL_MASK, cf, 1), pc = __shfl_up_sync(PDEC_FULL_MASK, cc, 1);
            ordered = ordered && (lane == 0 || (cf > pf && cc > pc));
        }
    }
    if (__all_sync(PDEC_FULL_MASK, ordered)) {
        // strictly increasing floor(xi) over the lanes: the in-domain targets of one warp instruction are distinct (floor targets
        // among themselves, ceil targets among themselves); out-of-domain lanes all hit the scratch words
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float* af = ptr_off(sacc, S.offF[k]);
            *af = fmaf(-q[k], w1[k], *af);
            __syncwarp();
            float* ac = ptr_off(sacc, S.offC[k]);
            *ac = fmaf(-q[k], w0[k], *ac);
            __syncwarp();
        }
    } else
    {
        // folded back-trace (a shock steeper than one cell per cell): lanes that share a floor index are ranked by lane id
        // (match_any) and take turns by rank -- same sums in a fixed order, as many rounds as the largest group
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const bool inside = PAD != PDEC_MODE_CONSTANT || (unsigned)S.bi[k] < 127u;
            // group key of the floor tap; 'replicate' clamps floor and ceil taps separately, so the ceil taps get groups of their own
            const int keyF = PAD == PDEC_MODE_CIRCULAR ? (S.bi[k] & 127) : (PAD == PDEC_MODE_REPLICATE ? min(max(S.bi[k], 0), 127) : (inside ? S.bi[k] : -1 - lane));
            const unsigned peers = __match_any_sync(PDEC_FULL_MASK, keyF);
            const int rank = __popc(peers & ((1u << lane) - 1u));
            int rankC = rank;
            if (PAD == PDEC_MODE_REPLICATE) {
                const unsigned peersC = __match_any_sync(PDEC_FULL_MASK, min(max(S.bi[k], -1), 126));
                rankC = __popc(peersC & ((1u << lane) - 1u));
            }
            for (int r = 0; __any_sync(PDEC_FULL_MASK, inside && (rank >= r || rankC >= r)); ++r) {
                if (inside && rank == r) {
                    float* af = ptr_off(sacc, S.offF[k]);
                    *af = fmaf(-q[k], w1[k], *af);
                }
                __syncwarp();
                if (inside && rankC == r) {
                    float* ac = ptr_off(sacc, S.offC[k]);
                    *ac = fmaf(-q[k], w0[k], *ac);
                }
                __syncwarp();
            }
        }
    }
    // transpose of the gather (own four cells) plus the derivative through xi(u)
    const float o0 = mine_acc[0], o1 = mine_acc[32], o2 = mine_acc[64], o3 = mine_acc[96];
    mine_acc[0] = 0.0f; mine_acc[32] = 0.0f; mine_acc[64] = 0.0f; mine_acc[96] = 0.0f;
    const float2 dt2 = f2(p.dt128, p.dt128);
    const float2 e01 = __fmul2_rn(f2(S.du[0], S.du[1]), dt2), e23 = __fmul2_rn(f2(S.du[2], S.du[3]), dt2);   // -du * dxi/du
    gin01 = __ffma2_rn(q01, e01, f2(o0, o1));
    gin23 = __ffma2_rn(q23, e23, f2(o2, o3));
    __syncwarp();
}

__device__ __forceinline__ float sgn(float d) { return (d > 0.0f) ? 1.0f : ((d < 0.0f) ? -1.0f : 0.0f); }

// Work = the (row, pair) space B x [t0, t1) in row-major order (t0 = 0, t1 = T: the whole sequence; a narrower window serves the
// frame-sliced host pipeline), cut into gridDim.x * F2_WARPS equal contiguous ranges, one per warp.  A warp walks its range one ROW
// SEGMENT at a time: the frames of the segment's pairs sit in two register sets that take turns as (frame, next frame), and the set of
// the frame just consumed is reloaded a few instructions into its forward step with the frame after next -- no per-pair cursor, no
// register moves (round 2: 299 -> 267 instructions per pair, 72 registers -> 14 warps x 2 CTAs per SM).  A segment that does not
// start at pair 0 first walks the pair before it "carry-only" (forward step only: its residual gradient w.r.t. its target frame is
// the carry the first real pair needs); the first frame of the next row is requested during the segment's last pair.
template <bool L1TERM, bool GSTRIDE, int PAD, int WARPS, int MINB>
__global__ void __launch_bounds__(32 * WARPS, MINB) burgers_pair_loss_unit_kernel(Unit2P p, const float* __restrict__ s, float* __restrict__ grad_s,
                                                                            float2* partials, int T, int B, int t0, int t1, long long per_warp,
                                                                            long long gf4, unsigned* done_counter, float* __restrict__ losses, float inv_B) {
    __shared__ float smem[WARPS][2][F2_ROW];
    __shared__ float red2[WARPS], red1[WARPS];
    __shared__ unsigned s_last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float* su = smem[warp][0];
    float* sacc = smem[warp][1];
    for (int q = lane; q < F2_ROW; q += 32) { su[q] = 0.0f; sacc[q] = 0.0f; }
    __syncwarp();
    float* mine = su + lane;
    float* mine_acc = sacc + lane;
#ifdef F2_TIMING
    // diagnostic build (scripts/diag/burgers_tail.py): per warp (SM id, start, end) in ns of the global timer, kept in the unused part of `partials`
    unsigned long long t_begin;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_begin));
#endif
    const float2 X01 = f2((float)(4 * lane) + 0.5f, (float)(4 * lane) + 1.5f), X23 = f2((float)(4 * lane) + 2.5f, (float)(4 * lane) + 3.5f);
    typedef unsigned off_t32;                                          // float4 offsets of s fit 32 bits (checked by the launcher): one-instruction address updates
    const off_t32 f4 = (off_t32)B * 32;                               // float4 stride between frames
    const float4* s4 = reinterpret_cast<const float4*>(s) + lane;
    // frame (b, f) of the gradient lives at g4 + f * gf4 + b * 32: its frame stride may differ from s's (a row chunk written straight
    // into the caller's full-batch array, which may be page-locked host memory)
    float4* g4 = grad_s ? reinterpret_cast<float4*>(grad_s) + lane : nullptr;
    const float2 gl2v = f2(p.gl2, p.gl2);
    const int Cw = t1 - t0;                                            // pairs per row in the window
    const long long total = (long long)B * Cw;
    float2 acc2 = f2(0.0f, 0.0f);
    float acc1 = 0.0f;
    do {                                                               // this warp's range: per_warp pairs from P0
        const long long P0 = ((long long)blockIdx.x * WARPS + warp) * per_warp;
        const long long P1 = min(total, P0 + per_warp);
        if (P0 >= P1) break;
        int left = (int)(P1 - P0);                                     // real pairs of the range not yet handed to a row segment
        int b = (int)(P0 / Cw);
        int pa = t0 + (int)(P0 - (long long)b * Cw);                   // first real pair of the current row
        const bool want_g = g4 != nullptr;
        const int lead = want_g && t0 > 0 ? 1 : 0;                     // later rows start one frame early (carry-only pair)
        off_t32 off = (off_t32)(pa - (want_g && pa > 0 ? 1 : 0)) * f4 + (off_t32)b * 32;      // float4 offset of the first frame walked
        // two frame registers, A and B, take turns as (u, target): u's registers are dead a few instructions into the forward step and
        // are reloaded there with the frame after the target (or, at the last pair of the row, the first frame of the next row)
        float4 A = s4[off], Bq = __ldcs(s4 + off + f4);
        Step2 S;
        while (true) {                                                 // one row segment per turn
            int cnt = min(left, t1 - pa);                              // its real pairs: pa .. pa + cnt - 1
            left -= cnt;
            const bool ends_at_T = pa + cnt == T;
            const off_t32 next_first = (off_t32)(t0 - lead) * f4 + (off_t32)(b + 1) * 32;
            const bool more = left > 0;
            float2 carry01 = f2(0.0f, 0.0f), carry23 = f2(0.0f, 0.0f), v01, v23;
            if (want_g && pa > 0) {
                // carry-only pair (pa - 1, pa): forward step only, its residual gradient w.r.t. frame pa is the carry of the first real pair
                const float2 n01 = f2(Bq.x, Bq.y), n23 = f2(Bq.z, Bq.w);
                step2_forward<PAD>(p, X01, X23, f2(A.x, A.y), f2(A.z, A.w), su, mine, lane, v01, v23, S, [&] { A = __ldcs(s4 + off + 2 * f4); });
                const float2 d01 = sub2(n01, v01), d23 = sub2(n23, v23);
                carry01 = __fmul2_rn(d01, gl2v); carry23 = __fmul2_rn(d23, gl2v);
                if (L1TERM) {
                    carry01.x += p.gl1 * sgn(d01.x); carry01.y += p.gl1 * sgn(d01.y);
                    carry23.x += p.gl1 * sgn(d23.x); carry23.y += p.gl1 * sgn(d23.y);
                }
                __syncwarp();
                const float4 t = A; A = Bq; Bq = t;                    // (u, target) = (frame pa, frame pa + 1) again in (A, B)
                off += f4;
            }
            size_t goff = GSTRIDE ? (size_t)pa * gf4 + (size_t)b * 32 : 0;       // gradient frame of u when its stride differs from s's
            // one pair: (U, N) = (frame p, frame p + 1); U is reloaded inside; `off` = float4 offset of frame p
#define F2_PAIR(U, N)                                                                                                                    \
            {                                                                                                                            \
                const float2 n01 = f2(N.x, N.y), n23 = f2(N.z, N.w);                                                                     \
                step2_forward<PAD>(p, X01, X23, f2(U.x, U.y), f2(U.z, U.w), su, mine, lane, v01, v23, S, [&] {                           \
                    if (cnt > 1) U = __ldcs(s4 + off + 2 * f4); else if (more) U = __ldcs(s4 + next_first);                              \
                });                                                                                                                      \
                const float2 d01 = sub2(n01, v01), d23 = sub2(n23, v23);     /* burgers.py:29 */                                         \
                float2 gd01 = __fmul2_rn(d01, gl2v), gd23 = __fmul2_rn(d23, gl2v);                                                       \
                if (L1TERM) {                                                                                                            \
                    gd01.x += p.gl1 * sgn(d01.x); gd01.y += p.gl1 * sgn(d01.y);                                                          \
                    gd23.x += p.gl1 * sgn(d23.x); gd23.y += p.gl1 * sgn(d23.y);                                                          \
                }                                                                                                                        \
                acc2 = __ffma2_rn(d01, d01, acc2);                                                                                       \
                acc2 = __ffma2_rn(d23, d23, acc2);                                                                                       \
                acc1 += (fabsf(d01.x) + fabsf(d01.y)) + (fabsf(d23.x) + fabsf(d23.y));                                                   \
                if (want_g) {                                                                                                            \
                    float2 gin01, gin23;                                                                                                 \
                    F2_ADJOINT_CALL                                                                                                      \
                    gin01 = __fadd2_rn(gin01, carry01);                                                                                  \
                    gin23 = __fadd2_rn(gin23, carry23);                                                                                  \
                    __stcs(g4 + (GSTRIDE ? goff : (size_t)off), make_float4(gin01.x, gin01.y, gin23.x, gin23.y));   /* written once */   \
                    carry01 = gd01; carry23 = gd23;                                                                                      \
                } else {                                                                                                                 \
                    __syncwarp();                                                                                                        \
                }                                                                                                                        \
                off += f4;                                                                                                               \
                if (GSTRIDE) goff += (size_t)gf4;                                                                                        \
            }
#define F2_ADJOINT_CALL step2_adjoint<PAD>(p, S, gd01, gd23, sacc, mine_acc, lane, gin01, gin23);      /* dL/dv = -gd */
            bool in_b;                                                 // the row's last frame ended up in B (else in A); the other holds the next row's first
            while (true) {
                F2_PAIR(A, Bq)
                if (--cnt == 0) { in_b = true; break; }
                F2_PAIR(Bq, A)
                if (--cnt == 0) { in_b = false; break; }
            }
#undef F2_PAIR
#undef F2_ADJOINT_CALL
            // `off` = the row segment's last frame.  Frame T receives the residual gradient of the last pair; a window that ends earlier leaves
            // its last frame to the next window
            if (want_g && ends_at_T)
                __stcs(g4 + (GSTRIDE ? goff : (size_t)off), make_float4(carry01.x, carry01.y, carry23.x, carry23.y));
            if (!more) break;
            // next row: its first frame is in flight in the register set that does not hold the row's last frame
            if (in_b) { /* A holds it */ } else { A = Bq; }
            b += 1;
            pa = t0;
            off = next_first;
            Bq = __ldcs(s4 + off + f4);
        }
    } while (false);
#ifdef F2_TIMING
    if (lane == 0) {
        unsigned long long t_end;
        unsigned smid;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_end));
        asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
        uint4* dbg = reinterpret_cast<uint4*>(partials + 65536) + (size_t)blockIdx.x * WARPS + warp;
        *dbg = make_uint4(smid, (unsigned)(t_begin & 0xffffffffu), (unsigned)(t_end & 0xffffffffu), 0u);
    }
#endif
    {
        const float a2 = warp_sum(acc2.x + acc2.y);
        acc1 = warp_sum(acc1);
        if (lane == 0) partials[(size_t)blockIdx.x * WARPS + warp] = make_float2(a2, acc1);
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
        const int count = (int)gridDim.x * WARPS;
        for (int i = threadIdx.x; i < count; i += 32 * WARPS) {
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
            for (int w = 0; w < WARPS; ++w) { t2 += red2[w]; t1 += red1[w]; }
            losses[0] = t2 * 0.5f * inv_B;
            losses[1] = t1 * inv_B;
        }
    }
}

}  // namespace pdec

using namespace pdec;

// returns PDEC_OK if it handled the call, 1 if the case is not this kernel's (caller falls back to burgers_fast.cu).
// partials must hold max(32, B * T) float2 (pdec_burgers_pair_loss_workspace_bytes).  [t0, t1) = window of pairs (0, T = all):
// losses <- the window's share of (L2, L1); grad_s frames t0..t1-1 (and T when t1 == T) <- their complete gradient
// grad_frame_stride: floats between consecutive frames of grad_s (0: dense, B * 128)
int pdec_burgers_pair_loss_unit(const float* s, float* grad_s, float2* partials, int T, int B, int t0, int t1, int pad_mode, double dt, double viscosity,
                                float gl2, float gl1, long long grad_frame_stride, unsigned* done_counter, float* losses, cudaStream_t st) {
    if (t0 < 0 || t1 > T || t0 >= t1) return PDEC_ERR_INVALID_ARGUMENT;
    if ((long long)B * 32 * ((long long)T + 1) >= (1LL << 32)) return 1;       // float4 offsets are 32-bit in the kernel: larger arrays (> 64 GB) go to the v1 kernel
    if (grad_frame_stride != 0 && (grad_frame_stride < (long long)B * 128 || (grad_frame_stride & 3))) return PDEC_ERR_INVALID_ARGUMENT;
    const long long gf4 = grad_frame_stride ? grad_frame_stride / 4 : (long long)B * 32;
    Unit2P p;
    const float dtf = (float)dt;
    p.dt128 = dtf * 128.0f;                                   // exact (power-of-two scale)
    p.c = (float)(dt * viscosity) * 16384.0f;                 // a / dx^2, dx^2 = 2^-14: exact scale of float32(dt * nu)
    if (!(fabsf(p.dt128) < 3.0e38f) || !(fabsf(p.c) < 3.0e38f)) return 1;
    const float invB = 1.0f / (float)B;
    p.gl2 = gl2 * invB; p.gl1 = gl1 * invB;
    const long long total = (long long)B * (t1 - t0);
    const bool strided = gf4 != (long long)B * 32;
    const bool lean = gl1 == 0.0f && !strided && pad_mode == PDEC_MODE_CONSTANT;
    const int warps = lean ? F2_WARPS_LEAN : F2_WARPS;
    // one range per resident warp (148 SMs x F2_MINB CTAs x warps), at least 4 pairs per range
    long long grid = 148LL * (lean ? F2_MINB_LEAN : F2_MINB);
    if (grid * warps * 4 > total) grid = total / (4 * warps) > 0 ? total / (4 * warps) : 1;
    long long per_warp = (total + grid * warps - 1) / (grid * warps);
#define F2_LAUNCH(L1, GS, PER, W, MB) burgers_pair_loss_unit_kernel<L1, GS, PER, W, MB><<<(unsigned)grid, 32 * W, 0, st>>>(p, s, grad_s, partials, T, B, t0, t1, per_warp, gf4, done_counter, losses, invB)
#define F2_PICK(L1, GS) do { if (pad_mode == PDEC_MODE_CIRCULAR) F2_LAUNCH(L1, GS, PDEC_MODE_CIRCULAR, F2_WARPS, F2_MINB); else if (pad_mode == PDEC_MODE_REPLICATE) F2_LAUNCH(L1, GS, PDEC_MODE_REPLICATE, F2_WARPS, F2_MINB); else F2_LAUNCH(L1, GS, PDEC_MODE_CONSTANT, F2_WARPS, F2_MINB); } while (0)
    if (lean) F2_LAUNCH(false, false, PDEC_MODE_CONSTANT, F2_WARPS_LEAN, F2_MINB_LEAN);
    else if (gl1 != 0.0f) { if (strided) F2_PICK(true, true); else F2_PICK(true, false); }
    else { if (strided) F2_PICK(false, true); else F2_PICK(false, false); }
#undef F2_PICK
#undef F2_LAUNCH
    return cudaPeekAtLastError() == cudaSuccess ? PDEC_OK : PDEC_ERR_CUDA_LAUNCH;
}
