// cg_pair.cu -- the 128 x 128 Poisson CG with one batch element on a CLUSTER OF TWO CTAs (two SMs).
//
// cg_block_kernel gives one SM to one batch element, so BASELINE config 3 (B = 64) leaves 84 of the 148 SMs idle while the 64 busy
// ones run a latency-bound iteration (1.78 us at 128 x 128, 1.00 us at 64 x 128: profiles/r1n_cg_packed.md).  Here the element's
// rows are split between the two CTAs of a thread-block cluster (64 rows each, 8 warps, 8 x 4 cells per thread as in cg_block_kernel):
//   * the CTA's boundary row of m goes into the PEER's edge-row slot with a distributed-shared-memory store;
//   * every warp writes its reduction partials into both CTAs' arrays, indexed by its cluster-wide warp number, so both CTAs add the
//     same 16 partials in the same order and take the same branch;
//   * the three barriers per iteration are cluster barriers (arrive.release / wait.acquire, ~380 cycles), split so that independent
//     work sits between arrive and wait: the x update behind reduction B, the six interior rows of the matvec behind the publish.
// Same recurrence and operation order as cg_block_kernel<16, FULL, two reductions> (blas.py:5-64, sparse.py:157-221); the partial
// sums are grouped as in its 16-warp reduction, so the iterates are bit-identical to it (tests/test_fluid_gpu.py).
//
// MEASURED (profiles/r1p_cg_pair.md): 1.90 us per iteration with cluster barriers, 2.47 us with the mbarrier hand-shake, against 1.78 us
// for one CTA per element: three cross-SM synchronisations per iteration (~0.3 us each) cost more than halving the per-SM work saves
// (0.78 us).  Kept as an experimental variant of pdec_cg_poisson_2d (stop_mode | 6 << 8 / 7 << 8), not dispatched by default.
#include <cooperative_groups.h>

#include "pdec_common.cuh"
#include "cg_reduce.cuh"

namespace cg = cooperative_groups;

namespace pdec {

constexpr int CP_WARPS = 8;      // warps per CTA; the pair has 16

// first half of a two-value reduction over the pair: after the xor butterflies lane 0 holds the warp total of `a`, lane 16 of `b`;
// both go to slot value * 16 + (cluster-wide warp) of the local AND the peer's array
__device__ __forceinline__ void pair_reduce_begin(float a, float b, float* red, float* red_peer, int gw, const bool second_is_max) {
    const int lane = threadIdx.x & 31;
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
    if ((lane & 15) == 0) {
        red[(lane >> 4) * 16 + gw] = keep;
        red_peer[(lane >> 4) * 16 + gw] = keep;
    }
}
// second half, after the cluster barrier: one LDS + five shuffles per lane (the order of cg_block_kernel's 16-warp path)
__device__ __forceinline__ void pair_reduce_end(float& a, float& b, const float* red, const bool second_is_max) {
    const int lane = threadIdx.x & 31;
    const bool hi = lane >= 16;
    const bool mx = hi && second_is_max;
    float v = red[lane], recv;
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
        recv = __shfl_xor_sync(0xffffffffu, v, o);
        v = mx ? fmaxf(v, recv) : v + recv;
    }
    recv = __shfl_xor_sync(0xffffffffu, v, 16);
    a = hi ? recv : v;
    b = hi ? v : recv;
}

// Point-to-point synchronisation of the pair on mbarriers: every warp of both CTAs arrives once per phase on BOTH CTAs' barrier
// (the remote arrive travels the same DSMEM path as the data stored before it, release at cluster scope), and waits on its own
// CTA's copy only -- one one-way DSMEM latency instead of the hardware cluster barrier's round trip.
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned peer_u32(unsigned addr, unsigned peer) {
    unsigned r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(peer));
    return r;
}
__device__ __forceinline__ void pair_arrive(unsigned local, unsigned remote) {
    __syncwarp();                            // the lanes' stores are ordered before lane 0's release
    if ((threadIdx.x & 31) == 0) {
        asm volatile("mbarrier.arrive.release.cluster.shared::cta.b64 _, [%0];" ::"r"(local) : "memory");
        asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
    }
}
__device__ __forceinline__ void pair_wait(unsigned local, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "PAIR_WAIT_%=:\n"
        "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n"
        "@p bra PAIR_DONE_%=;\n"
        "bra PAIR_WAIT_%=;\n"
        "PAIR_DONE_%=:\n"
        "}\n" ::"r"(local), "r"(parity) : "memory");
}

template <bool P2P>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(32 * CP_WARPS, 1)
    cg_pair_kernel(const float* __restrict__ rhs, float* __restrict__ out, int* __restrict__ iters, float* __restrict__ resid, int closed,
                   float accuracy, int max_it) {
    constexpr int NX = 128, NY = 128, ROWS = 8 * CP_WARPS;
    __shared__ __align__(16) float4 sm_x[ROWS * 32];                    // this CTA's rows of x
    __shared__ __align__(16) float4 sm_e[(2 * CP_WARPS + 2) * 32];      // 0 = row above the CTA, 1+2w / 2+2w = first / last row of warp w, 17 = row below
    __shared__ __align__(16) float redA[32], redB[32];
    __shared__ __align__(8) unsigned long long bars[3];
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int tx = threadIdx.x & 31, w = threadIdx.x >> 5, gw = rank * CP_WARPS + w;
    const int i0 = ROWS * rank + 8 * w, j0 = 4 * tx;                   // global row of the thread's first cell
    const size_t base = (size_t)(blockIdx.x >> 1) * NY * NX;
    float* redA_peer = cluster.map_shared_rank(redA, rank ^ 1);
    float* redB_peer = cluster.map_shared_rank(redB, rank ^ 1);
    // the peer's edge slot that holds MY boundary row: rank 0's last row is "the row above" rank 1 (slot 0), rank 1's first row is
    // "the row below" rank 0 (slot 2 * CP_WARPS + 1)
    float4* edge_peer = cluster.map_shared_rank(sm_e, rank ^ 1) + (rank == 0 ? 0 : (2 * CP_WARPS + 1) * 32) + tx;
    const bool sends = rank == 0 ? w == CP_WARPS - 1 : w == 0;
    float r[32], am[32], m[32];
    float dcol[4], drow[8];
    const float cf = closed ? 1.0f : 0.0f;
#pragma unroll
    for (int kx = 0; kx < 4; ++kx) dcol[kx] = -4.0f + cf * ((j0 + kx == 0 ? 1.0f : 0.0f) + (j0 + kx == NX - 1 ? 1.0f : 0.0f));
#pragma unroll
    for (int ky = 0; ky < 8; ++ky) drow[ky] = cf * ((i0 + ky == 0 ? 1.0f : 0.0f) + (i0 + ky == NY - 1 ? 1.0f : 0.0f));
    const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    for (int q = threadIdx.x; q < (2 * CP_WARPS + 2) * 32; q += 32 * CP_WARPS) sm_e[q] = zero4;
#pragma unroll
    for (int ky = 0; ky < 8; ++ky) {
        const float4 r4 = *reinterpret_cast<const float4*>(rhs + base + (size_t)(i0 + ky) * NX + j0);
        r[ky * 4 + 0] = r4.x; r[ky * 4 + 1] = r4.y; r[ky * 4 + 2] = r4.z; r[ky * 4 + 3] = r4.w;
        sm_x[(8 * w + ky) * 32 + tx] = zero4;
    }
    const unsigned barA = smem_u32(&bars[0]), barB = smem_u32(&bars[1]), barC = smem_u32(&bars[2]);
    const unsigned barA_peer = peer_u32(barA, rank ^ 1), barB_peer = peer_u32(barB, rank ^ 1), barC_peer = peer_u32(barC, rank ^ 1);
    unsigned parA = 0u, parB = 0u, parC = 0u;
    if (P2P && threadIdx.x == 0) {
#pragma unroll
        for (int q = 0; q < 3; ++q) asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bars[q])), "r"(2 * CP_WARPS) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    cluster.sync();                          // both CTAs' edge rows are zero-filled (and barriers initialised) before anyone publishes
    auto arrive = [&](const unsigned bar, const unsigned bar_peer) {
        if (P2P) pair_arrive(bar, bar_peer);
        else cluster.barrier_arrive();
    };
    auto wait = [&](const unsigned bar, unsigned& parity) {
        if (P2P) { pair_wait(bar, parity); parity ^= 1u; }
        else cluster.barrier_wait();
    };
    auto publish = [&]() {                   // first and last row of the warp's block: local slots, and the peer's slot at the seam
        const float4 first = make_float4(m[0], m[1], m[2], m[3]), last = make_float4(m[28], m[29], m[30], m[31]);
        sm_e[(1 + 2 * w) * 32 + tx] = first;
        sm_e[(2 + 2 * w) * 32 + tx] = last;
        if (sends) *edge_peer = rank == 0 ? last : first;
    };
    float4 u4, d4;
    auto matvec_row = [&](const int ky) {
        float left = __shfl_up_sync(0xffffffffu, m[ky * 4 + 3], 1);
        float right = __shfl_down_sync(0xffffffffu, m[ky * 4 + 0], 1);
        if (tx == 0) left = 0.0f;
        if (tx == 31) right = 0.0f;
        const float upv[4] = {u4.x, u4.y, u4.z, u4.w}, dnv[4] = {d4.x, d4.y, d4.z, d4.w};
#pragma unroll
        for (int kx = 0; kx < 4; ++kx) {
            const int c = ky * 4 + kx;
            const float u = ky > 0 ? m[c >= 4 ? c - 4 : 0] : upv[kx];
            const float d = ky < 7 ? m[c + 4 <= 31 ? c + 4 : 31] : dnv[kx];
            const float l = kx > 0 ? m[c >= 1 ? c - 1 : 0] : left;
            const float rr = kx < 3 ? m[c + 1 <= 31 ? c + 1 : 31] : right;
            float dg;
            if (kx > 0 && kx < 3 && ky > 0 && ky < 7) dg = -4.0f;
            else if (kx > 0 && kx < 3) dg = -4.0f + drow[ky];
            else if (ky > 0 && ky < 7) dg = dcol[kx];
            else dg = dcol[kx] + drow[ky];
            am[c] = (u + l) + (rr + d) + dg * m[c];            // column order of the sorted sparse row (sparse.py:173-176)
        }
    };
    auto matvec_interior = [&]() {           // rows 1..6 of the thread's block: no edge rows needed
#pragma unroll
        for (int ky = 1; ky < 7; ++ky) matvec_row(ky);
    };
    auto matvec_edges = [&]() {
        u4 = sm_e[(2 * w) * 32 + tx];
        d4 = sm_e[(2 * w + 3) * 32 + tx];
        matvec_row(0);
        matvec_row(7);
    };
    float rmax = 0.0f, dummy = 0.0f;
    int it = 0;
    u4 = zero4; d4 = zero4;
#pragma unroll
    for (int c = 0; c < 32; ++c) { m[c] = r[c]; rmax = fmaxf(rmax, fabsf(r[c])); }
    publish();
    pair_reduce_begin(dummy, rmax, redB, redB_peer, gw, true);
    arrive(barB, barB_peer);
    matvec_interior();
    wait(barB, parB);
    pair_reduce_end(dummy, rmax, redB, true);
    matvec_edges();
    while (it < max_it && (accuracy <= 0.0f || rmax >= accuracy)) {
        float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
        for (int c = 0; c < 32; ++c) { s1 += m[c] * am[c]; s2 += m[c] * r[c]; }
        pair_reduce_begin(s1, s2, redA, redA_peer, gw, false);
        arrive(barA, barA_peer);
        wait(barA, parA);
        pair_reduce_end(s1, s2, redA, false);
        const float tmp = s1, a = div_no_nan_fast(s2, tmp);
        float s3 = 0.0f, mx = 0.0f;
#pragma unroll
        for (int c = 0; c < 32; ++c) {
            r[c] -= a * am[c];
            s3 += r[c] * am[c];
            mx = fmaxf(mx, fabsf(r[c]));
        }
        pair_reduce_begin(s3, mx, redB, redB_peer, gw, true);
        arrive(barB, barB_peer);
#pragma unroll
        for (int ky = 0; ky < 8; ++ky) {       // x += a m behind the barrier: needs only a and the old m
            float4 x4 = sm_x[(8 * w + ky) * 32 + tx];
            x4.x += a * m[ky * 4 + 0]; x4.y += a * m[ky * 4 + 1]; x4.z += a * m[ky * 4 + 2]; x4.w += a * m[ky * 4 + 3];
            sm_x[(8 * w + ky) * 32 + tx] = x4;
        }
        wait(barB, parB);
        pair_reduce_end(s3, mx, redB, true);
        const float beta = div_no_nan_fast(s3, tmp);
#pragma unroll
        for (int c = 0; c < 32; ++c) m[c] = r[c] - beta * m[c];
        publish();                           // every read of the old edge rows happened before this thread arrived at barrier A
        arrive(barC, barC_peer);
        matvec_interior();                   // behind the barrier: rows that need no edge row
        wait(barC, parC);
        matvec_edges();
        rmax = mx;
        ++it;
    }
#pragma unroll
    for (int ky = 0; ky < 8; ++ky)
        *reinterpret_cast<float4*>(out + base + (size_t)(i0 + ky) * NX + j0) = sm_x[(8 * w + ky) * 32 + tx];
    if (threadIdx.x == 0 && rank == 0) {
        if (iters) iters[blockIdx.x >> 1] = it;
        if (resid) resid[blockIdx.x >> 1] = rmax;
    }
    cluster.sync();                          // no CTA of the pair exits while the peer may still store into its shared memory
}

// 128 x 128, zero guess, per-element stop rule, B <= (SMs / 2): one CTA pair per element.  Returns PDEC_OK or a launch error.
int cg_pair_launch(const float* rhs, float* p, int* iterations, float* residual, int B, int closed, float accuracy, int max_iterations,
                   bool p2p, cudaStream_t st) {
    if (p2p) cg_pair_kernel<true><<<2 * B, 32 * CP_WARPS, 0, st>>>(rhs, p, iterations, residual, closed, accuracy, max_iterations);
    else cg_pair_kernel<false><<<2 * B, 32 * CP_WARPS, 0, st>>>(rhs, p, iterations, residual, closed, accuracy, max_iterations);
    return cudaPeekAtLastError() == cudaSuccess ? PDEC_OK : PDEC_ERR_CUDA_LAUNCH;
}

}  // namespace pdec
