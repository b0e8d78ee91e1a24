// cg_tiled.cu -- conjugate-gradient Poisson solve on a GRID OF TILES that stays on chip, on one GPU or on a row-slab
// decomposition over several GPUs (BASELINE config 5).
//
// Reference: poisson_solve / SparseCG / conjugate_gradient (PF/physics/pressuresolver/solver_api.py:132-158,
// sparse.py:73-221, PF/math/blas.py:5-64).  The reference streams every CG vector through memory every iteration (and its
// native twin, PF/tf/cuda/src/pressure_solve_op.cu.cc, launches one kernel per vector operation).  Here:
//   * one persistent CTA per tile of (8*WARPS rows x 128 columns); r, A m and m of the tile live in REGISTERS, x in shared
//     memory (the register/shared-memory layout of cg_block_kernel, cg.cu) -- HBM is touched once (rhs in, p out);
//   * ONE grid-wide barrier per iteration.  With q = A m, all dot products are taken before the update
//         a = (m.r)/(m.q),   beta = ((r.q) - a (q.q))/(m.q)          [ = (r'.q)/(m.q) with r' = r - a q ]
//     and every tile publishes the edge rows/columns of (r, q, m) BEFORE the barrier.  After it, each tile knows a and beta
//     and forms its neighbours' new search direction on the halo itself: m' = (r - a q) - beta m.  The barrier therefore
//     carries the reduction AND the halo exchange;
//   * on several GPUs (row slabs) the same barrier spans the GPUs over NVLink peer memory: edge tiles store their rows straight
//     into the neighbour GPU's inbox, the last CTA to arrive on each GPU pushes that GPU's five partial sums into every peer's
//     mailbox and raises a flag; no NCCL call, no host round trip, nothing but the barrier between two iterations.
// Sums are formed in a fixed order (tile order, then rank order), so every CTA on every GPU computes bit-identical a and beta.
// Spin loops are bounded (clock64 budget); on a time-out every participant leaves and the iteration count reads -1.
#include "pdec_common.cuh"
#include "cg_reduce.cuh"
#include <string.h>

namespace pdec {

typedef unsigned long long u64;

enum { SIDE_TOP = 0, SIDE_BOTTOM = 1, SIDE_LEFT = 2, SIDE_RIGHT = 3 };
enum { FROM_ABOVE = 0, FROM_BELOW = 1 };
#define PDEC_SLAB_MAX_RANKS 8
// layout of the peer-visible communication buffer (bytes): [flags u64[8] | done u64[8] | abort u32 .. pad to 256]
//   [epoch u64 (local use) .. pad to 512] [mail float 2 x 8 x B x 8] [inbox float 2 x B x 2 x tiles_x x 3 x 128]
#define COMM_FLAGS 0
#define COMM_DONE 64
#define COMM_ABORT 128
#define COMM_EPOCH 256
#define COMM_MAIL 512

struct CgTiled {
    const float* rhs; float* out; int* iters; float* resid;
    int B, ny_loc, nx, ny_glob, row0, closed, max_it;
    float accuracy;
    int tiles_y, tiles_x, tile_rows;
    int rank, nranks;
    // local workspace
    unsigned* cnt; unsigned* release; unsigned* abort_local;
    float* totals;      // [2][B][8]
    float* partials;    // [2][B*tiles][8]
    float* edges;       // [2][B*tiles][4][3][128]
    // communication buffers (this rank's and the peers' mappings of theirs)
    char* comm; char* peer[PDEC_SLAB_MAX_RANKS];
    long long spin_budget;      // clock64 ticks a wait may take before the solve is abandoned
};

__device__ __forceinline__ unsigned ld_volatile_u32(const unsigned* p) { return *reinterpret_cast<const volatile unsigned*>(p); }
__device__ __forceinline__ u64 ld_volatile_u64(const u64* p) { return *reinterpret_cast<const volatile u64*>(p); }
__device__ __forceinline__ void st_volatile_u32(unsigned* p, unsigned v) { *reinterpret_cast<volatile unsigned*>(p) = v; }
__device__ __forceinline__ void st_volatile_u64(u64* p, u64 v) { *reinterpret_cast<volatile u64*>(p) = v; }

__device__ __forceinline__ float* comm_mail(char* comm, int B, int par, int src_rank, int b) {
    return reinterpret_cast<float*>(comm + COMM_MAIL) + (((size_t)par * PDEC_SLAB_MAX_RANKS + src_rank) * B + b) * 8;
}
__device__ __forceinline__ float* comm_inbox(char* comm, int B, int tiles_x, int par, int b, int dir, int tx, int vec) {
    float* base = reinterpret_cast<float*>(comm + COMM_MAIL) + (size_t)2 * PDEC_SLAB_MAX_RANKS * B * 8;
    return base + (((((size_t)par * B + b) * 2 + dir) * tiles_x + tx) * 3 + vec) * 128;
}
__device__ __forceinline__ float* edge_ptr(const CgTiled& P, int par, int tile_global, int side, int vec) {
    return P.edges + ((((size_t)par * P.B * P.tiles_y * P.tiles_x + tile_global) * 4 + side) * 3 + vec) * 128;
}

// raise the abort flag everywhere (local + peers); idempotent
__device__ __forceinline__ void raise_abort(const CgTiled& P) {
    st_volatile_u32(P.abort_local, 1u);
    for (int p = 0; p < P.nranks; ++p) st_volatile_u32(reinterpret_cast<unsigned*>(P.peer[p] + COMM_ABORT), 1u);
}
__device__ __forceinline__ bool aborted(const CgTiled& P) {
    return ld_volatile_u32(P.abort_local) != 0u || ld_volatile_u32(reinterpret_cast<const unsigned*>(P.comm + COMM_ABORT)) != 0u;
}

// Grid-wide (and, with nranks > 1, GPU-wide) barrier number k of this launch.  Before the call every CTA has written its
// partials(par, tile) and its edges; on return totals(par, b)[0..4] hold the five global reductions of every batch element.
// Returns false when the solve was abandoned (time-out somewhere).
template <int THREADS>
__device__ bool tiled_barrier(const CgTiled& P, unsigned k, u64 seq, unsigned* s_flag) {
    const int tid = threadIdx.x, par = k & 1;
    const int tiles = P.tiles_y * P.tiles_x;
    __syncthreads();
    if (tid == 0) {
        if (P.nranks > 1) __threadfence_system(); else __threadfence();
        const unsigned old = atomicAdd(P.cnt, 1u);
        s_flag[0] = (old == (k + 1u) * gridDim.x - 1u) ? 1u : 0u;
        s_flag[1] = 0u;
    }
    __syncthreads();
    if (s_flag[0]) {                                     // last CTA of this GPU to arrive
        __threadfence();
        const int lane = tid & 31, w = tid >> 5;
        for (int pair = w; pair < P.B * 5; pair += THREADS / 32) {
            const int b = pair / 5, v = pair - 5 * b;
            const float* src = P.partials + ((size_t)par * P.B * tiles + (size_t)b * tiles) * 8 + v;
            float acc = 0.0f;
            for (int t = lane; t < tiles; t += 32) {
                const float val = __ldcg(src + (size_t)t * 8);
                acc = (v == 4) ? fmaxf(acc, val) : acc + val;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float other = __shfl_xor_sync(0xffffffffu, acc, o);
                acc = (v == 4) ? fmaxf(acc, other) : acc + other;
            }
            if (lane == 0) {
                if (P.nranks == 1) P.totals[((size_t)par * P.B + b) * 8 + v] = acc;
                else
                    for (int p = 0; p < P.nranks; ++p) comm_mail(P.peer[p], P.B, par, P.rank, b)[v] = acc;
            }
        }
        if (P.nranks > 1) {
            __syncthreads();
            if (tid == 0) {
                __threadfence_system();
                for (int p = 0; p < P.nranks; ++p)
                    if (p != P.rank) st_volatile_u64(reinterpret_cast<u64*>(P.peer[p] + COMM_FLAGS) + P.rank, seq);
                const long long t0 = clock64();
                bool ok = true;
                for (int r = 0; r < P.nranks && ok; ++r) {
                    if (r == P.rank) continue;
                    while (ld_volatile_u64(reinterpret_cast<const u64*>(P.comm + COMM_FLAGS) + r) < seq) {
                        if (aborted(P) || clock64() - t0 > P.spin_budget) { ok = false; break; }
                    }
                }
                if (!ok) { raise_abort(P); s_flag[1] = 1u; }
                __threadfence_system();
            }
            __syncthreads();
            for (int pair = tid; pair < P.B * 5; pair += THREADS) {
                const int b = pair / 5, v = pair - 5 * b;
                float acc = 0.0f;
                for (int r = 0; r < P.nranks; ++r) {
                    const float val = __ldcg(comm_mail(P.comm, P.B, par, r, b) + v);
                    acc = (v == 4) ? fmaxf(acc, val) : acc + val;
                }
                P.totals[((size_t)par * P.B + b) * 8 + v] = acc;
            }
        }
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            st_volatile_u32(P.release, k + 1u);
        }
    } else if (tid == 0) {
        const long long t0 = clock64();
        while (ld_volatile_u32(P.release) < k + 1u) {
            if (aborted(P) || clock64() - t0 > P.spin_budget) { raise_abort(P); s_flag[1] = 1u; break; }
        }
        __threadfence();
    }
    __syncthreads();
    if (tid == 0 && aborted(P)) s_flag[1] = 1u;
    __syncthreads();
    return s_flag[1] == 0u;
}

template <int WARPS>
__global__ void __launch_bounds__(32 * WARPS, 1) cg_tiled_kernel(CgTiled P) {
    constexpr int THREADS = 32 * WARPS, ROWS = 8 * WARPS;
    extern __shared__ __align__(16) float smem_dyn[];
    __shared__ __align__(16) float red5[2][80];
    __shared__ unsigned s_flag[2];
    __shared__ int s_done;
    float4* sm_x = reinterpret_cast<float4*>(smem_dyn);          // [ROWS][32]
    float4* sm_e = sm_x + ROWS * 32;                            // [2*WARPS + 2][32] edge rows of m (0 / last used = halos)
    float* sm_lr = reinterpret_cast<float*>(sm_e + (2 * WARPS + 2) * 32);      // [2][ROWS] left / right halo column of m
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int tiles = P.tiles_y * P.tiles_x;
    const int b = blockIdx.x / tiles, t = blockIdx.x - b * tiles;
    const int ty = t / P.tiles_x, txt = t - ty * P.tiles_x;
    const int wu = P.tile_rows >> 3;                             // warps that own rows
    const int i0 = ty * P.tile_rows + 8 * w, j0 = txt * 128 + 4 * lane;
    const size_t base = (size_t)b * P.ny_loc * P.nx;
    const int nx = P.nx;
    float r[32], am[32], m[32];
    unsigned valid = 0u;
    float dcol[4], drow[8];
    const float cf = P.closed ? 1.0f : 0.0f;
#pragma unroll
    for (int kx = 0; kx < 4; ++kx) dcol[kx] = -4.0f + cf * ((j0 + kx == 0 ? 1.0f : 0.0f) + (j0 + kx == nx - 1 ? 1.0f : 0.0f));
#pragma unroll
    for (int ky = 0; ky < 8; ++ky) {
        const int gi = P.row0 + i0 + ky;
        drow[ky] = cf * ((gi == 0 ? 1.0f : 0.0f) + (gi == P.ny_glob - 1 ? 1.0f : 0.0f));
    }
    const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    for (int q = tid; q < (2 * WARPS + 2) * 32; q += THREADS) sm_e[q] = zero4;
    for (int q = tid; q < 2 * ROWS; q += THREADS) sm_lr[q] = 0.0f;
    const bool vec_ok = (nx & 3) == 0 && (reinterpret_cast<uintptr_t>(P.rhs) & 15) == 0 && (reinterpret_cast<uintptr_t>(P.out) & 15) == 0;
#pragma unroll
    for (int ky = 0; ky < 8; ++ky) {
        float rv[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        if (w < wu) {
            if (vec_ok && j0 + 3 < nx) {
                const float4 r4 = *reinterpret_cast<const float4*>(P.rhs + base + (size_t)(i0 + ky) * nx + j0);
                rv[0] = r4.x; rv[1] = r4.y; rv[2] = r4.z; rv[3] = r4.w;
                valid |= 15u << (ky * 4);
            } else {
#pragma unroll
                for (int kx = 0; kx < 4; ++kx)
                    if (j0 + kx < nx) { valid |= 1u << (ky * 4 + kx); rv[kx] = P.rhs[base + (size_t)(i0 + ky) * nx + j0 + kx]; }
            }
        }
#pragma unroll
        for (int kx = 0; kx < 4; ++kx) { r[ky * 4 + kx] = rv[kx]; m[ky * 4 + kx] = 0.0f; am[ky * 4 + kx] = 0.0f; }
        sm_x[(8 * w + ky) * 32 + lane] = zero4;
    }
    // neighbours
    const bool has_up = P.row0 + ty * P.tile_rows > 0, has_dn = P.row0 + (ty + 1) * P.tile_rows < P.ny_glob;
    const bool has_l = txt > 0, has_r = txt < P.tiles_x - 1;
    const bool up_remote = has_up && ty == 0, dn_remote = has_dn && ty == P.tiles_y - 1;
    const int tg = b * tiles + t;
    u64* epoch_ptr = reinterpret_cast<u64*>(P.comm + COMM_EPOCH);
    const u64 seq_base = ld_volatile_u64(epoch_ptr);
    // a peer may still be reading what the PREVIOUS solve put into its mailbox / inbox: wait for its "done" stamp
    if (P.nranks > 1 && tid == 0) {
        const long long t0 = clock64();
        for (int rk = 0; rk < P.nranks; ++rk) {
            if (rk == P.rank) continue;
            while (ld_volatile_u64(reinterpret_cast<const u64*>(P.comm + COMM_DONE) + rk) + 1 < seq_base) {
                if (aborted(P) || clock64() - t0 > P.spin_budget) { raise_abort(P); break; }
            }
        }
    }
    __syncthreads();
    auto publish_local = [&]() {
        if (w < wu) {
            sm_e[(1 + 2 * w) * 32 + lane] = make_float4(m[0], m[1], m[2], m[3]);
            sm_e[(2 + 2 * w) * 32 + lane] = make_float4(m[28], m[29], m[30], m[31]);
        }
    };
    auto matvec = [&]() {
        // warp w reads the last row of warp w-1 (or the top halo in row 0) and the first row of warp w+1 (or the bottom halo,
        // which the halo loader put right after the last used warp: row 2*wu + 1)
        const float4 u4 = sm_e[(2 * w) * 32 + lane], d4 = sm_e[(w == wu - 1 ? 2 * wu + 1 : 2 * w + 3) * 32 + lane];
        const float upv[4] = {u4.x, u4.y, u4.z, u4.w}, dnv[4] = {d4.x, d4.y, d4.z, d4.w};
        const float4 l0 = *reinterpret_cast<const float4*>(sm_lr + 8 * w), l1 = *reinterpret_cast<const float4*>(sm_lr + 8 * w + 4);
        const float4 r0 = *reinterpret_cast<const float4*>(sm_lr + ROWS + 8 * w), r1 = *reinterpret_cast<const float4*>(sm_lr + ROWS + 8 * w + 4);
        const float lh[8] = {l0.x, l0.y, l0.z, l0.w, l1.x, l1.y, l1.z, l1.w}, rh[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
        for (int ky = 0; ky < 8; ++ky) {
            float left = __shfl_up_sync(0xffffffffu, m[ky * 4 + 3], 1);
            float right = __shfl_down_sync(0xffffffffu, m[ky * 4 + 0], 1);
            if (lane == 0) left = lh[ky];
            if (lane == 31) right = rh[ky];
#pragma unroll
            for (int kx = 0; kx < 4; ++kx) {
                const int c = ky * 4 + kx;
                const float u = ky > 0 ? m[c >= 4 ? c - 4 : 0] : upv[kx];
                const float d = ky < 7 ? m[c + 4 <= 31 ? c + 4 : 31] : dnv[kx];
                const float l = kx > 0 ? m[c >= 1 ? c - 1 : 0] : left;
                const float rr = kx < 3 ? m[c + 1 <= 31 ? c + 1 : 31] : right;
                const float val = (u + l) + (rr + d) + (dcol[kx] + drow[ky]) * m[c];
                am[c] = (valid & (1u << c)) ? val : 0.0f;
            }
        }
    };
    bool done_me = false, ok = true;
    int my_iters = 0;
    float my_res = 0.0f;
    unsigned k = 0;
    for (;; ++k) {
        const int par = k & 1;
        // ---- 1. the five partial reductions of this tile ----
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
        block_reduce5<WARPS>(s1, s2, s4, s5, mx, red5[par]);
        // ---- 2. publish the tile's edges of (r, A m, m) and its partials ----
        if (!done_me) {
            auto publish_edges = [&](const float (&v)[32], const int vec) {
                if (w == 0) {
                    const float4 row = make_float4(v[0], v[1], v[2], v[3]);
                    if (has_up && !up_remote) *reinterpret_cast<float4*>(edge_ptr(P, par, tg, SIDE_TOP, vec) + 4 * lane) = row;
                    if (up_remote) *reinterpret_cast<float4*>(comm_inbox(P.peer[P.rank - 1], P.B, P.tiles_x, par, b, FROM_BELOW, txt, vec) + 4 * lane) = row;
                }
                if (w == wu - 1) {
                    const float4 row = make_float4(v[28], v[29], v[30], v[31]);
                    if (has_dn && !dn_remote) *reinterpret_cast<float4*>(edge_ptr(P, par, tg, SIDE_BOTTOM, vec) + 4 * lane) = row;
                    if (dn_remote) *reinterpret_cast<float4*>(comm_inbox(P.peer[P.rank + 1], P.B, P.tiles_x, par, b, FROM_ABOVE, txt, vec) + 4 * lane) = row;
                }
                if (w < wu && lane == 0 && has_l) {
                    float* e = edge_ptr(P, par, tg, SIDE_LEFT, vec) + 8 * w;
                    *reinterpret_cast<float4*>(e) = make_float4(v[0], v[4], v[8], v[12]);
                    *reinterpret_cast<float4*>(e + 4) = make_float4(v[16], v[20], v[24], v[28]);
                }
                if (w < wu && lane == 31 && has_r) {
                    float* e = edge_ptr(P, par, tg, SIDE_RIGHT, vec) + 8 * w;
                    *reinterpret_cast<float4*>(e) = make_float4(v[3], v[7], v[11], v[15]);
                    *reinterpret_cast<float4*>(e + 4) = make_float4(v[19], v[23], v[27], v[31]);
                }
            };
            publish_edges(r, 0);
            publish_edges(am, 1);
            publish_edges(m, 2);
        }
        if (tid == 0) {
            float* pp = P.partials + ((size_t)par * P.B * tiles + tg) * 8;
            pp[0] = s1; pp[1] = s2; pp[2] = s4; pp[3] = s5; pp[4] = mx;
        }
        // ---- 3. the one barrier of the iteration ----
        ok = tiled_barrier<THREADS>(P, k, seq_base + k, s_flag);
        if (!ok) break;
        // ---- 4. stop tests (per batch element; everybody leaves when every element is done) ----
        const int it = k == 0 ? 0 : (int)k - 1;                // real updates applied so far
        const bool budget = it < P.max_it;
        if (tid == 0) s_done = 1;
        __syncthreads();
        for (int q = tid; q < P.B; q += THREADS) {
            const float rq = __ldcg(P.totals + ((size_t)par * P.B + q) * 8 + 4);
            if (budget && (P.accuracy <= 0.0f || rq >= P.accuracy)) s_done = 0;      // benign race: all writers store 0
        }
        const float* tot = P.totals + ((size_t)par * P.B + b) * 8;
        const float S1 = __ldcg(tot), S2 = __ldcg(tot + 1), S4 = __ldcg(tot + 2), S5 = __ldcg(tot + 3), MX = __ldcg(tot + 4);
        __syncthreads();
        const bool all_done = s_done != 0;
        if (!done_me && !(budget && (P.accuracy <= 0.0f || MX >= P.accuracy))) { done_me = true; my_iters = it; my_res = MX; }
        if (all_done) break;
        if (done_me) continue;
        // ---- 5. update (pass 0 has m = A m = 0, hence a = beta = 0: it only turns r into the first search direction) ----
        const float a = div_no_nan(S2, S1);
        const float beta = div_no_nan(S4 - a * S5, S1);
#pragma unroll
        for (int ky = 0; ky < 8; ++ky) {
            float4 x4 = sm_x[(8 * w + ky) * 32 + lane];
            x4.x += a * m[ky * 4 + 0]; x4.y += a * m[ky * 4 + 1]; x4.z += a * m[ky * 4 + 2]; x4.w += a * m[ky * 4 + 3];
            sm_x[(8 * w + ky) * 32 + lane] = x4;
#pragma unroll
            for (int kx = 0; kx < 4; ++kx) {
                const int c = ky * 4 + kx;
                r[c] -= a * am[c];
                m[c] = r[c] - beta * m[c];
            }
        }
        // ---- 6. the neighbours' new search direction on the halo, from their published (r, A m, m) ----
        for (int q = tid; q < 512; q += THREADS) {
            const int side = q >> 7, idx = q & 127;
            const float *er = nullptr, *ea = nullptr, *em = nullptr;
            bool have = false;
            if (side == SIDE_TOP && has_up) {
                have = true;
                if (up_remote) { er = comm_inbox(P.comm, P.B, P.tiles_x, par, b, FROM_ABOVE, txt, 0); ea = er + 128; em = ea + 128; }
                else { er = edge_ptr(P, par, tg - P.tiles_x, SIDE_BOTTOM, 0); ea = er + 128; em = ea + 128; }
            } else if (side == SIDE_BOTTOM && has_dn) {
                have = true;
                if (dn_remote) { er = comm_inbox(P.comm, P.B, P.tiles_x, par, b, FROM_BELOW, txt, 0); ea = er + 128; em = ea + 128; }
                else { er = edge_ptr(P, par, tg + P.tiles_x, SIDE_TOP, 0); ea = er + 128; em = ea + 128; }
            } else if (side == SIDE_LEFT && has_l) {
                have = idx < P.tile_rows;
                er = edge_ptr(P, par, tg - 1, SIDE_RIGHT, 0); ea = er + 128; em = ea + 128;
            } else if (side == SIDE_RIGHT && has_r) {
                have = idx < P.tile_rows;
                er = edge_ptr(P, par, tg + 1, SIDE_LEFT, 0); ea = er + 128; em = ea + 128;
            }
            float hv = 0.0f;
            if (have) {
                const float rn = __ldcg(er + idx) - a * __ldcg(ea + idx);
                hv = rn - beta * __ldcg(em + idx);
            }
            if (side == SIDE_TOP) reinterpret_cast<float*>(sm_e)[idx] = hv;
            else if (side == SIDE_BOTTOM) reinterpret_cast<float*>(sm_e + (2 * wu + 1) * 32)[idx] = hv;
            else if (side == SIDE_LEFT) { if (idx < ROWS) sm_lr[idx] = hv; }
            else if (idx < ROWS) sm_lr[ROWS + idx] = hv;
        }
        publish_local();
        __syncthreads();
        matvec();
    }
    // ---- epilogue ----
    if (w < wu) {
#pragma unroll
        for (int ky = 0; ky < 8; ++ky) {
            const float4 x4 = sm_x[(8 * w + ky) * 32 + lane];
            if (vec_ok && j0 + 3 < nx) {
                *reinterpret_cast<float4*>(P.out + base + (size_t)(i0 + ky) * nx + j0) = x4;
            } else {
                const float xv[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
                for (int kx = 0; kx < 4; ++kx)
                    if (valid & (1u << (ky * 4 + kx))) P.out[base + (size_t)(i0 + ky) * nx + j0 + kx] = xv[kx];
            }
        }
    }
    if (t == 0 && tid == 0) {
        if (P.iters) P.iters[b] = ok ? my_iters : -1;
        if (P.resid) P.resid[b] = my_res;
    }
    if (blockIdx.x == 0 && tid == 0) {
        // the next solve starts max_it + 2 sequence numbers later on every rank; tell the peers this rank has finished reading
        const u64 next = seq_base + (u64)P.max_it + 2;
        st_volatile_u64(epoch_ptr, next);
        if (P.nranks > 1) {
            __threadfence_system();
            for (int p = 0; p < P.nranks; ++p)
                if (p != P.rank) st_volatile_u64(reinterpret_cast<u64*>(P.peer[p] + COMM_DONE) + P.rank, next - 1);
        }
    }
}

}  // namespace pdec

// ================================================== host side ==================================================
using namespace pdec;

namespace {

struct Tiling { int warps, tiles_y, tiles_x, tile_rows; };

size_t tiled_smem(int warps) { return (size_t)(8 * warps * 32 + (2 * warps + 2) * 32) * sizeof(float4) + (size_t)2 * 8 * warps * sizeof(float); }

template <int WARPS>
int resident_ctas(int sms) {
    auto kernel = cg_tiled_kernel<WARPS>;
    const size_t smem = tiled_smem(WARPS);
    if (smem > 48 * 1024 && cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return 0;
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, 32 * WARPS, smem) != cudaSuccess) return 0;
    return per_sm * sms;
}

// Pick the tile height: ny_loc = tiles_y * tile_rows, tile_rows a multiple of 8 and <= 128, every CTA resident at once
// (cooperative launch); among the feasible ones minimise (waves per SM) x (measured cost of one tile iteration).
bool choose_tiling(int B, int ny_loc, int nx, Tiling* out) {
    if (ny_loc < 8 || (ny_loc & 7) != 0 || nx < 2) return false;
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return false;
    const int tiles_x = (nx + 127) / 128;
    const int cap[3] = {resident_ctas<4>(sms), resident_ctas<8>(sms), resident_ctas<16>(sms)};
    double best = 1e30;
    bool found = false;
    for (int tiles_y = 1; tiles_y <= ny_loc / 8; ++tiles_y) {
        if (ny_loc % tiles_y) continue;
        const int rows = ny_loc / tiles_y;
        if ((rows & 7) != 0 || rows > 128) continue;
        const int wi = rows <= 32 ? 0 : (rows <= 64 ? 1 : 2);
        const long long ctas = (long long)B * tiles_y * tiles_x;
        if (cap[wi] <= 0 || ctas > cap[wi]) continue;
        const double waves = (double)((ctas + sms - 1) / sms);
        const double c = waves * (0.4 + 1.5 * rows / 128.0);      // us per iteration of one tile: 0.8 / 1.15 / 1.9 at 32 / 64 / 128 rows
        if (c < best) { best = c; found = true; out->warps = wi == 0 ? 4 : (wi == 1 ? 8 : 16); out->tiles_y = tiles_y; out->tiles_x = tiles_x; out->tile_rows = rows; }
    }
    return found;
}

size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

}  // namespace

// bytes of the peer-visible communication buffer of one rank
extern "C" size_t pdec_cg_slab_comm_bytes(int B, int nx) {
    const size_t b = B > 0 ? B : 1, tiles_x = nx > 0 ? (nx + 127) / 128 : 1;
    return COMM_MAIL + ((size_t)2 * PDEC_SLAB_MAX_RANKS * b * 8 + (size_t)2 * b * 2 * tiles_x * 3 * 128) * sizeof(float);
}

// local (not peer-visible) scratch: sync words, a single-GPU stand-in for the comm header, totals, partials, edges
extern "C" size_t pdec_cg_slab_workspace_bytes(int B, int ny_local, int nx) {
    const size_t b = B > 0 ? B : 1;
    size_t bt = b * (size_t)((ny_local + 7) / 8) * ((nx + 127) / 128);            // upper bound: 8-row tiles ...
    if (bt > 1024) bt = 1024;                                                       // ... all resident at once
    return 256 + COMM_MAIL + align256(2 * b * 8 * sizeof(float)) + align256(2 * bt * 8 * sizeof(float)) +
           align256(2 * bt * 4 * 3 * 128 * sizeof(float));
}

extern "C" int pdec_cg_slab_comm_init(void* comm, int B, int nx, void* stream) {
    if (!comm) return PDEC_ERR_INVALID_ARGUMENT;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(comm, 0, pdec_cg_slab_comm_bytes(B, nx), st) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
    const unsigned long long one = 1ull;          // first sequence number (flags start at 0)
    if (cudaMemcpyAsync((char*)comm + COMM_EPOCH, &one, sizeof(one), cudaMemcpyHostToDevice, st) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
    return PDEC_OK;
}

extern "C" int pdec_cg_poisson_2d_slab(const float* rhs, float* p, int* iterations, float* residual, int B, int ny_local, int nx,
                                       int ny_global, int row0, int closed, float accuracy, int max_iterations,
                                       const pdec_slab_comm* comm, void* workspace, size_t workspace_bytes, void* stream) {
    if (B < 0 || ny_local < 1 || nx < 2 || ny_global < ny_local || row0 < 0 || row0 + ny_local > ny_global || max_iterations < 0)
        return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!rhs || !p || !workspace) return PDEC_ERR_INVALID_ARGUMENT;
    const int nranks = comm ? comm->nranks : 1, rank = comm ? comm->rank : 0;
    if (nranks < 1 || nranks > PDEC_SLAB_MAX_RANKS || rank < 0 || rank >= nranks) return PDEC_ERR_INVALID_ARGUMENT;
    if (nranks == 1 && (row0 != 0 || ny_local != ny_global)) return PDEC_ERR_INVALID_ARGUMENT;
    if (nranks > 1 && (!comm->local || comm->peers[rank] != comm->local)) return PDEC_ERR_INVALID_ARGUMENT;
    if (workspace_bytes < pdec_cg_slab_workspace_bytes(B, ny_local, nx)) return PDEC_ERR_WORKSPACE_TOO_SMALL;
    Tiling tl;
    if (!choose_tiling(B, ny_local, nx, &tl) || (size_t)B * tl.tiles_y * tl.tiles_x > 1024) return PDEC_ERR_UNSUPPORTED_SIZE;
    cudaStream_t st = (cudaStream_t)stream;
    CgTiled P;
    P.rhs = rhs; P.out = p; P.iters = iterations; P.resid = residual;
    P.B = B; P.ny_loc = ny_local; P.nx = nx; P.ny_glob = ny_global; P.row0 = row0; P.closed = closed ? 1 : 0; P.max_it = max_iterations;
    P.accuracy = accuracy;
    P.tiles_y = tl.tiles_y; P.tiles_x = tl.tiles_x; P.tile_rows = tl.tile_rows;
    P.rank = rank; P.nranks = nranks;
    char* w = reinterpret_cast<char*>(workspace);
    P.cnt = reinterpret_cast<unsigned*>(w); P.release = P.cnt + 1; P.abort_local = P.cnt + 2;
    char* fake_comm = w + 256;
    const size_t tiles = (size_t)tl.tiles_y * tl.tiles_x;
    char* q = w + 256 + COMM_MAIL;
    P.totals = reinterpret_cast<float*>(q); q += align256((size_t)2 * B * 8 * sizeof(float));
    P.partials = reinterpret_cast<float*>(q); q += align256((size_t)2 * B * tiles * 8 * sizeof(float));
    P.edges = reinterpret_cast<float*>(q);
    for (int i = 0; i < PDEC_SLAB_MAX_RANKS; ++i) P.peer[i] = nullptr;
    if (nranks == 1) {
        // single GPU: the sync words AND the stand-in comm header are reset by this launch (nobody else writes them)
        if (cudaMemsetAsync(w, 0, 256 + COMM_MAIL, st) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
        P.comm = fake_comm; P.peer[0] = fake_comm;
    } else {
        if (cudaMemsetAsync(w, 0, 256, st) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
        P.comm = reinterpret_cast<char*>(comm->local);
        for (int i = 0; i < nranks; ++i) {
            if (!comm->peers[i]) return PDEC_ERR_INVALID_ARGUMENT;
            P.peer[i] = reinterpret_cast<char*>(comm->peers[i]);
        }
    }
    int dev = 0, khz = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    P.spin_budget = (long long)(khz > 0 ? khz : 1500000) * 1000LL * (nranks > 1 ? 20 : 5);     // 20 s across GPUs, 5 s inside one
    const unsigned grid = (unsigned)((size_t)B * tiles);
    void* args[] = {&P};
    const void* fn = tl.warps == 4 ? (const void*)cg_tiled_kernel<4> : (tl.warps == 8 ? (const void*)cg_tiled_kernel<8> : (const void*)cg_tiled_kernel<16>);
    if (cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(32 * tl.warps), args, tiled_smem(tl.warps), st) != cudaSuccess) {
        cudaGetLastError();
        return PDEC_ERR_CUDA_LAUNCH;
    }
    return PDEC_OK;
}

// ---- peer-visible allocations: cudaMalloc + CUDA IPC handles (one process per GPU, torchrun) ----
extern "C" int pdec_comm_alloc(size_t bytes, void** ptr, unsigned char* handle64) {
    if (!ptr || !handle64 || bytes == 0) return PDEC_ERR_INVALID_ARGUMENT;
    void* d = nullptr;
    if (cudaMalloc(&d, bytes) != cudaSuccess) { cudaGetLastError(); return PDEC_ERR_CUDA_LAUNCH; }
    cudaIpcMemHandle_t h;
    if (cudaIpcGetMemHandle(&h, d) != cudaSuccess) { cudaGetLastError(); cudaFree(d); return PDEC_ERR_CUDA_LAUNCH; }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    memcpy(handle64, &h, 64);
    *ptr = d;
    return PDEC_OK;
}
extern "C" int pdec_comm_open(const unsigned char* handle64, void** ptr) {
    if (!ptr || !handle64) return PDEC_ERR_INVALID_ARGUMENT;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    if (cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); return PDEC_ERR_CUDA_LAUNCH; }
    return PDEC_OK;
}
extern "C" int pdec_comm_close(void* ptr) {
    if (ptr && cudaIpcCloseMemHandle(ptr) != cudaSuccess) { cudaGetLastError(); return PDEC_ERR_CUDA_LAUNCH; }
    return PDEC_OK;
}
extern "C" int pdec_comm_free(void* ptr) {
    if (ptr && cudaFree(ptr) != cudaSuccess) { cudaGetLastError(); return PDEC_ERR_CUDA_LAUNCH; }
    return PDEC_OK;
}
