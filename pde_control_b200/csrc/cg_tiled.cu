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
//     into the neighbour GPU's inbox, every tile stores its five partial sums into EVERY rank's copy of the partials table and
//     then adds 1 to every rank's arrival counter (a fire-and-forget reduction over NVLink); a tile proceeds when its local
//     counter shows that all tiles of all ranks have arrived.  No NCCL call, no host round trip, no designated reducer.
// The barrier is a monotonic counting barrier: sequence numbers continue across launches (a device-resident epoch), and a tile
// that leaves early (convergence) adds the arrivals it did not use, so the counter of every rank ends each solve at
// (epoch + max_iterations + 2) x tiles, whatever the iteration count was.
// Sums are formed in a fixed order (rank, then tile) by every CTA itself, so every CTA on every GPU computes bit-identical
// a and beta.  Spin loops are bounded (clock64 budget); on a time-out every participant leaves and the iteration count reads -1.
#include "pdec_common.cuh"
#include "cg_reduce.cuh"
#include <string.h>

namespace pdec {

typedef unsigned long long u64;

enum { SIDE_TOP = 0, SIDE_BOTTOM = 1, SIDE_LEFT = 2, SIDE_RIGHT = 3 };
enum { FROM_ABOVE = 0, FROM_BELOW = 1 };
#define PDEC_SLAB_MAX_RANKS 8
// layout of the peer-visible communication buffer (bytes): [arrival counter u64 @0] [this rank's own arrivals u64 @64, local
// use] [abort u32 @128] [epoch u64 @256, local use] pad to 512 | partials float [2 parities][8 ranks][PDEC_SLAB_MAX_CTAS][8] | inbox float [2][B][2][tiles_x][3][128]
#define COMM_COUNTER 0
#define COMM_LOCAL_ARRIVALS 64
#define COMM_ABORT 128
#define COMM_EPOCH 256
#define COMM_PARTIALS 512
#define PDEC_SLAB_MAX_CTAS 1024
#define COMM_INBOX (COMM_PARTIALS + 2 * PDEC_SLAB_MAX_RANKS * PDEC_SLAB_MAX_CTAS * 8 * 4)

struct CgTiled {
    const float* rhs; float* out; int* iters; float* resid;
    int B, ny_loc, nx, ny_glob, row0, closed, max_it;
    float accuracy;
    int tiles_y, tiles_x, tile_rows;
    int rank, nranks;
    // local workspace
    unsigned* abort_local;
    float* edges;       // [2][B*tiles][4][3][128]
    // communication buffers (this rank's and the peers' mappings of theirs)
    char* comm; char* peer[PDEC_SLAB_MAX_RANKS];
    long long spin_budget;      // clock64 ticks a wait may take before the solve is abandoned
};

__device__ __forceinline__ unsigned ld_volatile_u32(const unsigned* p) { return *reinterpret_cast<const volatile unsigned*>(p); }
__device__ __forceinline__ u64 ld_volatile_u64(const u64* p) { return *reinterpret_cast<const volatile u64*>(p); }
__device__ __forceinline__ void st_volatile_u32(unsigned* p, unsigned v) { *reinterpret_cast<volatile unsigned*>(p) = v; }
__device__ __forceinline__ void st_volatile_u64(u64* p, u64 v) { *reinterpret_cast<volatile u64*>(p) = v; }

__device__ __forceinline__ float* comm_partial(char* comm, int ctas, int par, int src_rank, int cta) {
    return reinterpret_cast<float*>(comm + COMM_PARTIALS) + (((size_t)par * PDEC_SLAB_MAX_RANKS + src_rank) * ctas + cta) * 8;
}
__device__ __forceinline__ float* comm_inbox(char* comm, int B, int tiles_x, int par, int b, int dir, int tx, int vec) {
    float* base = reinterpret_cast<float*>(comm + COMM_INBOX);
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

// Arrival at barrier `seq` (global sequence number): everything this CTA stored before (edges, inbox rows, partials) is made
// visible, then the arrival is counted.  One GPU: +1 on the counter.  Several GPUs: +1 on this rank's own arrival count; the CTA
// that completes it adds the whole rank's arrivals to EVERY rank's counter (one NVLink reduction per peer and iteration instead
// of one per tile).  Wait: until the local counter shows all CTAs of all ranks.  Returns false when the solve was abandoned.
__device__ __forceinline__ void count_arrivals(const CgTiled& P, u64 n, u64 rank_total_after) {
    if (P.nranks > 1) {
        __threadfence_system();
        const u64 old = atomicAdd(reinterpret_cast<u64*>(P.comm + COMM_LOCAL_ARRIVALS), n);
        if (old + n == rank_total_after) {                  // last tile of this rank for this barrier
            __threadfence_system();
            const u64 all = n * gridDim.x;
            for (int p = 0; p < P.nranks; ++p) atomicAdd_system(reinterpret_cast<u64*>(P.peer[p] + COMM_COUNTER), all);
        }
    } else {
        __threadfence();
        atomicAdd(reinterpret_cast<u64*>(P.comm + COMM_COUNTER), n);
    }
}

__device__ bool tiled_barrier(const CgTiled& P, u64 seq, unsigned* s_flag) {
    const int tid = threadIdx.x;
    __syncthreads();
    if (tid == 0) {
        count_arrivals(P, 1ull, (seq + 1) * gridDim.x);
        const u64 target = (seq + 1) * (u64)P.nranks * gridDim.x;
        const long long t0 = clock64();
        unsigned bad = 0u;
        while (ld_volatile_u64(reinterpret_cast<const u64*>(P.comm + COMM_COUNTER)) < target) {
            if (aborted(P) || clock64() - t0 > P.spin_budget) { raise_abort(P); bad = 1u; break; }
        }
        if (P.nranks > 1) __threadfence_system(); else __threadfence();
        s_flag[0] = bad | (aborted(P) ? 1u : 0u);
    }
    __syncthreads();
    return s_flag[0] == 0u;
}

template <int WARPS>
__global__ void __launch_bounds__(32 * WARPS, 1) cg_tiled_kernel(CgTiled P) {
    constexpr int THREADS = 32 * WARPS, ROWS = 8 * WARPS, NH = 512 / THREADS;
    extern __shared__ __align__(16) float smem_dyn[];
    __shared__ __align__(16) float red5[2][80], red5t[2][80];
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
    const int ctas = gridDim.x;                                  // tiles of this rank (all ranks run the same grid)
    // entry: nobody publishes before every rank has finished the previous solve on this communication buffer (its tiles have
    // added their unused arrivals, so the counter stands at epoch x tiles-of-all-ranks)
    if (tid == 0) {
        const u64 target = seq_base * (u64)P.nranks * gridDim.x;
        const long long t0 = clock64();
        while (ld_volatile_u64(reinterpret_cast<const u64*>(P.comm + COMM_COUNTER)) < target) {
            if (aborted(P) || clock64() - t0 > P.spin_budget) { raise_abort(P); break; }
        }
        if (P.nranks > 1) __threadfence_system(); else __threadfence();
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
        if (tid < P.nranks) {                                   // this tile's partials into every rank's copy of the table
            float* pp = comm_partial(P.peer[tid], ctas, par, P.rank, tg);
            *reinterpret_cast<float4*>(pp) = make_float4(s1, s2, s4, s5);
            pp[4] = mx;
        }
        // ---- 3. the one barrier of the iteration ----
        ok = tiled_barrier(P, seq_base + k, s_flag);
        if (!ok) break;
        // the neighbours' published edges (needed in step 6): issue the loads now, they overlap the reduction of the partials
        float hr[NH], ha[NH], hm[NH];
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            const int q = tid + j * THREADS, side = q >> 7, idx = q & 127;
            const float* er = nullptr;
            if (side == SIDE_TOP && has_up)
                er = up_remote ? comm_inbox(P.comm, P.B, P.tiles_x, par, b, FROM_ABOVE, txt, 0) : edge_ptr(P, par, tg - P.tiles_x, SIDE_BOTTOM, 0);
            else if (side == SIDE_BOTTOM && has_dn)
                er = dn_remote ? comm_inbox(P.comm, P.B, P.tiles_x, par, b, FROM_BELOW, txt, 0) : edge_ptr(P, par, tg + P.tiles_x, SIDE_TOP, 0);
            else if (side == SIDE_LEFT && has_l && idx < P.tile_rows)
                er = edge_ptr(P, par, tg - 1, SIDE_RIGHT, 0);
            else if (side == SIDE_RIGHT && has_r && idx < P.tile_rows)
                er = edge_ptr(P, par, tg + 1, SIDE_LEFT, 0);
            hr[j] = er ? __ldcg(er + idx) : 0.0f;
            ha[j] = er ? __ldcg(er + 128 + idx) : 0.0f;
            hm[j] = er ? __ldcg(er + 256 + idx) : 0.0f;
        }
        // ---- 4. stop tests (per batch element; everybody leaves when every element is done) ----
        const int it = k == 0 ? 0 : (int)k - 1;                // real updates applied so far
        const bool budget = it < P.max_it;
        if (tid == 0) s_done = 1;
        __syncthreads();
        // some element still above the accuracy  <=>  some tile partial of max|r| still above it
        if (budget && P.accuracy > 0.0f) {
            for (int rk = 0; rk < P.nranks; ++rk)
                for (int q = tid; q < ctas; q += THREADS)
                    if (__ldcg(comm_partial(P.comm, ctas, par, rk, q) + 4) >= P.accuracy) s_done = 0;   // benign race: all store 0
        } else if (budget && tid == 0) {
            s_done = 0;
        }
        // the five sums of THIS element, every CTA by itself, in a fixed order (rank, tile): entry e = rank * tiles + tile
        float S1 = 0.0f, S2 = 0.0f, S4 = 0.0f, S5 = 0.0f, MX = 0.0f;
        for (int e = tid; e < P.nranks * tiles; e += THREADS) {
            const int rk = e / tiles, tt = e - rk * tiles;
            const float* pp = comm_partial(P.comm, ctas, par, rk, b * tiles + tt);
            const float4 v = __ldcg(reinterpret_cast<const float4*>(pp));
            S1 += v.x; S2 += v.y; S4 += v.z; S5 += v.w;
            MX = fmaxf(MX, __ldcg(pp + 4));
        }
        block_reduce5<WARPS>(S1, S2, S4, S5, MX, red5t[par]);    // its barrier also publishes s_done
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
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            const int q = tid + j * THREADS, side = q >> 7, idx = q & 127;
            const float hv = (hr[j] - a * ha[j]) - beta * hm[j];
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
    // every solve consumes max_it + 2 sequence numbers on every rank: hand in the arrivals this tile did not use (after its
    // last read of peer-written data), so the next solve's entry test sees when everybody has left this one
    __syncthreads();
    if (tid == 0 && ok) {
        const u64 unused = (u64)P.max_it + 2 - ((u64)k + 1);          // the same on every tile: all leave at the same barrier
        if (unused > 0) count_arrivals(P, unused, (seq_base + (u64)P.max_it + 2) * gridDim.x);
        if (blockIdx.x == 0) st_volatile_u64(epoch_ptr, seq_base + (u64)P.max_it + 2);
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
    // resident-CTA capacities are a property of (kernel, device): queried once per device (benign race: same values)
    static int cached[64][3];
    static bool have[64];
    int cap[3];
    if (dev >= 0 && dev < 64 && have[dev]) {
        cap[0] = cached[dev][0]; cap[1] = cached[dev][1]; cap[2] = cached[dev][2];
    } else {
        cap[0] = resident_ctas<4>(sms); cap[1] = resident_ctas<8>(sms); cap[2] = resident_ctas<16>(sms);
        if (dev >= 0 && dev < 64) { cached[dev][0] = cap[0]; cached[dev][1] = cap[1]; cached[dev][2] = cap[2]; have[dev] = true; }
    }
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
    return COMM_INBOX + (size_t)2 * b * 2 * tiles_x * 3 * 128 * sizeof(float);
}

// local (not peer-visible) scratch: sync words, a single-GPU stand-in for the comm header, totals, partials, edges
extern "C" size_t pdec_cg_slab_workspace_bytes(int B, int ny_local, int nx) {
    const size_t b = B > 0 ? B : 1;
    size_t bt = b * (size_t)((ny_local + 7) / 8) * ((nx + 127) / 128);            // upper bound: 8-row tiles ...
    if (bt > PDEC_SLAB_MAX_CTAS) bt = PDEC_SLAB_MAX_CTAS;                                                     // ... all resident at once
    return 256 + COMM_INBOX + align256(2 * bt * 4 * 3 * 128 * sizeof(float));
}

extern "C" int pdec_cg_slab_comm_init(void* comm, int B, int nx, void* stream) {
    if (!comm) return PDEC_ERR_INVALID_ARGUMENT;
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(comm, 0, pdec_cg_slab_comm_bytes(B, nx), st) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
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
    if (!choose_tiling(B, ny_local, nx, &tl) || (size_t)B * tl.tiles_y * tl.tiles_x > PDEC_SLAB_MAX_CTAS) return PDEC_ERR_UNSUPPORTED_SIZE;
    cudaStream_t st = (cudaStream_t)stream;
    CgTiled P;
    P.rhs = rhs; P.out = p; P.iters = iterations; P.resid = residual;
    P.B = B; P.ny_loc = ny_local; P.nx = nx; P.ny_glob = ny_global; P.row0 = row0; P.closed = closed ? 1 : 0; P.max_it = max_iterations;
    P.accuracy = accuracy;
    P.tiles_y = tl.tiles_y; P.tiles_x = tl.tiles_x; P.tile_rows = tl.tile_rows;
    P.rank = rank; P.nranks = nranks;
    char* w = reinterpret_cast<char*>(workspace);
    P.abort_local = reinterpret_cast<unsigned*>(w);
    char* fake_comm = w + 256;
    const size_t tiles = (size_t)tl.tiles_y * tl.tiles_x;
    P.edges = reinterpret_cast<float*>(w + 256 + COMM_INBOX);
    for (int i = 0; i < PDEC_SLAB_MAX_RANKS; ++i) P.peer[i] = nullptr;
    if (nranks == 1) {
        // single GPU: counter, epoch and abort words live in the scratch and are reset by this launch (nobody else writes them)
        if (cudaMemsetAsync(w, 0, 256 + COMM_PARTIALS, st) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
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
