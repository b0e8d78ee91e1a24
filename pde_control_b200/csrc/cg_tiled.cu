// cg_tiled.cu -- conjugate-gradient Poisson solve on a GRID OF TILES that stays on chip, on one GPU or on a row-slab
// decomposition over several GPUs (BASELINE config 5).
//
// Reference: poisson_solve / SparseCG / conjugate_gradient (PF/physics/pressuresolver/solver_api.py:132-158,
// sparse.py:73-221, PF/math/blas.py:5-64).  The reference streams every CG vector through memory every iteration (and its
// native twin, PF/tf/cuda/src/pressure_solve_op.cu.cc, launches one kernel per vector operation).  Here:
//   * one persistent CTA per tile of (8*WARPS rows x 128 columns); r, A m and m of the tile live in REGISTERS, x in shared
//     memory (the register/shared-memory layout of cg_block_kernel, cg.cu) -- HBM is touched once (rhs in, p out);
//   * ONE exchange per iteration.  With q = A m, all dot products are taken before the update
//         a = (m.r)/(m.q),   beta = ((r.q) - a (q.q))/(m.q)          [ = (r'.q)/(m.q) with r' = r - a q ]
//     and every tile publishes the edge rows/columns of (r, q, m) together with its five partial sums.  Once a tile has
//     everybody's partials it knows a and beta and forms its neighbours' new search direction on the halo itself:
//     m' = (r - a q) - beta m.  The exchange therefore carries the reduction AND the halo;
//   * the exchange needs no fence, no atomic and no barrier: every published float travels in an 8-byte word together with the
//     32-bit sequence number of the iteration (single-copy atomic), and a consumer polls the words it needs until they carry
//     the current number.  Measured on this pool (scripts/diag/p2p_latency.cu): a flag hop over NVLink costs 1.15 us, the same
//     hop behind a system-scope fence 4.4 us -- so the data IS the flag;
//   * the reduction is a two-hop tree, because 128 tiles all polling 128 partials hot-spot the L2 (measured: 5 us): the tiles of
//     a batch element post their partials, the element's first tile (its leader) sums them in tile order and posts the result;
//   * on several GPUs (row slabs) edge tiles store their rows straight into the neighbour GPU's inbox and each leader stores its
//     rank's sums into EVERY rank's table over NVLink peer memory; every tile then adds the ranks' sums in rank order.  No NCCL
//     call, no host round trip.
// Everybody needs everybody's sums before it can publish again, which bounds the skew to one iteration: two parity buffers
// suffice, and sequence numbers simply continue across solves (a device-resident epoch).
// Sums are formed in a fixed order (rank, then tile) by every CTA itself, so every CTA on every GPU computes bit-identical
// a and beta.  Polls are bounded (clock64 budget); on a time-out every participant leaves and the iteration count reads -1.
#include "pdec_common.cuh"
#include "cg_reduce.cuh"
#include <string.h>

namespace pdec {

typedef unsigned long long u64;

enum { SIDE_TOP = 0, SIDE_BOTTOM = 1, SIDE_LEFT = 2, SIDE_RIGHT = 3 };
enum { FROM_ABOVE = 0, FROM_BELOW = 1 };
#define PDEC_SLAB_MAX_RANKS 8
#define PDEC_SLAB_MAX_CTAS 1024
// peer-visible communication buffer (bytes): [abort u32 @128] [epoch u64 @256, local use] pad to 512
//   | rank sums u64 [2 parities][8 ranks][PDEC_SLAB_MAX_BATCH][8 words, 5 used] | inbox u64 [2][B][2][tiles_x][3][128]
#define COMM_ABORT 128
#define COMM_EPOCH 256
#define COMM_PARTIALS 512
#define PDEC_SLAB_MAX_BATCH 64
#define COMM_INBOX (COMM_PARTIALS + (size_t)2 * PDEC_SLAB_MAX_RANKS * PDEC_SLAB_MAX_BATCH * 8 * 8)

struct CgTiled {
    const float* rhs; float* out; int* iters; float* resid;
    int B, ny_loc, nx, ny_glob, row0, closed, max_it;
    float accuracy;
    int tiles_y, tiles_x, tile_rows;
    int rank, nranks;
    unsigned* abort_local;      // local scratch
    u64* edges;                 // local scratch [2][B*tiles][4 sides][3 vecs][128] tagged words
    u64* tile_sums;             // local scratch [2][B*tiles][8] tagged words: the tiles' partial sums, read by the element's leader tile
    char* comm; char* peer[PDEC_SLAB_MAX_RANKS];      // this rank's communication buffer and everybody's, as mapped here
    long long spin_budget;      // clock64 ticks a poll may take before the solve is abandoned
};

__device__ __forceinline__ unsigned ld_volatile_u32(const unsigned* p) { return *reinterpret_cast<const volatile unsigned*>(p); }
__device__ __forceinline__ u64 ld_volatile_u64(const u64* p) { return *reinterpret_cast<const volatile u64*>(p); }
__device__ __forceinline__ void st_volatile_u32(unsigned* p, unsigned v) { *reinterpret_cast<volatile unsigned*>(p) = v; }
__device__ __forceinline__ void st_volatile_u64(u64* p, u64 v) { *reinterpret_cast<volatile u64*>(p) = v; }

// a float and the sequence number it belongs to, in one single-copy-atomic word
__device__ __forceinline__ u64 tagged(float v, unsigned tag) { return ((u64)tag << 32) | (u64)__float_as_uint(v); }
// strong (volatile) vector stores: a weak store may sit in the SM's write path until something flushes it, and here nothing does
__device__ __forceinline__ void st_volatile_v2(u64* p, u64 a, u64 b) {
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(a), "l"(b) : "memory");
}
__device__ __forceinline__ void st_tagged4(u64* p, float a, float b, float c, float d, unsigned tag) {       // p 16-byte aligned
    st_volatile_v2(p, tagged(a, tag), tagged(b, tag));
    st_volatile_v2(p + 2, tagged(c, tag), tagged(d, tag));
}

__device__ __forceinline__ u64* comm_rank_sums(char* comm, int par, int src_rank, int b) {
    return reinterpret_cast<u64*>(comm + COMM_PARTIALS) + (((size_t)par * PDEC_SLAB_MAX_RANKS + src_rank) * PDEC_SLAB_MAX_BATCH + b) * 8;
}
__device__ __forceinline__ u64* comm_inbox(char* comm, int B, int tiles_x, int par, int b, int dir, int tx, int vec) {
    u64* base = reinterpret_cast<u64*>(comm + COMM_INBOX);
    return base + (((((size_t)par * B + b) * 2 + dir) * tiles_x + tx) * 3 + vec) * 128;
}
__device__ __forceinline__ u64* edge_ptr(const CgTiled& P, int par, int tile_global, int side, int vec) {
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

__device__ __forceinline__ ulonglong2 ld_volatile_v2(const u64* p) {       // p 16-byte aligned; each half is single-copy atomic
    ulonglong2 v;
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ bool tag_is(u64 w, unsigned tag) { return (unsigned)(w >> 32) == tag; }
__device__ __forceinline__ float word_value(u64 w) { return __uint_as_float((unsigned)w); }

// Bounded polling: all the words a thread waits for are (re)loaded back to back -- one memory round trip per attempt, not one
// per word.  (A __nanosleep between attempts was measured slower: 7.4 vs 5.7 us per iteration at 1024^2.)
struct PollBudget {
    long long t0; unsigned spins;
    __device__ __forceinline__ PollBudget() : t0(0), spins(0u) {}
    __device__ __forceinline__ bool again(const CgTiled& P) {              // false: give up
        if (spins == 0u) t0 = clock64();
        if ((++spins & 31u) == 0u && (aborted(P) || clock64() - t0 > P.spin_budget)) { raise_abort(P); return false; }
        return true;
    }
};
// the five partial sums of one tile (words 0-3 only when `sums`), tag-checked
__device__ __forceinline__ bool poll_partials(const CgTiled& P, const u64* pp, unsigned tag, bool sums, float (&v)[5]) {
    PollBudget bud;
    for (;;) {
        ulonglong2 a = make_ulonglong2(0ull, 0ull), c = make_ulonglong2(0ull, 0ull);
        if (sums) { a = ld_volatile_v2(pp); c = ld_volatile_v2(pp + 2); }
        const u64 e = ld_volatile_u64(pp + 4);
        if (tag_is(e, tag) && (!sums || (tag_is(a.x, tag) && tag_is(a.y, tag) && tag_is(c.x, tag) && tag_is(c.y, tag)))) {
            v[0] = word_value(a.x); v[1] = word_value(a.y); v[2] = word_value(c.x); v[3] = word_value(c.y); v[4] = word_value(e);
            return true;
        }
        if (!bud.again(P)) return false;
    }
}

#ifdef PDEC_TILED_PROFILE
__device__ long long g_tiled_prof[8];      // CTA 0, thread 0: cycles in [dots+reduce, publish, collect, reduce totals, update+halo+matvec]
#define PROF_STAMP(i) do { const long long now_ = clock64(); prof_acc[i] += now_ - prof_t; prof_t = now_; } while (0)
#else
#define PROF_STAMP(i) do { } while (0)
#endif

template <int WARPS>
__global__ void __launch_bounds__(32 * WARPS, 1) cg_tiled_kernel(CgTiled P) {
    constexpr int THREADS = 32 * WARPS, ROWS = 8 * WARPS, NH = 512 / THREADS;
    extern __shared__ __align__(16) float smem_dyn[];
    __shared__ __align__(16) float red5[2][80], red5l[2][80];
    __shared__ int s_done, s_bad;
    __shared__ float s_tot[PDEC_SLAB_MAX_RANKS * PDEC_SLAB_MAX_BATCH][5];
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
    // P.closed = 4-bit mask of solid sides of the GLOBAL domain (pdec_closed_sides)
#pragma unroll
    for (int kx = 0; kx < 4; ++kx) dcol[kx] = -4.0f + ((j0 + kx == 0 && (P.closed & 4)) ? 1.0f : 0.0f) + ((j0 + kx == nx - 1 && (P.closed & 8)) ? 1.0f : 0.0f);
#pragma unroll
    for (int ky = 0; ky < 8; ++ky) {
        const int gi = P.row0 + i0 + ky;
        drow[ky] = ((gi == 0 && (P.closed & 1)) ? 1.0f : 0.0f) + ((gi == P.ny_glob - 1 && (P.closed & 2)) ? 1.0f : 0.0f);
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
    if (tid == 0) s_bad = 0;
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
    // every address of the exchange, once, for parity 0 (64-bit address arithmetic inside the iteration cost more than the
    // stores themselves); parity 1 lies one fixed stride further
    const unsigned edge_stride = (unsigned)ctas * 4u * 3u * 128u, inbox_stride = (unsigned)P.B * 2u * (unsigned)P.tiles_x * 3u * 128u;
    const unsigned sums_stride = (unsigned)ctas * 8u;
    u64* const own_edges = edge_ptr(P, 0, tg, 0, 0);
    u64* const own_sums = P.tile_sums + (size_t)tg * 8;
    u64* const up_target = !has_up ? nullptr
                         : (up_remote ? comm_inbox(P.peer[P.rank - 1], P.B, P.tiles_x, 0, b, FROM_BELOW, txt, 0) : own_edges + (SIDE_TOP * 3) * 128);
    u64* const dn_target = !has_dn ? nullptr
                         : (dn_remote ? comm_inbox(P.peer[P.rank + 1], P.B, P.tiles_x, 0, b, FROM_ABOVE, txt, 0) : own_edges + (SIDE_BOTTOM * 3) * 128);
    const unsigned up_stride = up_remote ? inbox_stride : edge_stride, dn_stride = dn_remote ? inbox_stride : edge_stride;
    const u64* halo_src[NH];
    unsigned halo_stride[NH];
#pragma unroll
    for (int j = 0; j < NH; ++j) {
        const int q = tid + j * THREADS, side = q >> 7, idx = q & 127;
        const u64* er = nullptr;
        halo_stride[j] = edge_stride;
        if (side == SIDE_TOP && has_up) {
            er = up_remote ? comm_inbox(P.comm, P.B, P.tiles_x, 0, b, FROM_ABOVE, txt, 0) : edge_ptr(P, 0, tg - P.tiles_x, SIDE_BOTTOM, 0);
            if (up_remote) halo_stride[j] = inbox_stride;
        } else if (side == SIDE_BOTTOM && has_dn) {
            er = dn_remote ? comm_inbox(P.comm, P.B, P.tiles_x, 0, b, FROM_BELOW, txt, 0) : edge_ptr(P, 0, tg + P.tiles_x, SIDE_TOP, 0);
            if (dn_remote) halo_stride[j] = inbox_stride;
        } else if (side == SIDE_LEFT && has_l && idx < P.tile_rows) {
            er = edge_ptr(P, 0, tg - 1, SIDE_RIGHT, 0);
        } else if (side == SIDE_RIGHT && has_r && idx < P.tile_rows) {
            er = edge_ptr(P, 0, tg + 1, SIDE_LEFT, 0);
        }
        halo_src[j] = er ? er + idx : nullptr;
    }
    const bool col_lane = w < wu && ((lane == 0 && has_l) || (lane == 31 && has_r));
    const bool left_lane = lane == 0;
    bool done_me = false, ok = true;
    int my_iters = 0;
    float my_res = 0.0f;
    unsigned k = 0;
#ifdef PDEC_TILED_PROFILE
    long long prof_t = clock64(), prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#endif
    for (;; ++k) {
        const u64 seq = seq_base + k;
        const int par = (int)(seq & 1ull);
        const unsigned tag = (unsigned)(seq + 1ull);               // 0 is "never written"
        // ---- 1. the five partial reductions of this tile ----
        constexpr int ACC = WARPS >= 16 ? 1 : 4;               // independent accumulation chains (the 16-warp tile has no registers to spare)
        float p1[ACC], p2[ACC], p4[ACC], p5[ACC], pm[ACC];
#pragma unroll
        for (int i = 0; i < ACC; ++i) { p1[i] = 0.0f; p2[i] = 0.0f; p4[i] = 0.0f; p5[i] = 0.0f; pm[i] = 0.0f; }
#pragma unroll
        for (int c = 0; c < 32; ++c) {
            p1[c % ACC] += m[c] * am[c];
            p2[c % ACC] += m[c] * r[c];
            p4[c % ACC] += r[c] * am[c];
            p5[c % ACC] += am[c] * am[c];
            pm[c % ACC] = fmaxf(pm[c % ACC], fabsf(r[c]));
        }
        float s1 = p1[0], s2 = p2[0], s4 = p4[0], s5 = p5[0], mx = pm[0];
        if (ACC == 4) {
            s1 = (p1[0] + p1[1 % ACC]) + (p1[2 % ACC] + p1[3 % ACC]); s2 = (p2[0] + p2[1 % ACC]) + (p2[2 % ACC] + p2[3 % ACC]);
            s4 = (p4[0] + p4[1 % ACC]) + (p4[2 % ACC] + p4[3 % ACC]); s5 = (p5[0] + p5[1 % ACC]) + (p5[2 % ACC] + p5[3 % ACC]);
            mx = fmaxf(fmaxf(pm[0], pm[1 % ACC]), fmaxf(pm[2 % ACC], pm[3 % ACC]));
        }
        block_reduce5<WARPS>(s1, s2, s4, s5, mx, red5[par]);
        PROF_STAMP(0);
        // ---- 2. publish the tile's edges of (r, A m, m) and its partials ----
        if (!done_me) {
            u64* const eo = own_edges + (size_t)par * edge_stride;
            u64* const upt = up_target ? up_target + (size_t)par * up_stride : nullptr;
            u64* const dnt = dn_target ? dn_target + (size_t)par * dn_stride : nullptr;
            auto publish_edges = [&](const float (&v)[32], const int vec) {
                if (w == 0 && upt) st_tagged4(upt + vec * 128 + 4 * lane, v[0], v[1], v[2], v[3], tag);
                if (w == wu - 1 && dnt) st_tagged4(dnt + vec * 128 + 4 * lane, v[28], v[29], v[30], v[31], tag);
                if (col_lane) {                                   // lane 0: first column -> left edge; lane 31: last column -> right edge
                    u64* e = eo + ((left_lane ? SIDE_LEFT : SIDE_RIGHT) * 3 + vec) * 128 + 8 * w;
                    st_tagged4(e, left_lane ? v[0] : v[3], left_lane ? v[4] : v[7], left_lane ? v[8] : v[11], left_lane ? v[12] : v[15], tag);
                    st_tagged4(e + 4, left_lane ? v[16] : v[19], left_lane ? v[20] : v[23], left_lane ? v[24] : v[27], left_lane ? v[28] : v[31], tag);
                }
            };
            publish_edges(r, 0);
            publish_edges(am, 1);
            publish_edges(m, 2);
        }
        PROF_STAMP(5);
        if (tid == 0) {                                         // this tile's partials, for the leader tile of its batch element
            u64* pp = own_sums + (size_t)par * sums_stride;
            st_tagged4(pp, s1, s2, s4, s5, tag);
            st_volatile_u64(pp + 4, tagged(mx, tag));
        }
        PROF_STAMP(1);
        // ---- 3. collect: everybody's partials (the implicit barrier) and the neighbours' edges ----
        const int it = k == 0 ? 0 : (int)k - 1;                // real updates applied so far
        const bool budget = it < P.max_it;
        if (tid == 0) s_done = 1;
        __syncthreads();
        bool good = true;
        // (a) some element still above the accuracy  <=>  some tile partial of max|r| still above it
        // (b) the five sums of THIS element, every CTA by itself, in a fixed order (rank, tile): entry e = rank * tiles + tile
        float S1 = 0.0f, S2 = 0.0f, S4 = 0.0f, S5 = 0.0f, MX = 0.0f;
        bool any_above = false;
        // first look at the neighbours' published edges (needed in step 6): these loads are in flight while the partials are
        // collected; the tags are checked (and stragglers re-read) afterwards
        const u64* hp[NH];
        u64 hw[NH][3];
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            hp[j] = (done_me || !halo_src[j]) ? nullptr : halo_src[j] + (size_t)par * halo_stride[j];
            hw[j][0] = 0ull; hw[j][1] = 0ull; hw[j][2] = 0ull;
            if (hp[j]) { hw[j][0] = ld_volatile_u64(hp[j]); hw[j][1] = ld_volatile_u64(hp[j] + 128); hw[j][2] = ld_volatile_u64(hp[j] + 256); }
        }
        PROF_STAMP(6);
        if (t == 0) {
            // leader of element b on this rank: the tiles' partials in tile order -> this rank's sums -> every rank's table
            float L[5] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
            for (int e = tid; e < tiles && good; e += THREADS) {
                float v[5];
                good = poll_partials(P, P.tile_sums + ((size_t)par * ctas + b * tiles + e) * 8, tag, true, v);
                L[0] += v[0]; L[1] += v[1]; L[2] += v[2]; L[3] += v[3]; L[4] = fmaxf(L[4], v[4]);
            }
            block_reduce5<WARPS>(L[0], L[1], L[2], L[3], L[4], red5l[par]);
            if (tid < P.nranks) {
                u64* pp = comm_rank_sums(P.peer[tid], par, P.rank, b);
                st_tagged4(pp, L[0], L[1], L[2], L[3], tag);
                st_volatile_u64(pp + 4, tagged(L[4], tag));
            }
        }
        // every tile: the ranks' sums of every element (all elements for the stop test, its own for a and beta)
        for (int e = tid; e < P.nranks * P.B && good; e += THREADS) {
            const int rk = e / P.B, q = e - rk * P.B;
            float v[5];
            good = poll_partials(P, comm_rank_sums(P.comm, par, rk, q), tag, true, v);
            if (P.accuracy <= 0.0f || v[4] >= P.accuracy) any_above = true;
#pragma unroll
            for (int i = 0; i < 5; ++i) s_tot[e][i] = v[i];
        }
        if (budget && any_above) s_done = 0;                    // benign race: all writers store 0
        PROF_STAMP(7);
        float hr[NH], ha[NH], hm[NH];
#pragma unroll
        for (int j = 0; j < NH; ++j) {
            if (hp[j] && good) {
                PollBudget bud;
                while (!(tag_is(hw[j][0], tag) && tag_is(hw[j][1], tag) && tag_is(hw[j][2], tag))) {
                    if (!bud.again(P)) { good = false; break; }
                    hw[j][0] = ld_volatile_u64(hp[j]); hw[j][1] = ld_volatile_u64(hp[j] + 128); hw[j][2] = ld_volatile_u64(hp[j] + 256);
                }
            }
            hr[j] = word_value(hw[j][0]); ha[j] = word_value(hw[j][1]); hm[j] = word_value(hw[j][2]);
        }
        PROF_STAMP(2);
        if (!good) s_bad = 1;
        __syncthreads();                                         // publishes s_tot / s_done / s_bad
        for (int rk = 0; rk < P.nranks; ++rk) {                  // rank order: the same bits on every tile of every rank
            const float* v = s_tot[rk * P.B + b];
            S1 += v[0]; S2 += v[1]; S4 += v[2]; S5 += v[3]; MX = fmaxf(MX, v[4]);
        }
        PROF_STAMP(3);
        if (s_bad) { ok = false; break; }
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
        PROF_STAMP(4);
    }
#ifdef PDEC_TILED_PROFILE
    if (blockIdx.x == gridDim.x / 2 && tid == 0) for (int i = 0; i < 8; ++i) g_tiled_prof[i] = prof_acc[i];
#endif
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
    // the next solve on this communication buffer continues the sequence (identically on every rank: all left at barrier k)
    if (blockIdx.x == 0 && tid == 0 && ok) st_volatile_u64(epoch_ptr, seq_base + (u64)k + 1);
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
    return COMM_INBOX + (size_t)2 * b * 2 * tiles_x * 3 * 128 * sizeof(u64);
}

// local (not peer-visible) scratch: abort word, a single-GPU stand-in for the communication buffer, the tiles' edge words
extern "C" size_t pdec_cg_slab_workspace_bytes(int B, int ny_local, int nx) {
    const size_t b = B > 0 ? B : 1;
    size_t bt = b * (size_t)((ny_local + 7) / 8) * ((nx + 127) / 128);            // upper bound: 8-row tiles ...
    if (bt > PDEC_SLAB_MAX_CTAS) bt = PDEC_SLAB_MAX_CTAS;                                                     // ... all resident at once
    return 256 + COMM_INBOX + align256(2 * bt * 8 * sizeof(u64)) + align256(2 * bt * 4 * 3 * 128 * sizeof(u64));
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
    if (B > PDEC_SLAB_MAX_BATCH || !choose_tiling(B, ny_local, nx, &tl) || (size_t)B * tl.tiles_y * tl.tiles_x > PDEC_SLAB_MAX_CTAS)
        return PDEC_ERR_UNSUPPORTED_SIZE;
    cudaStream_t st = (cudaStream_t)stream;
    CgTiled P;
    P.rhs = rhs; P.out = p; P.iters = iterations; P.resid = residual;
    P.B = B; P.ny_loc = ny_local; P.nx = nx; P.ny_glob = ny_global; P.row0 = row0; P.closed = (int)pdec_closed_sides(closed); P.max_it = max_iterations;
    P.accuracy = accuracy;
    P.tiles_y = tl.tiles_y; P.tiles_x = tl.tiles_x; P.tile_rows = tl.tile_rows;
    P.rank = rank; P.nranks = nranks;
    char* w = reinterpret_cast<char*>(workspace);
    P.abort_local = reinterpret_cast<unsigned*>(w);
    char* fake_comm = w + 256;
    const size_t tiles = (size_t)tl.tiles_y * tl.tiles_x, ctas = (size_t)B * tiles;
    P.tile_sums = reinterpret_cast<u64*>(w + 256 + COMM_INBOX);
    P.edges = P.tile_sums + align256((size_t)2 * ctas * 8 * sizeof(u64)) / sizeof(u64);
    for (int i = 0; i < PDEC_SLAB_MAX_RANKS; ++i) P.peer[i] = nullptr;
    // the tiles' sums and edge words are local to this launch: clear them (scratch is shared with other operators, stale bits
    // must not look like a sequence number)
    if (cudaMemsetAsync(P.tile_sums, 0, align256((size_t)2 * ctas * 8 * sizeof(u64)) + (size_t)2 * ctas * 4 * 3 * 128 * sizeof(u64), st) != cudaSuccess)
        return PDEC_ERR_CUDA_LAUNCH;
    if (nranks == 1) {
        // single GPU: abort word, epoch and the partials table live in the scratch too and restart from zero with every launch
        if (cudaMemsetAsync(w, 0, 256 + COMM_INBOX, st) != cudaSuccess)
            return PDEC_ERR_CUDA_LAUNCH;
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
    const unsigned grid = (unsigned)ctas;
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

#ifdef PDEC_TILED_PROFILE
extern "C" int pdec_debug_tiled_profile(long long* out8) {
    return cudaMemcpyFromSymbol(out8, pdec::g_tiled_prof, 8 * sizeof(long long)) == cudaSuccess ? 0 : -3;
}
#endif
