// poisson_direct.cu -- EXACT Poisson pressure solve by fast diagonalisation, one CTA per batch element, everything on chip.
//
// Replaces, as an opt-in solver next to the CG of cg.cu, the reference's direct solver class SparseSciPy
// (PF/physics/pressuresolver/sparse.py:14-37: scipy.sparse.linalg.spsolve on the matrix of sparse.py:114-221 -- the default solver on
// the reference's NumPy backend, solver_api.py:144-148, "the fastest single-grid solver", documentation/Pressure_Solvers.md:40).
//
// The matrix of a rectangular domain without obstacles is a Kronecker sum  A = T_y (x) I + I (x) T_x  of two 1-D operators
// (tridiagonal 1, d, 1 with d = -2; on a CLOSED side the end diagonal is -1: one in-domain neighbour fewer, sparse.py:151-221), and the
// 1-D operators have closed-form orthonormal eigenvectors (sine basis for OPEN ends, cosine basis for CLOSED ends).  With
// T = E L E^T per axis:       p = E_y [ (E_y^T rhs E_x) / (l_y(i) + l_x(j)) ] E_x^T
// -- four 128 x 128 x 128 fp32 matrix products per element (16.8 MFLOP, against ~300 CG iterations of 0.2 MFLOP each with two block
// barriers per iteration), no iteration count, no accuracy knob, error at fp32 rounding level.  On a CLOSED domain A is singular
// (constants); the zero eigenvalue's reciprocal is 0: the mean-free solution CG converges to from a zero guess.
//
// Kernel: 512 threads, thread (warp w, lane tx) owns the 8 x 4 output block rows 8w.., columns 4tx.. of every product (the layout of
// cg_block_kernel).  Each product is a LEFT multiplication C = M B with B in shared memory (row-major, LDS.128 per k, conflict free)
// and M staged k-major into a two-slot shared-memory ring by cp.async, 32 rows at a time (two broadcast LDS.128 per k); right
// multiplications are left multiplications of the transpose, so stages 1 and 3 store their result transposed.  6 shared-memory
// wavefronts per 32 FFMA per warp and k: FFMA-bound.
#include "pdec_common.cuh"

namespace pdec {

constexpr int DS_N = 128;                 // tile size: grids up to 128 x 128, zero padded
constexpr int DS_PITCH = 132;             // shared row pitch (floats): rows stay 16-byte aligned, B-row reads are conflict free
constexpr int DS_MAT = DS_N * DS_N;       // floats per spectral matrix

// spectral block layout (floats): [0] Ey  (Ey[k][m], columns = eigenvectors of T_y)   [1] Ex   [2] Ex^T   [3] Ey^T
//                                 [4] inv (inv[j][i] = 1 / (l_x(j) + l_y(i)), 0 for padding and for the singular mode)

// 1-D eigenpairs of tridiag(1, d, 1), d = -2 except -1 at an end whose side is SOLID (one accessible neighbour fewer); an OPEN end sees
// a zero outside.  ends = (lower solid) | (upper solid) << 1.  Orthonormal closed forms: open-open: sine basis; solid-solid: cosine basis
// (DCT-II); mixed: quarter-wave sines anchored at the open end.
__device__ __forceinline__ double eig_vec(int i, int k, int n, int ends) {
    const double pi = 3.14159265358979323846;
    if (ends == 3) {
        const double c = k == 0 ? sqrt(1.0 / n) : sqrt(2.0 / n);
        return c * cos(pi * (i + 0.5) * k / n);
    }
    if (ends == 0) return sqrt(2.0 / (n + 1)) * sin(pi * (i + 1.0) * (k + 1.0) / (n + 1));
    const double theta = pi * (2.0 * k + 1.0) / (2.0 * n + 1.0), c = sqrt(4.0 / (2.0 * n + 1.0));
    return c * sin((ends == 2 ? (i + 1.0) : (double)(n - i)) * theta);        // ends == 2: open below, solid above; ends == 1: mirrored
}
__device__ __forceinline__ double eig_val(int k, int n, int ends) {
    const double pi = 3.14159265358979323846;
    if (ends == 3) return -2.0 + 2.0 * cos(pi * k / n);
    if (ends == 0) return -2.0 + 2.0 * cos(pi * (k + 1.0) / (n + 1));
    return -2.0 + 2.0 * cos(pi * (2.0 * k + 1.0) / (2.0 * n + 1.0));
}

__global__ void __launch_bounds__(256) poisson_direct_setup_kernel(float* __restrict__ spec, int ny, int nx, int closed) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= DS_MAT) return;
    const int a = idx / DS_N, b = idx % DS_N;
    // E[k][m]: row = cell index k, column = eigenvector m
    const int ends_y = closed & 3, ends_x = (closed >> 2) & 3;          // `closed` = 4-bit side mask (pdec_closed_sides)
    const float ey = (a < ny && b < ny) ? (float)eig_vec(a, b, ny, ends_y) : 0.0f;
    const float ex = (a < nx && b < nx) ? (float)eig_vec(a, b, nx, ends_x) : 0.0f;
    spec[0 * DS_MAT + idx] = ey;
    spec[1 * DS_MAT + idx] = ex;
    spec[2 * DS_MAT + b * DS_N + a] = ex;
    spec[3 * DS_MAT + b * DS_N + a] = ey;
    float inv = 0.0f;
    if (a < nx && b < ny) {              // inv[j = a][i = b]
        const double l = eig_val(a, nx, ends_x) + eig_val(b, ny, ends_y);
        // the one singular mode (constants on a closed domain) -> 0; every other eigenvalue sum is <= -2 (1 - cos(pi / 128)) ~ -6e-4
        inv = (fabs(l) < 1e-9) ? 0.0f : (float)(1.0 / l);
    }
    spec[4 * DS_MAT + idx] = inv;
}

constexpr int DS_CHUNK = 32;              // rows of M per staged chunk (16 KB); 4 chunks per product, 16 per solve
constexpr int DS_CHUNKS = 4 * (DS_N / DS_CHUNK);

// chunk g of the solve = rows [c * 32, c * 32 + 32) of spectral matrix `stage` (g = stage * 4 + c), contiguous in global memory:
// every thread copies two 16-byte pieces into the ring slot g & 1 (cp.async, L2 -> shared without registers)
__device__ __forceinline__ void stage_chunk(const float* __restrict__ spec, float* ring, int g) {
    if (g < DS_CHUNKS) {
        const float* src = spec + (size_t)(g / (DS_N / DS_CHUNK)) * DS_MAT + (size_t)(g % (DS_N / DS_CHUNK)) * DS_CHUNK * DS_N;
        float* dst = ring + (g & 1) * DS_CHUNK * DS_N;
        const int q = threadIdx.x * 8;
        cp_async16(dst + q, src + q);
        cp_async16(dst + q + 4, src + q + 4);
    }
    cp_async_commit();                      // an empty group at the tail keeps the wait count uniform
}

// C(8 x 4 block of this thread) = sum_k M[k][8w .. 8w+7] * Bs[k][4tx .. 4tx+3], M streamed through the two-slot ring:
// while chunk g is multiplied, chunk g + 1 is in flight (ncu on the first version, M read with LDG: 4.1 long-scoreboard stalls per
// issue, 49 % issue slots used -- the broadcast loads missed L1 two times in three).  g0 = first chunk of this product.
__device__ __forceinline__ void left_multiply(const float* __restrict__ spec, const float* ring, const float* Bs, int K, int g0, int w, int tx,
                                              float (&acc)[8][4]) {
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[r][c] = 0.0f;
    const float4* b4 = reinterpret_cast<const float4*>(Bs) + tx;
    for (int c = 0; c < DS_N / DS_CHUNK; ++c) {
        const int g = g0 + c;
        cp_async_wait<1>();                 // chunk g has landed (this thread's pieces) ...
        __syncthreads();                    // ... and everybody else's
        const float4* m4 = reinterpret_cast<const float4*>(ring + (g & 1) * DS_CHUNK * DS_N) + 2 * w;
        const int k0 = c * DS_CHUNK, kn = min(DS_CHUNK, K - k0);
#pragma unroll 4
        for (int kk = 0; kk < kn; ++kk) {
            const float4 a0 = m4[kk * (DS_N / 4)], a1 = m4[kk * (DS_N / 4) + 1];          // broadcast LDS.128
            const float4 b = b4[(k0 + kk) * (DS_PITCH / 4)];
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int cc = 0; cc < 4; ++cc) acc[r][cc] = fmaf(av[r], bv[cc], acc[r][cc]);
        }
        __syncthreads();                    // the slot is free again
        stage_chunk(spec, const_cast<float*>(ring), g + 2);
    }
}

// dst[4tx + c][8w + r] = acc[r][c]: the transposed block is 4 rows x 8 consecutive floats (two STS.128 per row)
__device__ __forceinline__ void store_transposed(float* dst, int w, int tx, const float (&acc)[8][4]) {
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        float4* row = reinterpret_cast<float4*>(dst + (4 * tx + c) * DS_PITCH + 8 * w);
        row[0] = make_float4(acc[0][c], acc[1][c], acc[2][c], acc[3][c]);
        row[1] = make_float4(acc[4][c], acc[5][c], acc[6][c], acc[7][c]);
    }
}

__global__ void __launch_bounds__(512, 1) poisson_direct_kernel(const float* __restrict__ rhs, float* __restrict__ out, const float* __restrict__ spec,
                                                               int ny, int nx) {
    extern __shared__ __align__(16) float smem[];
    float* buf0 = smem;
    float* buf1 = smem + DS_N * DS_PITCH;
    float* ring = smem + 2 * DS_N * DS_PITCH;          // two 16 KB slots for the spectral matrices
    stage_chunk(spec, ring, 0);
    stage_chunk(spec, ring, 1);
    const int tx = threadIdx.x & 31, w = threadIdx.x >> 5;
    const size_t base = (size_t)blockIdx.x * ny * nx;
    const bool vec = (nx & 3) == 0 && ((reinterpret_cast<uintptr_t>(rhs) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    // rhs -> buf0 (row-major, zero padded to 128 x 128)
    for (int r = 0; r < 8; ++r) {
        const int i = 8 * w + r, j = 4 * tx;
        float4 v = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (i < ny) {
            if (vec && j < nx) v = *reinterpret_cast<const float4*>(rhs + base + (size_t)i * nx + j);
            else {
                float t[4] = {0.0f, 0.0f, 0.0f, 0.0f};
                for (int c = 0; c < 4; ++c)
                    if (j + c < nx) t[c] = rhs[base + (size_t)i * nx + j + c];
                v = make_float4(t[0], t[1], t[2], t[3]);
            }
        }
        *reinterpret_cast<float4*>(buf0 + i * DS_PITCH + j) = v;
    }
    __syncthreads();
    float acc[8][4];
    // 1: U = Ey^T R            -> stored transposed: buf1 = U^T [nx][ny]
    left_multiply(spec, ring, buf0, ny, 0 * (DS_N / DS_CHUNK), w, tx, acc);
    store_transposed(buf1, w, tx, acc);
    __syncthreads();
    // 2: V^T = Ex^T U^T, scaled by 1 / (l_x + l_y)  -> buf0 = S^T [nx][ny]
    left_multiply(spec, ring, buf1, nx, 1 * (DS_N / DS_CHUNK), w, tx, acc);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const float4 s = __ldg(reinterpret_cast<const float4*>(spec + 4 * DS_MAT + (8 * w + r) * DS_N) + tx);
        *reinterpret_cast<float4*>(buf0 + (8 * w + r) * DS_PITCH + 4 * tx) =
            make_float4(acc[r][0] * s.x, acc[r][1] * s.y, acc[r][2] * s.z, acc[r][3] * s.w);
    }
    __syncthreads();
    // 3: W^T = Ex S^T          -> stored transposed: buf1 = W [ny][nx]
    left_multiply(spec, ring, buf0, nx, 2 * (DS_N / DS_CHUNK), w, tx, acc);
    store_transposed(buf1, w, tx, acc);
    __syncthreads();
    // 4: P = Ey W              -> global
    left_multiply(spec, ring, buf1, ny, 3 * (DS_N / DS_CHUNK), w, tx, acc);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const int i = 8 * w + r, j = 4 * tx;
        if (i >= ny) continue;
        if (vec && j < nx) *reinterpret_cast<float4*>(out + base + (size_t)i * nx + j) = make_float4(acc[r][0], acc[r][1], acc[r][2], acc[r][3]);
        else
            for (int c = 0; c < 4; ++c)
                if (j + c < nx) out[base + (size_t)i * nx + j + c] = acc[r][c];
    }
}


// =====================================================================================================================================
// The same solve on the 5th-generation tensor cores (tcgen05, accumulators in tensor memory): each of the four 128^3 products is 16
// K-steps of `tcgen05.mma.cta_group::1.kind::tf32` (M = N = 128, K = 8), issued by one thread, both operands K-major in shared memory
// in the canonical no-swizzle layout (8-row x 16-byte core matrices; scripts/diag/umma_probe.cu pins the descriptors and the layout
// against the CPU).  TF32 keeps 11 significant bits, so every product runs as THREE tensor-core products on exact splits
// x = hi + lo (hi = x rounded to TF32, lo = the remainder rounded to TF32: together 22 significant bits):
// A B ~ A_hi B_hi + A_lo B_hi + A_hi B_lo, accumulated in fp32 in tensor memory; the dropped A_lo B_lo term is 2^-24 relative.
//
// The running matrix is always the A operand and each epilogue writes the next A straight from the accumulator rows (one thread = one
// tensor-memory lane = one row), either as it is (next product multiplies from the right again) or TRANSPOSED (the left
// multiplications by E_y, taken as right multiplications of the transpose):
//   1  T1[y][kx]   = R[y][:] Ex           A = R            B(n = kx, k = x)  = Ex^T      epilogue: transposed
//   2  T2^T[kx][ky] = T1^T[kx][:] Ey      A = T1^T         B(n = ky, k = y)  = Ey^T      epilogue: scale by 1 / (l_x + l_y), as it is
//   3  T3^T[kx][y]  = S^T[kx][:] Ey^T     A = S^T          B(n = y, k = ky)  = Ey        epilogue: transposed
//   4  P[y][x]      = T3[y][:] Ex^T       A = T3           B(n = x, k = kx)  = Ex        epilogue: to global memory
// A lives in shared memory whole (hi and lo, 2 x 72 KB; the step along K is padded to 144 bytes so that the transposed 4-byte stores
// of a warp fall into 32 different banks), the constant B matrices stream through a ring in chunks (hi + lo =
// 16 KB per K = 16, pre-split and pre-arranged by the set-up kernel so that a chunk is ONE bulk copy: cp.async.bulk completing on an
// mbarrier, five slots).  Warp-specialised: warps 0-7 write operands and run the epilogues, one thread of warp 8 issues the copies, one
// thread of warp 9 the tensor-core instructions; slots are handed back by tcgen05.commit.
constexpr unsigned TC_A_LBO = 144, TC_A_SBO = 32 * TC_A_LBO;           // bytes: next 16-byte unit along K / next 8 rows
constexpr unsigned TC_A_BYTES = 16 * TC_A_SBO;                         // one 128 x 128 A operand (hi or lo): 73 728 bytes
constexpr int TC_KC = 16;                                              // K per streamed chunk of B
constexpr unsigned TC_B_LBO = 128, TC_B_SBO = (TC_KC / 4) * TC_B_LBO;  // inside a chunk
constexpr unsigned TC_B_PART = 16 * TC_B_SBO;                          // hi or lo part of a chunk: 8 KB
constexpr unsigned TC_B_CHUNK = 2 * TC_B_PART;                         // 16 KB, one bulk copy
constexpr int TC_CHUNKS_PER_PRODUCT = DS_N / TC_KC;                    // 8
constexpr int TC_CHUNKS = 4 * TC_CHUNKS_PER_PRODUCT;                   // 32 per solve
constexpr int TC_SLOTS = 5;                                            // ring depth: four copies in flight while one chunk is multiplied
constexpr size_t TC_INV_OFFSET = (size_t)TC_CHUNKS * TC_B_CHUNK;       // byte offset of inv, stored [ky][kx] (plain 128 x 128 floats), in the block
constexpr int TC_EPI_THREADS = 256;                                    // warps 0-7: operand writers / epilogue
constexpr int TC_THREADS = TC_EPI_THREADS + 64;                        // warp 8: copy issuer, warp 9: tensor-core issuer

__device__ __forceinline__ unsigned tc_canon(unsigned r, unsigned k, unsigned lbo, unsigned sbo) {
    return (r >> 3) * sbo + (k >> 2) * lbo + (r & 7u) * 16u + (k & 3u) * 4u;
}
// x rounded to TF32 (round to nearest, ties away: 10 mantissa bits, the low 13 bits of the word are zero, so the tensor core -- which
// ignores them -- reads exactly this value).  x = hi + lo with hi = tf32(x), lo = tf32(x - hi): |x - hi - lo| <= 2^-24 |x|
__device__ __forceinline__ float tf32_hi(float x) {
    unsigned r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}
__device__ __forceinline__ float tf32_lo(float x, float hi) { return tf32_hi(x - hi); }

// spectral block of the tensor-core solver: chunk q = 8 g + c of product g holds B_g(n, k), k in [16 c, 16 c + 16), hi part then lo part
__global__ void __launch_bounds__(256) poisson_direct_tc_setup_kernel(unsigned char* __restrict__ spec, int ny, int nx, int closed) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 4 * DS_MAT) return;
    const int g = idx / DS_MAT, n = (idx % DS_MAT) / DS_N, k = idx % DS_N;
    const int ends_y = closed & 3, ends_x = (closed >> 2) & 3;
    // E[cell][mode].  B_0(n, k) = Ex[k][n], B_1(n, k) = Ey[k][n], B_2(n, k) = Ey[n][k], B_3(n, k) = Ex[n][k]
    float v;
    if (g == 0) v = (k < nx && n < nx) ? (float)eig_vec(k, n, nx, ends_x) : 0.0f;
    else if (g == 1) v = (k < ny && n < ny) ? (float)eig_vec(k, n, ny, ends_y) : 0.0f;
    else if (g == 2) v = (n < ny && k < ny) ? (float)eig_vec(n, k, ny, ends_y) : 0.0f;
    else v = (n < nx && k < nx) ? (float)eig_vec(n, k, nx, ends_x) : 0.0f;
    const float hi = tf32_hi(v), lo = tf32_lo(v, hi);
    unsigned char* chunk = spec + (size_t)(TC_CHUNKS_PER_PRODUCT * g + k / TC_KC) * TC_B_CHUNK;
    const unsigned at = tc_canon((unsigned)n, (unsigned)(k % TC_KC), TC_B_LBO, TC_B_SBO);
    *reinterpret_cast<float*>(chunk + at) = hi;
    *reinterpret_cast<float*>(chunk + TC_B_PART + at) = lo;
    if (g == 0) {                                                      // inv[kx = n][ky = k]
        float inv = 0.0f;
        if (n < nx && k < ny) {
            const double l = eig_val(n, nx, ends_x) + eig_val(k, ny, ends_y);
            inv = (fabs(l) < 1e-9) ? 0.0f : (float)(1.0 / l);
        }
        reinterpret_cast<float*>(spec + TC_INV_OFFSET)[k * DS_N + n] = inv;       // stored [ky][kx]: a warp (32 rows kx) reads one line per ky
    }
}


// ---- tcgen05 / mbarrier wrappers (PTX as in the CUTLASS sm_100 headers; encodings pinned by scripts/diag/umma_probe.cu) ----
__device__ __forceinline__ unsigned tc_smem(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ unsigned long long tc_desc(unsigned saddr, unsigned lbo, unsigned sbo) {
    // shared-memory matrix descriptor: start, leading (K) and stride (M/N) byte offsets in 16-byte units, version 1, no swizzle
    return (unsigned long long)((saddr >> 4) & 0x3fffu) | ((unsigned long long)((lbo >> 4) & 0x3fffu) << 16) |
           ((unsigned long long)((sbo >> 4) & 0x3fffu) << 32) | (1ull << 46);
}
// instruction descriptor: D fp32 (bits 4-5 = 1), A and B tf32 (bits 7-9, 10-12 = 2), both K-major, N = 128 (>> 3 at bit 17), M = 128 (>> 4 at bit 24)
constexpr unsigned TC_IDESC = (1u << 4) | (2u << 7) | (2u << 10) | (16u << 17) | (8u << 24);
__device__ __forceinline__ void tc_mma(unsigned tmem_d, unsigned long long da, unsigned long long db, unsigned accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(da), "l"(db), "r"(TC_IDESC), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_commit(unsigned bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// bounded wait for phase `parity` of an mbarrier; a tensor-core product that never completes must not hang the GPU
__device__ __forceinline__ void tc_wait(unsigned bar, unsigned parity) {
    unsigned done = 0;
    for (int spin = 0; !done && spin < (1 << 22); ++spin)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (!done) __trap();
}
// 32 consecutive accumulator columns of this thread's lane
__device__ __forceinline__ void tc_ld32(unsigned taddr, float (&v)[32]) {
    unsigned r[32];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, "
                 "%21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
                   "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
                   "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
}

__device__ __forceinline__ void tc_bar_init(unsigned bar, unsigned count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory"); }
__device__ __forceinline__ void tc_bar_arrive(unsigned bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
// one bulk copy global -> shared (16-byte aligned, a multiple of 16 bytes), completion counted in bytes on `bar`
__device__ __forceinline__ void tc_bulk_load(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// dynamic shared memory: A_hi | A_lo | ring of TC_SLOTS chunks | barriers: full[TC_SLOTS], empty[TC_SLOTS], accumulator complete, A ready | tmem address
constexpr unsigned TC_SMEM_RING = 2 * TC_A_BYTES, TC_SMEM_BARS = TC_SMEM_RING + TC_SLOTS * TC_B_CHUNK;
constexpr unsigned TC_SMEM_BYTES = TC_SMEM_BARS + 128;

#ifdef TC_TIMING
__device__ unsigned long long g_tc_times[16];                          // diagnostic build: phase time stamps of CTA 0 (ns)
#define TC_STAMP(i) do { if (blockIdx.x == 0 && tid == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_)); g_tc_times[i] = t_; } } while (0)
#else
#define TC_STAMP(i) do { } while (0)
#endif

__global__ void __launch_bounds__(TC_THREADS, 1) poisson_direct_tc_kernel(const float* __restrict__ rhs, float* __restrict__ out,
                                                                         const unsigned char* __restrict__ spec, int ny, int nx) {
    extern __shared__ __align__(128) unsigned char tc_smem_raw[];
    unsigned char* a_hi = tc_smem_raw;
    unsigned char* a_lo = tc_smem_raw + TC_A_BYTES;
    const unsigned s0 = tc_smem(tc_smem_raw);
    const unsigned bar_full = s0 + TC_SMEM_BARS, bar_empty = bar_full + 8 * TC_SLOTS, bar_acc = bar_empty + 8 * TC_SLOTS, bar_a = bar_acc + 8;
    volatile unsigned* tmem_slot = reinterpret_cast<volatile unsigned*>(tc_smem_raw + TC_SMEM_BARS + 8 * (2 * TC_SLOTS + 2));
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int i = 0; i < TC_SLOTS; ++i) { tc_bar_init(bar_full + 8 * i, 1); tc_bar_init(bar_empty + 8 * i, 1); }
        tc_bar_init(bar_acc, 1);
        tc_bar_init(bar_a, TC_EPI_THREADS);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 9) {                                                   // 128 columns of tensor memory: one 128 x 128 fp32 accumulator
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 128;" ::"r"(s0 + TC_SMEM_BARS + 8 * (2 * TC_SLOTS + 2)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // the right-hand side is requested before anything else is waited for: 16 pieces of 16 bytes per operand-writer thread
    const size_t base = (size_t)blockIdx.x * ny * nx;
    const bool vec = (nx & 3) == 0 && ((reinterpret_cast<uintptr_t>(rhs) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
    float4 rq[DS_MAT / 4 / TC_EPI_THREADS];
    if (warp < 8) {
#pragma unroll
        for (int i = 0; i < DS_MAT / 4 / TC_EPI_THREADS; ++i) {
            const int piece = tid + i * TC_EPI_THREADS, row = piece >> 5, j = 4 * (piece & 31);
            float t[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            if (row < ny) {
                if (vec && j < nx) {
                    const float4 v = __ldcs(reinterpret_cast<const float4*>(rhs + base + (size_t)row * nx + j));
                    t[0] = v.x; t[1] = v.y; t[2] = v.z; t[3] = v.w;
                } else {
                    for (int c = 0; c < 4; ++c)
                        if (j + c < nx) t[c] = rhs[base + (size_t)row * nx + j + c];
                }
            }
            rq[i] = make_float4(t[0], t[1], t[2], t[3]);
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const unsigned tm = *tmem_slot;
    TC_STAMP(0);

    if (warp == 8) {
        // ---- copy issuer: the 32 chunks of the solve, in order, each into the slot its predecessor five chunks back has left ----
        if (lane == 0) {
            for (int q = 0; q < TC_CHUNKS; ++q) {
                const int slot = q % TC_SLOTS, use = q / TC_SLOTS;
                if (use > 0) tc_wait(bar_empty + 8 * slot, (unsigned)(use - 1) & 1u);
                tc_bulk_load(s0 + TC_SMEM_RING + (unsigned)slot * TC_B_CHUNK, spec + (size_t)q * TC_B_CHUNK, TC_B_CHUNK, bar_full + 8 * slot);
            }
        }
        __syncwarp();
    } else if (warp == 9) {
        // ---- tensor-core issuer ----
        if (lane == 0) {
            for (int g = 0; g < 4; ++g) {
                tc_wait(bar_a, (unsigned)g & 1u);                      // operand A of product g is in shared memory, the accumulator has been read
                tc_fence_after();
                for (int c = 0; c < TC_CHUNKS_PER_PRODUCT; ++c) {
                    const int q = TC_CHUNKS_PER_PRODUCT * g + c, slot = q % TC_SLOTS, use = q / TC_SLOTS;
#ifndef TC_DIAG_NO_B_WAIT                                               // diagnostic: tensor-core time alone (products of whatever is in the ring)
                    tc_wait(bar_full + 8 * slot, (unsigned)use & 1u);
#endif
                    tc_fence_after();
                    const unsigned b_s = s0 + TC_SMEM_RING + (unsigned)slot * TC_B_CHUNK;
#pragma unroll
                    for (int ks = 0; ks < TC_KC / 8; ++ks) {            // K = 8 per instruction: two 16-byte units of either operand
                        const unsigned ka = (unsigned)((TC_KC / 8) * c + ks) * 2u * TC_A_LBO, kb = (unsigned)ks * 2u * TC_B_LBO;
                        const unsigned long long dah = tc_desc(s0 + ka, TC_A_LBO, TC_A_SBO), dal = tc_desc(s0 + TC_A_BYTES + ka, TC_A_LBO, TC_A_SBO);
                        const unsigned long long dbh = tc_desc(b_s + kb, TC_B_LBO, TC_B_SBO), dbl = tc_desc(b_s + TC_B_PART + kb, TC_B_LBO, TC_B_SBO);
                        tc_mma(tm, dah, dbh, (c | ks) != 0 ? 1u : 0u);
                        tc_mma(tm, dal, dbh, 1u);
                        tc_mma(tm, dah, dbl, 1u);
                    }
                    tc_commit(bar_empty + 8 * slot);                   // the slot is free once these products have read it
                }
                tc_commit(bar_acc);
            }
        }
        __syncwarp();
    } else {
        // ---- operand writers / epilogue warps ----
        // R -> A of product 1 (rows y, K = x), zero padded, split into hi + lo.  One 16-byte piece per thread and turn: coalesced global
        // reads; the 144-byte K step puts the eight pieces of a quarter warp into 32 different banks
#pragma unroll
        for (int i = 0; i < DS_MAT / 4 / TC_EPI_THREADS; ++i) {
            const int piece = tid + i * TC_EPI_THREADS, row = piece >> 5, kq = piece & 31;
            const float4 t = rq[i];
            const float4 h = make_float4(tf32_hi(t.x), tf32_hi(t.y), tf32_hi(t.z), tf32_hi(t.w));
            const unsigned at = (unsigned)(row >> 3) * TC_A_SBO + (unsigned)kq * TC_A_LBO + (unsigned)(row & 7) * 16u;
            *reinterpret_cast<float4*>(a_hi + at) = h;
            *reinterpret_cast<float4*>(a_lo + at) = make_float4(tf32_lo(t.x, h.x), tf32_lo(t.y, h.y), tf32_lo(t.z, h.z), tf32_lo(t.w, h.w));
        }
        tc_fence_async_smem();                                         // generic-proxy stores -> visible to the tensor core
        tc_bar_arrive(bar_a);
        TC_STAMP(1);
        const unsigned m = 32u * (unsigned)(warp & 3) + (unsigned)lane; // this thread's accumulator row (tensor-memory lane)
        const unsigned col0 = 64u * (unsigned)(warp >> 2);              // and its half of the columns
#pragma unroll 1
        for (int g = 0; g < 4; ++g) {
            // the eigenvalue reciprocals of this thread's 64 accumulator entries are fetched while product 2 runs
            float sc[64];
            if (g == 1) {
                const float* iv = reinterpret_cast<const float*>(spec + TC_INV_OFFSET) + (size_t)col0 * DS_N + m;
#pragma unroll
                for (int j = 0; j < 64; ++j) sc[j] = __ldg(iv + (size_t)j * DS_N);
            }
            tc_wait(bar_acc, (unsigned)g & 1u);
            tc_fence_after();
            TC_STAMP(2 + 2 * g);
#pragma unroll
            for (int turn = 0; turn < 2; ++turn) {
                const unsigned n0 = col0 + 32u * (unsigned)turn;
                float v[32];
                tc_ld32(tm + ((32u * (unsigned)(warp & 3)) << 16) + n0, v);
                if (g == 1) {                                          // divide by the eigenvalue sums: inv[kx = m][ky = n]
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] *= sc[32 * turn + j];
                }
                if (g == 3) {                                          // P[y = m][x = n] -> plain rows in shared memory (the A operand is dead)
                    float* prow = reinterpret_cast<float*>(tc_smem_raw) + m * DS_PITCH + n0;
#pragma unroll
                    for (int j4 = 0; j4 < 8; ++j4) *reinterpret_cast<float4*>(prow + 4 * j4) = make_float4(v[4 * j4], v[4 * j4 + 1], v[4 * j4 + 2], v[4 * j4 + 3]);
                } else if (g == 1) {                                   // next A(row = m, K = n): as it is, 16 bytes at a time
#pragma unroll
                    for (int j4 = 0; j4 < 8; ++j4) {
                        const float4 h = make_float4(tf32_hi(v[4 * j4]), tf32_hi(v[4 * j4 + 1]), tf32_hi(v[4 * j4 + 2]), tf32_hi(v[4 * j4 + 3]));
                        const unsigned at = (m >> 3) * TC_A_SBO + ((n0 >> 2) + (unsigned)j4) * TC_A_LBO + (m & 7u) * 16u;
                        *reinterpret_cast<float4*>(a_hi + at) = h;
                        *reinterpret_cast<float4*>(a_lo + at) =
                            make_float4(tf32_lo(v[4 * j4], h.x), tf32_lo(v[4 * j4 + 1], h.y), tf32_lo(v[4 * j4 + 2], h.z), tf32_lo(v[4 * j4 + 3], h.w));
                    }
                } else {                                               // next A(row = n, K = m): transposed, 4 bytes at a time
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const unsigned n = n0 + (unsigned)j;
                        const unsigned at = (n >> 3) * TC_A_SBO + (m >> 2) * TC_A_LBO + (n & 7u) * 16u + (m & 3u) * 4u;
                        const float h = tf32_hi(v[j]);
                        *reinterpret_cast<float*>(a_hi + at) = h;
                        *reinterpret_cast<float*>(a_lo + at) = tf32_lo(v[j], h);
                    }
                }
            }
            if (g < 3) {
                tc_fence_async_smem();
                tc_fence_before();                                     // the accumulator has been read: the next product may overwrite it
                tc_bar_arrive(bar_a);
            } else {
                // coalesced copy-out of P by the 256 epilogue threads (named barrier 1: the copy and tensor-core warps are not part of it)
                asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI_THREADS) : "memory");
                const float* ps = reinterpret_cast<const float*>(tc_smem_raw);
#pragma unroll 4
                for (int i = 0; i < DS_MAT / 4 / TC_EPI_THREADS; ++i) {
                    const int piece = tid + i * TC_EPI_THREADS, row = piece >> 5, x = 4 * (piece & 31);
                    if (row >= ny || x >= nx) continue;
                    const float4 pv = *reinterpret_cast<const float4*>(ps + row * DS_PITCH + x);
                    float* dst = out + base + (size_t)row * nx + x;
                    if (vec) __stcs(reinterpret_cast<float4*>(dst), pv);
                    else {
                        const float t[4] = {pv.x, pv.y, pv.z, pv.w};
                        for (int cc = 0; cc < 4; ++cc)
                            if (x + cc < nx) dst[cc] = t[cc];
                    }
                }
            }
            TC_STAMP(3 + 2 * g);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 128;" ::"r"(tm) : "memory");
}

}  // namespace pdec

using namespace pdec;

extern "C" size_t pdec_poisson_direct_2d_spectral_bytes(void) { return (size_t)5 * DS_MAT * sizeof(float); }

extern "C" int pdec_poisson_direct_2d_supported(int ny, int nx) { return ny >= 2 && nx >= 2 && ny <= DS_N && nx <= DS_N; }

extern "C" int pdec_poisson_direct_2d_setup(void* spectral, size_t spectral_bytes, int ny, int nx, int closed, void* stream) {
    if (!spectral) return PDEC_ERR_INVALID_ARGUMENT;
    if (!pdec_poisson_direct_2d_supported(ny, nx)) return PDEC_ERR_UNSUPPORTED_SIZE;
    if (spectral_bytes < pdec_poisson_direct_2d_spectral_bytes()) return PDEC_ERR_WORKSPACE_TOO_SMALL;
    poisson_direct_setup_kernel<<<(DS_MAT + 255) / 256, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<float*>(spectral), ny, nx, (int)pdec_closed_sides(closed));
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" int pdec_poisson_direct_2d(const float* rhs, float* p, int B, int ny, int nx, const void* spectral, void* stream) {
    if (B < 0) return PDEC_ERR_INVALID_ARGUMENT;
    if (!pdec_poisson_direct_2d_supported(ny, nx)) return PDEC_ERR_UNSUPPORTED_SIZE;
    if (B == 0) return PDEC_OK;
    if (!rhs || !p || !spectral) return PDEC_ERR_INVALID_ARGUMENT;
    const size_t smem = ((size_t)2 * DS_N * DS_PITCH + 2 * DS_CHUNK * DS_N) * sizeof(float);
    if (cudaFuncSetAttribute(poisson_direct_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
    poisson_direct_kernel<<<B, 512, smem, (cudaStream_t)stream>>>(rhs, p, reinterpret_cast<const float*>(spectral), ny, nx);
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

// ---- tensor-core variant (tcgen05, 3 x TF32): same interface, its own spectral block ----
extern "C" size_t pdec_poisson_direct_tc_2d_spectral_bytes(void) { return TC_INV_OFFSET + (size_t)DS_MAT * sizeof(float); }

extern "C" int pdec_poisson_direct_tc_2d_setup(void* spectral, size_t spectral_bytes, int ny, int nx, int closed, void* stream) {
    if (!spectral) return PDEC_ERR_INVALID_ARGUMENT;
    if (!pdec_poisson_direct_2d_supported(ny, nx)) return PDEC_ERR_UNSUPPORTED_SIZE;
    if (spectral_bytes < pdec_poisson_direct_tc_2d_spectral_bytes()) return PDEC_ERR_WORKSPACE_TOO_SMALL;
    if ((reinterpret_cast<uintptr_t>(spectral) & 15) != 0) return PDEC_ERR_INVALID_ARGUMENT;
    poisson_direct_tc_setup_kernel<<<(4 * DS_MAT + 255) / 256, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<unsigned char*>(spectral), ny, nx,
                                                                                              (int)pdec_closed_sides(closed));
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" int pdec_poisson_direct_tc_2d(const float* rhs, float* p, int B, int ny, int nx, const void* spectral, void* stream) {
    if (B < 0) return PDEC_ERR_INVALID_ARGUMENT;
    if (!pdec_poisson_direct_2d_supported(ny, nx)) return PDEC_ERR_UNSUPPORTED_SIZE;
    if (B == 0) return PDEC_OK;
    if (!rhs || !p || !spectral || (reinterpret_cast<uintptr_t>(spectral) & 15) != 0) return PDEC_ERR_INVALID_ARGUMENT;
    const size_t smem = TC_SMEM_BYTES;
    if (cudaFuncSetAttribute(poisson_direct_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
    poisson_direct_tc_kernel<<<B, TC_THREADS, smem, (cudaStream_t)stream>>>(rhs, p, reinterpret_cast<const unsigned char*>(spectral), ny, nx);
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

#ifdef TC_TIMING
extern "C" int pdec_debug_tc_times(unsigned long long* host16) {
    return cudaMemcpyFromSymbol(host16, g_tc_times, sizeof(unsigned long long) * 16) == cudaSuccess ? PDEC_OK : PDEC_ERR_CUDA_LAUNCH;
}
#endif
