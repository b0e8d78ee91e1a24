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
