// resample.cu -- the backend operator the whole path is built on, as an operator of its own: multilinear resampling of a grid at
// arbitrary points and its reverse mode.
//
// Replaces  backend.resample(inputs, sample_coords, interpolation='linear', boundary)   PF/tf/tf_backend.py:145-154, 451-502
//           native twins  Resample / ResampleGradient                                  PF/tf/cuda/src/resample.cc:7-23,
//                                                                                       resample_gradient.cc:6-13, .cu.cc:80-134
// Semantics are the TF-CPU ones the fused kernels follow (tf_backend.py:462-476): 'constant' clamps both corner indices and takes
// the weights from the CLAMPED indices, so a coordinate outside [0, N-1) along any axis gives exactly 0; 'replicate' clamps the
// indices only; 'circular' wraps.  Combination order of the corners: over axis 0 first, then axis 1 (:491-497).
// Layout: data [B, ny, nx] (rank 2) or [B, nx] (rank 1), one channel; coords [Bc, P, rank] in INDEX space of data (what
// CenteredGrid.sample_at passes, grid.py:101-107), Bc = 1 (shared by the batch) or B; out [B, P].
// Reverse mode: g_coords [B, P, rank] per batch element; g_data is the transpose of a gather at ARBITRARY points, so there is no window
// to bound a gather-transpose with (the fused kernels of the hot path have one).  It is accumulated DETERMINISTICALLY all the same, without
// float atomics (the reference's ResampleGradient, resample_gradient.cu.cc:80-134, uses them): every contribution is converted to 64-bit
// fixed point with a power-of-two scale taken from max|g_out| (an order-independent integer atomicMax) and added with integer atomics --
// integer addition is associative, the sum does not depend on the order -- then converted back.  Resolution: max|g_out| * 2^-40 per term.
#include "pdec_common.cuh"

namespace pdec {

template <int MODE, int RANK>
__global__ void __launch_bounds__(256) resample_fwd_kernel(const float* __restrict__ data, const float* __restrict__ coords, float* __restrict__ out,
                                                           int ny, int nx, long long P, int shared_coords) {
    const long long p = (long long)blockIdx.x * 256 + threadIdx.x;
    const int b = blockIdx.y;
    if (p >= P) return;
    const float* c = coords + ((shared_coords ? 0 : (long long)b * P) + p) * RANK;
    const float* f = data + (size_t)b * ny * nx;
    int fx, cx;
    float w0x, w1x;
    interp_setup<MODE>(c[RANK - 1], nx, fx, cx, w0x, w1x);
    float v;
    if (RANK == 1) {
        v = __ldg(f + fx) * w1x + __ldg(f + cx) * w0x;
    } else {
        int fy, cy;
        float w0y, w1y;
        interp_setup<MODE>(c[0], ny, fy, cy, w0y, w1y);
        const float f0 = __ldg(f + (size_t)fy * nx + fx) * w1y + __ldg(f + (size_t)cy * nx + fx) * w0y;
        const float f1 = __ldg(f + (size_t)fy * nx + cx) * w1y + __ldg(f + (size_t)cy * nx + cx) * w0y;
        v = f0 * w1x + f1 * w0x;
    }
    out[(long long)b * P + p] = v;
}

// workspace header: word 0 = bits of max|g_out| (non-negative floats order like their bit patterns)
__global__ void __launch_bounds__(256) resample_absmax_kernel(const float* __restrict__ g, long long n, unsigned* __restrict__ header) {
    float m = 0.0f;
    for (long long q = (long long)blockIdx.x * 256 + threadIdx.x; q < n; q += (long long)gridDim.x * 256) {
        const float v = fabsf(g[q]);
        if (v == v && v < 3.0e38f) m = fmaxf(m, v);                 // NaN / inf gradients do not set the scale
    }
    m = warp_max(m);
    __shared__ float sm[8];
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x < 8) {
        m = sm[threadIdx.x];
        for (int o = 4; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffu, m, o));
        if (threadIdx.x == 0) atomicMax(header, __float_as_uint(m));
    }
}


template <int MODE, int RANK>
__global__ void __launch_bounds__(256) resample_bwd_kernel(const float* __restrict__ data, const float* __restrict__ coords, const float* __restrict__ g_out,
                                                           unsigned long long* __restrict__ acc, float* __restrict__ g_coords, int ny, int nx, long long P,
                                                           int shared_coords, const unsigned* __restrict__ header, int bits) {
    const long long p = (long long)blockIdx.x * 256 + threadIdx.x;
    const int b = blockIdx.y;
    if (p >= P) return;
    const float* c = coords + ((shared_coords ? 0 : (long long)b * P) + p) * RANK;
    const float* f = data + (size_t)b * ny * nx;
    const float g = g_out[(long long)b * P + p];
    int fx, cx, fy = 0, cy = 0;
    float w0x, w1x, w0y = 0.0f, w1y = 1.0f;
    interp_setup<MODE>(c[RANK - 1], nx, fx, cx, w0x, w1x);
    if (RANK == 2) interp_setup<MODE>(c[0], ny, fy, cy, w0y, w1y);
    if (acc != nullptr) {
        const float scale = fixed_scale(header, bits);
        unsigned long long* gd = acc + (size_t)b * ny * nx;
        auto add = [&](size_t idx, float v) {                       // two's-complement add: order independent
            const float sv = v * scale;
            if (sv == sv && fabsf(sv) < 9.0e18f) atomicAdd(gd + idx, (unsigned long long)__float2ll_rn(sv));
        };
        if (scale > 0.0f) {
            if (RANK == 1) {
                add(fx, g * w1x);
                add(cx, g * w0x);
            } else {
                add((size_t)fy * nx + fx, g * w1x * w1y);
                add((size_t)cy * nx + fx, g * w1x * w0y);
                add((size_t)fy * nx + cx, g * w0x * w1y);
                add((size_t)cy * nx + cx, g * w0x * w0y);
            }
        }
    }
    if (g_coords != nullptr) {
        float* gc = g_coords + ((long long)b * P + p) * RANK;
        if (RANK == 1) {
            gc[0] = g * (__ldg(f + cx) - __ldg(f + fx));
        } else {
            const float s00 = __ldg(f + (size_t)fy * nx + fx), s01 = __ldg(f + (size_t)fy * nx + cx);
            const float s10 = __ldg(f + (size_t)cy * nx + fx), s11 = __ldg(f + (size_t)cy * nx + cx);
            const float f0 = s00 * w1y + s10 * w0y, f1 = s01 * w1y + s11 * w0y;
            gc[0] = g * ((s10 - s00) * w1x + (s11 - s01) * w0x);
            gc[1] = g * (f1 - f0);
        }
    }
}

__global__ void __launch_bounds__(256) resample_unfix_kernel(const unsigned long long* __restrict__ acc, float* __restrict__ g_data, long long n,
                                                             const unsigned* __restrict__ header, int bits) {
    const float inv = fixed_inv_scale(header, bits);
    for (long long q = (long long)blockIdx.x * 256 + threadIdx.x; q < n; q += (long long)gridDim.x * 256)
        g_data[q] = fixed_to_float(acc[q], inv);
}

}  // namespace pdec

using namespace pdec;

static bool resample_args_ok(const void* data, const void* coords, int B, int rank, int ny, int nx, int coords_batch, long long P, int mode) {
    if (!data || !coords || B < 0 || P < 0 || nx < 1 || (rank != 1 && rank != 2) || (rank == 2 && ny < 1)) return false;
    if (coords_batch != 1 && coords_batch != B) return false;
    return mode == PDEC_MODE_CONSTANT || mode == PDEC_MODE_REPLICATE || mode == PDEC_MODE_CIRCULAR;
}

extern "C" int pdec_resample_fwd(const float* data, const float* coords, float* out, int B, int rank, int ny, int nx, int coords_batch,
                                 long long npoints, int mode, void* stream) {
    if (!resample_args_ok(data, coords, B, rank, ny, nx, coords_batch, npoints, mode) || !out) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0 || npoints == 0) return PDEC_OK;
    if (B > 65535) return PDEC_ERR_UNSUPPORTED_SIZE;
    cudaStream_t st = (cudaStream_t)stream;
    const dim3 grid((unsigned)((npoints + 255) / 256), (unsigned)B);
    const int shared = coords_batch == 1 ? 1 : 0;
    if (rank == 1) ny = 1;
#define RS_FWD(M)                                                                                                    \
    do {                                                                                                             \
        if (rank == 1) resample_fwd_kernel<M, 1><<<grid, 256, 0, st>>>(data, coords, out, ny, nx, npoints, shared);   \
        else resample_fwd_kernel<M, 2><<<grid, 256, 0, st>>>(data, coords, out, ny, nx, npoints, shared);             \
    } while (0)
    if (mode == PDEC_MODE_CONSTANT) RS_FWD(PDEC_MODE_CONSTANT);
    else if (mode == PDEC_MODE_REPLICATE) RS_FWD(PDEC_MODE_REPLICATE);
    else RS_FWD(PDEC_MODE_CIRCULAR);
#undef RS_FWD
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" size_t pdec_resample_bwd_workspace_bytes(int B, int rank, int ny, int nx) {
    if (rank == 1) ny = 1;
    return 256 + sizeof(unsigned long long) * (size_t)(B > 0 ? B : 1) * (size_t)(ny > 0 ? ny : 1) * (size_t)(nx > 0 ? nx : 1);
}

extern "C" int pdec_resample_bwd(const float* data, const float* coords, const float* g_out, float* g_data, float* g_coords, int B, int rank,
                                 int ny, int nx, int coords_batch, long long npoints, int mode, void* workspace, size_t workspace_bytes,
                                 void* stream) {
    if (!resample_args_ok(data, coords, B, rank, ny, nx, coords_batch, npoints, mode) || !g_out) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (B > 65535) return PDEC_ERR_UNSUPPORTED_SIZE;
    cudaStream_t st = (cudaStream_t)stream;
    if (rank == 1) ny = 1;
    const long long cells = (long long)B * ny * nx;
    if (g_data) {
        if (!workspace) return PDEC_ERR_INVALID_ARGUMENT;
        if (workspace_bytes < pdec_resample_bwd_workspace_bytes(B, rank, ny, nx)) return PDEC_ERR_WORKSPACE_TOO_SMALL;
        if (cudaMemsetAsync(workspace, 0, pdec_resample_bwd_workspace_bytes(B, rank, ny, nx), st) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
    }
    if (npoints == 0 || (!g_data && !g_coords)) {
        if (g_data && cudaMemsetAsync(g_data, 0, sizeof(float) * (size_t)cells, st) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
        return PDEC_OK;
    }
    unsigned* header = reinterpret_cast<unsigned*>(workspace);
    unsigned long long* acc = g_data ? reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(workspace) + 256) : nullptr;
    // up to 4 * npoints terms can land on one cell: keep the sum below 2^62
    int bits = 40, terms = 2;
    for (long long t = 4 * npoints; t > 1 && terms < 62; t >>= 1) ++terms;
    if (bits > 62 - terms) bits = 62 - terms;
    const long long total = (long long)B * npoints;
    if (g_data) {
        const int blocks = (int)((total + 255) / 256 > 1184 ? 1184 : (total + 255) / 256);
        resample_absmax_kernel<<<blocks, 256, 0, st>>>(g_out, total, header);
    }
    const dim3 grid((unsigned)((npoints + 255) / 256), (unsigned)B);
    const int shared = coords_batch == 1 ? 1 : 0;
#define RS_BWD(M)                                                                                                                                      \
    do {                                                                                                                                               \
        if (rank == 1) resample_bwd_kernel<M, 1><<<grid, 256, 0, st>>>(data, coords, g_out, acc, g_coords, ny, nx, npoints, shared, header, bits);      \
        else resample_bwd_kernel<M, 2><<<grid, 256, 0, st>>>(data, coords, g_out, acc, g_coords, ny, nx, npoints, shared, header, bits);                \
    } while (0)
    if (mode == PDEC_MODE_CONSTANT) RS_BWD(PDEC_MODE_CONSTANT);
    else if (mode == PDEC_MODE_REPLICATE) RS_BWD(PDEC_MODE_REPLICATE);
    else RS_BWD(PDEC_MODE_CIRCULAR);
#undef RS_BWD
    if (g_data) {
        const int blocks = (int)((cells + 255) / 256 > 1184 ? 1184 : (cells + 255) / 256);
        resample_unfix_kernel<<<blocks, 256, 0, st>>>(acc, g_data, cells, header, bits);
    }
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}
