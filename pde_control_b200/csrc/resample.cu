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
// Reverse mode: g_data by atomicAdd of the four corner weights (like the reference's ResampleGradient: the order of float additions
// is not fixed; the fused kernels of the hot path use their own deterministic transposes), g_coords [B, P, rank] per batch element.
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

template <int MODE, int RANK>
__global__ void __launch_bounds__(256) resample_bwd_kernel(const float* __restrict__ data, const float* __restrict__ coords, const float* __restrict__ g_out,
                                                           float* __restrict__ g_data, float* __restrict__ g_coords, int ny, int nx, long long P,
                                                           int shared_coords) {
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
    if (g_data != nullptr) {
        float* gd = g_data + (size_t)b * ny * nx;
        if (RANK == 1) {
            atomicAdd(gd + fx, g * w1x);
            atomicAdd(gd + cx, g * w0x);
        } else {
            atomicAdd(gd + (size_t)fy * nx + fx, g * w1x * w1y);
            atomicAdd(gd + (size_t)cy * nx + fx, g * w1x * w0y);
            atomicAdd(gd + (size_t)fy * nx + cx, g * w0x * w1y);
            atomicAdd(gd + (size_t)cy * nx + cx, g * w0x * w0y);
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

extern "C" int pdec_resample_bwd(const float* data, const float* coords, const float* g_out, float* g_data, float* g_coords, int B, int rank,
                                 int ny, int nx, int coords_batch, long long npoints, int mode, void* stream) {
    if (!resample_args_ok(data, coords, B, rank, ny, nx, coords_batch, npoints, mode) || !g_out) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (B > 65535) return PDEC_ERR_UNSUPPORTED_SIZE;
    cudaStream_t st = (cudaStream_t)stream;
    if (rank == 1) ny = 1;
    if (g_data && cudaMemsetAsync(g_data, 0, sizeof(float) * (size_t)B * ny * nx, st) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
    if (npoints == 0 || (!g_data && !g_coords)) return PDEC_OK;
    const dim3 grid((unsigned)((npoints + 255) / 256), (unsigned)B);
    const int shared = coords_batch == 1 ? 1 : 0;
#define RS_BWD(M)                                                                                                                          \
    do {                                                                                                                                   \
        if (rank == 1) resample_bwd_kernel<M, 1><<<grid, 256, 0, st>>>(data, coords, g_out, g_data, g_coords, ny, nx, npoints, shared);     \
        else resample_bwd_kernel<M, 2><<<grid, 256, 0, st>>>(data, coords, g_out, g_data, g_coords, ny, nx, npoints, shared);               \
    } while (0)
    if (mode == PDEC_MODE_CONSTANT) RS_BWD(PDEC_MODE_CONSTANT);
    else if (mode == PDEC_MODE_REPLICATE) RS_BWD(PDEC_MODE_REPLICATE);
    else RS_BWD(PDEC_MODE_CIRCULAR);
#undef RS_BWD
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}
