// laplace_matrix.cu -- the stencil of the pressure matrix on a domain with obstacles, built on the device from the cell mask.
//
// Replaces  LaplaceMatrix (PF/tf/cuda/src/laplace_op.cc:10-17, laplace_op.cu.cc:47-145: banded int8 matrix [N, 2D+1] from the active /
//           fluid masks) and PoissonDomain's mask logic (PF/physics/pressuresolver/solver_api.py:51-121, sparse.py:157-221).
// active [ny, nx]: 1 = fluid, 0 = obstacle (the obstacle geometries sampled at the cell centres, fluid.py:228-233).  Outputs
//   code [ny, nx] uint8: bits 0..3 = coupled to the neighbour above (i-1) / below (i+1) / left (j-1) / right (j+1), i.e.
//                        active_self * active_neighbour (0 outside the domain);  bits 4..7 = minus the diagonal = the number of
//                        ACCESSIBLE neighbours (outside the domain: 1 on open, 0 on closed boundaries), at least 1  (sparse.py:186-213)
//   fy [ny+1, nx], fx [ny, nx+1]: face masks = min of the two adjacent accessible cells, the accessible grid extended by its own
//                        extrapolation: replicate on open, 0 on closed domains (_frictionless_velocity_mask, solver_api.py:111-121)
// One byte per cell instead of the reference's 2D+1: the CG kernels decode it (cg.cu, pdec_cg_poisson_2d_masked).
#include "pdec_common.cuh"

namespace pdec {

__global__ void __launch_bounds__(256) laplace_matrix_kernel(const float* __restrict__ active, unsigned char* __restrict__ code, float* __restrict__ fy,
                                                             float* __restrict__ fx, int ny, int nx, int closed) {
    const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y;
    if (i > ny || j > nx) return;
    auto act = [&](int a, int b) { return (a >= 0 && a < ny && b >= 0 && b < nx) ? active[(size_t)a * nx + b] : 0.0f; };
    const float outside = closed ? 0.0f : 1.0f;
    auto acc = [&](int a, int b) { return (a >= 0 && a < ny && b >= 0 && b < nx) ? active[(size_t)a * nx + b] : outside; };
    if (i < ny && j < nx && code != nullptr) {
        const float self = act(i, j);
        const unsigned up = act(i - 1, j) * self != 0.0f, dn = act(i + 1, j) * self != 0.0f;
        const unsigned lf = act(i, j - 1) * self != 0.0f, rt = act(i, j + 1) * self != 0.0f;
        const float d = fmaxf(acc(i - 1, j) + acc(i + 1, j) + acc(i, j - 1) + acc(i, j + 1), 1.0f);
        code[(size_t)i * nx + j] = (unsigned char)(up | (dn << 1) | (lf << 2) | (rt << 3) | ((unsigned)d << 4));
    }
    // face (i, j) of vy lies between cells (i-1, j) and (i, j); the mask grid is extended by replicate (open) or 0 (closed)
    if (fy != nullptr && j < nx) {
        const float a = i > 0 ? active[(size_t)(i - 1) * nx + j] : (closed ? 0.0f : active[j]);
        const float b = i < ny ? active[(size_t)i * nx + j] : (closed ? 0.0f : active[(size_t)(ny - 1) * nx + j]);
        fy[(size_t)i * nx + j] = fminf(a, b);
    }
    if (fx != nullptr && i < ny) {
        const float a = j > 0 ? active[(size_t)i * nx + j - 1] : (closed ? 0.0f : active[(size_t)i * nx]);
        const float b = j < nx ? active[(size_t)i * nx + j] : (closed ? 0.0f : active[(size_t)i * nx + nx - 1]);
        fx[(size_t)i * (nx + 1) + j] = fminf(a, b);
    }
}

}  // namespace pdec

extern "C" int pdec_laplace_matrix_2d(const float* active, unsigned char* code, float* fy, float* fx, int ny, int nx, int closed, void* stream) {
    if (!active || ny < 1 || nx < 1) return PDEC_ERR_INVALID_ARGUMENT;
    const dim3 grid((unsigned)(nx / 32 + 1), (unsigned)(ny / 8 + 1));
    pdec::laplace_matrix_kernel<<<grid, dim3(32, 8), 0, (cudaStream_t)stream>>>(active, code, fy, fx, ny, nx, closed ? 1 : 0);
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}
