// fluid.cu -- 2-D incompressible-flow stencil / gather kernels for sm_100a:
// staggered curl, semi-Lagrangian advection of a centred field and of the staggered velocity, masked
// divergence, pressure-gradient subtraction, density normalisation, and the reverse-mode adjoint of each.
//
// Compile with -fmad=false: back-traced sample coordinates follow the reference's fp32 operation order
// (the 'constant' extrapolation rule is discontinuous in the coordinate).
//
// Layouts (channel dropped, contiguous fp32): rho[B,ny,nx], vy[B,ny+1,nx], vx[B,ny,nx+1]; axis 0 = y.
// Thread mapping: one thread per output sample, blockDim (32, 8), grid (x tiles, y tiles, batch): every warp
// reads/writes 128 contiguous bytes of a row; gathers go through the read-only L1 path (__ldg) -- at the
// configured sizes all fields are L2 resident (128x128 x 64 = 4 MiB per field, L2 = 126 MB).
//
// Adjoint of the bilinear gather = a scatter made ORDER-INDEPENDENT (no float atomics): pass A adds every sample's four weighted
// shares of the output gradient to 64-bit fixed-point accumulators with integer atomics and stores the sample's coordinate
// gradients; pass B converts the sums and brings the coordinate gradients to the velocity faces (fixed small stencils).
#include "pdec_common.cuh"

namespace pdec {

struct FluidGeo {
    int ny, nx;
    float lo[2], sz[2];          // domain box (centred grids)
    float clo[2][2], csz[2][2];  // staggered component boxes [component][axis]  (staggered_grid.py:49-53)
    float inv[2], cinv[2][2];    // exact reciprocals of sz / csz where those are powers of two, else 0
    float c_wu[2], c_wl[2];      // component c -> cell centres along axis c      (grid.py:112-120)
    float s_wu[2][2], s_wl[2][2];// component 1-c -> face centres of component c, [c][axis]
    float dt;
    float gfac[3][2];            // [variant][axis]: d(xi)/d(u) = -(dt * resolution) / size of the variant's grid
};

enum { V_CENTERED = 0, V_COMP_Y = 1, V_COMP_X = 2 };

template <int MODE>
__device__ __forceinline__ int map_index(int i, int n) {   // pad-mode index map; -1 = outside ('constant')
    if (MODE == PDEC_MODE_CONSTANT) return (i < 0 || i >= n) ? -1 : i;
    if (MODE == PDEC_MODE_REPLICATE) return min(max(i, 0), n - 1);
    int m = i % n;
    return m < 0 ? m + n : m;
}
template <int MODE>
__device__ __forceinline__ float load_mapped(const float* __restrict__ f, int i, int j, int ny, int nx) {
    const int a = map_index<MODE>(i, ny), b = map_index<MODE>(j, nx);
    if (MODE == PDEC_MODE_CONSTANT && (a < 0 || b < 0)) return 0.0f;
    return __ldg(f + (size_t)a * nx + b);
}

__device__ __forceinline__ float point_coord(int i, int n, float lower, float size) {
    return cell_center_local(i, n) * size + lower;      // box.local_to_global (geom/_box.py:54-56)
}
__device__ __forceinline__ float index_coord(float x, float lower, float size, float resf) {
    return ((x - lower) / size) * resf - 0.5f;          // _box.py:50-52, grid.py:104-105
}
// the same with the division replaced by the exact reciprocal when `inv` != 0 (size a power of two: x / 2^k == x * 2^-k bit for bit)
__device__ __forceinline__ float index_coord_fast(float x, float lower, float size, float inv, float resf) {
    const float local = inv != 0.0f ? (x - lower) * inv : (x - lower) / size;
    return local * resf - 0.5f;
}

// Sample-point coordinates of this block's 8 rows and 32 columns, computed ONCE per block into shared memory.  point_coord evaluates
// numpy.linspace in float64 (two double divisions per call): per thread that was ~40 % of the advection kernels' instructions
// (ncu, profiles/r2n_solver_kernels_summary.md: 400 thread instructions per cell).  Same function, same bits.
__device__ __forceinline__ void block_points(int ry, int rx, const float* lo, const float* sz, float* s_py, float* s_px) {
    const int t = threadIdx.y * 32 + threadIdx.x;
    if (t < 8) {
        const int i = blockIdx.y * 8 + t;
        s_py[t] = i < ry ? point_coord(i, ry, lo[0], sz[0]) : 0.0f;
    } else if (t >= 32 && t < 64) {
        const int j = blockIdx.x * 32 + t - 32;
        s_px[t - 32] = j < rx ? point_coord(j, rx, lo[1], sz[1]) : 0.0f;
    }
    __syncthreads();
}

// Resolution of the grid a VARIANT advects (= samples from and writes to).
template <int VARIANT>
__device__ __host__ __forceinline__ void variant_res(int ny, int nx, int& ry, int& rx) {
    ry = ny + (VARIANT == V_COMP_Y ? 1 : 0);
    rx = nx + (VARIANT == V_COMP_X ? 1 : 0);
}

// Velocity (uy, ux) at sample point (i, j) of the VARIANT's grid.  VMODE = extrapolation of the velocity field.
template <int VARIANT, int VMODE>
__device__ __forceinline__ void velocity_at(const FluidGeo& g, const float* __restrict__ vy, const float* __restrict__ vx,
                                            int i, int j, float& uy, float& ux) {
    const int ny = g.ny, nx = g.nx;
    if (VARIANT == V_CENTERED) {            // StaggeredGrid.at(centres): upper*wu + lower*wl per component
        uy = __ldg(vy + (size_t)(i + 1) * nx + j) * g.c_wu[0] + __ldg(vy + (size_t)i * nx + j) * g.c_wl[0];
        ux = __ldg(vx + (size_t)i * (nx + 1) + j + 1) * g.c_wu[1] + __ldg(vx + (size_t)i * (nx + 1) + j) * g.c_wl[1];
    } else if (VARIANT == V_COMP_Y) {       // vy faces: uy = vy; ux = vx padded along y, lerp y then x
        uy = __ldg(vy + (size_t)i * nx + j);
        const float a0 = load_mapped<VMODE>(vx, i, j, ny, nx + 1) * g.s_wu[0][0] + load_mapped<VMODE>(vx, i - 1, j, ny, nx + 1) * g.s_wl[0][0];
        const float a1 = load_mapped<VMODE>(vx, i, j + 1, ny, nx + 1) * g.s_wu[0][0] + load_mapped<VMODE>(vx, i - 1, j + 1, ny, nx + 1) * g.s_wl[0][0];
        ux = a1 * g.s_wu[0][1] + a0 * g.s_wl[0][1];
    } else {                                // vx faces: ux = vx; uy = vy padded along x, lerp y then x
        ux = __ldg(vx + (size_t)i * (nx + 1) + j);
        const float a0 = load_mapped<VMODE>(vy, i + 1, j - 1, ny + 1, nx) * g.s_wu[1][0] + load_mapped<VMODE>(vy, i, j - 1, ny + 1, nx) * g.s_wl[1][0];
        const float a1 = load_mapped<VMODE>(vy, i + 1, j, ny + 1, nx) * g.s_wu[1][0] + load_mapped<VMODE>(vy, i, j, ny + 1, nx) * g.s_wl[1][0];
        uy = a1 * g.s_wu[1][1] + a0 * g.s_wl[1][1];
    }
}

// Back-traced index coordinates of sample (i, j) of the VARIANT's grid.
template <int VARIANT, int VMODE>
__device__ __forceinline__ void backtrace(const FluidGeo& g, const float* __restrict__ vy, const float* __restrict__ vx,
                                          int i, int j, int ry, int rx, float py, float px, float& xiy, float& xix) {
    float uy, ux;
    velocity_at<VARIANT, VMODE>(g, vy, vx, i, j, uy, ux);
    const float *lo = (VARIANT == V_CENTERED) ? g.lo : g.clo[VARIANT - 1];
    const float *sz = (VARIANT == V_CENTERED) ? g.sz : g.csz[VARIANT - 1];
    const float *inv = (VARIANT == V_CENTERED) ? g.inv : g.cinv[VARIANT - 1];
    const float y = py - uy * g.dt;                                  // advect.py:35; (py, px) = point_coord of (i, j), from block_points
    const float x = px - ux * g.dt;
    xiy = index_coord_fast(y, lo[0], sz[0], inv[0], (float)ry);
    xix = index_coord_fast(x, lo[1], sz[1], inv[1], (float)rx);
}

// Bilinear sample, TF op order (tf_backend.py:491-497): combine over y first, then x.
// Optionally returns d/dxi_y and d/dxi_x.
template <int MODE, bool DERIV>
__device__ __forceinline__ float sample2d(const float* __restrict__ f, int ry, int rx, float xiy, float xix, float& dy, float& dx) {
    int fy, cy, fx, cx;
    float w0y, w1y, w0x, w1x;
    interp_setup<MODE>(xiy, ry, fy, cy, w0y, w1y);
    interp_setup<MODE>(xix, rx, fx, cx, w0x, w1x);
    const float s00 = __ldg(f + (size_t)fy * rx + fx), s01 = __ldg(f + (size_t)fy * rx + cx);
    const float s10 = __ldg(f + (size_t)cy * rx + fx), s11 = __ldg(f + (size_t)cy * rx + cx);
    const float f0 = s00 * w1y + s10 * w0y;
    const float f1 = s01 * w1y + s11 * w0y;
    if (DERIV) {
        dy = (s10 - s00) * w1x + (s11 - s01) * w0x;
        dx = f1 - f0;
    }
    return f0 * w1x + f1 * w0x;
}

// ---- forward advection ---------------------------------------------------------------------------
template <int VARIANT, int FMODE, int VMODE>
__global__ void __launch_bounds__(256) advect_fwd_kernel(FluidGeo g, const float* __restrict__ field, const float* __restrict__ vy,
                                                         const float* __restrict__ vx, float* __restrict__ out) {
    int ry, rx;
    variant_res<VARIANT>(g.ny, g.nx, ry, rx);
    const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y, b = blockIdx.z;
    __shared__ float s_py[8], s_px[32];
    block_points(ry, rx, (VARIANT == V_CENTERED) ? g.lo : g.clo[VARIANT - 1], (VARIANT == V_CENTERED) ? g.sz : g.csz[VARIANT - 1], s_py, s_px);
    if (i >= ry || j >= rx) return;
    const size_t sy = (size_t)(g.ny + 1) * g.nx, sx = (size_t)g.ny * (g.nx + 1), sf = (size_t)ry * rx;
    float xiy, xix, d0, d1;
    backtrace<VARIANT, VMODE>(g, vy + b * sy, vx + b * sx, i, j, ry, rx, s_py[threadIdx.y], s_px[threadIdx.x], xiy, xix);
    out[b * sf + (size_t)i * rx + j] = sample2d<FMODE, false>(field + b * sf, ry, rx, xiy, xix, d0, d1);
}

// ---- fused staggered force loss (control/incompressible_flow total_force_loss: l2_loss(semi_lagrangian(v1, v1, dt) - v2)) ----------------
// One launch per component: back-trace, sample, subtract the next frame's velocity, keep the residual (the adjoint's seed) and the block's
// sum of squares.  The advected field itself never reaches HBM.  Partials are re-reduced in a fixed order (no float atomics).
__device__ __forceinline__ float block_sum_2d(float v, float* sh) {
    const int t = threadIdx.y * 32 + threadIdx.x;
    v = warp_sum(v);
    if ((t & 31) == 0) sh[t >> 5] = v;
    __syncthreads();
    float r = 0.0f;
    if (t < 32) {
        r = t < 8 ? sh[t] : 0.0f;
        r = warp_sum(r);
    }
    return r;                                                         // valid in thread 0
}

template <int VARIANT, int M>
__global__ void __launch_bounds__(256) advect_residual_kernel(FluidGeo g, const float* __restrict__ vy, const float* __restrict__ vx,
                                                              const float* __restrict__ target, float* __restrict__ resid,
                                                              float* __restrict__ partials) {
    int ry, rx;
    variant_res<VARIANT>(g.ny, g.nx, ry, rx);
    const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y, b = blockIdx.z;
    __shared__ float s_py[8], s_px[32], sh[8];
    block_points(ry, rx, g.clo[VARIANT - 1], g.csz[VARIANT - 1], s_py, s_px);
    float d = 0.0f;
    if (i < ry && j < rx) {
        const size_t sy = (size_t)(g.ny + 1) * g.nx, sx = (size_t)g.ny * (g.nx + 1), sf = (size_t)ry * rx;
        const float* field = (VARIANT == V_COMP_Y ? vy + b * sy : vx + b * sx);
        float xiy, xix, d0, d1;
        backtrace<VARIANT, M>(g, vy + b * sy, vx + b * sx, i, j, ry, rx, s_py[threadIdx.y], s_px[threadIdx.x], xiy, xix);
        const size_t at = b * sf + (size_t)i * rx + j;
        d = sample2d<M, false>(field, ry, rx, xiy, xix, d0, d1) - __ldg(target + at);
        resid[at] = d;
    }
    const float s = block_sum_2d(d * d, sh);
    if (threadIdx.x == 0 && threadIdx.y == 0) partials[((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = s;
}

// loss = sum(partials of component y) / 2 / B + sum(partials of component x) / 2 / B, each sum in float64 and a fixed order
__global__ void __launch_bounds__(1024) force_loss_finish_kernel(const float* __restrict__ partials, int n0, int n1, int B, float* __restrict__ loss) {
    __shared__ double sh[2][1024];
    double a0 = 0.0, a1 = 0.0;
    for (int i = threadIdx.x; i < n0; i += 1024) a0 += (double)partials[i];
    for (int i = threadIdx.x; i < n1; i += 1024) a1 += (double)partials[n0 + i];
    sh[0][threadIdx.x] = a0; sh[1][threadIdx.x] = a1;
    __syncthreads();
    for (int w = 512; w > 0; w >>= 1) {
        if (threadIdx.x < w) { sh[0][threadIdx.x] += sh[0][threadIdx.x + w]; sh[1][threadIdx.x] += sh[1][threadIdx.x + w]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) loss[0] = (float)(sh[0][0] / 2.0 / (double)B) + (float)(sh[1][0] / 2.0 / (double)B);
}

// seeds of the adjoint: g_a = resid * (g_loss / B) for both components, the target's gradient -g_a (optional), and max|g_a| per
// component into header[0], header[1] (the fixed-point scale of the adjoint that follows: saves its own reduction pass)
__global__ void __launch_bounds__(256) force_loss_seed_kernel(const float* __restrict__ ry_, const float* __restrict__ rx_, const float* __restrict__ g_loss,
                                                              float inv_B, float* __restrict__ gy, float* __restrict__ gx, float* __restrict__ g_ty,
                                                              float* __restrict__ g_tx, size_t n0, size_t n1, unsigned* __restrict__ header) {
    const float k = g_loss[0] * inv_B;
    float m0 = 0.0f, m1 = 0.0f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n0 + n1; i += (size_t)gridDim.x * blockDim.x) {
        if (i < n0) {
            const float v = ry_[i] * k;
            gy[i] = v;
            if (g_ty) g_ty[i] = -v;
            const float a = fabsf(v);
            if (a == a && a < 3.0e38f) m0 = fmaxf(m0, a);
        } else {
            const float v = rx_[i - n0] * k;
            gx[i - n0] = v;
            if (g_tx) g_tx[i - n0] = -v;
            const float a = fabsf(v);
            if (a == a && a < 3.0e38f) m1 = fmaxf(m1, a);
        }
    }
    m0 = warp_max(m0); m1 = warp_max(m1);
    __shared__ float sm[2][8];
    if ((threadIdx.x & 31) == 0) { sm[0][threadIdx.x >> 5] = m0; sm[1][threadIdx.x >> 5] = m1; }
    __syncthreads();
    if (threadIdx.x < 8) {
        m0 = sm[0][threadIdx.x]; m1 = sm[1][threadIdx.x];
        for (int o = 4; o > 0; o >>= 1) { m0 = fmaxf(m0, __shfl_xor_sync(0xffu, m0, o)); m1 = fmaxf(m1, __shfl_xor_sync(0xffu, m1, o)); }
        if (threadIdx.x == 0) { atomicMax(header, __float_as_uint(m0)); atomicMax(header + 1, __float_as_uint(m1)); }
    }
}

// ---- adjoint pass A: scatter of the output gradient through the sampling, coordinate gradients ----------------------------
// The transpose of the bilinear sampling is a scatter to data-dependent cells.  Every sample adds its four weighted shares of g_out
// to 64-bit FIXED-POINT accumulators with integer atomics (pdec_common.cuh: scale from max|g_out|, integer adds commute): the sums do
// not depend on the order the atomics land in -- bit-reproducible, no float atomics, no bound on the displacement, one pass.
// (Rounds 1-2a gathered instead: coordinates written out, then every cell summed hat weights over a window bounded by the largest
// displacement; 2.2 x the time of this at 128 x 128 x 64, profiles/r2af_advection_adjoint.md.)
constexpr int ADV_BITS = 40;                 // |share| <= max|g_out| -> below 2^40; 2^22 shares per cell still fit the 64-bit sum

// max |g| of up to two arrays -> header[0], header[1] (bits of non-negative floats order like unsigned integers)
__global__ void __launch_bounds__(256) advect_absmax_kernel(const float* __restrict__ g0, long long n0, const float* __restrict__ g1, long long n1,
                                                            unsigned* __restrict__ header) {
    // NaN / inf gradients do not set the scale.  16 bytes per load where the array allows it (the max does not depend on the order)
    auto scan = [&](const float* __restrict__ g, long long n) -> float {
        float m = 0.0f;
        auto take = [&](float x) { const float v = fabsf(x); if (v == v && v < 3.0e38f) m = fmaxf(m, v); };
        const long long tid = (long long)blockIdx.x * 256 + threadIdx.x, step = (long long)gridDim.x * 256;
        long long done = 0;
        if ((reinterpret_cast<uintptr_t>(g) & 15) == 0) {
            const float4* g4 = reinterpret_cast<const float4*>(g);
            for (long long q = tid; q < n / 4; q += step) { const float4 v = g4[q]; take(v.x); take(v.y); take(v.z); take(v.w); }
            done = (n / 4) * 4;
        }
        for (long long q = done + tid; q < n; q += step) take(g[q]);
        return m;
    };
    float m0 = scan(g0, n0), m1 = g1 != nullptr ? scan(g1, n1) : 0.0f;
    m0 = warp_max(m0); m1 = warp_max(m1);
    __shared__ float sm[2][8];
    if ((threadIdx.x & 31) == 0) { sm[0][threadIdx.x >> 5] = m0; sm[1][threadIdx.x >> 5] = m1; }
    __syncthreads();
    if (threadIdx.x < 8) {
        m0 = sm[0][threadIdx.x]; m1 = sm[1][threadIdx.x];
        for (int o = 4; o > 0; o >>= 1) { m0 = fmaxf(m0, __shfl_xor_sync(0xffu, m0, o)); m1 = fmaxf(m1, __shfl_xor_sync(0xffu, m1, o)); }
        if (threadIdx.x == 0) {
            atomicMax(header, __float_as_uint(m0));
            if (g1 != nullptr) atomicMax(header + 1, __float_as_uint(m1));
        }
    }
}


// acc (may be null: field gradient not wanted) <- shares of g_out; GY / GX <- g_out * d(sample)/d(xi) * d(xi)/d(u): the gradient w.r.t.
// the velocity (uy, ux) AT the sample point, brought to the faces by pass B
template <int VARIANT, int FMODE, int VMODE>
__device__ __forceinline__ void advect_bwd_scatter_body(const FluidGeo& g, const float* __restrict__ field, const float* __restrict__ vy,
                                                        const float* __restrict__ vx, const float* __restrict__ gout,
                                                        unsigned long long* __restrict__ acc, float* __restrict__ GY, float* __restrict__ GX,
                                                        const unsigned* __restrict__ header, int b, float* s_py, float* s_px) {
    int ry, rx;
    variant_res<VARIANT>(g.ny, g.nx, ry, rx);
    const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y;
    const size_t sy = (size_t)(g.ny + 1) * g.nx, sx = (size_t)g.ny * (g.nx + 1), sf = (size_t)ry * rx;
    block_points(ry, rx, (VARIANT == V_CENTERED) ? g.lo : g.clo[VARIANT - 1], (VARIANT == V_CENTERED) ? g.sz : g.csz[VARIANT - 1], s_py, s_px);
    if (i >= ry || j >= rx) return;
    float xiy, xix;
    backtrace<VARIANT, VMODE>(g, vy + b * sy, vx + b * sx, i, j, ry, rx, s_py[threadIdx.y], s_px[threadIdx.x], xiy, xix);
    int fy, cy, fx, cx;
    float w0y, w1y, w0x, w1x;
    interp_setup<FMODE>(xiy, ry, fy, cy, w0y, w1y);
    interp_setup<FMODE>(xix, rx, fx, cx, w0x, w1x);
    const float* f = field + b * sf;
    const float s00 = __ldg(f + (size_t)fy * rx + fx), s01 = __ldg(f + (size_t)fy * rx + cx);
    const float s10 = __ldg(f + (size_t)cy * rx + fx), s11 = __ldg(f + (size_t)cy * rx + cx);
    const float dy = (s10 - s00) * w1x + (s11 - s01) * w0x;           // as sample2d<FMODE, true>
    const float dx = (s01 * w1y + s11 * w0y) - (s00 * w1y + s10 * w0y);
    const size_t o = b * sf + (size_t)i * rx + j;
    const float go = gout[o];
    const float *gf = g.gfac[VARIANT];                               // -(dt * res) / size per axis: through xi = ((p - u*dt - lower)/size)*res - 0.5
    GY[o] = go * dy * gf[0];
    GX[o] = go * dx * gf[1];
    if (acc != nullptr) {
        const float scale = fixed_scale(header, ADV_BITS);
        if (scale > 0.0f) {
            // transpose of out = (s00 w1y + s10 w0y) w1x + (s01 w1y + s11 w0y) w0x; an out-of-domain 'constant' sample has both taps of
            // an axis on one cell with weights w and -w: its two shares cancel exactly in fixed point, like its value is exactly +0
            unsigned long long* a = acc + b * sf;
            fixed_add(a + (size_t)fy * rx + fx, go * w1x * w1y, scale);
            fixed_add(a + (size_t)cy * rx + fx, go * w1x * w0y, scale);
            fixed_add(a + (size_t)fy * rx + cx, go * w0x * w1y, scale);
            fixed_add(a + (size_t)cy * rx + cx, go * w0x * w0y, scale);
        }
    }
}

template <int VARIANT, int FMODE, int VMODE>
__global__ void __launch_bounds__(256) advect_bwd_scatter_kernel(FluidGeo g, const float* __restrict__ field, const float* __restrict__ vy,
                                                                 const float* __restrict__ vx, const float* __restrict__ gout,
                                                                 unsigned long long* __restrict__ acc, float* __restrict__ GY,
                                                                 float* __restrict__ GX, const unsigned* __restrict__ header) {
    __shared__ float s_py[8], s_px[32];
    advect_bwd_scatter_body<VARIANT, FMODE, VMODE>(g, field, vy, vx, gout, acc, GY, GX, header, blockIdx.z, s_py, s_px);
}

// both components of the staggered self-advection in one launch: blockIdx.z < B -> the vy grid, else the vx grid (the tile grid covers
// (ny + 1) x (nx + 1); blocks beyond a component's own extent leave at once)
template <int M>
__global__ void __launch_bounds__(256) advect_bwd_scatter_staggered_kernel(FluidGeo g, const float* __restrict__ vy, const float* __restrict__ vx,
                                                                           const float* __restrict__ g_ay, const float* __restrict__ g_ax,
                                                                           unsigned long long* __restrict__ acc0, unsigned long long* __restrict__ acc1,
                                                                           float* __restrict__ G0, float* __restrict__ G1,
                                                                           const unsigned* __restrict__ header, int B) {
    __shared__ float s_py[8], s_px[32];
    const size_t sy = (size_t)B * (g.ny + 1) * g.nx, sx = (size_t)B * g.ny * (g.nx + 1);
    if ((int)blockIdx.z < B) advect_bwd_scatter_body<V_COMP_Y, M, M>(g, vy, vy, vx, g_ay, acc0, G0, G0 + sy, header, blockIdx.z, s_py, s_px);
    else advect_bwd_scatter_body<V_COMP_X, M, M>(g, vx, vy, vx, g_ax, acc1, G1, G1 + sx, header + 1, (int)blockIdx.z - B, s_py, s_px);
}

// transpose of "component brought to the other component's faces" (two nested 2-point lerps with pad mode)
// G0X: coordinate gradient d/d(ux) at vy faces [ny+1, nx]  ->  contribution to g_vx[r, c]  (vx is [ny, nx+1])
template <int VMODE>
__device__ __forceinline__ float cross_to_vx(const FluidGeo& g, const float* __restrict__ G0X, int r, int c) {
    const int ny = g.ny, nx = g.nx;
    // g_t[i, c] = wu_x * G0X[i, c-1] + wl_x * G0X[i, c]   (i in [0, ny], c in [0, nx])
    auto gt = [&](int i) -> float {
        float v = 0.0f;
        if (c >= 1) v += g.s_wu[0][1] * __ldg(G0X + (size_t)i * nx + c - 1);
        if (c <= nx - 1) v += g.s_wl[0][1] * __ldg(G0X + (size_t)i * nx + c);
        return v;
    };
    // t[i, c] = vx(i, c) * wu_y + vx(i-1, c) * wl_y with the row index mapped by VMODE
    float acc = 0.0f;
    if (VMODE == PDEC_MODE_CIRCULAR) {
        acc = g.s_wu[0][0] * gt(r) + g.s_wl[0][0] * gt(r + 1);
        if (r == 0) acc += g.s_wu[0][0] * gt(ny);          // i = ny reads vx(ny) -> row 0
        if (r == ny - 1) acc += g.s_wl[0][0] * gt(0);      // i = 0 reads vx(-1) -> row ny-1
        return acc;
    }
    acc = g.s_wu[0][0] * gt(r) + g.s_wl[0][0] * gt(r + 1);
    if (VMODE == PDEC_MODE_REPLICATE) {
        if (r == 0) acc += g.s_wl[0][0] * gt(0);           // vx(-1) -> row 0
        if (r == ny - 1) acc += g.s_wu[0][0] * gt(ny);     // vx(ny) -> row ny-1
    }
    return acc;
}
// G1Y: coordinate gradient d/d(uy) at vx faces [ny, nx+1]  ->  contribution to g_vy[r, c]  (vy is [ny+1, nx])
template <int VMODE>
__device__ __forceinline__ float cross_to_vy(const FluidGeo& g, const float* __restrict__ G1Y, int r, int c) {
    const int ny = g.ny, nx = g.nx;
    // uy(i, j) = T(i, j) * wu_x + T(i, j-1) * wl_x,  T(i, c) = vy[i+1, c] * wu_y + vy[i, c] * wl_y, column mapped by VMODE
    // g_T[i, c] = wu_x * G1Y[i, c] + wl_x * G1Y[i, c+1] (+ mode terms)
    auto gT = [&](int i) -> float {
        float v = g.s_wu[1][1] * __ldg(G1Y + (size_t)i * (nx + 1) + c) + g.s_wl[1][1] * __ldg(G1Y + (size_t)i * (nx + 1) + c + 1);
        if (VMODE == PDEC_MODE_REPLICATE) {
            if (c == 0) v += g.s_wl[1][1] * __ldg(G1Y + (size_t)i * (nx + 1));               // j = 0 reads T(i, -1) -> col 0
            if (c == nx - 1) v += g.s_wu[1][1] * __ldg(G1Y + (size_t)i * (nx + 1) + nx);     // j = nx reads T(i, nx) -> col nx-1
        } else if (VMODE == PDEC_MODE_CIRCULAR) {
            if (c == nx - 1) v += g.s_wl[1][1] * __ldg(G1Y + (size_t)i * (nx + 1));          // j = 0 reads T(i, -1) -> col nx-1
            if (c == 0) v += g.s_wu[1][1] * __ldg(G1Y + (size_t)i * (nx + 1) + nx);          // j = nx reads T(i, nx) -> col 0
        }
        return v;
    };
    float acc = 0.0f;
    if (r >= 1) acc += g.s_wu[1][0] * gT(r - 1);
    if (r <= ny - 1) acc += g.s_wl[1][0] * gT(r);
    return acc;
}

// centred advection, pass B: g_rho out of the fixed-point sums, and the velocity gradient from the coordinate gradients
template <int FMODE>
__global__ void __launch_bounds__(256) advect_centered_bwd_kernel(FluidGeo g, const unsigned long long* __restrict__ acc, const float* __restrict__ GY,
                                                                  const float* __restrict__ GX, const unsigned* __restrict__ header,
                                                                  float* __restrict__ g_rho, float* g_vy, float* g_vx,
                                                                  const float* add_vy, const float* add_vx) {
    // add_vy / add_vx (optional, same shapes as g_vy / g_vx, may alias them): gradients that reached the velocity by other routes
    const int ny = g.ny, nx = g.nx;
    const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y, b = blockIdx.z;
    const size_t sf = (size_t)ny * nx;
    if (g_rho != nullptr && i < ny && j < nx) {
        const size_t at = b * sf + (size_t)i * nx + j;
        g_rho[at] = fixed_to_float(acc[at], fixed_inv_scale(header, ADV_BITS));
    }
    // velocity at centres = vy[i+1]*wu + vy[i]*wl  ->  g_vy[i] = wu * GY[i-1] + wl * GY[i]
    if (g_vy != nullptr && i <= ny && j < nx) {
        float v = 0.0f;
        if (i >= 1) v += g.c_wu[0] * __ldg(GY + b * sf + (size_t)(i - 1) * nx + j);
        if (i <= ny - 1) v += g.c_wl[0] * __ldg(GY + b * sf + (size_t)i * nx + j);
        const size_t at = b * (size_t)(ny + 1) * nx + (size_t)i * nx + j;
        g_vy[at] = add_vy != nullptr ? add_vy[at] + v : v;
    }
    if (g_vx != nullptr && i < ny && j <= nx) {
        float v = 0.0f;
        if (j >= 1) v += g.c_wu[1] * __ldg(GX + b * sf + (size_t)i * nx + j - 1);
        if (j <= nx - 1) v += g.c_wl[1] * __ldg(GX + b * sf + (size_t)i * nx + j);
        const size_t at = b * (size_t)ny * (nx + 1) + (size_t)i * (nx + 1) + j;
        g_vx[at] = add_vx != nullptr ? add_vx[at] + v : v;
    }
}

// staggered self-advection, pass B: the field gradient out of the fixed-point sums plus the three velocity routes
template <int VMODE>
__global__ void __launch_bounds__(256) advect_staggered_bwd_kernel(FluidGeo g, const unsigned long long* __restrict__ acc0, const unsigned long long* __restrict__ acc1,
                                                                   const float* __restrict__ G0, const float* __restrict__ G1,
                                                                   const unsigned* __restrict__ header, float* __restrict__ g_vy,
                                                                   float* __restrict__ g_vx, int accumulate, int B) {
    const int ny = g.ny, nx = g.nx;
    const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y, b = blockIdx.z;
    const size_t sy = (size_t)(ny + 1) * nx, sx = (size_t)ny * (nx + 1);
    // G0 = {GY, GX} of the vy grid, G1 of the vx grid (arrays of B elements each)
    const float *GY0 = G0 + b * sy, *GX0 = G0 + (size_t)B * sy + b * sy, *GY1 = G1 + b * sx, *GX1 = G1 + (size_t)B * sx + b * sx;
    if (i <= ny && j < nx) {
        float v = fixed_to_float(acc0[b * sy + (size_t)i * nx + j], fixed_inv_scale(header, ADV_BITS));
        v += __ldg(GY0 + (size_t)i * nx + j);               // uy = vy itself at vy faces
        v += cross_to_vy<VMODE>(g, GY1, i, j);              // vy averaged to vx faces
        float* p = g_vy + b * sy + (size_t)i * nx + j;
        *p = accumulate ? *p + v : v;
    }
    if (i < ny && j <= nx) {
        float v = fixed_to_float(acc1[b * sx + (size_t)i * (nx + 1) + j], fixed_inv_scale(header + 1, ADV_BITS));
        v += __ldg(GX1 + (size_t)i * (nx + 1) + j);
        v += cross_to_vx<VMODE>(g, GX0, i, j);
        float* p = g_vx + b * sx + (size_t)i * (nx + 1) + j;
        *p = accumulate ? *p + v : v;
    }
}

// ---- staggered curl (util.py:60-67) -------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(256) curl_fwd_kernel(const float* __restrict__ phi, float* __restrict__ vy, float* __restrict__ vx, int ny, int nx) {
    const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y, b = blockIdx.z;
    if (i > ny || j > nx) return;
    const float* f = phi + (size_t)b * ny * nx;
    const float c = load_mapped<MODE>(f, i, j, ny, nx);
    if (j < nx) vy[(size_t)b * (ny + 1) * nx + (size_t)i * nx + j] = c - load_mapped<MODE>(f, i, j + 1, ny, nx);
    if (i < ny) vx[(size_t)b * ny * (nx + 1) + (size_t)i * (nx + 1) + j] = load_mapped<MODE>(f, i + 1, j, ny, nx) - c;
}

// staggered_curl_2d + advect.semi_lagrangian(density, curl, dt) in one launch (ControlledIncompressibleFlow.step,
// src/control/pde/incompressible_flow.py:53-59): a thread owns lattice point (i, j), writes the two face velocities it owns and, inside
// the domain, advects the cell: the four face values of the cell come from the potential directly (same expressions as curl_fwd_kernel,
// same bits), so the velocity is written for the later force loss / adjoint but never read back here.
template <int PMODE, int FMODE>
__global__ void __launch_bounds__(256) curl_advect_fwd_kernel(FluidGeo g, const float* __restrict__ rho, const float* __restrict__ phi,
                                                              float* __restrict__ vy, float* __restrict__ vx, float* __restrict__ out) {
    const int ny = g.ny, nx = g.nx;
    const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y, b = blockIdx.z;
    __shared__ float s_py[8], s_px[32];
    block_points(ny, nx, g.lo, g.sz, s_py, s_px);
    if (i > ny || j > nx) return;
    const float* f = phi + (size_t)b * ny * nx;
    const float c = load_mapped<PMODE>(f, i, j, ny, nx);
    const float pr = load_mapped<PMODE>(f, i, j + 1, ny, nx), pd = load_mapped<PMODE>(f, i + 1, j, ny, nx);
    const float vy0 = c - pr, vx0 = pd - c;
    if (j < nx) vy[(size_t)b * (ny + 1) * nx + (size_t)i * nx + j] = vy0;
    if (i < ny) vx[(size_t)b * ny * (nx + 1) + (size_t)i * (nx + 1) + j] = vx0;
    if (i >= ny || j >= nx) return;
    const float pdr = load_mapped<PMODE>(f, i + 1, j + 1, ny, nx);
    const float vy1 = pd - pdr, vx1 = pdr - pr;                        // vy[i + 1][j], vx[i][j + 1]
    const float uy = vy1 * g.c_wu[0] + vy0 * g.c_wl[0];                // StaggeredGrid.at(centres), as velocity_at<V_CENTERED>
    const float ux = vx1 * g.c_wu[1] + vx0 * g.c_wl[1];
    const float y = s_py[threadIdx.y] - uy * g.dt, x = s_px[threadIdx.x] - ux * g.dt;
    const float xiy = index_coord_fast(y, g.lo[0], g.sz[0], g.inv[0], (float)ny);
    const float xix = index_coord_fast(x, g.lo[1], g.sz[1], g.inv[1], (float)nx);
    float d0, d1;
    const size_t sf = (size_t)ny * nx;
    out[b * sf + (size_t)i * nx + j] = sample2d<FMODE, false>(rho + b * sf, ny, nx, xiy, xix, d0, d1);
}

template <int MODE>
__global__ void __launch_bounds__(256) curl_bwd_kernel(const float* __restrict__ g_vy, const float* __restrict__ g_vx, float* __restrict__ g_phi, int ny, int nx) {
    const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y, b = blockIdx.z;
    if (i >= ny || j >= nx) return;
    const float* gy = g_vy + (size_t)b * (ny + 1) * nx;
    const float* gx = g_vx + (size_t)b * ny * (nx + 1);
    // vy[a, c] = phi(m(a), m(c)) - phi(m(a), m(c+1)),  a in [0, ny], c in [0, nx-1]
    // vx[a, c] = phi(m(a+1), m(c)) - phi(m(a), m(c)),  a in [0, ny-1], c in [0, nx]
    float acc = 0.0f;
    for (int pass = 0; pass < 2; ++pass) {
        // rows a of vy whose mapped row is i: a == i always; a == ny if replicate and i == ny-1 / circular and i == 0
        int a = i;
        if (pass == 1) {
            if ((MODE == PDEC_MODE_REPLICATE && i == ny - 1) || (MODE == PDEC_MODE_CIRCULAR && i == 0)) a = ny; else break;
        }
        acc += __ldg(gy + (size_t)a * nx + j);                                                          // + term, c = j
        if (j >= 1) acc -= __ldg(gy + (size_t)a * nx + j - 1);                                          // - term, c + 1 = j
        if (MODE == PDEC_MODE_REPLICATE && j == nx - 1) acc -= __ldg(gy + (size_t)a * nx + nx - 1);     // c + 1 = nx -> nx-1
        if (MODE == PDEC_MODE_CIRCULAR && j == 0) acc -= __ldg(gy + (size_t)a * nx + nx - 1);           // c + 1 = nx -> 0
    }
    for (int pass = 0; pass < 2; ++pass) {
        // columns c of vx whose mapped column is j: c == j always; c == nx if replicate and j == nx-1 / circular and j == 0
        int c = j;
        if (pass == 1) {
            if ((MODE == PDEC_MODE_REPLICATE && j == nx - 1) || (MODE == PDEC_MODE_CIRCULAR && j == 0)) c = nx; else break;
        }
        acc -= __ldg(gx + (size_t)i * (nx + 1) + c);                                // - term, a = i
        if (i >= 1) acc += __ldg(gx + (size_t)(i - 1) * (nx + 1) + c);              // + term, a + 1 = i
        if (MODE == PDEC_MODE_REPLICATE && i == ny - 1) acc += __ldg(gx + (size_t)(ny - 1) * (nx + 1) + c);   // a + 1 = ny -> ny-1
        if (MODE == PDEC_MODE_CIRCULAR && i == 0) acc += __ldg(gx + (size_t)(ny - 1) * (nx + 1) + c);        // a + 1 = ny -> 0
    }
    g_phi[(size_t)b * ny * nx + (size_t)i * nx + j] = acc;
}

// ---- masked divergence (solver_api.py:111-121, staggered_grid.py:201-209) ----------------------------
// boundary faces of a SOLID side are zeroed (with_hard_boundary_conditions, solver_api.py:111-121); `closed` = 4-bit side mask
// (pdec_closed_sides): bits 0 / 1 = lower / upper side along y, bits 2 / 3 along x
__device__ __forceinline__ float face_mask_y(int idx, int n, unsigned closed) { return ((idx == 0 && (closed & 1u)) || (idx == n && (closed & 2u))) ? 0.0f : 1.0f; }
__device__ __forceinline__ float face_mask_x(int idx, int n, unsigned closed) { return ((idx == 0 && (closed & 4u)) || (idx == n && (closed & 8u))) ? 0.0f : 1.0f; }

__global__ void __launch_bounds__(256) mask_divergence_fwd_kernel(const float* __restrict__ vy, const float* __restrict__ vx, float* __restrict__ div,
                                                                  int ny, int nx, unsigned closed) {
    const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y, b = blockIdx.z;
    if (i >= ny || j >= nx) return;
    const float* y = vy + (size_t)b * (ny + 1) * nx;
    const float* x = vx + (size_t)b * ny * (nx + 1);
    const float dy = y[(size_t)(i + 1) * nx + j] * face_mask_y(i + 1, ny, closed) - y[(size_t)i * nx + j] * face_mask_y(i, ny, closed);
    const float dx = x[(size_t)i * (nx + 1) + j + 1] * face_mask_x(j + 1, nx, closed) - x[(size_t)i * (nx + 1) + j] * face_mask_x(j, nx, closed);
    div[(size_t)b * ny * nx + (size_t)i * nx + j] = dy + dx;
}

__global__ void __launch_bounds__(256) mask_divergence_bwd_kernel(const float* __restrict__ g_div, float* __restrict__ g_vy, float* __restrict__ g_vx,
                                                                  int ny, int nx, unsigned closed, int accumulate) {
    const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y, b = blockIdx.z;
    const float* d = g_div + (size_t)b * ny * nx;
    if (i <= ny && j < nx) {
        const float lo = (i >= 1) ? d[(size_t)(i - 1) * nx + j] : 0.0f, hi = (i < ny) ? d[(size_t)i * nx + j] : 0.0f;
        const float v = (lo - hi) * face_mask_y(i, ny, closed);
        float* p = g_vy + (size_t)b * (ny + 1) * nx + (size_t)i * nx + j;
        *p = accumulate ? *p + v : v;
    }
    if (i < ny && j <= nx) {
        const float lo = (j >= 1) ? d[(size_t)i * nx + j - 1] : 0.0f, hi = (j < nx) ? d[(size_t)i * nx + j] : 0.0f;
        const float v = (lo - hi) * face_mask_x(j, nx, closed);
        float* p = g_vx + (size_t)b * ny * (nx + 1) + (size_t)i * (nx + 1) + j;
        *p = accumulate ? *p + v : v;
    }
}

// ---- pressure-gradient subtraction (fluid.py:240-242, staggered_grid.py:219-231) ---------------------
__global__ void __launch_bounds__(256) subtract_gradient_fwd_kernel(const float* __restrict__ vy, const float* __restrict__ vx, const float* __restrict__ p,
                                                                    float* __restrict__ oy, float* __restrict__ ox, int ny, int nx, unsigned closed,
                                                                    float dx0, float dxy, float dxx, float iy, float ix) {
    // iy / ix: exact reciprocal of dxy / dxx when that is a power of two (x / 2^k == x * 2^-k bit for bit), else 0 -> divide
    const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y, b = blockIdx.z;
    const float* q = p + (size_t)b * ny * nx;
    if (i <= ny && j < nx) {
        const size_t o = (size_t)b * (ny + 1) * nx + (size_t)i * nx + j;
        const float m = face_mask_y(i, ny, closed);
        float gr = 0.0f;                                   // replicate pad: boundary-face gradient is 0
        if (i >= 1 && i < ny) { const float d = q[(size_t)i * nx + j] * dx0 - q[(size_t)(i - 1) * nx + j] * dx0; gr = iy != 0.0f ? d * iy : d / dxy; }
        oy[o] = vy[o] * m - gr * m;
    }
    if (i < ny && j <= nx) {
        const size_t o = (size_t)b * ny * (nx + 1) + (size_t)i * (nx + 1) + j;
        const float m = face_mask_x(j, nx, closed);
        float gr = 0.0f;
        if (j >= 1 && j < nx) { const float d = q[(size_t)i * nx + j] * dx0 - q[(size_t)i * nx + j - 1] * dx0; gr = ix != 0.0f ? d * ix : d / dxx; }
        ox[o] = vx[o] * m - gr * m;
    }
}

__global__ void __launch_bounds__(256) subtract_gradient_bwd_kernel(const float* __restrict__ g_oy, const float* __restrict__ g_ox, float* __restrict__ g_vy,
                                                                    float* __restrict__ g_vx, float* __restrict__ g_p, int ny, int nx, unsigned closed,
                                                                    float dx0, float dxy, float dxx, int accumulate, float iy, float ix) {
    const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y, b = blockIdx.z;
    const float* gy = g_oy + (size_t)b * (ny + 1) * nx;
    const float* gx = g_ox + (size_t)b * ny * (nx + 1);
    if (g_vy != nullptr && i <= ny && j < nx) {
        const float v = gy[(size_t)i * nx + j] * face_mask_y(i, ny, closed);
        float* p = g_vy + (size_t)b * (ny + 1) * nx + (size_t)i * nx + j;
        *p = accumulate ? *p + v : v;
    }
    if (g_vx != nullptr && i < ny && j <= nx) {
        const float v = gx[(size_t)i * (nx + 1) + j] * face_mask_x(j, nx, closed);
        float* p = g_vx + (size_t)b * ny * (nx + 1) + (size_t)i * (nx + 1) + j;
        *p = accumulate ? *p + v : v;
    }
    if (g_p != nullptr && i < ny && j < nx) {
        // interior faces only: face i (1..ny-1) contributes -g/dx to p[i] and +g/dx to p[i-1]
        float acc = 0.0f;
        auto by = [&](float v) { return iy != 0.0f ? v * iy : v / dxy; };
        auto bx = [&](float v) { return ix != 0.0f ? v * ix : v / dxx; };
        if (i >= 1) acc -= by(gy[(size_t)i * nx + j]);
        if (i + 1 < ny) acc += by(gy[(size_t)(i + 1) * nx + j]);
        if (j >= 1) acc -= bx(gx[(size_t)i * (nx + 1) + j]);
        if (j + 1 < nx) acc += bx(gx[(size_t)i * (nx + 1) + j + 1]);
        g_p[(size_t)b * ny * nx + (size_t)i * nx + j] = acc * dx0;
    }
}

// ---- density normalisation (grid.py:205-209, nd.py:28-40) ---------------------------------------------
// stage 1: per (block, b) partial sums of up to three products, fixed order; stage 2 re-reduces the partials.
__device__ __forceinline__ float block_sum_256(float v, float* sh) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5;
    if ((threadIdx.x & 31) == 0) sh[w] = v;
    __syncthreads();
    float r = 0.0f;
    if (threadIdx.x < 32) {
        r = (threadIdx.x < (blockDim.x >> 5)) ? sh[threadIdx.x] : 0.0f;
        r = warp_sum(r);
        if (threadIdx.x == 0) sh[0] = r;
    }
    __syncthreads();
    r = sh[0];
    __syncthreads();
    return r;
}

// partials[(k * B + b) * nblk + blk], k = 0: sum a, 1: sum c, 2: sum a*c   (c may be null)
__global__ void __launch_bounds__(256) sums_kernel(const float* __restrict__ a, const float* __restrict__ c, float* __restrict__ partials, int n, int B) {
    __shared__ float sh[32];
    const int b = blockIdx.y, nblk = gridDim.x;
    const float* pa = a + (size_t)b * n;
    const float* pc = c ? c + (size_t)b * n : nullptr;
    float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += nblk * blockDim.x) {
        const float x = pa[i], y = pc ? pc[i] : 0.0f;
        s0 += x; s1 += y; s2 += x * y;
    }
    s0 = block_sum_256(s0, sh); s1 = block_sum_256(s1, sh); s2 = block_sum_256(s2, sh);
    if (threadIdx.x == 0) {
        partials[((size_t)0 * B + b) * nblk + blockIdx.x] = s0;
        partials[((size_t)1 * B + b) * nblk + blockIdx.x] = s1;
        partials[((size_t)2 * B + b) * nblk + blockIdx.x] = s2;
    }
}
__device__ __forceinline__ float reduce_partials(const float* __restrict__ partials, int k, int B, int b, int nblk) {
    float s = 0.0f;
    for (int i = 0; i < nblk; ++i) s += partials[((size_t)k * B + b) * nblk + i];
    return s;
}
// out = target * (sum(source) / max(sum(target), eps));  partials from sums_kernel(a = target, c = source)
__global__ void __launch_bounds__(256) normalize_fwd_kernel(const float* __restrict__ target, const float* __restrict__ partials, float* __restrict__ out,
                                                            int n, int B, int nblk, float eps) {
    const int b = blockIdx.y;
    const float T = reduce_partials(partials, 0, B, b, nblk), S = reduce_partials(partials, 1, B, b, nblk);
    const float scale = S / fmaxf(T, eps);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[(size_t)b * n + i] = target[(size_t)b * n + i] * scale;
}
// p1 = sums(target, source) ; p2 = sums(g_out, target)[k=2] = sum(g_out * target)
__global__ void __launch_bounds__(256) normalize_bwd_kernel(const float* __restrict__ g_out, const float* __restrict__ p1, const float* __restrict__ p2,
                                                            float* __restrict__ g_target, float* __restrict__ g_source, int n, int B, int nblk, float eps) {
    const int b = blockIdx.y;
    const float T = reduce_partials(p1, 0, B, b, nblk), S = reduce_partials(p1, 1, B, b, nblk), D = reduce_partials(p2, 2, B, b, nblk);
    const float den = fmaxf(T, eps);
    const float scale = S / den;
    const float dT = (T > eps) ? -(D * S) / (den * den) : 0.0f;
    const float dS = D / den;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        if (g_target) g_target[(size_t)b * n + i] = g_out[(size_t)b * n + i] * scale + dT;
        if (g_source) g_source[(size_t)b * n + i] = dS;
    }
}

}  // namespace pdec

// ======================================================================================================
// host side
// ======================================================================================================
using namespace pdec;

// upper_weight of CenteredGrid.at's fast path (grid.py:112-119), computed like the reference:
// float32 box arithmetic, float64 product with the resolution, % 1.0; weights cast to float32.
static void half_weights(float self_lower, float self_size, int self_res, float other_lower, float& wu, float& wl) {
    const float origin32 = (other_lower - self_lower) / self_size;
    const double origin = (double)origin32 * (double)self_res;
    double w = fmod(origin, 1.0);
    if (w < 0) w += 1.0;
    wu = (float)w;
    wl = (float)(1.0 - w);
}

// 1 / v when v is a power of two (then x / v == x * (1 / v) bit for bit), else 0
static float pow2_reciprocal(float v) {
    int e;
    return (v > 0.0f && frexpf(v, &e) == 0.5f && e > -100 && e < 100) ? 1.0f / v : 0.0f;
}

static int make_geo(FluidGeo& g, int B, int ny, int nx, double dt, float lower_y, float lower_x, float size_y, float size_x) {
    if (B < 0 || ny < 2 || nx < 2 || !(size_y > 0.0f) || !(size_x > 0.0f)) return PDEC_ERR_INVALID_ARGUMENT;
    if (B > 65535) return PDEC_ERR_UNSUPPORTED_SIZE;
    g.ny = ny; g.nx = nx; g.dt = (float)dt;
    g.lo[0] = lower_y; g.lo[1] = lower_x; g.sz[0] = size_y; g.sz[1] = size_x;
    const int res[2] = {ny, nx};
    double dx[2];
    for (int a = 0; a < 2; ++a) dx[a] = (double)g.sz[a] / (double)res[a];
    for (int c = 0; c < 2; ++c)
        for (int a = 0; a < 2; ++a) {
            const double unit = (a == c) ? dx[a] : 0.0;
            const float up = g.lo[a] + g.sz[a];
            const float lo_c = (float)((double)g.lo[a] - unit / 2), up_c = (float)((double)up + unit / 2);
            g.clo[c][a] = lo_c;
            g.csz[c][a] = up_c - lo_c;
        }
    auto exact_inv = [](float v) { return pow2_reciprocal(v); };
    for (int a = 0; a < 2; ++a) {
        g.inv[a] = exact_inv(g.sz[a]);
        for (int c = 0; c < 2; ++c) g.cinv[c][a] = exact_inv(g.csz[c][a]);
    }
    for (int v = 0; v < 3; ++v)
        for (int a = 0; a < 2; ++a) {
            const int r = res[a] + ((v == 1 && a == 0) || (v == 2 && a == 1) ? 1 : 0);
            const float size = v == 0 ? g.sz[a] : g.csz[v - 1][a];
            g.gfac[v][a] = -(g.dt * (float)r) / size;                // the expression the kernels evaluated per thread
        }
    for (int c = 0; c < 2; ++c) half_weights(g.clo[c][c], g.csz[c][c], res[c] + 1, g.lo[c], g.c_wu[c], g.c_wl[c]);
    for (int c = 0; c < 2; ++c) {
        const int d = 1 - c;   // the other component, padded by one cell on both sides along axis c
        for (int a = 0; a < 2; ++a) {
            float lo_p = g.clo[d][a], sz_p = g.csz[d][a];
            int res_p = res[a] + (a == d ? 1 : 0);
            if (a == c) {
                lo_p = (float)((double)g.clo[d][a] - dx[a]);
                sz_p = (float)((double)g.csz[d][a] + 2 * dx[a]);
                res_p += 2;
            }
            half_weights(lo_p, sz_p, res_p, g.clo[c][a], g.s_wu[c][a], g.s_wl[c][a]);
        }
    }
    return PDEC_OK;
}

static dim3 tiles(int ry, int rx, int B) { return dim3((rx + 31) / 32, (ry + 7) / 8, B); }
static const dim3 kBlock(32, 8, 1);

#define MODE_SWITCH(mode, CALL)                                   \
    do {                                                          \
        if ((mode) == PDEC_MODE_CONSTANT) { CALL(PDEC_MODE_CONSTANT); }       \
        else if ((mode) == PDEC_MODE_REPLICATE) { CALL(PDEC_MODE_REPLICATE); } \
        else { CALL(PDEC_MODE_CIRCULAR); }                        \
    } while (0)
static bool bad_mode(int m) { return m < 0 || m > 2; }

extern "C" int pdec_curl_2d_fwd(const float* phi, float* vy, float* vx, int B, int ny, int nx, int mode, void* stream) {
    if (B < 0 || ny < 2 || nx < 2 || bad_mode(mode) || B > 65535) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!phi || !vy || !vx) return PDEC_ERR_INVALID_ARGUMENT;
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(M) curl_fwd_kernel<M><<<tiles(ny + 1, nx + 1, B), kBlock, 0, st>>>(phi, vy, vx, ny, nx)
    MODE_SWITCH(mode, CALL);
#undef CALL
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" int pdec_curl_2d_bwd(const float* g_vy, const float* g_vx, float* g_phi, int B, int ny, int nx, int mode, void* stream) {
    if (B < 0 || ny < 2 || nx < 2 || bad_mode(mode) || B > 65535) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!g_vy || !g_vx || !g_phi) return PDEC_ERR_INVALID_ARGUMENT;
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(M) curl_bwd_kernel<M><<<tiles(ny, nx, B), kBlock, 0, st>>>(g_vy, g_vx, g_phi, ny, nx)
    MODE_SWITCH(mode, CALL);
#undef CALL
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" int pdec_advect_centered_2d_fwd(const float* rho, const float* vy, const float* vx, float* rho_out, int B, int ny, int nx,
                                           int mode, double dt, float lower_y, float lower_x, float size_y, float size_x, void* stream) {
    FluidGeo g;
    int rc = make_geo(g, B, ny, nx, dt, lower_y, lower_x, size_y, size_x);
    if (rc != PDEC_OK) return rc;
    if (bad_mode(mode)) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!rho || !vy || !vx || !rho_out) return PDEC_ERR_INVALID_ARGUMENT;
    cudaStream_t st = (cudaStream_t)stream;
    // the velocity is never extrapolated for centred sample points -> VMODE irrelevant (use 0)
#define CALL(M) advect_fwd_kernel<V_CENTERED, M, 0><<<tiles(ny, nx, B), kBlock, 0, st>>>(g, rho, vy, vx, rho_out)
    MODE_SWITCH(mode, CALL);
#undef CALL
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" int pdec_advect_staggered_2d_fwd(const float* vy, const float* vx, float* ay, float* ax, int B, int ny, int nx, int mode,
                                            double dt, float lower_y, float lower_x, float size_y, float size_x, void* stream) {
    FluidGeo g;
    int rc = make_geo(g, B, ny, nx, dt, lower_y, lower_x, size_y, size_x);
    if (rc != PDEC_OK) return rc;
    if (bad_mode(mode)) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!vy || !vx || !ay || !ax) return PDEC_ERR_INVALID_ARGUMENT;
    cudaStream_t st = (cudaStream_t)stream;
#define CALL(M)                                                                                              \
    advect_fwd_kernel<V_COMP_Y, M, M><<<tiles(ny + 1, nx, B), kBlock, 0, st>>>(g, vy, vy, vx, ay);           \
    advect_fwd_kernel<V_COMP_X, M, M><<<tiles(ny, nx + 1, B), kBlock, 0, st>>>(g, vx, vy, vx, ax)
    MODE_SWITCH(mode, CALL);
#undef CALL
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

// workspace layout for the advection adjoints: [header: max|g_out| of the two grids, 256-byte multiple]
// [fixed-point sums: u64 x B*(ny+1)*nx, then u64 x B*ny*(nx+1)] [GY, GX of the vy grid] [GY, GX of the vx grid]  (16 bytes per sample;
// the centred variant uses one grid of B*ny*nx: sums, then GY, GX)
static size_t ws_header(int B) { return (((size_t)4 * (B > 0 ? B : 1) * sizeof(int)) + 255) & ~(size_t)255; }
static int absmax_blocks(long long n) { const long long b = (n / 4 + 255) / 256; return (int)(b > 148 * 8 ? 148 * 8 : (b < 1 ? 1 : b)); }

extern "C" size_t pdec_advect_2d_workspace_bytes(int B, int ny, int nx) {
    const size_t b = B > 0 ? B : 1;
    return ws_header(B) + 4 * sizeof(float) * b * ((size_t)(ny + 1) * nx + (size_t)ny * (nx + 1));
}

extern "C" int pdec_advect_centered_2d_bwd(const float* rho, const float* vy, const float* vx, const float* g_out, float* g_rho,
                                           float* g_vy, float* g_vx, int accumulate_v, int B, int ny, int nx, int mode, double dt,
                                           float lower_y, float lower_x, float size_y, float size_x, void* workspace,
                                           size_t workspace_bytes, void* stream) {
    FluidGeo g;
    int rc = make_geo(g, B, ny, nx, dt, lower_y, lower_x, size_y, size_x);
    if (rc != PDEC_OK) return rc;
    if (bad_mode(mode)) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!rho || !vy || !vx || !g_out || !workspace) return PDEC_ERR_INVALID_ARGUMENT;
    if (workspace_bytes < pdec_advect_2d_workspace_bytes(B, ny, nx)) return PDEC_ERR_WORKSPACE_TOO_SMALL;
    cudaStream_t st = (cudaStream_t)stream;
    unsigned* header = reinterpret_cast<unsigned*>(workspace);
    const size_t sf = (size_t)B * ny * nx;
    unsigned long long* acc = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(workspace) + ws_header(B));
    float* G = reinterpret_cast<float*>(acc + sf);                     // GY, GX
    if (cudaMemsetAsync(workspace, 0, ws_header(B) + (g_rho ? sf * sizeof(unsigned long long) : 0), st) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
    if (g_rho) advect_absmax_kernel<<<absmax_blocks((long long)sf), 256, 0, st>>>(g_out, (long long)sf, nullptr, 0, header);
#define CALL(M)                                                                                                                     \
    advect_bwd_scatter_kernel<V_CENTERED, M, 0><<<tiles(ny, nx, B), kBlock, 0, st>>>(g, rho, vy, vx, g_out, g_rho ? acc : nullptr, G, G + sf, header); \
    advect_centered_bwd_kernel<M><<<tiles(ny + 1, nx + 1, B), kBlock, 0, st>>>(g, acc, G, G + sf, header, g_rho, g_vy, g_vx, accumulate_v ? g_vy : nullptr, accumulate_v ? g_vx : nullptr)
    MODE_SWITCH(mode, CALL);
#undef CALL
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

// header_ready: the caller has cleared header + sums and put max|g_ay|, max|g_ax| into the header already
static int advect_staggered_bwd_impl(const float* vy, const float* vx, const float* g_ay, const float* g_ax, float* g_vy, float* g_vx, int accumulate,
                                     int B, int ny, int nx, int mode, double dt, float lower_y, float lower_x, float size_y, float size_x,
                                     void* workspace, size_t workspace_bytes, void* stream, bool header_ready) {
    FluidGeo g;
    int rc = make_geo(g, B, ny, nx, dt, lower_y, lower_x, size_y, size_x);
    if (rc != PDEC_OK) return rc;
    if (bad_mode(mode)) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!vy || !vx || !g_ay || !g_ax || !g_vy || !g_vx || !workspace) return PDEC_ERR_INVALID_ARGUMENT;
    if (workspace_bytes < pdec_advect_2d_workspace_bytes(B, ny, nx)) return PDEC_ERR_WORKSPACE_TOO_SMALL;
    cudaStream_t st = (cudaStream_t)stream;
    unsigned* header = reinterpret_cast<unsigned*>(workspace);
    const size_t sy = (size_t)B * (ny + 1) * nx, sx = (size_t)B * ny * (nx + 1);
    unsigned long long* acc0 = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(workspace) + ws_header(B));
    unsigned long long* acc1 = acc0 + sy;
    float* G0 = reinterpret_cast<float*>(acc1 + sx);                   // GY, GX of the vy grid
    float* G1 = G0 + 2 * sy;                                           // GY, GX of the vx grid
    if (!header_ready) {
        if (cudaMemsetAsync(workspace, 0, ws_header(B) + (sy + sx) * sizeof(unsigned long long), st) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
        advect_absmax_kernel<<<absmax_blocks((long long)(sy > sx ? sy : sx)), 256, 0, st>>>(g_ay, (long long)sy, g_ax, (long long)sx, header);
    }
    const bool one_launch = 2 * B <= 65535;                            // grid.z limit
#define CALL(M)                                                                                                                              \
    if (one_launch) advect_bwd_scatter_staggered_kernel<M><<<tiles(ny + 1, nx + 1, 2 * B), kBlock, 0, st>>>(g, vy, vx, g_ay, g_ax, acc0, acc1, G0, G1, header, B); \
    else {                                                                                                                                   \
        advect_bwd_scatter_kernel<V_COMP_Y, M, M><<<tiles(ny + 1, nx, B), kBlock, 0, st>>>(g, vy, vy, vx, g_ay, acc0, G0, G0 + sy, header);  \
        advect_bwd_scatter_kernel<V_COMP_X, M, M><<<tiles(ny, nx + 1, B), kBlock, 0, st>>>(g, vx, vy, vx, g_ax, acc1, G1, G1 + sx, header + 1); \
    }                                                                                                                                        \
    advect_staggered_bwd_kernel<M><<<tiles(ny + 1, nx + 1, B), kBlock, 0, st>>>(g, acc0, acc1, G0, G1, header, g_vy, g_vx, accumulate, B)
    MODE_SWITCH(mode, CALL);
#undef CALL
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" int pdec_advect_staggered_2d_bwd(const float* vy, const float* vx, const float* g_ay, const float* g_ax, float* g_vy,
                                            float* g_vx, int accumulate, int B, int ny, int nx, int mode, double dt, float lower_y,
                                            float lower_x, float size_y, float size_x, void* workspace, size_t workspace_bytes,
                                            void* stream) {
    return advect_staggered_bwd_impl(vy, vx, g_ay, g_ax, g_vy, g_vx, accumulate, B, ny, nx, mode, dt, lower_y, lower_x, size_y, size_x, workspace,
                                     workspace_bytes, stream, false);
}

// ---- fused curl + density advection ----------------------------------------------------------------------
template <int PM>
static void launch_curl_advect(int fmode, const FluidGeo& g, const float* rho, const float* phi, float* vy, float* vx, float* out, int B, cudaStream_t st) {
#define CALL(M) curl_advect_fwd_kernel<PM, M><<<tiles(g.ny + 1, g.nx + 1, B), kBlock, 0, st>>>(g, rho, phi, vy, vx, out)
    MODE_SWITCH(fmode, CALL);
#undef CALL
}

extern "C" int pdec_curl_advect_2d_fwd(const float* rho, const float* phi, float* vy, float* vx, float* rho_out, int B, int ny, int nx,
                                       int potential_mode, int mode, double dt, float lower_y, float lower_x, float size_y, float size_x,
                                       void* stream) {
    FluidGeo g;
    int rc = make_geo(g, B, ny, nx, dt, lower_y, lower_x, size_y, size_x);
    if (rc != PDEC_OK) return rc;
    if (bad_mode(mode) || bad_mode(potential_mode)) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!rho || !phi || !vy || !vx || !rho_out) return PDEC_ERR_INVALID_ARGUMENT;
    cudaStream_t st = (cudaStream_t)stream;
    if (potential_mode == PDEC_MODE_CONSTANT) launch_curl_advect<PDEC_MODE_CONSTANT>(mode, g, rho, phi, vy, vx, rho_out, B, st);
    else if (potential_mode == PDEC_MODE_REPLICATE) launch_curl_advect<PDEC_MODE_REPLICATE>(mode, g, rho, phi, vy, vx, rho_out, B, st);
    else launch_curl_advect<PDEC_MODE_CIRCULAR>(mode, g, rho, phi, vy, vx, rho_out, B, st);
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

// adjoint: g_rho <- transpose of the sampling; the velocity gradient of the advection PLUS add_vy / add_vx (what reached the curl's
// velocity by other routes: the force-loss terms; either may be NULL) goes through the transpose of the curl into g_phi.  The summed
// face gradients live in the workspace (pdec_advect_2d_workspace_bytes has the room behind the four centred arrays).
extern "C" int pdec_curl_advect_2d_bwd(const float* rho, const float* vy, const float* vx, const float* g_out, const float* add_vy,
                                       const float* add_vx, float* g_rho, float* g_phi, int B, int ny, int nx, int potential_mode, int mode,
                                       double dt, float lower_y, float lower_x, float size_y, float size_x, void* workspace,
                                       size_t workspace_bytes, void* stream) {
    FluidGeo g;
    int rc = make_geo(g, B, ny, nx, dt, lower_y, lower_x, size_y, size_x);
    if (rc != PDEC_OK) return rc;
    if (bad_mode(mode) || bad_mode(potential_mode)) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!rho || !vy || !vx || !g_out || !g_phi || !workspace) return PDEC_ERR_INVALID_ARGUMENT;
    if (workspace_bytes < pdec_advect_2d_workspace_bytes(B, ny, nx)) return PDEC_ERR_WORKSPACE_TOO_SMALL;
    cudaStream_t st = (cudaStream_t)stream;
    unsigned* header = reinterpret_cast<unsigned*>(workspace);
    const size_t sf = (size_t)B * ny * nx;
    unsigned long long* acc = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(workspace) + ws_header(B));
    float* G = reinterpret_cast<float*>(acc + sf);                     // GY, GX
    float* t_vy = G + 2 * sf;                                          // B * (ny + 1) * nx
    float* t_vx = t_vy + (size_t)B * (ny + 1) * nx;                    // B * ny * (nx + 1)
    if (cudaMemsetAsync(workspace, 0, ws_header(B) + (g_rho ? sf * sizeof(unsigned long long) : 0), st) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
    if (g_rho) advect_absmax_kernel<<<absmax_blocks((long long)sf), 256, 0, st>>>(g_out, (long long)sf, nullptr, 0, header);
#define CALL(M)                                                                                                                     \
    advect_bwd_scatter_kernel<V_CENTERED, M, 0><<<tiles(ny, nx, B), kBlock, 0, st>>>(g, rho, vy, vx, g_out, g_rho ? acc : nullptr, G, G + sf, header); \
    advect_centered_bwd_kernel<M><<<tiles(ny + 1, nx + 1, B), kBlock, 0, st>>>(g, acc, G, G + sf, header, g_rho, t_vy, t_vx, add_vy, add_vx)
    MODE_SWITCH(mode, CALL);
#undef CALL
#define CALL(M) curl_bwd_kernel<M><<<tiles(ny, nx, B), kBlock, 0, st>>>(t_vy, t_vx, g_phi, ny, nx)
    MODE_SWITCH(potential_mode, CALL);
#undef CALL
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

// ---- fused staggered force loss ------------------------------------------------------------------------
static size_t force_loss_blocks(int B, int ny, int nx, size_t& n0) {
    const dim3 t0 = tiles(ny + 1, nx, B), t1 = tiles(ny, nx + 1, B);
    n0 = (size_t)t0.x * t0.y * t0.z;
    return n0 + (size_t)t1.x * t1.y * t1.z;
}

extern "C" size_t pdec_staggered_force_loss_workspace_bytes(int B, int ny, int nx) {
    size_t n0;
    return force_loss_blocks(B > 0 ? B : 1, ny, nx, n0) * sizeof(float);
}

extern "C" int pdec_staggered_force_loss_fwd(const float* vy, const float* vx, const float* ty, const float* tx, float* resid_y, float* resid_x,
                                             float* loss, int B, int ny, int nx, int mode, double dt, float lower_y, float lower_x, float size_y,
                                             float size_x, void* workspace, size_t workspace_bytes, void* stream) {
    FluidGeo g;
    int rc = make_geo(g, B, ny, nx, dt, lower_y, lower_x, size_y, size_x);
    if (rc != PDEC_OK) return rc;
    if (bad_mode(mode)) return PDEC_ERR_INVALID_ARGUMENT;
    if (!loss) return PDEC_ERR_INVALID_ARGUMENT;
    cudaStream_t st = (cudaStream_t)stream;
    if (B == 0) return cudaMemsetAsync(loss, 0, sizeof(float), st) == cudaSuccess ? PDEC_OK : PDEC_ERR_CUDA_LAUNCH;
    if (!vy || !vx || !ty || !tx || !resid_y || !resid_x || !workspace) return PDEC_ERR_INVALID_ARGUMENT;
    if (workspace_bytes < pdec_staggered_force_loss_workspace_bytes(B, ny, nx)) return PDEC_ERR_WORKSPACE_TOO_SMALL;
    size_t n0;
    const size_t n = force_loss_blocks(B, ny, nx, n0);
    if (n > 0x7fffffffULL) return PDEC_ERR_INVALID_ARGUMENT;
    float* partials = reinterpret_cast<float*>(workspace);
#define CALL(M)                                                                                                                \
    advect_residual_kernel<V_COMP_Y, M><<<tiles(ny + 1, nx, B), kBlock, 0, st>>>(g, vy, vx, ty, resid_y, partials);            \
    advect_residual_kernel<V_COMP_X, M><<<tiles(ny, nx + 1, B), kBlock, 0, st>>>(g, vy, vx, tx, resid_x, partials + n0)
    MODE_SWITCH(mode, CALL);
#undef CALL
    force_loss_finish_kernel<<<1, 1024, 0, st>>>(partials, (int)n0, (int)(n - n0), B, loss);
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" int pdec_staggered_force_loss_bwd(const float* vy, const float* vx, const float* resid_y, const float* resid_x, const float* g_loss,
                                             float* seed_y, float* seed_x, float* g_vy, float* g_vx, float* g_ty, float* g_tx, int B, int ny, int nx,
                                             int mode, double dt, float lower_y, float lower_x, float size_y, float size_x, void* workspace,
                                             size_t workspace_bytes, void* stream) {
    if (B < 0) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!resid_y || !resid_x || !g_loss || !seed_y || !seed_x) return PDEC_ERR_INVALID_ARGUMENT;
    const size_t n0 = (size_t)B * (ny + 1) * nx, n1 = (size_t)B * ny * (nx + 1);
    size_t blocks = (n0 + n1 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (!workspace || ny < 2 || nx < 2) return PDEC_ERR_INVALID_ARGUMENT;
    if (workspace_bytes < pdec_advect_2d_workspace_bytes(B, ny, nx)) return PDEC_ERR_WORKSPACE_TOO_SMALL;
    if (cudaMemsetAsync(workspace, 0, ws_header(B) + (n0 + n1) * sizeof(unsigned long long), (cudaStream_t)stream) != cudaSuccess) return PDEC_ERR_CUDA_LAUNCH;
    force_loss_seed_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(resid_y, resid_x, g_loss, 1.0f / (float)B, seed_y, seed_x, g_ty, g_tx, n0, n1,
                                                                               reinterpret_cast<unsigned*>(workspace));
    PDEC_CHECK_LAUNCH();
    return advect_staggered_bwd_impl(vy, vx, seed_y, seed_x, g_vy, g_vx, 0, B, ny, nx, mode, dt, lower_y, lower_x, size_y, size_x, workspace,
                                     workspace_bytes, stream, true);
}

extern "C" int pdec_mask_divergence_2d_fwd(const float* vy, const float* vx, float* div, int B, int ny, int nx, int closed, void* stream) {
    if (B < 0 || ny < 2 || nx < 2 || B > 65535) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!vy || !vx || !div) return PDEC_ERR_INVALID_ARGUMENT;
    mask_divergence_fwd_kernel<<<tiles(ny, nx, B), kBlock, 0, (cudaStream_t)stream>>>(vy, vx, div, ny, nx, pdec_closed_sides(closed));
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" int pdec_mask_divergence_2d_bwd(const float* g_div, float* g_vy, float* g_vx, int accumulate, int B, int ny, int nx, int closed,
                                           void* stream) {
    if (B < 0 || ny < 2 || nx < 2 || B > 65535) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!g_div || !g_vy || !g_vx) return PDEC_ERR_INVALID_ARGUMENT;
    mask_divergence_bwd_kernel<<<tiles(ny + 1, nx + 1, B), kBlock, 0, (cudaStream_t)stream>>>(g_div, g_vy, g_vx, ny, nx, pdec_closed_sides(closed), accumulate);
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" int pdec_subtract_pressure_gradient_2d_fwd(const float* vy, const float* vx, const float* p, float* vy_out, float* vx_out, int B,
                                                      int ny, int nx, int closed, float size_y, float size_x, void* stream) {
    if (B < 0 || ny < 2 || nx < 2 || B > 65535 || !(size_y > 0.0f) || !(size_x > 0.0f)) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!vy || !vx || !p || !vy_out || !vx_out) return PDEC_ERR_INVALID_ARGUMENT;
    const float dxy = (float)((double)size_y / ny), dxx = (float)((double)size_x / nx);
    subtract_gradient_fwd_kernel<<<tiles(ny + 1, nx + 1, B), kBlock, 0, (cudaStream_t)stream>>>(vy, vx, p, vy_out, vx_out, ny, nx, pdec_closed_sides(closed),
                                                                                                 dxy, dxy, dxx, pow2_reciprocal(dxy), pow2_reciprocal(dxx));
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" int pdec_subtract_pressure_gradient_2d_bwd(const float* g_vy_out, const float* g_vx_out, float* g_vy, float* g_vx, float* g_p,
                                                      int accumulate_v, int B, int ny, int nx, int closed, float size_y, float size_x,
                                                      void* stream) {
    if (B < 0 || ny < 2 || nx < 2 || B > 65535 || !(size_y > 0.0f) || !(size_x > 0.0f)) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!g_vy_out || !g_vx_out) return PDEC_ERR_INVALID_ARGUMENT;
    const float dxy = (float)((double)size_y / ny), dxx = (float)((double)size_x / nx);
    subtract_gradient_bwd_kernel<<<tiles(ny + 1, nx + 1, B), kBlock, 0, (cudaStream_t)stream>>>(g_vy_out, g_vx_out, g_vy, g_vx, g_p, ny, nx,
                                                                                                 pdec_closed_sides(closed), dxy, dxy, dxx, accumulate_v,
                                                                                                 pow2_reciprocal(dxy), pow2_reciprocal(dxx));
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

static int norm_blocks(int n) {
    int nb = (n + 256 * 16 - 1) / (256 * 16);
    return nb < 1 ? 1 : (nb > 64 ? 64 : nb);
}
extern "C" size_t pdec_normalize_2d_workspace_bytes(int B, int ny, int nx) {
    return (size_t)2 * 3 * (B > 0 ? B : 1) * norm_blocks(ny * nx) * sizeof(float);
}

extern "C" int pdec_normalize_2d_fwd(const float* target, const float* source, float* out, int B, int ny, int nx, float epsilon,
                                     void* workspace, size_t workspace_bytes, void* stream) {
    if (B < 0 || ny < 1 || nx < 1 || B > 65535) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!target || !source || !out || !workspace) return PDEC_ERR_INVALID_ARGUMENT;
    if (workspace_bytes < pdec_normalize_2d_workspace_bytes(B, ny, nx)) return PDEC_ERR_WORKSPACE_TOO_SMALL;
    const int n = ny * nx, nb = norm_blocks(n);
    cudaStream_t st = (cudaStream_t)stream;
    float* p1 = reinterpret_cast<float*>(workspace);
    sums_kernel<<<dim3(nb, B), 256, 0, st>>>(target, source, p1, n, B);
    normalize_fwd_kernel<<<dim3(nb, B), 256, 0, st>>>(target, p1, out, n, B, nb, epsilon);
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" int pdec_normalize_2d_bwd(const float* target, const float* source, const float* g_out, float* g_target, float* g_source, int B,
                                     int ny, int nx, float epsilon, void* workspace, size_t workspace_bytes, void* stream) {
    if (B < 0 || ny < 1 || nx < 1 || B > 65535) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!target || !source || !g_out || !workspace) return PDEC_ERR_INVALID_ARGUMENT;
    if (workspace_bytes < pdec_normalize_2d_workspace_bytes(B, ny, nx)) return PDEC_ERR_WORKSPACE_TOO_SMALL;
    const int n = ny * nx, nb = norm_blocks(n);
    cudaStream_t st = (cudaStream_t)stream;
    float* p1 = reinterpret_cast<float*>(workspace);
    float* p2 = p1 + (size_t)3 * B * nb;
    sums_kernel<<<dim3(nb, B), 256, 0, st>>>(target, source, p1, n, B);
    sums_kernel<<<dim3(nb, B), 256, 0, st>>>(g_out, target, p2, n, B);
    normalize_bwd_kernel<<<dim3(nb, B), 256, 0, st>>>(g_out, p1, p2, g_target, g_source, n, B, nb, epsilon);
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}
