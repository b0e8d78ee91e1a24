// burgers.cu -- Burgers 1-D forward step, multi-step trajectory, reverse-mode adjoint and the fused
// force-loss (pair loss) kernel for sm_100a.  SURVEY rows B1-B6.
//
// Compile with -fmad=false: the back-traced index xi and the diffusion update follow the reference's fp32
// operation order (no FMA contraction), because the 'constant' boundary rule is discontinuous in xi.
//
// Design: ONE WARP OWNS ONE ROW.  The row (n floats) lives in shared memory for the whole life of the
// trajectory / pair sequence; global memory is touched exactly once per frame in each direction
// (cp.async 16-byte loads, float4 stores).  Gathers, the 3-point stencil and the transpose of the gather all
// work on the shared-memory copy.  The advection adjoint w.r.t. the data is the transpose of a gather; it is
// accumulated into a warp-private shared row in a fixed order (pdec::warp_ordered_scatter_add) -- no float
// atomics, bit-reproducible run to run.
#include "pdec_common.cuh"

namespace pdec {

struct BurgersP {
    int n, nrow;  // cells per row, shared-memory row stride (multiple of 4)
    int mode, substeps, resubstep, effect_mode;
    float dt, lower, size, nf;
    float a;       // float32(dt*viscosity/substeps)       (PF/physics/field/util.py:43)
    float dx2;     // float32(dx^2), dx = size/n           (PF/physics/field/grid.py:190)
    float dxi_du;  // d xi / d u = -dt * n / size
};

// back-traced index: advect.py:35, _box.py:50-52, grid.py:104-105
__device__ __forceinline__ float backtrace_xi(const BurgersP& p, float x0, float u) {
    const float x = x0 - u * p.dt;
    const float local = (x - p.lower) / p.size;
    return local * p.nf - 0.5f;
}

template <int MODE>
__device__ __forceinline__ float row_at(const float* s, int j, int n) {
    if (MODE == PDEC_MODE_CONSTANT) return (j < 0 || j >= n) ? 0.0f : s[j];
    if (MODE == PDEC_MODE_REPLICATE) return s[min(max(j, 0), n - 1)];
    return s[j < 0 ? j + n : (j >= n ? j - n : j)];
}

// raw 3-point laplacian (upper + lower - 2*center), PF/math/nd.py:281-282
template <int MODE>
__device__ __forceinline__ float lap_raw(const float* s, int i, int n) {
    return row_at<MODE>(s, i + 1, n) + row_at<MODE>(s, i - 1, n) - 2.0f * s[i];
}

// ---- forward step of one row held in shared memory ------------------------------------------------
// su: input row; sa, sb: scratch rows; sf: force row or nullptr; so: output row (distinct from su/sa/sb).
template <int MODE>
__device__ void warp_burgers_fwd(const BurgersP& p, const float* sx0, const float* su, float* sa, float* sb,
                                 const float* sf, float* so) {
    const int lane = lane_id();
    const int n = p.n;
    for (int i = lane; i < n; i += 32) {
        const float xi = backtrace_xi(p, sx0[i], su[i]);
        int fi, ci;
        float w0, w1;
        interp_setup<MODE>(xi, n, fi, ci, w0, w1);
        sa[i] = su[fi] * w1 + su[ci] * w0;
    }
    __syncwarp();
    const float* src = sa;
    if (p.resubstep) {  // NumPy semantics: every substep sees the updated field
        float* dst = sb;
        for (int k = 0; k < p.substeps; ++k) {
            for (int i = lane; i < n; i += 32) dst[i] = src[i] + p.a * (lap_raw<MODE>(src, i, n) / p.dx2);
            __syncwarp();
            const float* t = src;
            src = dst;
            dst = const_cast<float*>(t);
        }
    }
    for (int i = lane; i < n; i += 32) {
        float v = src[i];
        if (!p.resubstep) {  // TF semantics: the same laplacian added `substeps` times
            const float lap = lap_raw<MODE>(src, i, n) / p.dx2;
            for (int k = 0; k < p.substeps; ++k) v = v + p.a * lap;
        }
        if (p.effect_mode == PDEC_EFFECT_GROW) v = v + sf[i] * p.dt;
        else if (p.effect_mode == PDEC_EFFECT_ADD) v = v + sf[i];
        so[i] = v;
    }
    __syncwarp();
}

// ---- reverse-mode adjoint of one step of one row ---------------------------------------------------
// su: saved input row; sg: gradient w.r.t. the step output (clobbered); st: scratch;
// sacc: receives the gradient w.r.t. the input row; sgf: force-gradient accumulator row or nullptr.
template <int MODE>
__device__ void warp_burgers_bwd(const BurgersP& p, const float* sx0, const float* su, float* sg, float* st,
                                 float* sacc, float* sgf) {
    const int lane = lane_id();
    const int n = p.n;
    // effect adjoint (effect.py:48-53): d/df of u + f*dt (GROW) or u + f (ADD)
    if (sgf != nullptr) {
        for (int i = lane; i < n; i += 32) sgf[i] += (p.effect_mode == PDEC_EFFECT_GROW) ? sg[i] * p.dt : sg[i];
    }
    // diffusion adjoint: the 3-point stencil matrix is symmetric for all three pad modes
    float* g1 = sg;
    if (p.resubstep) {
        float* dst = st;
        for (int k = 0; k < p.substeps; ++k) {
            for (int i = lane; i < n; i += 32) {
                const float up = row_at<MODE>(g1, i + 1, n), lo = row_at<MODE>(g1, i - 1, n), c = g1[i];
                dst[i] = c + ((p.a * up) / p.dx2 + (p.a * lo) / p.dx2 - 2.0f * ((p.a * c) / p.dx2));
            }
            __syncwarp();
            float* t = g1;
            g1 = dst;
            dst = t;
        }
    } else {
        const float ka = p.a * (float)p.substeps;
        for (int i = lane; i < n; i += 32) {
            const float up = row_at<MODE>(sg, i + 1, n), lo = row_at<MODE>(sg, i - 1, n), c = sg[i];
            st[i] = c + ((ka * up) / p.dx2 + (ka * lo) / p.dx2 - 2.0f * ((ka * c) / p.dx2));
        }
        g1 = st;
        __syncwarp();
    }
    // advection adjoint: transpose of the gather (ordered scatter) + derivative through the back-trace
    for (int i = lane; i < n; i += 32) sacc[i] = 0.0f;
    __syncwarp();
    const int iters = (n + 31) >> 5;
    for (int k = 0; k < iters; ++k) {
        const int i = lane + (k << 5);
        const bool in = i < n;
        int fi = 0, ci = 0;
        float w0 = 0.0f, w1 = 0.0f, g = 0.0f;
        if (in) {
            const float xi = backtrace_xi(p, sx0[i], su[i]);
            interp_setup<MODE>(xi, n, fi, ci, w0, w1);
            g = g1[i];
        }
        // 'constant': an out-of-range sample reads the same cell with weights +w and -w: exactly zero
        const bool contributes = in && !(MODE == PDEC_MODE_CONSTANT && fi == ci);
        warp_ordered_scatter_add(sacc, fi, g * w1, contributes);
        warp_ordered_scatter_add(sacc, ci, g * w0, contributes);
        if (in) g1[i] = g * (su[ci] - su[fi]) * p.dxi_du;  // through xi = xi(u_i)
    }
    __syncwarp();
    for (int i = lane; i < n; i += 32) sacc[i] += g1[i];
    __syncwarp();
}

__device__ __forceinline__ void fill_x0(const BurgersP& p, float* sx0) {
    for (int i = threadIdx.x; i < p.n; i += blockDim.x) sx0[i] = cell_center_local(i, p.n) * p.size + p.lower;
    __syncthreads();
}

// ---- kernel: nsteps forward steps, row resident in shared memory ----------------------------------
template <int MODE>
__global__ void __launch_bounds__(128) burgers_trajectory_fwd_kernel(BurgersP p, const float* __restrict__ u0,
                                                                     const float* __restrict__ force,
                                                                     float* __restrict__ frames, int B, int nsteps,
                                                                     int store_all, int vec4) {
    extern __shared__ __align__(16) float smem[];
    float* sx0 = smem;
    const int warp = threadIdx.x >> 5, wpc = blockDim.x >> 5;
    float* base = smem + p.nrow + (size_t)warp * 5 * p.nrow;
    float *su = base, *so = base + p.nrow, *sa = base + 2 * p.nrow, *sb = base + 3 * p.nrow, *sf = base + 4 * p.nrow;
    fill_x0(p, sx0);
    const size_t rowsz = (size_t)p.n;
    for (int b = blockIdx.x * wpc + warp; b < B; b += gridDim.x * wpc) {
        warp_row_load_async(su, u0 + b * rowsz, p.n, vec4);
        if (p.effect_mode != PDEC_EFFECT_NONE) warp_row_load_async(sf, force + b * rowsz, p.n, vec4);
        cp_async_commit();
        cp_async_wait<0>();
        __syncwarp();
        for (int t = 0; t < nsteps; ++t) {
            warp_burgers_fwd<MODE>(p, sx0, su, sa, sb, sf, so);
            if (store_all) warp_row_store(frames + ((size_t)t * B + b) * rowsz, so, p.n, vec4);
            else if (t == nsteps - 1) warp_row_store(frames + b * rowsz, so, p.n, vec4);
            float* tmp = su;
            su = so;
            so = tmp;
        }
        __syncwarp();
    }
}

// ---- kernel: adjoint of the trajectory -------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(128) burgers_trajectory_bwd_kernel(BurgersP p, const float* __restrict__ u0,
                                                                     const float* __restrict__ frames,
                                                                     const float* __restrict__ g_frames,
                                                                     float* __restrict__ g_u0, float* __restrict__ g_force,
                                                                     int B, int nsteps, int g_all, int vec4) {
    extern __shared__ __align__(16) float smem[];
    float* sx0 = smem;
    const int warp = threadIdx.x >> 5, wpc = blockDim.x >> 5, lane = lane_id();
    float* base = smem + p.nrow + (size_t)warp * 6 * p.nrow;
    float *su = base, *sg = base + p.nrow, *st = base + 2 * p.nrow, *sacc = base + 3 * p.nrow,
          *sgf = base + 4 * p.nrow, *sgx = base + 5 * p.nrow;
    fill_x0(p, sx0);
    const size_t rowsz = (size_t)p.n;
    const bool want_gf = (g_force != nullptr) && p.effect_mode != PDEC_EFFECT_NONE;
    for (int b = blockIdx.x * wpc + warp; b < B; b += gridDim.x * wpc) {
        // lambda_{nsteps} = gradient w.r.t. the last frame
        const float* glast = g_all ? g_frames + ((size_t)(nsteps - 1) * B + b) * rowsz : g_frames + b * rowsz;
        warp_row_load_async(sg, glast, p.n, vec4);
        for (int i = lane; i < p.n; i += 32) sgf[i] = 0.0f;
        float* lam = sg;
        float* out = sacc;
        for (int t = nsteps - 1; t >= 0; --t) {
            const float* uin = (t == 0) ? u0 + b * rowsz : frames + ((size_t)(t - 1) * B + b) * rowsz;
            warp_row_load_async(su, uin, p.n, vec4);
            const bool extra = g_all && t > 0;  // the loss may also read frame t (output of step t-1)
            if (extra) warp_row_load_async(sgx, g_frames + ((size_t)(t - 1) * B + b) * rowsz, p.n, vec4);
            cp_async_commit();
            cp_async_wait<0>();
            __syncwarp();
            warp_burgers_bwd<MODE>(p, sx0, su, lam, st, out, want_gf ? sgf : nullptr);
            if (extra) {
                for (int i = lane; i < p.n; i += 32) out[i] += sgx[i];
                __syncwarp();
            }
            float* tmp = lam;
            lam = out;
            out = tmp;
        }
        warp_row_store(g_u0 + b * rowsz, lam, p.n, vec4);
        if (g_force != nullptr) warp_row_store(g_force + b * rowsz, sgf, p.n, vec4);
        __syncwarp();
    }
}

// =====================================================================================================
// Latency-bound shape (BASELINE config 1: B = 8 rows, 32 sequential steps): ONE CTA PER ROW, ONE THREAD PER CELL.
// With a warp per row a step is a chain of ~600 dependent warp instructions on one scheduler (10.7 us per step forward +
// adjoint at n = 128).  Here the row's cells spread over n / 32 warps on all four schedulers: a thread keeps its cell in a
// register, neighbours and gather taps come from a shared-memory copy of the row (ping-pong buffers, one block barrier per
// stencil pass), frames are written / read straight from the owning thread (coalesced).  Same expressions, same operation order
// (this file compiles with -fmad=false) => the forward is bit-identical to warp_burgers_fwd and to the pair-loss kernels.
// IEEE divisions by box size and dx^2 become exact multiplications when those are powers of two (the unit box at n = 128).
// The adjoint's transpose of the gather stays deterministic: every warp scatters its 32 cells into a PRIVATE shared row in a
// fixed order (warp_ordered_scatter_add) and the rows are summed in warp order -- no float atomics.
// =====================================================================================================
struct CtaFast {
    float inv_size, inv_dx2;       // used when the flags say the reciprocal is exact
    int size_pow2, dx2_pow2;
};

__device__ __forceinline__ float cta_backtrace_xi(const BurgersP& p, const CtaFast& f, float x0, float u) {
    const float x = x0 - u * p.dt;
    const float local = f.size_pow2 ? (x - p.lower) * f.inv_size : (x - p.lower) / p.size;
    return local * p.nf - 0.5f;
}
__device__ __forceinline__ float cta_div_dx2(const BurgersP& p, const CtaFast& f, float v) { return f.dx2_pow2 ? v * f.inv_dx2 : v / p.dx2; }

template <int MODE>
__global__ void __launch_bounds__(1024) burgers_trajectory_fwd_cta_kernel(BurgersP p, CtaFast f, const float* __restrict__ u0,
                                                                          const float* __restrict__ force, float* __restrict__ frames,
                                                                          int B, int nsteps, int store_all) {
    extern __shared__ __align__(16) float smem[];
    const int n = p.n, i = threadIdx.x, b = blockIdx.x;
    const bool in = i < n;
    float* su[2] = {smem, smem + p.nrow};
    float* sa[2] = {smem + 2 * p.nrow, smem + 3 * p.nrow};
    const size_t row = (size_t)b * n;
    float u = in ? u0[row + i] : 0.0f;
    const float fo = (in && p.effect_mode != PDEC_EFFECT_NONE) ? force[row + i] : 0.0f;
    const float x0 = in ? cell_center_local(i, n) * p.size + p.lower : 0.0f;
    if (in) su[0][i] = u;
    __syncthreads();
    for (int t = 0; t < nsteps; ++t) {
        const float* cur = su[t & 1];
        float v = 0.0f;
        if (in) {
            const float xi = cta_backtrace_xi(p, f, x0, u);
            int fi, ci;
            float w0, w1;
            interp_setup<MODE>(xi, n, fi, ci, w0, w1);
            v = cur[fi] * w1 + cur[ci] * w0;
        }
        if (p.resubstep) {          // NumPy semantics: every substep sees the updated field
            for (int k = 0; k < p.substeps; ++k) {
                float* buf = sa[k & 1];
                if (in) buf[i] = v;
                __syncthreads();
                if (in) v = v + p.a * cta_div_dx2(p, f, lap_raw<MODE>(buf, i, n));
            }
        } else {                    // TF semantics: the same laplacian added `substeps` times
            if (in) sa[0][i] = v;
            __syncthreads();
            if (in) {
                const float lap = cta_div_dx2(p, f, lap_raw<MODE>(sa[0], i, n));
                for (int k = 0; k < p.substeps; ++k) v = v + p.a * lap;
            }
        }
        if (in) {
            if (p.effect_mode == PDEC_EFFECT_GROW) v = v + fo * p.dt;
            else if (p.effect_mode == PDEC_EFFECT_ADD) v = v + fo;
            u = v;
            if (store_all) frames[((size_t)t * B + b) * n + i] = v;
            else if (t == nsteps - 1) frames[row + i] = v;
            su[(t + 1) & 1][i] = v;
        }
        __syncthreads();
    }
}

template <int MODE>
__global__ void __launch_bounds__(1024) burgers_trajectory_bwd_cta_kernel(BurgersP p, CtaFast f, const float* __restrict__ u0,
                                                                          const float* __restrict__ frames, const float* __restrict__ g_frames,
                                                                          float* __restrict__ g_u0, float* __restrict__ g_force, int B, int nsteps,
                                                                          int g_all) {
    extern __shared__ __align__(16) float smem[];
    const int n = p.n, i = threadIdx.x, b = blockIdx.x, lane = i & 31, w = i >> 5, nwarps = blockDim.x >> 5;
    const bool in = i < n;
    float* su = smem;
    float* sg[2] = {smem + p.nrow, smem + 2 * p.nrow};
    float* sacc = smem + 3 * p.nrow;                        // [nwarps][nrow] private scatter rows
    float* mine = sacc + (size_t)w * p.nrow;
    const size_t row = (size_t)b * n;
    const float x0 = in ? cell_center_local(i, n) * p.size + p.lower : 0.0f;
    const bool want_gf = (g_force != nullptr) && p.effect_mode != PDEC_EFFECT_NONE;
    float gf = 0.0f;
    float lam = in ? (g_all ? g_frames[((size_t)(nsteps - 1) * B + b) * n + i] : g_frames[row + i]) : 0.0f;
    // the saved input row of step t and the extra gradient of frame t are fetched one step ahead
    float u_next = in ? (nsteps == 1 ? u0[row + i] : frames[((size_t)(nsteps - 2) * B + b) * n + i]) : 0.0f;
    float gx_next = (in && g_all && nsteps > 1) ? g_frames[((size_t)(nsteps - 2) * B + b) * n + i] : 0.0f;
    for (int t = nsteps - 1; t >= 0; --t) {
        const float u = u_next, gextra = gx_next;
        if (in && t > 0) {
            u_next = (t == 1) ? u0[row + i] : frames[((size_t)(t - 2) * B + b) * n + i];
            gx_next = (g_all && t > 1) ? g_frames[((size_t)(t - 2) * B + b) * n + i] : 0.0f;
        }
        float g = lam;
        if (want_gf) gf += (p.effect_mode == PDEC_EFFECT_GROW) ? g * p.dt : g;
        // diffusion adjoint: the 3-point stencil matrix is symmetric for all three pad modes (same expression as warp_burgers_bwd)
        const int passes = p.resubstep ? p.substeps : 1;
        const float ka = p.resubstep ? p.a : p.a * (float)p.substeps;
        for (int k = 0; k < passes; ++k) {
            float* buf = sg[k & 1];
            if (in) buf[i] = g;
            if (k == 0 && in) su[i] = u;                    // the saved row rides on the first barrier
            __syncthreads();
            if (in) {
                const float up = row_at<MODE>(buf, i + 1, n), lo = row_at<MODE>(buf, i - 1, n);
                g = g + (cta_div_dx2(p, f, ka * up) + cta_div_dx2(p, f, ka * lo) - 2.0f * cta_div_dx2(p, f, ka * g));
            }
        }
        // advection adjoint: transpose of the gather into this warp's private row + derivative through the back-trace
        for (int q = lane; q < n; q += 32) mine[q] = 0.0f;
        __syncwarp();
        int fi = 0, ci = 0;
        float w0 = 0.0f, w1 = 0.0f, gxi = 0.0f;
        if (in) {
            const float xi = cta_backtrace_xi(p, f, x0, u);
            interp_setup<MODE>(xi, n, fi, ci, w0, w1);
            gxi = g * (su[ci] - su[fi]) * p.dxi_du;
        }
        // 'constant': an out-of-range sample reads the same cell with weights +w and -w: exactly zero
        const bool contributes = in && !(MODE == PDEC_MODE_CONSTANT && fi == ci);
        warp_ordered_scatter_add(mine, fi, g * w1, contributes);
        warp_ordered_scatter_add(mine, ci, g * w0, contributes);
        __syncthreads();
        if (in) {
            float acc = 0.0f;
            for (int q = 0; q < nwarps; ++q) acc += sacc[(size_t)q * p.nrow + i];      // fixed warp order
            lam = acc + gxi;
            if (g_all && t > 0) lam += gextra;
        }
        // no barrier needed here: the next step's first shared write (sg / su) is ordered behind this step's reads of sacc by its own
        // barrier only for OTHER buffers -- sacc is rewritten (zeroed) before that barrier, so close the step explicitly
        __syncthreads();
    }
    if (in) {
        g_u0[row + i] = lam;
        if (g_force != nullptr) g_force[row + i] = gf;
    }
}

// ---- kernel: fused force loss + gradient over T frame pairs ------------------
// One warp walks one batch row b through t = 0..T-1.  Every frame s[t,b,:] is read from HBM exactly once
// (3-slot cp.async ring: current, next, prefetch) and every gradient row is written exactly once:
//   grad[t] = (+gd_{t-1}: frame t is the target of pair t-1)  +  (-step^T gd_t: frame t is the input of pair t)
// Loss partials: per-row (L2, L1) sums written to `partials[b]`; a second single-block kernel reduces them in
// a fixed tree order (deterministic two-stage reduction).
#define PL_ROWS 10
template <int MODE>
__global__ void __launch_bounds__(128) burgers_pair_loss_kernel(BurgersP p, const float* __restrict__ s,
                                                                float* __restrict__ grad_s, float2* __restrict__ partials,
                                                                int T, int B, float gl2_over_B, float gl1_over_B,
                                                                int vec4) {
    extern __shared__ __align__(16) float smem[];
    float* sx0 = smem;
    const int warp = threadIdx.x >> 5, wpc = blockDim.x >> 5, lane = lane_id();
    float* base = smem + p.nrow + (size_t)warp * PL_ROWS * p.nrow;
    float* ring[3] = {base, base + p.nrow, base + 2 * p.nrow};
    float *sa = base + 3 * p.nrow, *sb = base + 4 * p.nrow, *so = base + 5 * p.nrow, *sg = base + 6 * p.nrow,
          *st = base + 7 * p.nrow, *sacc = base + 8 * p.nrow, *carry = base + 9 * p.nrow;
    fill_x0(p, sx0);
    const size_t rowsz = (size_t)p.n;
    const size_t fstride = (size_t)B * rowsz;
    const bool want_grad = grad_s != nullptr;
    for (int b = blockIdx.x * wpc + warp; b < B; b += gridDim.x * wpc) {
        const float* srow = s + b * rowsz;
        float* grow = want_grad ? grad_s + b * rowsz : nullptr;
        warp_row_load_async(ring[0], srow, p.n, vec4);
        cp_async_commit();
        warp_row_load_async(ring[1], srow + fstride, p.n, vec4);
        cp_async_commit();
        for (int i = lane; i < p.n; i += 32) carry[i] = 0.0f;
        float acc2 = 0.0f, acc1 = 0.0f;
        for (int t = 0; t < T; ++t) {
            float* cur = ring[t % 3];
            float* nxt = ring[(t + 1) % 3];
            if (t + 2 <= T) warp_row_load_async(ring[(t + 2) % 3], srow + (size_t)(t + 2) * fstride, p.n, vec4);
            cp_async_commit();      // (possibly empty) group keeps the wait count uniform
            cp_async_wait<1>();     // frames t and t+1 have landed
            __syncwarp();
            warp_burgers_fwd<MODE>(p, sx0, cur, sa, sb, nullptr, so);
            for (int i = lane; i < p.n; i += 32) {
                const float d = nxt[i] - so[i];
                acc2 += d * d;
                acc1 += fabsf(d);
                const float sgn = (d > 0.0f) ? 1.0f : ((d < 0.0f) ? -1.0f : 0.0f);
                const float gd = gl2_over_B * d + gl1_over_B * sgn;  // dL/dd
                so[i] = gd;
                sg[i] = -gd;  // d = s[t+1] - step(s[t])
            }
            __syncwarp();
            if (want_grad) {
                warp_burgers_bwd<MODE>(p, sx0, cur, sg, st, sacc, nullptr);
                for (int i = lane; i < p.n; i += 32) {
                    st[i] = sacc[i] + carry[i];
                    carry[i] = so[i];
                }
                __syncwarp();
                warp_row_store(grow + (size_t)t * fstride, st, p.n, vec4);
                __syncwarp();
            }
        }
        cp_async_wait<0>();
        if (want_grad) {
            warp_row_store(grow + (size_t)T * fstride, carry, p.n, vec4);
        }
        acc2 = warp_sum(acc2);
        acc1 = warp_sum(acc1);
        if (lane == 0) partials[b] = make_float2(acc2, acc1);
        __syncwarp();
    }
}

// second stage: fixed-order reduction of the per-row partials -> losses[0] = L2, losses[1] = L1
__global__ void __launch_bounds__(1024) burgers_loss_reduce_kernel(const float2* __restrict__ partials, int count,
                                                                    float inv_B, float* __restrict__ losses) {
    __shared__ float s2[32], s1[32];
    float a2 = 0.0f, a1 = 0.0f;
    for (int i = threadIdx.x; i < count; i += blockDim.x) {
        const float2 v = partials[i];
        a2 += v.x;
        a1 += v.y;
    }
    a2 = warp_sum(a2);
    a1 = warp_sum(a1);
    if ((threadIdx.x & 31) == 0) {
        s2[threadIdx.x >> 5] = a2;
        s1[threadIdx.x >> 5] = a1;
    }
    __syncthreads();
    if (threadIdx.x < 32) {
        a2 = (threadIdx.x < (blockDim.x >> 5)) ? s2[threadIdx.x] : 0.0f;
        a1 = (threadIdx.x < (blockDim.x >> 5)) ? s1[threadIdx.x] : 0.0f;
        a2 = warp_sum(a2);
        a1 = warp_sum(a1);
        if (threadIdx.x == 0) {
            losses[0] = a2 * 0.5f * inv_B;  // l2_loss: sum(d^2)/2/B  (PF/math/nd.py:115-121)
            losses[1] = a1 * inv_B;         // l1_loss: sum|d|/B      (PF/math/nd.py:96-108)
        }
    }
}

}  // namespace pdec

// ======================================================================================================
// host side: C-ABI entry points
// ======================================================================================================
using namespace pdec;

static int make_params(BurgersP& p, int n, int mode, double dt, double viscosity, float lower, float size, int substeps,
                       int resubstep, int effect_mode) {
    if (n < 2 || substeps < 1 || mode < 0 || mode > 2 || effect_mode < 0 || effect_mode > 2 || !(size > 0.0f))
        return PDEC_ERR_INVALID_ARGUMENT;
    p.n = n;
    p.nrow = (n + 3) & ~3;
    p.mode = mode;
    p.substeps = substeps;
    p.resubstep = resubstep ? 1 : 0;
    p.effect_mode = effect_mode;
    p.dt = (float)dt;
    p.lower = lower;
    p.size = size;
    p.nf = (float)n;
    p.a = (float)(dt * viscosity / (double)substeps);
    const double dx = (double)size / (double)n;
    p.dx2 = (float)(dx * dx);
    p.dxi_du = -(p.nf / size) * p.dt;
    return PDEC_OK;
}

static bool aligned16(const void* ptr) { return (reinterpret_cast<uintptr_t>(ptr) & 15u) == 0; }

static int sm_count() {
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return sms > 0 ? sms : 148;
}

// one CTA per row / one thread per cell pays while every row's CTA is resident at once with room to spare (few rows: the
// latency-bound regime); beyond that the warp-per-row kernels have the better throughput
static bool cta_shape(int B, int n) {
    const int warps = (n + 31) / 32;
    return n <= 1024 && (long long)B * warps <= (long long)sm_count() * 8;
}
static bool is_pow2(float v) {
    int e;
    return v > 0.0f && frexpf(v, &e) == 0.5f && e > -100 && e < 100;
}
static CtaFast cta_fast(const BurgersP& p) {
    CtaFast f;
    f.size_pow2 = is_pow2(p.size) ? 1 : 0;
    f.dx2_pow2 = is_pow2(p.dx2) ? 1 : 0;
    f.inv_size = f.size_pow2 ? 1.0f / p.size : 0.0f;      // exact reciprocals: x / 2^k == x * 2^-k bit for bit (no overflow at these scales)
    f.inv_dx2 = f.dx2_pow2 ? 1.0f / p.dx2 : 0.0f;
    return f;
}

// grid for a warp-per-row kernel with `wpc` warps per CTA and `smem` bytes per CTA
static int rows_grid(int B, int wpc, size_t smem) {
    const int ctas_needed = (B + wpc - 1) / wpc;
    size_t per_sm = (size_t)(220 * 1024) / (smem + 1024);
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 16) per_sm = 16;
    const int cap = sm_count() * (int)per_sm;
    return ctas_needed < cap ? ctas_needed : cap;
}

#define PDEC_MAX_DYN_SMEM (227 * 1024)

template <typename K>
static int set_smem(K kernel, size_t smem) {
    if (smem > PDEC_MAX_DYN_SMEM) return PDEC_ERR_UNSUPPORTED_SIZE;
    if (smem > 48 * 1024) {
        if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
            return PDEC_ERR_CUDA_LAUNCH;
    }
    return PDEC_OK;
}

// warps per CTA such that the per-CTA shared memory fits; rows of up to ~11k cells are supported
static int pick_wpc(int nrow, int rows_per_warp) {
    for (int wpc = 4; wpc >= 1; wpc >>= 1) {
        const size_t smem = ((size_t)nrow * (1 + (size_t)wpc * rows_per_warp)) * sizeof(float);
        if (smem <= PDEC_MAX_DYN_SMEM) return wpc;
    }
    return 0;
}

extern "C" int pdec_burgers_trajectory_fwd(const float* u0, const float* force, float* frames, int B, int n, int nsteps,
                                           int store_all, int mode, double dt, double viscosity, float box_lower,
                                           float box_size, int substeps, int resubstep, int effect_mode, void* stream) {
    BurgersP p;
    int rc = make_params(p, n, mode, dt, viscosity, box_lower, box_size, substeps, resubstep, effect_mode);
    if (rc != PDEC_OK) return rc;
    if (B < 0 || nsteps < 1) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!u0 || !frames || (effect_mode != PDEC_EFFECT_NONE && !force)) return PDEC_ERR_INVALID_ARGUMENT;
    cudaStream_t st = (cudaStream_t)stream;
    if (cta_shape(B, n)) {          // few rows: one CTA per row, one thread per cell (latency-bound regime)
        const CtaFast f = cta_fast(p);
        const int threads = (n + 31) & ~31;
        const size_t csmem = (size_t)4 * p.nrow * sizeof(float);
#define LAUNCH_FWD_CTA(M) burgers_trajectory_fwd_cta_kernel<M><<<B, threads, csmem, st>>>(p, f, u0, force, frames, B, nsteps, store_all)
        if (mode == PDEC_MODE_CONSTANT) LAUNCH_FWD_CTA(PDEC_MODE_CONSTANT);
        else if (mode == PDEC_MODE_REPLICATE) LAUNCH_FWD_CTA(PDEC_MODE_REPLICATE);
        else LAUNCH_FWD_CTA(PDEC_MODE_CIRCULAR);
#undef LAUNCH_FWD_CTA
        PDEC_CHECK_LAUNCH();
        return PDEC_OK;
    }
    const int wpc = pick_wpc(p.nrow, 5);
    if (wpc == 0) return PDEC_ERR_UNSUPPORTED_SIZE;
    const size_t smem = ((size_t)p.nrow * (1 + (size_t)wpc * 5)) * sizeof(float);
    const int vec4 = (n % 4 == 0) && aligned16(u0) && aligned16(frames) && (force == nullptr || aligned16(force));
    const int grid = rows_grid(B, wpc, smem);
#define LAUNCH_FWD(M)                                                                                          \
    do {                                                                                                       \
        rc = set_smem(burgers_trajectory_fwd_kernel<M>, smem);                                                 \
        if (rc != PDEC_OK) return rc;                                                                          \
        burgers_trajectory_fwd_kernel<M><<<grid, wpc * 32, smem, st>>>(p, u0, force, frames, B, nsteps, store_all, vec4); \
    } while (0)
    if (mode == PDEC_MODE_CONSTANT) LAUNCH_FWD(PDEC_MODE_CONSTANT);
    else if (mode == PDEC_MODE_REPLICATE) LAUNCH_FWD(PDEC_MODE_REPLICATE);
    else LAUNCH_FWD(PDEC_MODE_CIRCULAR);
#undef LAUNCH_FWD
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

extern "C" int pdec_burgers_trajectory_bwd(const float* u0, const float* frames, const float* g_frames, float* g_u0,
                                           float* g_force, int B, int n, int nsteps, int g_all, int mode, double dt,
                                           double viscosity, float box_lower, float box_size, int substeps,
                                           int resubstep, int effect_mode, void* stream) {
    BurgersP p;
    int rc = make_params(p, n, mode, dt, viscosity, box_lower, box_size, substeps, resubstep, effect_mode);
    if (rc != PDEC_OK) return rc;
    if (B < 0 || nsteps < 1) return PDEC_ERR_INVALID_ARGUMENT;
    if (B == 0) return PDEC_OK;
    if (!u0 || !g_frames || !g_u0 || (nsteps > 1 && !frames)) return PDEC_ERR_INVALID_ARGUMENT;
    cudaStream_t st = (cudaStream_t)stream;
    if (cta_shape(B, n)) {
        const CtaFast f = cta_fast(p);
        const int threads = (n + 31) & ~31;
        const size_t csmem = (size_t)(3 + threads / 32) * p.nrow * sizeof(float);
#define LAUNCH_BWD_CTA(M)                                                                                      \
    do {                                                                                                       \
        rc = set_smem(burgers_trajectory_bwd_cta_kernel<M>, csmem);                                            \
        if (rc != PDEC_OK) return rc;                                                                          \
        burgers_trajectory_bwd_cta_kernel<M><<<B, threads, csmem, st>>>(p, f, u0, frames, g_frames, g_u0, g_force, B, nsteps, g_all); \
    } while (0)
        if (mode == PDEC_MODE_CONSTANT) LAUNCH_BWD_CTA(PDEC_MODE_CONSTANT);
        else if (mode == PDEC_MODE_REPLICATE) LAUNCH_BWD_CTA(PDEC_MODE_REPLICATE);
        else LAUNCH_BWD_CTA(PDEC_MODE_CIRCULAR);
#undef LAUNCH_BWD_CTA
        PDEC_CHECK_LAUNCH();
        return PDEC_OK;
    }
    const int wpc = pick_wpc(p.nrow, 6);
    if (wpc == 0) return PDEC_ERR_UNSUPPORTED_SIZE;
    const size_t smem = ((size_t)p.nrow * (1 + (size_t)wpc * 6)) * sizeof(float);
    const int vec4 = (n % 4 == 0) && aligned16(u0) && aligned16(g_frames) && aligned16(g_u0) &&
                     (frames == nullptr || aligned16(frames)) && (g_force == nullptr || aligned16(g_force));
    const int grid = rows_grid(B, wpc, smem);
#define LAUNCH_BWD(M)                                                                                          \
    do {                                                                                                       \
        rc = set_smem(burgers_trajectory_bwd_kernel<M>, smem);                                                 \
        if (rc != PDEC_OK) return rc;                                                                          \
        burgers_trajectory_bwd_kernel<M><<<grid, wpc * 32, smem, st>>>(p, u0, frames, g_frames, g_u0, g_force, B, nsteps, g_all, vec4); \
    } while (0)
    if (mode == PDEC_MODE_CONSTANT) LAUNCH_BWD(PDEC_MODE_CONSTANT);
    else if (mode == PDEC_MODE_REPLICATE) LAUNCH_BWD(PDEC_MODE_REPLICATE);
    else LAUNCH_BWD(PDEC_MODE_CIRCULAR);
#undef LAUNCH_BWD
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}

// burgers_fast.cu: register-resident kernel for n == 128 (returns 1 when the shape is not eligible)
int pdec_burgers_pair_loss_fast(const float* s, float* grad_s, float2* partials, int T, int B, int n, int mode, double dt, double viscosity,
                                float lower, float size, int substeps, int resubstep, float gl2, float gl1, unsigned* done_counter,
                                float* losses, cudaStream_t st);

static size_t pair_loss_partials_bytes(int B, int T) {       // one loss partial per (row, chunk of pairs), rounded to 256 bytes
    const size_t raw = (size_t)(B > 0 ? B : 1) * (size_t)(T > 0 ? T : 1) * sizeof(float2);
    return (raw + 255) & ~(size_t)255;
}

extern "C" size_t pdec_burgers_pair_loss_workspace_bytes(int B, int n, int T) {
    (void)n;
    return pair_loss_partials_bytes(B, T) + 256;                                   // + the "CTAs done" counter of the fused reduction
}

extern "C" int pdec_burgers_pair_loss(const float* s, float* grad_s, float* losses, int T, int B, int n, int mode,
                                      double dt, double viscosity, float box_lower, float box_size, int substeps,
                                      int resubstep, float gl2, float gl1, void* workspace, size_t workspace_bytes,
                                      void* stream) {
    BurgersP p;
    int rc = make_params(p, n, mode, dt, viscosity, box_lower, box_size, substeps, resubstep, PDEC_EFFECT_NONE);
    if (rc != PDEC_OK) return rc;
    if (B < 1 || T < 1 || !s || !losses || !workspace) return PDEC_ERR_INVALID_ARGUMENT;
    if (workspace_bytes < pdec_burgers_pair_loss_workspace_bytes(B, n, T)) return PDEC_ERR_WORKSPACE_TOO_SMALL;
    cudaStream_t st = (cudaStream_t)stream;
    float2* partials = reinterpret_cast<float2*>(workspace);
    const float invB = 1.0f / (float)B;
    {
        // n == 128: register-resident kernel, loss reduction fused into the same launch (counter word behind the partials)
        unsigned* done_counter = reinterpret_cast<unsigned*>(reinterpret_cast<char*>(workspace) + pair_loss_partials_bytes(B, T));
        rc = pdec_burgers_pair_loss_fast(s, grad_s, partials, T, B, n, mode, dt, viscosity, box_lower, box_size, substeps, resubstep, gl2,
                                         gl1, done_counter, losses, st);
        if (rc <= 0) return rc;
    }
    const int wpc = pick_wpc(p.nrow, PL_ROWS);
    if (wpc == 0) return PDEC_ERR_UNSUPPORTED_SIZE;
    const size_t smem = ((size_t)p.nrow * (1 + (size_t)wpc * PL_ROWS)) * sizeof(float);
    const int vec4 = (n % 4 == 0) && aligned16(s) && (grad_s == nullptr || aligned16(grad_s));
    const int grid = rows_grid(B, wpc, smem);
#define LAUNCH_PL(M)                                                                                           \
    do {                                                                                                       \
        rc = set_smem(burgers_pair_loss_kernel<M>, smem);                                                      \
        if (rc != PDEC_OK) return rc;                                                                          \
        burgers_pair_loss_kernel<M><<<grid, wpc * 32, smem, st>>>(p, s, grad_s, partials, T, B, gl2 * invB, gl1 * invB, vec4); \
    } while (0)
    if (mode == PDEC_MODE_CONSTANT) LAUNCH_PL(PDEC_MODE_CONSTANT);
    else if (mode == PDEC_MODE_REPLICATE) LAUNCH_PL(PDEC_MODE_REPLICATE);
    else LAUNCH_PL(PDEC_MODE_CIRCULAR);
#undef LAUNCH_PL
    PDEC_CHECK_LAUNCH();
    burgers_loss_reduce_kernel<<<1, 1024, 0, st>>>(partials, B, invB, losses);
    PDEC_CHECK_LAUNCH();
    return PDEC_OK;
}
