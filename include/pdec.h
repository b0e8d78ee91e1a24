/* pdec.h -- C-ABI of the B200-native PDE-Control physics path (libpdec.so, sm_100a).
 *
 * This is the operator-level drop-in boundary.  Each entry point replaces one of the
 * reference's backend calls / native TF custom ops; the reference interface it replaces is cited per function
 * (paths relative to /root/reference, PF/ = PhiFlow/phi/).
 *
 * Conventions
 *  - every pointer is a DEVICE pointer owned by the caller unless the name ends in `_host`;
 *  - fp32, contiguous, channels-last with the (size-1) channel dropped: 1-D fields [B, N],
 *    2-D centred fields [B, ny, nx], staggered velocity vy[B, ny+1, nx], vx[B, ny, nx+1];
 *  - no allocation, no device synchronisation, no default-stream use inside: every call only enqueues work on
 *    `stream` (a cudaStream_t passed as void*), so all calls are CUDA-graph-capture safe;
 *  - returns PDEC_OK (0) or a negative error code; never exits, never throws (the reference's native ops
 *    call exit(1) on CUDA errors: PF/tf/cuda/src/helpers.h:56-65);
 *  - no mutable global state; re-entrant for distinct streams.
 */
#ifndef PDEC_H
#define PDEC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PDEC_VERSION 100

/* error codes */
#define PDEC_OK 0
#define PDEC_ERR_INVALID_ARGUMENT (-1)
#define PDEC_ERR_UNSUPPORTED_SIZE (-2)
#define PDEC_ERR_CUDA_LAUNCH (-3)
#define PDEC_ERR_WORKSPACE_TOO_SMALL (-4)
#define PDEC_ERR_NO_DEVICE (-5)
#define PDEC_ERR_NOT_CONVERGED (-6)
#define PDEC_ERR_STATIC_SCRATCH_TOO_SMALL (-7) /* pdec_burgers_pair_loss_static: B * T beyond PDEC_STATIC_MAX_PAIRS */

/* boundary / extrapolation modes.  Reference: PF/physics/field/grid.py:240-249 (_pad_mode_str)
 * 'constant' -> zero outside, 'boundary' -> replicate, 'periodic' -> circular;
 * resample semantics follow the TF CPU path PF/tf/tf_backend.py:451-538. */
#define PDEC_MODE_CONSTANT 0
#define PDEC_MODE_REPLICATE 1
#define PDEC_MODE_CIRCULAR 2

/* The `closed` argument of the 2-D entry points (which sides of the domain are SOLID; Domain boundaries, PF/physics/domain.py:13-35 with
 * Material.solid, material.py:8-58): 0 = every side OPEN, 1 = every side CLOSED, or per side -- the reference's
 * Domain(..., boundaries=[(lower_y, upper_y), (lower_x, upper_x)]) -- with PDEC_CLOSED_SIDES(ylo, yhi, xlo, xhi), each 0 / 1.
 * A solid side zeroes its boundary faces (solver_api.py:111-121) and does not count as an accessible neighbour in the matrix diagonal
 * (sparse.py:186-213); an open side counts. */
#define PDEC_CLOSED_SIDES(ylo, yhi, xlo, xhi) (0x10 | ((ylo) ? 1 : 0) | ((yhi) ? 2 : 0) | ((xlo) ? 4 : 0) | ((xhi) ? 8 : 0))

/* FieldEffect modes, PF/physics/field/effect.py:48-53 */
#define PDEC_EFFECT_NONE 0
#define PDEC_EFFECT_GROW 1 /* u += f*dt */
#define PDEC_EFFECT_ADD 2  /* u += f    */

int pdec_version(void);
const char* pdec_error_string(int code);
/* number of CUDA devices visible (0 if none / no driver); never fails */
int pdec_device_count(void);

/* ------------------------------------------------------------------------------------------------
 * Burgers 1-D  
 * Replaces: Burgers.step_velocity  PF/physics/burgers.py:42-48
 *             = advect.semi_lagrangian PF/physics/field/advect.py:24-37  (-> tf_backend.resample :145-154)
 *             + diffuse explicit branch PF/physics/field/util.py:38-44   (-> math.laplace PF/math/nd.py:204-283)
 *             + effect_applied         PF/physics/field/effect.py:48-53
 * and, for the adjoint, TF autodiff of those ops (native analogue ResampleGradient,
 * PF/tf/cuda/src/resample_gradient.cu.cc:80-134, which uses float atomics; this library does not).
 *
 * Common physics arguments:
 *   n            cells per row;  mode  PDEC_MODE_*;  dt  time increment;  viscosity  nu
 *   box_lower, box_size   AABox of the 1-D domain (PF/geom/_box.py)
 *   substeps     diffusion_substeps; resubstep = 0: TF semantics (k substeps collapse to k adds of the same
 *                laplacian), 1: NumPy semantics (true substepping) -- SURVEY fact 5
 *   effect_mode  PDEC_EFFECT_*; force [B, n] or NULL (required unless PDEC_EFFECT_NONE)
 * ------------------------------------------------------------------------------------------------ */

/* nsteps sequential steps of every row, the row staying in shared memory across steps.
 * u0 [B,n] -> frames: store_all=1: [nsteps,B,n] (frames[k] = state after k+1 steps); store_all=0: [B,n] (final). */
int pdec_burgers_trajectory_fwd(const float* u0, const float* force, float* frames, int B, int n, int nsteps,
                                int store_all, int mode, double dt, double viscosity, float box_lower,
                                float box_size, int substeps, int resubstep, int effect_mode, void* stream);

/* reverse-mode adjoint of pdec_burgers_trajectory_fwd.
 * frames: the store_all=1 output of the forward (only frames[0..nsteps-2] are read; may be NULL if nsteps==1).
 * g_frames: gradient w.r.t. the forward output: g_all=1: [nsteps,B,n]; g_all=0: [B,n] (final frame only).
 * g_u0 [B,n] (written); g_force [B,n] (written) or NULL. */
int pdec_burgers_trajectory_bwd(const float* u0, const float* frames, const float* g_frames, float* g_u0,
                                float* g_force, int B, int n, int nsteps, int g_all, int mode, double dt,
                                double viscosity, float box_lower, float box_size, int substeps, int resubstep,
                                int effect_mode, void* stream);

/* Fused force loss + its gradient over T consecutive frame pairs.
 * Replaces BurgersPDE.total_force_loss src/control/pde/burgers.py:24-35 with math.l2_loss / l1_loss
 * PF/math/nd.py:96-124 and the TF autodiff of the whole expression:
 *   d_t = s[t+1] - step(s[t]);  L2 = sum_t sum(d_t^2)/2/B;  L1 = sum_t sum|d_t|/B.
 * s [T+1,B,n].  losses[2] <- (L2, L1).  grad_s [T+1,B,n] or NULL <- gl2*dL2/ds + gl1*dL1/ds.
 * workspace: pdec_burgers_pair_loss_workspace_bytes(B, n, T) bytes of device scratch (loss partials). */
size_t pdec_burgers_pair_loss_workspace_bytes(int B, int n, int T);
int pdec_burgers_pair_loss(const float* s, float* grad_s, float* losses, int T, int B, int n, int mode, double dt,
                           double viscosity, float box_lower, float box_size, int substeps, int resubstep,
                           float gl2, float gl1, void* workspace, size_t workspace_bytes, void* stream);

/* A WINDOW of frame pairs [t0, t1) of the same loss (0 <= t0 < t1 <= T): losses[2] <- the window's share of (L2, L1) (the shares
 * of disjoint windows add up to the whole), grad_s frames t0 .. t1-1 (and frame T when t1 == T) <- their COMPLETE gradient
 * (frame t0 includes the residual term of the pair before the window, recomputed from frame t0-1, which must be present).
 * Only frames t0-1 .. t1 of s are read: the host pipeline below runs window k while frames of window k+1 are still in flight.
 * Supported for the fused unit-box kernel only (n = 128, any PDEC_MODE_*, box [0, 1], one substep):
 * pdec_burgers_pair_loss_window_supported() != 0, else PDEC_ERR_UNSUPPORTED_SIZE.  Workspace as for pdec_burgers_pair_loss. */
int pdec_burgers_pair_loss_window_supported(int n, int mode, float box_lower, float box_size, int substeps);
int pdec_burgers_pair_loss_window(const float* s, float* grad_s, float* losses, int T, int B, int n, int t0, int t1, int mode, double dt,
                                  double viscosity, float box_lower, float box_size, int substeps, int resubstep, float gl2, float gl1,
                                  void* workspace, size_t workspace_bytes, void* stream);
/* The same with a frame stride for the gradient: frame f of row b is written at grad_s + f * grad_frame_stride + b * n (floats;
 * a multiple of 4, at least B * n; 0 = dense).  grad_s may then be the rows [b0, b0 + B) of a larger batch's array, and it may be
 * page-locked HOST memory that the device can address (cudaHostAlloc / torch pin_memory under unified addressing): the kernel's
 * stores then travel over PCIe themselves, 512 contiguous bytes per warp, and no device-to-host copy follows. */
int pdec_burgers_pair_loss_window_strided(const float* s, float* grad_s, float* losses, int T, int B, int n, int t0, int t1, int mode,
                                          double dt, double viscosity, float box_lower, float box_size, int substeps, int resubstep,
                                          float gl2, float gl1, long long grad_frame_stride, void* workspace, size_t workspace_bytes,
                                          void* stream);

/* The same loss and gradient as pdec_burgers_pair_loss for callers that CANNOT ALLOCATE (code under CUDA-graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The scratch (loss partials + the "CTAs done" counter) is a file-scope `static __device__` array of
 * PDEC_STATIC_SCRATCH_BYTES inside the library (csrc/noalloc.cu); every GPU that loads libpdec.so holds it until the process
 * exits.  The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine) and only enqueues on `stream`.
 * Limit: B * T <= PDEC_STATIC_MAX_PAIRS frame pairs (= BASELINE config 2, s[33, 4096, 128]); beyond it the call enqueues nothing
 * and returns PDEC_ERR_STATIC_SCRATCH_TOO_SMALL.  Every call on a device shares the one array, which keeps the last call's
 * contents (the call clears the counter it needs at zero): calls that overlap on the GPU -- two streams, or two graph branches --
 * RACE; serialise them.  Results are bit-identical to pdec_burgers_pair_loss on the same inputs. */
#define PDEC_STATIC_MAX_PAIRS 131072
#define PDEC_STATIC_SCRATCH_BYTES (PDEC_STATIC_MAX_PAIRS * 8 + 256)
int pdec_burgers_pair_loss_static(const float* s, float* grad_s, float* losses, int T, int B, int n, int mode, double dt,
                                  double viscosity, float box_lower, float box_size, int substeps, int resubstep, float gl2,
                                  float gl1, void* stream);

/* The same force loss fed from HOST arrays: what the reference does every optimisation step through
 * session.run(fetches, feed_dict) (src/control/control_training.py:113-142, PF/tf/session.py:30-70 copies every fed array to
 * the device and every fetched one back).  s_host [T+1,B,n], grad_host [T+1,B,n] or NULL and losses_host[2] are HOST pointers
 * (page-locked for the copies to be asynchronous).  The batch is cut into chunks of `chunk_rows` rows; H2D of chunk c+1, the
 * fused kernel of chunk c and D2H of chunk c-1 overlap on three streams owned by the pipeline object, which also owns the
 * device staging (allocated by _create, never by the call).  The call is ordered on `stream` like a kernel launch: it starts
 * after the work already enqueued there and `stream` waits for the last copy; the host buffers are valid after
 * synchronising `stream`.  Results equal pdec_burgers_pair_loss on the whole batch (the gradient bit for bit when
 * B / chunk_rows is a power of two; the loss scalars are summed chunk by chunk in ascending order).
 * max_chunks bounds B <= chunk_rows * max_chunks.
 * frames_per_slice > 0 additionally allocates whole-batch staging ((T+1) * chunk_rows * max_chunks * n floats, twice) and, when
 * pdec_burgers_pair_loss_window_supported(), cuts the work along TIME instead (slices of that many frame pairs of all rows):
 * every copy is then one contiguous block.  Same results contract.
 * Direct gradient (the default where it applies: fused unit-box kernel, grad_host addressable by the device, frames_per_slice == 0,
 * PDEC_PIPE_STAGED_D2H unset): the kernel of every row chunk stores its gradient straight into grad_host
 * (pdec_burgers_pair_loss_window_strided) instead of into device staging followed by a copy: no D2H leg, nothing to drain
 * (profiles/r1q_host_pipeline.md). */
typedef struct pdec_host_pipeline pdec_host_pipeline;
int pdec_host_pipeline_create(pdec_host_pipeline** pipeline, int chunk_rows, int n, int T, int max_chunks, int frames_per_slice);
int pdec_host_pipeline_destroy(pdec_host_pipeline* pipeline);
int pdec_burgers_pair_loss_host(pdec_host_pipeline* pipeline, const float* s_host, float* grad_host, float* losses_host, int T,
                                int B, int n, int mode, double dt, double viscosity, float box_lower, float box_size,
                                int substeps, int resubstep, float gl2, float gl1, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Resampling as an operator of its own  (operator-level boundary)
 * Replaces backend.resample(inputs, sample_coords, 'linear', boundary) (PF/tf/tf_backend.py:145-154, 451-502) and the native ops
 * Resample / ResampleGradient (PF/tf/cuda/src/resample.cc:7-23, resample_gradient.cc:6-13) for one-channel grids of rank 1 or 2.
 * data [B, ny, nx] (rank 2) or [B, nx] (rank 1, ny ignored); coords [coords_batch, npoints, rank] in INDEX space of data (axis 0
 * first), coords_batch = 1 (shared by the batch) or B; out [B, npoints].  mode = PDEC_MODE_* with the TF-CPU semantics of the fused
 * kernels ('constant': exactly 0 outside [0, N-1) along any axis).
 * _bwd: g_data [B, ny, nx] and/or g_coords [B, npoints, rank] (per batch element; sum over B for shared coordinates); either may be
 * NULL.  g_data is accumulated WITHOUT float atomics (the reference's ResampleGradient uses them): 64-bit fixed-point integer atomics
 * with a power-of-two scale from max|g_out| -- the sum does not depend on the order, two runs are bit-identical; workspace =
 * pdec_resample_bwd_workspace_bytes() bytes of device scratch (needed when g_data != NULL). */
int pdec_resample_fwd(const float* data, const float* coords, float* out, int B, int rank, int ny, int nx, int coords_batch,
                      long long npoints, int mode, void* stream);
size_t pdec_resample_bwd_workspace_bytes(int B, int rank, int ny, int nx);
int pdec_resample_bwd(const float* data, const float* coords, const float* g_out, float* g_data, float* g_coords, int B, int rank,
                      int ny, int nx, int coords_batch, long long npoints, int mode, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * 2-D incompressible flow  
 * Layouts: rho[B,ny,nx], vy[B,ny+1,nx], vx[B,ny,nx+1] (PF/physics/field/staggered_grid.py:33-39), axis 0 = y.
 * Box arguments: lower/size of the DOMAIN box per axis (PF/geom/_box.py); the staggered component boxes are
 * derived as in staggered_grid.py:49-53.  `mode` = PDEC_MODE_* of the sampled field
 * (Material.extrapolation_mode / vector_extrapolation_mode, PF/physics/material.py:66-98).
 * Adjoint entry points take the forward INPUTS (nothing else is saved) and a caller-owned workspace.
 * `accumulate` != 0 adds into the velocity-gradient outputs instead of overwriting them.
 * ------------------------------------------------------------------------------------------------ */

/* staggered_curl_2d, PF/physics/field/util.py:60-67: vy[i,j] = phi(i,j)-phi(i,j+1), vx[i,j] = phi(i+1,j)-phi(i,j),
 * phi extended by `mode`.  bwd: transpose (replaces TF conv2d gradient). */
int pdec_curl_2d_fwd(const float* phi, float* vy, float* vx, int B, int ny, int nx, int mode, void* stream);
int pdec_curl_2d_bwd(const float* g_vy, const float* g_vx, float* g_phi, int B, int ny, int nx, int mode, void* stream);

/* advect.semi_lagrangian(density, velocity, dt), PF/physics/field/advect.py:24-37 with StaggeredGrid.at(centres)
 * staggered_grid.py:126-137 and CenteredGrid.sample_at grid.py:101-107 (-> tf_backend.resample :145-154; native twin
 * Resample, PF/tf/cuda/src/resample.cu.cc:112-141). */
int pdec_advect_centered_2d_fwd(const float* rho, const float* vy, const float* vx, float* rho_out, int B, int ny, int nx,
                                int mode, double dt, float lower_y, float lower_x, float size_y, float size_x, void* stream);
/* advect.semi_lagrangian(v, v, dt) on a StaggeredGrid (advect.py:38-40; grid.py:112-123). */
int pdec_advect_staggered_2d_fwd(const float* vy, const float* vx, float* ay, float* ax, int B, int ny, int nx, int mode,
                                 double dt, float lower_y, float lower_x, float size_y, float size_x, void* stream);
/* adjoints (replace TF autodiff of gather_nd; native twin ResampleGradient, resample_gradient.cu.cc:138-226, which uses
 * float atomics): every sample's shares are added to 64-bit fixed-point accumulators with integer atomics -- order-independent, so
 * the result is bit-reproducible -- then converted and combined with the fixed face stencils of the velocity routes.
 * g_rho / g_vy / g_vx of the centred variant may be NULL (not wanted). */
size_t pdec_advect_2d_workspace_bytes(int B, int ny, int nx);
int pdec_advect_centered_2d_bwd(const float* rho, const float* vy, const float* vx, const float* g_out, float* g_rho,
                                float* g_vy, float* g_vx, int accumulate_v, int B, int ny, int nx, int mode, double dt,
                                float lower_y, float lower_x, float size_y, float size_x, void* workspace,
                                size_t workspace_bytes, void* stream);
int pdec_advect_staggered_2d_bwd(const float* vy, const float* vx, const float* g_ay, const float* g_ax, float* g_vy,
                                 float* g_vx, int accumulate, int B, int ny, int nx, int mode, double dt, float lower_y,
                                 float lower_x, float size_y, float size_x, void* workspace, size_t workspace_bytes,
                                 void* stream);

/* ControlledIncompressibleFlow.step fused (src/control/pde/incompressible_flow.py:53-59): velocity = staggered_curl_2d(potential)
 * (PF/physics/field/util.py:60-67) and next density = advect.semi_lagrangian(density, velocity, dt) (advect.py:24-37) in one launch;
 * vy / vx are written (the force loss and the adjoint read them) but not read back.  Same bits as pdec_curl_2d_fwd followed by
 * pdec_advect_centered_2d_fwd.  potential_mode = extrapolation of the potential, mode = of the density.
 * _bwd: g_rho <- adjoint of the sampling (may be NULL), g_phi <- transpose of the curl applied to the velocity gradient of the
 * advection PLUS add_vy / add_vx (gradients that reached the velocity by other routes, e.g. the force-loss terms; either may be
 * NULL).  Workspace: pdec_advect_2d_workspace_bytes. */
int pdec_curl_advect_2d_fwd(const float* rho, const float* phi, float* vy, float* vx, float* rho_out, int B, int ny, int nx,
                            int potential_mode, int mode, double dt, float lower_y, float lower_x, float size_y, float size_x, void* stream);
int pdec_curl_advect_2d_bwd(const float* rho, const float* vy, const float* vx, const float* g_out, const float* add_vy,
                            const float* add_vx, float* g_rho, float* g_phi, int B, int ny, int nx, int potential_mode, int mode,
                            double dt, float lower_y, float lower_x, float size_y, float size_x, void* workspace,
                            size_t workspace_bytes, void* stream);

/* One term of IncompFlowControl.total_force_loss fused (src/control/pde/incompressible_flow.py:31-43:
 * l2_loss(advect.semi_lagrangian(v1, v1, dt) - v2) summed over both staggered components, l2_loss = sum(d^2) / 2 / batch):
 * loss[0] = sum(resid_y^2) / 2 / B + sum(resid_x^2) / 2 / B with resid = advected(v) - (ty, tx); the residuals are kept for the adjoint,
 * the advected field is never stored.  Deterministic: block partials re-reduced in float64 in a fixed order.
 * _bwd: seed = resid * (g_loss[0] / B) (g_loss is a DEVICE scalar, so the call can sit in a CUDA graph), then the adjoint of the
 * staggered advection into g_vy / g_vx; g_ty / g_tx (= -seed) may be NULL.  workspace of _bwd: pdec_advect_2d_workspace_bytes. */
size_t pdec_staggered_force_loss_workspace_bytes(int B, int ny, int nx);
int pdec_staggered_force_loss_fwd(const float* vy, const float* vx, const float* ty, const float* tx, float* resid_y, float* resid_x,
                                  float* loss, int B, int ny, int nx, int mode, double dt, float lower_y, float lower_x, float size_y,
                                  float size_x, void* workspace, size_t workspace_bytes, void* stream);
int pdec_staggered_force_loss_bwd(const float* vy, const float* vx, const float* resid_y, const float* resid_x, const float* g_loss,
                                  float* seed_y, float* seed_x, float* g_vy, float* g_vx, float* g_ty, float* g_tx, int B, int ny, int nx,
                                  int mode, double dt, float lower_y, float lower_x, float size_y, float size_x, void* workspace,
                                  size_t workspace_bytes, void* stream);

/* PoissonDomain.with_hard_boundary_conditions (PF/physics/pressuresolver/solver_api.py:111-121, no obstacles: faces on a
 * solid domain boundary are zeroed when closed != 0) fused with StaggeredGrid.divergence(physical_units=False)
 * (staggered_grid.py:201-209). */
int pdec_mask_divergence_2d_fwd(const float* vy, const float* vx, float* div, int B, int ny, int nx, int closed, void* stream);
int pdec_mask_divergence_2d_bwd(const float* g_div, float* g_vy, float* g_vx, int accumulate, int B, int ny, int nx, int closed,
                                void* stream);

/* divergence_free tail, PF/physics/fluid.py:237-242: v_out = mask*v - mask*grad(p*dx0)/dx with replicate-padded
 * gradient (staggered_grid.py:219-231: boundary-face gradient 0).  bwd outputs may be NULL. */
int pdec_subtract_pressure_gradient_2d_fwd(const float* vy, const float* vx, const float* p, float* vy_out, float* vx_out, int B,
                                           int ny, int nx, int closed, float size_y, float size_x, void* stream);
int pdec_subtract_pressure_gradient_2d_bwd(const float* g_vy_out, const float* g_vx_out, float* g_vy, float* g_vx, float* g_p,
                                           int accumulate_v, int B, int ny, int nx, int closed, float size_y, float size_x,
                                           void* stream);

/* CenteredGrid.normalized -> math.normalize_to (PF/physics/field/grid.py:205-209, PF/math/nd.py:28-40):
 * out = target * (sum(source) / max(sum(target), epsilon)) per batch element; deterministic two-stage sums. */
size_t pdec_normalize_2d_workspace_bytes(int B, int ny, int nx);
int pdec_normalize_2d_fwd(const float* target, const float* source, float* out, int B, int ny, int nx, float epsilon,
                          void* workspace, size_t workspace_bytes, void* stream);
int pdec_normalize_2d_bwd(const float* target, const float* source, const float* g_out, float* g_target, float* g_source, int B,
                          int ny, int nx, float epsilon, void* workspace, size_t workspace_bytes, void* stream);

/* poisson_solve with SparseCG (PF/physics/pressuresolver/solver_api.py:132-158, sparse.py:73-111) = conjugate_gradient
 * (PF/math/blas.py:5-64) on the matrix-free 5-point operator of sparse.py:114-221 without obstacles (off-diagonals 1 towards
 * in-domain neighbours; diagonal -4 when open, -(in-domain neighbours) when closed).  Native twin: PressureSolve,
 * PF/tf/cuda/src/pressure_solve_op.cu.cc:19-227.  The SAME entry serves the adjoint: dL/d(rhs) = solve(dL/dp)
 * (sparse.py:93-94; tf_cuda_pressuresolver.py:43-50).
 * rhs, p [B,ny,nx]; guess [B,ny,nx] or NULL (zero); iterations [B] int32 (device) or NULL; residual [B] (device, final
 * max|r| per element) or NULL.  accuracy <= 0: run exactly max_iterations.
 * stop_mode PDEC_CG_STOP_PER_ELEMENT: every batch element stops on its own max|r| < accuracy;
 * PDEC_CG_STOP_WHOLE_BATCH: all elements iterate until the batch-wide max|r| < accuracy (blas.py:33-35). */
#define PDEC_CG_STOP_PER_ELEMENT 0
#define PDEC_CG_STOP_WHOLE_BATCH 1
size_t pdec_cg_poisson_2d_workspace_bytes(int B, int ny, int nx);
int pdec_cg_poisson_2d(const float* rhs, const float* guess, float* p, int* iterations, float* residual, int B, int ny, int nx,
                       int closed, float accuracy, int max_iterations, int stop_mode, void* workspace, size_t workspace_bytes,
                       void* stream);

/* The WHOLE pressure projection in one launch (divergence_free, PF/physics/fluid.py:214-243 = with_hard_boundary_conditions
 * solver_api.py:111-121 -> StaggeredGrid.divergence staggered_grid.py:201-209 -> poisson_solve / SparseCG sparse.py:73-111 ->
 * p *= dx, velocity -= mask * gradient(p) fluid.py:240-242) for grids up to 128 x 128 without obstacles: the on-chip CG of
 * pdec_cg_poisson_2d (one CTA per batch element, zero initial guess, per-element stop rule) forms its right-hand side from the
 * staggered field in its prologue and writes the corrected field in its epilogue, so the divergence and the pressure never make
 * a round trip through HBM and the two stencil launches around the solve disappear.  Bit-identical to the composition
 * pdec_mask_divergence_2d_fwd -> pdec_cg_poisson_2d -> pdec_subtract_pressure_gradient_2d_fwd.
 * p [B,ny,nx] and div [B,ny,nx] are optional outputs (NULL: not stored); iterations / residual as pdec_cg_poisson_2d.
 * _bwd is the adjoint of the whole triple (TF autodiff of those ops + the registered CG gradient sparse.py:93-94): given
 * dL/dv' (and optionally the upstream dL/dp, dL/ddiv, else NULL) it writes dL/dv, running the same CG on the symmetric matrix.
 * Larger grids: pdec_project_2d_supported() == 0 and the calls return PDEC_ERR_UNSUPPORTED_SIZE; compose the three entry points. */
int pdec_project_2d_supported(int ny, int nx);
int pdec_project_2d_fwd(const float* vy, const float* vx, float* vy_out, float* vx_out, float* p, float* div, int* iterations,
                        float* residual, int B, int ny, int nx, int closed, float size_y, float size_x, float accuracy,
                        int max_iterations, void* stream);
int pdec_project_2d_bwd(const float* g_vy_out, const float* g_vx_out, const float* g_p, const float* g_div, float* g_vy, float* g_vx,
                        int* iterations, float* residual, int B, int ny, int nx, int closed, float size_y, float size_x,
                        float accuracy, int max_iterations, void* stream);

/* EXACT pressure solve by fast diagonalisation (opt-in, next to the CG): replaces the reference's direct solver class SparseSciPy
 * (PF/physics/pressuresolver/sparse.py:14-37, scipy spsolve on the matrix of sparse.py:114-221; the reference's default on its NumPy
 * backend, solver_api.py:144-148).  Rectangular domains up to 128 x 128 without obstacles: A = T_y (x) I + I (x) T_x with closed-form
 * eigenvectors per axis (sine basis on OPEN, cosine basis on CLOSED domains), p = E_y [(E_y^T rhs E_x) / (l_y + l_x)] E_x^T as four
 * on-chip 128^3 fp32 products per batch element (one CTA each).  No iteration count, no accuracy knob; on a CLOSED domain the
 * mean-free solution (what CG reaches from a zero guess).  The same call is its own adjoint (A symmetric, sparse.py:93-94).
 * `spectral`: pdec_poisson_direct_2d_spectral_bytes() bytes of device memory owned by the caller, filled ONCE per (ny, nx, closed)
 * by pdec_poisson_direct_2d_setup (eigenvectors evaluated in fp64 on the device) and read-only afterwards (shareable by streams).
 * A binder with other boundary rules (per-side mixed) may fill the block itself: five [128][128] float matrices, zero padded --
 * E_y (E[k][m]: cell k, eigenvector m), E_x, E_x^T, E_y^T, inv[j][i] = 1 / (l_x(j) + l_y(i)) (0 for a singular mode). */
size_t pdec_poisson_direct_2d_spectral_bytes(void);
int pdec_poisson_direct_2d_supported(int ny, int nx);
int pdec_poisson_direct_2d_setup(void* spectral, size_t spectral_bytes, int ny, int nx, int closed, void* stream);
int pdec_poisson_direct_2d(const float* rhs, float* p, int B, int ny, int nx, const void* spectral, void* stream);

/* The same exact solve on the tensor cores (tcgen05.mma kind::tf32, accumulators in tensor memory): every 128^3 product is three TF32
 * products on exact hi / lo splits of both operands (A_hi B_hi + A_lo B_hi + A_hi B_lo, fp32 accumulation), so the result stays at
 * fp32 rounding level.  Same arguments and meaning as pdec_poisson_direct_2d*, but its own spectral block (the constant matrices
 * pre-split and laid out as tensor-core operands): pdec_poisson_direct_tc_2d_spectral_bytes() bytes, 16-byte aligned, filled by
 * pdec_poisson_direct_tc_2d_setup.  Replaces the same reference class (SparseSciPy, PF/physics/pressuresolver/sparse.py:14-37). */
size_t pdec_poisson_direct_tc_2d_spectral_bytes(void);
int pdec_poisson_direct_tc_2d_setup(void* spectral, size_t spectral_bytes, int ny, int nx, int closed, void* stream);
int pdec_poisson_direct_tc_2d(const float* rhs, float* p, int B, int ny, int nx, const void* spectral, void* stream);

/* The stencil of a domain with obstacles from its cell mask (replaces the LaplaceMatrix op, PF/tf/cuda/src/laplace_op.cc:10-17, and
 * PoissonDomain's mask logic, solver_api.py:51-121): active [ny*nx] floats (1 fluid, 0 obstacle) -> code [ny*nx] bytes as
 * pdec_cg_poisson_2d_masked reads them, and the face masks fy [(ny+1)*nx], fx [ny*(nx+1)] that multiply the velocity before the
 * divergence and the pressure gradient before the subtraction (_frictionless_velocity_mask).  Any output may be NULL. */
int pdec_laplace_matrix_2d(const float* active, unsigned char* code, float* fy, float* fx, int ny, int nx, int closed, void* stream);

/* The same solve on a domain with OBSTACLES (PoissonDomain active / accessible masks, solver_api.py:51-121; matrix
 * sparse.py:157-221): `code` [ny*nx] bytes (device, shared by the batch) gives every cell its stencil: bits 0..3 = coupled to the
 * neighbour above (i-1) / below (i+1) / left (j-1) / right (j+1), i.e. active_self * active_neighbour; bits 4..7 = minus the
 * diagonal = number of accessible neighbours clamped to >= 1 (domain boundaries count as accessible when open).  Grids up to
 * 128 x 128 with guess == NULL and the per-element stop rule run on chip (one CTA per batch element; rows without a coupled
 * neighbour are solved directly: p = rhs / diagonal); everything else runs on the streaming cooperative kernel (stop_mode | 0x100
 * forces it).  Workspace as pdec_cg_poisson_2d.  The adjoint is the same call (the masked matrix is symmetric). */
int pdec_cg_poisson_2d_masked(const float* rhs, const float* guess, float* p, int* iterations, float* residual,
                              const unsigned char* code, int B, int ny, int nx, float accuracy, int max_iterations, int stop_mode,
                              void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Tiled / slab-decomposed CG  (BASELINE config 5: 2048 x 2048, row slabs over 2/4/8 GPUs)
 * The same solve as pdec_cg_poisson_2d (zero initial guess, per-element stop rule), on a grid of on-chip tiles with ONE
 * grid-wide barrier per iteration; with nranks > 1 every rank passes its slab of rows [row0, row0 + ny_local) of the
 * [ny_global, nx] domain and the barrier spans the GPUs over peer memory (halo rows and the dot-product partials are stored
 * straight into the neighbours' communication buffers; no NCCL call on the data path).  Replaces, for the decomposed case,
 * the halo exchange + allreduce a distributed PF/math/blas.py:5-64 would need around sparse.py:105-111.
 * All ranks must call with the same B, nx, ny_global, accuracy, max_iterations, in the same order.
 * Requirements: ny_local a multiple of 8; B * tiles <= resident CTAs (else PDEC_ERR_UNSUPPORTED_SIZE).
 * iterations[b] = -1 if the solve was abandoned (a participant timed out).
 * Communication buffer: pdec_cg_slab_comm_bytes(B, nx) bytes per rank, peer-accessible (pdec_comm_alloc + pdec_comm_open on
 * the other ranks, or any other peer-mapped allocation), initialised ONCE by pdec_cg_slab_comm_init before the peers use it. */
typedef struct pdec_slab_comm {
    void* local;     /* this rank's communication buffer */
    void* peers[8];  /* every rank's buffer as mapped into THIS process (peers[rank] == local) */
    int rank, nranks;
} pdec_slab_comm;
size_t pdec_cg_slab_comm_bytes(int B, int nx);
size_t pdec_cg_slab_workspace_bytes(int B, int ny_local, int nx);
int pdec_cg_slab_comm_init(void* comm, int B, int nx, void* stream);
int pdec_cg_poisson_2d_slab(const float* rhs, float* p, int* iterations, float* residual, int B, int ny_local, int nx,
                            int ny_global, int row0, int closed, float accuracy, int max_iterations, const pdec_slab_comm* comm,
                            void* workspace, size_t workspace_bytes, void* stream);
/* Halo rows of row-slab decomposed fields, exchanged over NVLink peer memory (SURVEY 8e / 8b "halo pack/unpack for slabs").  The reference
 * has no distributed code; these entries are what a slab-decomposed IncompressibleFlow.step (PF/physics/fluid.py:90-114) needs around
 * advect.semi_lagrangian (advect.py:24-40: rows within max|v| dt / dx + 1 of a seam) and StaggeredGrid.gradient (staggered_grid.py:219-231:
 * one row of the pressure).  field [B, rows_total, row_len] (row_len = nx, or nx + 1 for the x-face component).
 * _push_rows STORES rows [row_first, row_first + nrows) of every batch element straight into the NEIGHBOUR's inbox (peer_inbox = the
 * neighbour's inbox as mapped into this process: pdec_comm_alloc / pdec_comm_open); _pull_rows waits for the rows the neighbour pushed
 * and writes (accumulate = 0) or adds (accumulate = 1: the adjoint of an exchange is the reverse exchange, summed) them to rows
 * [row_first, row_first + nrows) of its own field.  side: 0 = the rows travel UP (they arrive from the rank below and become a lower
 * halo), 1 = DOWN.  Both ends pass the same `seq` (>= 1, incremented by the host for every exchange on an inbox; a sender is never more
 * than one exchange ahead, which the two parity slots of the inbox absorb).  Every float travels in an 8-byte word with its sequence
 * number: no fence, no NCCL call, no host round trip.  status (device int, may be NULL) <- -1 if a word did not arrive within ~10 s.
 * An inbox holds pdec_halo_inbox_bytes(B, row_len, max_rows) bytes, zeroed ONCE by pdec_halo_inbox_init before any peer pushes. */
size_t pdec_halo_inbox_bytes(int B, int row_len, int max_rows);
int pdec_halo_inbox_init(void* inbox, int B, int row_len, int max_rows, void* stream);
int pdec_halo_push_rows(const float* field, int B, int rows_total, int row_len, int row_first, int nrows, void* peer_inbox, int side,
                        unsigned seq, int max_rows, void* stream);
int pdec_halo_pull_rows(float* field, int B, int rows_total, int row_len, int row_first, int nrows, const void* my_inbox, int side,
                        unsigned seq, int accumulate, int max_rows, int* status, void* stream);
/* peer-visible device allocations for the above: cudaMalloc + CUDA IPC handle (64 bytes, exchanged by the host program) */
int pdec_comm_alloc(size_t bytes, void** ptr, unsigned char* handle64);
int pdec_comm_open(const unsigned char* handle64, void** ptr);
int pdec_comm_close(void* ptr);
int pdec_comm_free(void* ptr);

#ifdef __cplusplus
}
#endif
#endif /* PDEC_H */
