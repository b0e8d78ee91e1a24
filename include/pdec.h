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

/* boundary / extrapolation modes.  Reference: PF/physics/field/grid.py:240-249 (_pad_mode_str)
 * 'constant' -> zero outside, 'boundary' -> replicate, 'periodic' -> circular;
 * resample semantics follow the TF CPU path PF/tf/tf_backend.py:451-538. */
#define PDEC_MODE_CONSTANT 0
#define PDEC_MODE_REPLICATE 1
#define PDEC_MODE_CIRCULAR 2

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

#ifdef __cplusplus
}
#endif
#endif /* PDEC_H */
