"""torch.autograd Functions over the C-ABI kernels of libpdec.so.

Every op runs on CUDA tensors on `torch.cuda.current_stream()`; none has a CPU fallback (a CPU tensor raises).
Reference semantics are cited per op (PF/ = /root/reference/PhiFlow/phi/).
"""
import torch

from . import _lib
from ._lib import MODE_CONSTANT, MODE_REPLICATE, MODE_CIRCULAR, EFFECT_NONE, EFFECT_GROW, EFFECT_ADD  # noqa: F401


def _req_cuda(*tensors):
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError("pde_control_b200 ops need CUDA tensors (there is no CPU fallback)")
        if t.dtype != torch.float32:
            raise RuntimeError("pde_control_b200 ops are fp32; got %s" % t.dtype)


def _c(t):
    return None if t is None else t.contiguous()


# ------------------------------------------------------------------------------------------------
# Burgers 1-D
# ------------------------------------------------------------------------------------------------

class _BurgersTrajectory(torch.autograd.Function):
    """nsteps of Burgers.step_velocity (PF/physics/burgers.py:42-48) on [B, n] rows; returns all frames
    [nsteps, B, n].  Backward = pdec_burgers_trajectory_bwd (deterministic, no float atomics)."""

    @staticmethod
    def forward(ctx, u0, force, nsteps, mode, dt, viscosity, lower, size, substeps, resubstep, effect_mode):
        _req_cuda(u0, force)
        u0 = _c(u0)
        force = _c(force)
        B, n = u0.shape
        frames = torch.empty((nsteps, B, n), device=u0.device, dtype=torch.float32)
        with torch.cuda.device(u0.device):
            rc = _lib.lib().pdec_burgers_trajectory_fwd(_lib.ptr(u0), _lib.ptr(force), _lib.ptr(frames), B, n, nsteps, 1,
                                                        mode, dt, viscosity, lower, size, substeps, int(resubstep),
                                                        effect_mode, _lib.stream_ptr())
        _lib.check(rc, "pdec_burgers_trajectory_fwd")
        ctx.save_for_backward(u0, frames, force)
        ctx.cfg = (nsteps, mode, dt, viscosity, lower, size, substeps, resubstep, effect_mode)
        return frames

    @staticmethod
    def backward(ctx, g_frames):
        u0, frames, force = ctx.saved_tensors
        nsteps, mode, dt, viscosity, lower, size, substeps, resubstep, effect_mode = ctx.cfg
        g_frames = _c(g_frames)
        B, n = u0.shape
        g_u0 = torch.empty_like(u0)
        want_gf = force is not None and ctx.needs_input_grad[1]
        g_force = torch.empty_like(u0) if want_gf else None
        with torch.cuda.device(u0.device):
            rc = _lib.lib().pdec_burgers_trajectory_bwd(_lib.ptr(u0), _lib.ptr(frames), _lib.ptr(g_frames), _lib.ptr(g_u0),
                                                        _lib.ptr(g_force), B, n, nsteps, 1, mode, dt, viscosity, lower,
                                                        size, substeps, int(resubstep), effect_mode, _lib.stream_ptr())
        _lib.check(rc, "pdec_burgers_trajectory_bwd")
        return (g_u0, g_force) + (None,) * 9


def burgers_trajectory(u0, nsteps, dt, viscosity, lower=0.0, size=1.0, mode=MODE_CONSTANT, force=None,
                       effect_mode=None, substeps=1, resubstep=False):
    """All frames 1..nsteps of the Burgers solver started at u0 [B, n]; differentiable w.r.t. u0 and force."""
    if effect_mode is None:
        effect_mode = EFFECT_NONE if force is None else EFFECT_GROW
    if force is None:
        effect_mode = EFFECT_NONE
    return _BurgersTrajectory.apply(u0, force, int(nsteps), int(mode), float(dt), float(viscosity), float(lower),
                                    float(size), int(substeps), bool(resubstep), int(effect_mode))


def burgers_step(u, dt, viscosity, **kw):
    """One Burgers.step_velocity; [B, n] -> [B, n]."""
    return burgers_trajectory(u, 1, dt, viscosity, **kw)[0]


class _BurgersPairLoss(torch.autograd.Function):
    """BurgersPDE.total_force_loss (src/control/pde/burgers.py:24-35) fused with its gradient: the forward
    launch already computes dL/ds for unit upstream gradients; backward only scales it."""

    @staticmethod
    def forward(ctx, s, mode, dt, viscosity, lower, size, substeps, resubstep):
        _req_cuda(s)
        s = _c(s)
        T1, B, n = s.shape
        T = T1 - 1
        need_grad = s.requires_grad
        losses = torch.empty(2, device=s.device, dtype=torch.float32)
        L = _lib.lib()
        ws_bytes = L.pdec_burgers_pair_loss_workspace_bytes(B, n, T)
        ws = torch.empty((ws_bytes + 3) // 4, device=s.device, dtype=torch.float32)
        grad_l2 = torch.empty_like(s) if need_grad else None
        with torch.cuda.device(s.device):
            rc = L.pdec_burgers_pair_loss(_lib.ptr(s), _lib.ptr(grad_l2), _lib.ptr(losses), T, B, n, mode, dt, viscosity,
                                          lower, size, substeps, int(resubstep), 1.0, 0.0, _lib.ptr(ws), ws_bytes,
                                          _lib.stream_ptr())
        _lib.check(rc, "pdec_burgers_pair_loss")
        ctx.cfg = (mode, dt, viscosity, lower, size, substeps, resubstep)
        ctx.save_for_backward(s, grad_l2)
        l2, l1 = losses[0], losses[1]
        ctx.mark_non_differentiable(l1)   # "Total Force" is a reported scalar (burgers.py:34), not optimised
        return l2, l1

    @staticmethod
    def backward(ctx, g_l2, g_l1):
        s, grad_l2 = ctx.saved_tensors
        return (grad_l2 * g_l2,) + (None,) * 7


def burgers_pair_loss(s, dt, viscosity, lower=0.0, size=1.0, mode=MODE_CONSTANT, substeps=1, resubstep=False):
    """(L2, L1) force losses of a frame sequence s [T+1, B, n]; L2 differentiable w.r.t. s."""
    return _BurgersPairLoss.apply(s, int(mode), float(dt), float(viscosity), float(lower), float(size), int(substeps),
                                  bool(resubstep))


def burgers_pair_loss_raw(s, grad_out, losses_out, workspace, dt, viscosity, lower=0.0, size=1.0, mode=MODE_CONSTANT,
                          substeps=1, resubstep=False, gl2=1.0, gl1=0.0):
    """Allocation-free form for benchmarks / CUDA graphs: writes grad_out [T+1,B,n] and losses_out [2]."""
    T1, B, n = s.shape
    rc = _lib.lib().pdec_burgers_pair_loss(_lib.ptr(s), _lib.ptr(grad_out), _lib.ptr(losses_out), T1 - 1, B, n, int(mode),
                                           float(dt), float(viscosity), float(lower), float(size), int(substeps),
                                           int(resubstep), float(gl2), float(gl1), _lib.ptr(workspace),
                                           workspace.numel() * workspace.element_size(), _lib.stream_ptr())
    _lib.check(rc, "pdec_burgers_pair_loss")
