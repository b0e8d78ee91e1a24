"""torch.autograd Functions over the C-ABI kernels of libpdec.so.

Every op runs on CUDA tensors on `torch.cuda.current_stream()`; none has a CPU fallback (a CPU tensor raises).
Reference semantics are cited per op (PF/ = /root/reference/PhiFlow/phi/).
"""
import os

import numpy as np
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
        need_grad = ctx.needs_input_grad[0]       # not s.requires_grad: forward runs with grad mode off and _c() may copy (non-contiguous s)
        s = _c(s)
        T1, B, n = s.shape
        T = T1 - 1
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


class HostPipeline(object):
    """Force loss fed from page-locked HOST arrays (the reference feeds and fetches host arrays every optimisation step:
    src/control/control_training.py:113-142 through PF/tf/session.py).  Owns three streams and the device staging of
    pdec_host_pipeline (include/pdec.h); `chunk_rows` rows of the batch travel per leg, so H2D, the fused kernel and D2H of
    neighbouring chunks overlap.  frames_per_slice > 0: when the fused unit-box kernel applies the work is cut along time instead
    (slices of that many frame pairs of all rows; contiguous copies both ways), at the price of whole-batch device staging for
    max_batch rows.  Measured on B200 (scripts/diag/pcie_probe.py, pcie_timeline.py): both cuts reach the same 1.84-1.86 ms for
    2 x 69 MB, within 8 % of what chunked full-duplex PCIe delivers (about 44 GB/s per direction), so rows are the default.
    Where the fused unit-box kernel applies and grad_host is addressable by the device (torch pin_memory), the kernel stores the
    gradient into grad_host itself and the D2H leg disappears (1.71-1.77 ms, profiles/r1q_host_pipeline.md)."""

    def __init__(self, n, T, chunk_rows=512, max_batch=4096, frames_per_slice=0, device=None):
        import ctypes
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.n, self.T, self.chunk_rows = int(n), int(T), int(chunk_rows)
        self.max_chunks = (int(max_batch) + self.chunk_rows - 1) // self.chunk_rows
        h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            rc = _lib.lib().pdec_host_pipeline_create(ctypes.byref(h), self.chunk_rows, self.n, self.T, self.max_chunks,
                                                      int(frames_per_slice))
        _lib.check(rc, "pdec_host_pipeline_create")
        self._h = h

    def pair_loss(self, s_host, grad_host, losses_host, dt, viscosity, lower=0.0, size=1.0, mode=MODE_CONSTANT, substeps=1,
                  resubstep=False, gl2=1.0, gl1=0.0):
        """s_host [T+1,B,n], grad_host (same shape, or None: losses only), losses_host [2]: contiguous fp32 CPU tensors,
        pinned.  Asynchronous on the current stream of the pipeline's device; synchronise it before reading the outputs."""
        for t in (s_host, grad_host, losses_host):
            if t is None:
                continue
            if t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous():
                raise ValueError("host pipeline takes contiguous fp32 CPU tensors")
            if not t.is_pinned():
                raise ValueError("host pipeline needs page-locked (pinned) host tensors")
        T1, B, n = s_host.shape
        if grad_host is not None and tuple(grad_host.shape) != (T1, B, n):
            raise ValueError("grad_host must have the shape of s_host")
        if losses_host.numel() != 2:
            raise ValueError("losses_host must hold 2 floats")
        import ctypes
        with torch.cuda.device(self.device):
            rc = _lib.lib().pdec_burgers_pair_loss_host(
                self._h, ctypes.c_void_p(s_host.data_ptr()), None if grad_host is None else ctypes.c_void_p(grad_host.data_ptr()),
                ctypes.c_void_p(losses_host.data_ptr()), T1 - 1, B, n, int(mode), float(dt), float(viscosity), float(lower),
                float(size), int(substeps), int(resubstep), float(gl2), float(gl1), _lib.stream_ptr())
        _lib.check(rc, "pdec_burgers_pair_loss_host")

    def close(self):
        if self._h is not None:
            with torch.cuda.device(self.device):
                torch.cuda.synchronize()
                _lib.lib().pdec_host_pipeline_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ------------------------------------------------------------------------------------------------
# 2-D incompressible flow.  Layouts: rho[B,ny,nx], vy[B,ny+1,nx], vx[B,ny,nx+1]; axis 0 = y.
# box = (lower_y, lower_x, size_y, size_x) of the domain; None -> AABox(0, resolution) (dx = 1).
# ------------------------------------------------------------------------------------------------

class _Resample(torch.autograd.Function):
    """backend.resample (tf_backend.py:145-154): out[b, p] = multilinear sample of data[b] at coords[b | 0, p]."""

    @staticmethod
    def forward(ctx, data, coords, mode):
        B = data.shape[0]
        rank = data.dim() - 1
        assert rank in (1, 2) and coords.shape[-1] == rank and coords.shape[0] in (1, B), (data.shape, coords.shape)
        data, coords = _c(data), _c(coords)
        _req_cuda(data, coords)
        P = int(np.prod(coords.shape[1:-1]))
        out = torch.empty((B,) + tuple(coords.shape[1:-1]), device=data.device, dtype=torch.float32)
        ny, nx = (1, data.shape[1]) if rank == 1 else (data.shape[1], data.shape[2])
        with torch.cuda.device(data.device):
            _call("pdec_resample_fwd", _lib.ptr(data), _lib.ptr(coords), _lib.ptr(out), B, rank, ny, nx, coords.shape[0], P, int(mode),
                  _lib.stream_ptr())
        ctx.save_for_backward(data, coords)
        ctx.meta = (B, rank, ny, nx, P, int(mode))
        return out

    @staticmethod
    def backward(ctx, g_out):
        data, coords = ctx.saved_tensors
        B, rank, ny, nx, P, mode = ctx.meta
        g_out = _c(g_out)
        g_data = torch.empty_like(data) if ctx.needs_input_grad[0] else None
        g_coords = torch.empty((B,) + tuple(coords.shape[1:]), device=data.device, dtype=torch.float32) if ctx.needs_input_grad[1] else None
        nbytes = _lib.lib().pdec_resample_bwd_workspace_bytes(B, rank, ny, nx)
        ws = _lib.workspace(data.device, nbytes)
        with torch.cuda.device(data.device):
            _call("pdec_resample_bwd", _lib.ptr(data), _lib.ptr(coords), _lib.ptr(g_out), _lib.ptr(g_data), _lib.ptr(g_coords), B, rank, ny,
                  nx, coords.shape[0], P, mode, _lib.ptr(ws), nbytes, _lib.stream_ptr())
        if g_coords is not None and coords.shape[0] == 1 and B > 1:
            g_coords = g_coords.sum(0, keepdim=True)
        return g_data, g_coords, None


def resample(data, coords, mode=MODE_CONSTANT):
    """Multilinear resampling of a one-channel grid data [B, ny, nx] | [B, nx] at coords [B | 1, ..., rank] (index space, axis 0 first)
    -> [B, ...]; differentiable w.r.t. data and coords.  The operator the reference's backend exposes (tf_backend.py:145-154)."""
    return _Resample.apply(data, coords, int(mode))


def closed_code(closed):
    """The C-ABI's `closed` argument: False / True (every side OPEN / CLOSED), an int code, or four flags (y-lower, y-upper, x-lower,
    x-upper side solid) -> PDEC_CLOSED_SIDES (include/pdec.h)."""
    if isinstance(closed, (tuple, list)):
        ylo, yhi, xlo, xhi = (bool(c) for c in closed)
        if ylo == yhi == xlo == xhi:
            return int(ylo)
        return 0x10 | int(ylo) | int(yhi) << 1 | int(xlo) << 2 | int(xhi) << 3
    return int(closed)


def _box(box, ny, nx):
    if box is None:
        return (0.0, 0.0, float(ny), float(nx))
    return tuple(float(v) for v in box)


def _call(name, *args):
    rc = getattr(_lib.lib(), name)(*args)
    _lib.check(rc, name)


class _Curl2D(torch.autograd.Function):
    """staggered_curl_2d (PF/physics/field/util.py:60-67)."""

    @staticmethod
    def forward(ctx, phi, mode):
        _req_cuda(phi)
        phi = _c(phi)
        B, ny, nx = phi.shape
        vy = torch.empty((B, ny + 1, nx), device=phi.device, dtype=torch.float32)
        vx = torch.empty((B, ny, nx + 1), device=phi.device, dtype=torch.float32)
        with torch.cuda.device(phi.device):
            _call("pdec_curl_2d_fwd", _lib.ptr(phi), _lib.ptr(vy), _lib.ptr(vx), B, ny, nx, mode, _lib.stream_ptr())
        ctx.cfg = (B, ny, nx, mode)
        return vy, vx

    @staticmethod
    def backward(ctx, g_vy, g_vx):
        B, ny, nx, mode = ctx.cfg
        g_vy, g_vx = _c(g_vy), _c(g_vx)
        g_phi = torch.empty((B, ny, nx), device=g_vy.device, dtype=torch.float32)
        with torch.cuda.device(g_vy.device):
            _call("pdec_curl_2d_bwd", _lib.ptr(g_vy), _lib.ptr(g_vx), _lib.ptr(g_phi), B, ny, nx, mode, _lib.stream_ptr())
        return g_phi, None


def curl_2d(phi, mode=MODE_CONSTANT):
    return _Curl2D.apply(phi, int(mode))


class _AdvectCentered(torch.autograd.Function):
    """advect.semi_lagrangian(density, velocity, dt) (PF/physics/field/advect.py:24-37)."""

    @staticmethod
    def forward(ctx, rho, vy, vx, dt, mode, box):
        _req_cuda(rho, vy, vx)
        rho, vy, vx = _c(rho), _c(vy), _c(vx)
        B, ny, nx = rho.shape
        assert vy.shape == (B, ny + 1, nx) and vx.shape == (B, ny, nx + 1), (rho.shape, vy.shape, vx.shape)
        out = torch.empty_like(rho)
        with torch.cuda.device(rho.device):
            _call("pdec_advect_centered_2d_fwd", _lib.ptr(rho), _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(out), B, ny, nx, mode, dt,
                  *box, _lib.stream_ptr())
        ctx.save_for_backward(rho, vy, vx)
        ctx.cfg = (dt, mode, box)
        return out

    @staticmethod
    def backward(ctx, g_out):
        rho, vy, vx = ctx.saved_tensors
        dt, mode, box = ctx.cfg
        B, ny, nx = rho.shape
        g_out = _c(g_out)
        need_v = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
        g_rho = torch.empty_like(rho) if ctx.needs_input_grad[0] else None
        g_vy = torch.empty_like(vy) if need_v else None
        g_vx = torch.empty_like(vx) if need_v else None
        L = _lib.lib()
        nbytes = L.pdec_advect_2d_workspace_bytes(B, ny, nx)
        ws = _lib.workspace(rho.device, nbytes)
        with torch.cuda.device(rho.device):
            _call("pdec_advect_centered_2d_bwd", _lib.ptr(rho), _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(g_out), _lib.ptr(g_rho),
                  _lib.ptr(g_vy), _lib.ptr(g_vx), 0, B, ny, nx, mode, dt, *box, _lib.ptr(ws), nbytes, _lib.stream_ptr())
        return g_rho, g_vy, g_vx, None, None, None


def advect_centered(rho, vy, vx, dt, mode=MODE_CONSTANT, box=None):
    return _AdvectCentered.apply(rho, vy, vx, float(dt), int(mode), _box(box, rho.shape[1], rho.shape[2]))


class _CurlAdvect(torch.autograd.Function):
    """velocity = staggered_curl_2d(potential); next = advect.semi_lagrangian(density, velocity, dt) in one launch
    (ControlledIncompressibleFlow.step, src/control/pde/incompressible_flow.py:53-59).  The backward takes the gradients that reached
    the velocity by other routes (the force-loss terms) and adds them inside the adjoint kernel instead of in separate launches."""

    @staticmethod
    def forward(ctx, rho, phi, dt, potential_mode, mode, box):
        _req_cuda(rho, phi)
        rho, phi = _c(rho), _c(phi)
        B, ny, nx = rho.shape
        assert phi.shape == rho.shape, (rho.shape, phi.shape)
        out = torch.empty_like(rho)
        vy = torch.empty((B, ny + 1, nx), device=rho.device, dtype=torch.float32)
        vx = torch.empty((B, ny, nx + 1), device=rho.device, dtype=torch.float32)
        with torch.cuda.device(rho.device):
            _call("pdec_curl_advect_2d_fwd", _lib.ptr(rho), _lib.ptr(phi), _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(out), B, ny, nx,
                  potential_mode, mode, dt, *box, _lib.stream_ptr())
        ctx.save_for_backward(rho, vy, vx)
        ctx.cfg = (dt, potential_mode, mode, box)
        ctx.set_materialize_grads(False)
        return out, vy, vx

    @staticmethod
    def backward(ctx, g_out, g_vy, g_vx):
        rho, vy, vx = ctx.saved_tensors
        dt, potential_mode, mode, box = ctx.cfg
        B, ny, nx = rho.shape
        if g_out is None:
            if g_vy is None and g_vx is None:
                return None, None, None, None, None, None
            g_out = torch.zeros_like(rho)
        g_out = _c(g_out)
        g_vy = _c(g_vy) if g_vy is not None else None
        g_vx = _c(g_vx) if g_vx is not None else None
        g_rho = torch.empty_like(rho) if ctx.needs_input_grad[0] else None
        g_phi = torch.empty_like(rho)
        nbytes = _lib.lib().pdec_advect_2d_workspace_bytes(B, ny, nx)
        ws = _lib.workspace(rho.device, nbytes)
        with torch.cuda.device(rho.device):
            _call("pdec_curl_advect_2d_bwd", _lib.ptr(rho), _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(g_out),
                  _lib.ptr(g_vy) if g_vy is not None else 0, _lib.ptr(g_vx) if g_vx is not None else 0,
                  _lib.ptr(g_rho) if g_rho is not None else 0, _lib.ptr(g_phi), B, ny, nx, potential_mode, mode, dt, *box,
                  _lib.ptr(ws), nbytes, _lib.stream_ptr())
        return g_rho, g_phi, None, None, None, None


def curl_advect(rho, phi, dt, potential_mode=MODE_CONSTANT, mode=MODE_CONSTANT, box=None):
    """(next density, vy, vx) of the controlled step; see _CurlAdvect."""
    return _CurlAdvect.apply(rho, phi, float(dt), int(potential_mode), int(mode), _box(box, rho.shape[1], rho.shape[2]))


class _AdvectStaggered(torch.autograd.Function):
    """advect.semi_lagrangian(v, v, dt) on a StaggeredGrid (advect.py:38-40)."""

    @staticmethod
    def forward(ctx, vy, vx, dt, mode, box):
        _req_cuda(vy, vx)
        vy, vx = _c(vy), _c(vx)
        B, ny1, nx = vy.shape
        ny = ny1 - 1
        assert vx.shape == (B, ny, nx + 1), (vy.shape, vx.shape)
        ay, ax = torch.empty_like(vy), torch.empty_like(vx)
        with torch.cuda.device(vy.device):
            _call("pdec_advect_staggered_2d_fwd", _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(ay), _lib.ptr(ax), B, ny, nx, mode, dt, *box,
                  _lib.stream_ptr())
        ctx.save_for_backward(vy, vx)
        ctx.cfg = (dt, mode, box)
        return ay, ax

    @staticmethod
    def backward(ctx, g_ay, g_ax):
        vy, vx = ctx.saved_tensors
        dt, mode, box = ctx.cfg
        B, ny1, nx = vy.shape
        ny = ny1 - 1
        g_ay, g_ax = _c(g_ay), _c(g_ax)
        g_vy, g_vx = torch.empty_like(vy), torch.empty_like(vx)
        nbytes = _lib.lib().pdec_advect_2d_workspace_bytes(B, ny, nx)
        ws = _lib.workspace(vy.device, nbytes)
        with torch.cuda.device(vy.device):
            _call("pdec_advect_staggered_2d_bwd", _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(g_ay), _lib.ptr(g_ax), _lib.ptr(g_vy),
                  _lib.ptr(g_vx), 0, B, ny, nx, mode, dt, *box, _lib.ptr(ws), nbytes, _lib.stream_ptr())
        return g_vy, g_vx, None, None, None


def advect_staggered(vy, vx, dt, mode=MODE_REPLICATE, box=None):
    return _AdvectStaggered.apply(vy, vx, float(dt), int(mode), _box(box, vx.shape[1], vy.shape[2]))


class _StaggeredForceLoss(torch.autograd.Function):
    """l2_loss(advect.semi_lagrangian(v1, v1, dt) - v2) over both staggered components, one fused term of total_force_loss
    (src/control/pde/incompressible_flow.py:31-43).  The residual is the only array kept for the adjoint."""

    @staticmethod
    def forward(ctx, vy, vx, ty, tx, dt, mode, box):
        _req_cuda(vy, vx, ty, tx)
        vy, vx, ty, tx = _c(vy), _c(vx), _c(ty), _c(tx)
        B, ny1, nx = vy.shape
        ny = ny1 - 1
        assert vx.shape == (B, ny, nx + 1) and ty.shape == vy.shape and tx.shape == vx.shape, (vy.shape, vx.shape, ty.shape, tx.shape)
        ry, rx = torch.empty_like(vy), torch.empty_like(vx)
        loss = torch.empty((), device=vy.device, dtype=torch.float32)
        nbytes = _lib.lib().pdec_staggered_force_loss_workspace_bytes(B, ny, nx)
        ws = _lib.workspace(vy.device, nbytes)
        with torch.cuda.device(vy.device):
            _call("pdec_staggered_force_loss_fwd", _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(ty), _lib.ptr(tx), _lib.ptr(ry), _lib.ptr(rx),
                  _lib.ptr(loss), B, ny, nx, mode, dt, *box, _lib.ptr(ws), nbytes, _lib.stream_ptr())
        ctx.save_for_backward(vy, vx, ry, rx)
        ctx.cfg = (dt, mode, box)
        return loss

    @staticmethod
    def backward(ctx, g_loss):
        vy, vx, ry, rx = ctx.saved_tensors
        dt, mode, box = ctx.cfg
        B, ny1, nx = vy.shape
        ny = ny1 - 1
        g_loss = _c(g_loss.to(torch.float32))
        want_t = ctx.needs_input_grad[2] or ctx.needs_input_grad[3]
        sy, sx = torch.empty_like(vy), torch.empty_like(vx)
        g_vy, g_vx = torch.empty_like(vy), torch.empty_like(vx)
        g_ty, g_tx = (torch.empty_like(vy), torch.empty_like(vx)) if want_t else (None, None)
        nbytes = _lib.lib().pdec_advect_2d_workspace_bytes(B, ny, nx)
        ws = _lib.workspace(vy.device, nbytes)
        with torch.cuda.device(vy.device):
            _call("pdec_staggered_force_loss_bwd", _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(ry), _lib.ptr(rx), _lib.ptr(g_loss), _lib.ptr(sy),
                  _lib.ptr(sx), _lib.ptr(g_vy), _lib.ptr(g_vx), _lib.ptr(g_ty) if want_t else 0, _lib.ptr(g_tx) if want_t else 0, B, ny, nx,
                  mode, dt, *box, _lib.ptr(ws), nbytes, _lib.stream_ptr())
        return g_vy, g_vx, g_ty, g_tx, None, None, None


def staggered_force_loss(vy, vx, ty, tx, dt, mode=MODE_REPLICATE, box=None):
    """sum((advected(vy) - ty)^2) / 2 / B + sum((advected(vx) - tx)^2) / 2 / B as a 0-d tensor."""
    return _StaggeredForceLoss.apply(vy, vx, ty, tx, float(dt), int(mode), _box(box, vx.shape[1], vy.shape[2]))


class _MaskDivergence(torch.autograd.Function):
    """with_hard_boundary_conditions + divergence(physical_units=False) (solver_api.py:111-121, staggered_grid.py:201-209)."""

    @staticmethod
    def forward(ctx, vy, vx, closed):
        _req_cuda(vy, vx)
        vy, vx = _c(vy), _c(vx)
        B, ny1, nx = vy.shape
        ny = ny1 - 1
        div = torch.empty((B, ny, nx), device=vy.device, dtype=torch.float32)
        with torch.cuda.device(vy.device):
            _call("pdec_mask_divergence_2d_fwd", _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(div), B, ny, nx, int(closed), _lib.stream_ptr())
        ctx.cfg = (B, ny, nx, int(closed))
        return div

    @staticmethod
    def backward(ctx, g_div):
        B, ny, nx, closed = ctx.cfg
        g_div = _c(g_div)
        g_vy = torch.empty((B, ny + 1, nx), device=g_div.device, dtype=torch.float32)
        g_vx = torch.empty((B, ny, nx + 1), device=g_div.device, dtype=torch.float32)
        with torch.cuda.device(g_div.device):
            _call("pdec_mask_divergence_2d_bwd", _lib.ptr(g_div), _lib.ptr(g_vy), _lib.ptr(g_vx), 0, B, ny, nx, closed, _lib.stream_ptr())
        return g_vy, g_vx, None


def mask_divergence(vy, vx, closed):
    return _MaskDivergence.apply(vy, vx, closed_code(closed))


CG_STOP_PER_ELEMENT, CG_STOP_WHOLE_BATCH = 0, 1


def cg_poisson_raw(rhs, out, closed, accuracy, max_iterations, guess=None, iterations=None, residual=None, stop_mode=0):
    """Allocation-light form: solves A out = rhs on the current stream."""
    B, ny, nx = rhs.shape
    L = _lib.lib()
    nbytes = L.pdec_cg_poisson_2d_workspace_bytes(B, ny, nx)
    ws = _lib.workspace(rhs.device, nbytes)
    with torch.cuda.device(rhs.device):
        _call("pdec_cg_poisson_2d", _lib.ptr(rhs), _lib.ptr(guess), _lib.ptr(out), _lib.ptr(iterations), _lib.ptr(residual), B, ny, nx,
              closed_code(closed), float(accuracy if accuracy is not None else 0.0), int(max_iterations), int(stop_mode), _lib.ptr(ws), nbytes,
              _lib.stream_ptr())


_SPECTRAL = {}


# PDEC_DIRECT_TENSOR_CORES=0 selects the fp32 FFMA kernel of the direct solver (A/B runs); default: the tcgen05 3 x TF32 kernel
_DIRECT_TC = os.environ.get("PDEC_DIRECT_TENSOR_CORES", "1") != "0"


def _spectral_block(device, ny, nx, closed, tensor_cores=None):
    """Eigenvector matrices + reciprocal eigenvalue sums of the (ny, nx, closed) Poisson matrix, built once on the device (one block per
    kernel flavour: plain fp32 matrices for the FFMA kernel; hi / lo TF32 splits in the tensor core's operand layout for the tcgen05 one)."""
    tc = _DIRECT_TC if tensor_cores is None else bool(tensor_cores)
    key = (device.index if device.index is not None else torch.cuda.current_device(), int(ny), int(nx), closed_code(closed), tc)
    blk = _SPECTRAL.get(key)
    if blk is None:
        L = _lib.lib()
        name = "pdec_poisson_direct_tc_2d" if tc else "pdec_poisson_direct_2d"
        nbytes = getattr(L, name + "_spectral_bytes")()
        blk = torch.empty(nbytes // 4, device=device, dtype=torch.float32)
        with torch.cuda.device(device):
            _call(name + "_setup", _lib.ptr(blk), nbytes, int(ny), int(nx), closed_code(closed), _lib.stream_ptr())
        # built on the current stream; other streams (graph branches) only ever read it after this one has been joined
        torch.cuda.current_stream(device).synchronize()
        _SPECTRAL[key] = blk
    return blk


_SPECTRAL_LARGE = {}
DIRECT_LARGE_MAX = 8192         # the eigenvector matrices are dense n x n: 256 MB each at 8192


def _eig_1d(n, ends, device):
    """Orthonormal eigenvectors E[cell, mode] and eigenvalues of tridiag(1, d, 1) with d = -2, -1 at a SOLID end (csrc/poisson_direct.cu
    eig_vec / eig_val: sine basis for two open ends, cosine basis for two solid ends, quarter-wave sines anchored at the open end for
    mixed ends), evaluated in fp64.  ends = (lower solid) | (upper solid) << 1."""
    import math
    i = torch.arange(n, device=device, dtype=torch.float64)[:, None]
    k = torch.arange(n, device=device, dtype=torch.float64)[None, :]
    if ends == 3:
        c = torch.full_like(k, math.sqrt(2.0 / n))
        c[:, 0] = math.sqrt(1.0 / n)
        return c * torch.cos(math.pi * (i + 0.5) * k / n), -2.0 + 2.0 * torch.cos(math.pi * k[0] / n)
    if ends == 0:
        return math.sqrt(2.0 / (n + 1)) * torch.sin(math.pi * (i + 1.0) * (k + 1.0) / (n + 1)), -2.0 + 2.0 * torch.cos(math.pi * (k[0] + 1.0) / (n + 1))
    theta = math.pi * (2.0 * k + 1.0) / (2.0 * n + 1.0)
    pos = (i + 1.0) if ends == 2 else (n - i)
    return math.sqrt(4.0 / (2.0 * n + 1.0)) * torch.sin(pos * theta), -2.0 + 2.0 * torch.cos(theta[0])


def _spectral_large(device, ny, nx, closed):
    """(Ey, Ey^T, Ex, Ex^T, 1 / (l_y + l_x)) in fp32 for grids beyond the on-chip kernels: the four products are plain library GEMMs."""
    code = closed_code(closed)
    key = (device.index if device.index is not None else torch.cuda.current_device(), int(ny), int(nx), code)
    blk = _SPECTRAL_LARGE.get(key)
    if blk is None:
        sides = (0xF if code == 1 else 0) if code < 0x10 else (code & 0xF)
        Ey, ly = _eig_1d(int(ny), sides & 3, device)
        Ex, lx = _eig_1d(int(nx), (sides >> 2) & 3, device)
        lam = ly[:, None] + lx[None, :]
        inv = torch.where(lam.abs() < 1e-9, torch.zeros_like(lam), 1.0 / lam)          # the singular mode of a closed domain -> 0
        blk = tuple(t.to(torch.float32).contiguous() for t in (Ey, Ey.T, Ex, Ex.T, inv))
        torch.cuda.current_stream(device).synchronize()
        _SPECTRAL_LARGE[key] = blk
    return blk


def poisson_direct_large_raw(rhs, out, closed):
    """The fast-diagonalisation solve for grids larger than 128 x 128: p = Ey [(Ey^T rhs Ex) / (l_y + l_x)] Ex^T with four library fp32
    GEMMs (cuBLAS SGEMM through torch.matmul, TF32 off).  2048 x 2048: 1.28 ms per solve, 6e-6 against the same formula in fp64, whose
    own residual is 1e-11 (scripts/diag/direct_large.py) -- against ~24 ms for the slab CG on 2-8 GPUs and 590 ms for a streaming CG."""
    Ey, EyT, Ex, ExT, inv = _spectral_large(rhs.device, rhs.shape[1], rhs.shape[2], closed)
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        t = torch.matmul(torch.matmul(EyT, rhs), Ex)
        t.mul_(inv)
        torch.matmul(torch.matmul(Ey, t), ExT, out=out)
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


def poisson_direct_raw(rhs, out, closed, tensor_cores=None):
    B, ny, nx = rhs.shape
    if not _lib.lib().pdec_poisson_direct_2d_supported(int(ny), int(nx)):
        assert max(ny, nx) <= DIRECT_LARGE_MAX, "direct solver: grid too large for dense eigenvector matrices"
        return poisson_direct_large_raw(rhs, out, closed)
    tc = _DIRECT_TC if tensor_cores is None else bool(tensor_cores)
    blk = _spectral_block(rhs.device, ny, nx, closed, tc)
    with torch.cuda.device(rhs.device):
        _call("pdec_poisson_direct_tc_2d" if tc else "pdec_poisson_direct_2d", _lib.ptr(rhs), _lib.ptr(out), B, ny, nx, _lib.ptr(blk),
              _lib.stream_ptr())


class _PressureSolveDirect(torch.autograd.Function):
    """Exact solve by fast diagonalisation (pdec_poisson_direct_2d; the reference's SparseSciPy, sparse.py:14-37).  Backward = the same
    solve of the incoming gradient (A symmetric, sparse.py:93-94)."""

    @staticmethod
    def forward(ctx, div, closed):
        _req_cuda(div)
        div = _c(div)
        p = torch.empty_like(div)
        poisson_direct_raw(div, p, closed)
        ctx.closed = closed
        return p

    @staticmethod
    def backward(ctx, g_p):
        g_p = _c(g_p)
        g_div = torch.empty_like(g_p)
        poisson_direct_raw(g_p, g_div, ctx.closed)
        return g_div, None


def pressure_solve_direct(div, closed):
    """Exact pressure [B, ny, nx] of a rectangular domain without obstacles (on-chip tensor-core kernel up to 128 x 128, library GEMMs
    beyond); differentiable w.r.t. div."""
    return _PressureSolveDirect.apply(div, closed_code(closed))


def direct_solver_supported(ny, nx):
    """On-chip kernels up to 128 x 128 (tensor cores), library GEMMs beyond (poisson_direct_large_raw)."""
    return min(int(ny), int(nx)) >= 2 and max(int(ny), int(nx)) <= DIRECT_LARGE_MAX


class SlabComm(object):
    """Peer-visible communication buffers of the slab-decomposed CG (include/pdec.h, pdec_slab_comm): one buffer per rank,
    allocated by the library (cudaMalloc), exchanged as CUDA IPC handles over `torch.distributed` and mapped into every
    process.  One process per GPU; `group` defaults to the world group.  With world size 1 no buffer is needed (pass
    comm=None to cg_poisson_slab_raw)."""

    def __init__(self, B, nx, group=None):
        import ctypes
        import torch.distributed as dist
        L = _lib.lib()
        self.rank, self.nranks = dist.get_rank(group), dist.get_world_size(group)
        assert 1 < self.nranks <= 8
        self.B, self.nx = B, nx
        nbytes = L.pdec_cg_slab_comm_bytes(B, nx)
        local = ctypes.c_void_p()
        handle = (ctypes.c_ubyte * 64)()
        _lib.check(L.pdec_comm_alloc(nbytes, ctypes.byref(local), handle), "pdec_comm_alloc")
        _lib.check(L.pdec_cg_slab_comm_init(local, B, nx, _lib.stream_ptr()), "pdec_cg_slab_comm_init")
        torch.cuda.synchronize()
        handles = [None] * self.nranks
        dist.all_gather_object(handles, bytes(handle), group=group)      # also orders every rank's init before any peer use
        self.struct = _lib.SlabCommStruct()
        self.struct.local = local
        self.struct.rank, self.struct.nranks = self.rank, self.nranks
        self._opened = []
        for r, h in enumerate(handles):
            if r == self.rank:
                self.struct.peers[r] = local
                continue
            mapped = ctypes.c_void_p()
            buf = (ctypes.c_ubyte * 64).from_buffer_copy(h)
            _lib.check(L.pdec_comm_open(buf, ctypes.byref(mapped)), "pdec_comm_open")
            self.struct.peers[r] = mapped
            self._opened.append(mapped)
        self._local = local
        dist.barrier(group=group)

    def close(self):
        L = _lib.lib()
        torch.cuda.synchronize()
        for m in self._opened:
            L.pdec_comm_close(m)
        self._opened = []
        if self._local is not None:
            L.pdec_comm_free(self._local)
            self._local = None


def cg_poisson_slab_raw(rhs, out, closed, accuracy, max_iterations, ny_global=None, row0=0, comm=None, iterations=None, residual=None):
    """Tiled on-chip CG of this rank's row slab rhs[B, ny_local, nx] of a [ny_global, nx] domain (whole domain if comm is None)."""
    import ctypes
    B, ny, nx = rhs.shape
    L = _lib.lib()
    nbytes = L.pdec_cg_slab_workspace_bytes(B, ny, nx)
    ws = _lib.workspace(rhs.device, nbytes)
    with torch.cuda.device(rhs.device):
        _call("pdec_cg_poisson_2d_slab", _lib.ptr(rhs), _lib.ptr(out), _lib.ptr(iterations), _lib.ptr(residual), B, ny, nx,
              int(ny if ny_global is None else ny_global), int(row0), closed_code(closed), float(accuracy if accuracy is not None else 0.0),
              int(max_iterations), None if comm is None else ctypes.cast(ctypes.pointer(comm.struct), ctypes.c_void_p),
              _lib.ptr(ws), nbytes, _lib.stream_ptr())


class _PressureSolve(torch.autograd.Function):
    """poisson_solve with SparseCG (solver_api.py:132-158, sparse.py:73-111, math/blas.py:5-64).  Backward = the same
    CG applied to the incoming gradient (sparse.py:93-94: A is symmetric)."""

    @staticmethod
    def forward(ctx, div, closed, accuracy, max_iterations, gradient_accuracy, max_gradient_iterations, guess, stop_mode):
        _req_cuda(div, guess)
        div = _c(div)
        B = div.shape[0]
        p = torch.empty_like(div)
        iters = torch.zeros(B, device=div.device, dtype=torch.int32)
        resid = torch.zeros(B, device=div.device, dtype=torch.float32)
        cg_poisson_raw(div, p, closed, accuracy, max_iterations, _c(guess), iters, resid, stop_mode)
        ctx.cfg = (closed, gradient_accuracy, max_gradient_iterations, stop_mode)
        ctx.mark_non_differentiable(iters, resid)
        return p, iters, resid

    @staticmethod
    def backward(ctx, g_p, _gi, _gr):
        closed, accuracy, max_it, stop_mode = ctx.cfg
        g_p = _c(g_p)
        g_div = torch.empty_like(g_p)
        cg_poisson_raw(g_p, g_div, closed, accuracy, max_it, None, None, None, stop_mode)
        return g_div, None, None, None, None, None, None, None


def pressure_solve(div, closed, accuracy=1e-5, max_iterations=2000, gradient_accuracy=None, max_gradient_iterations=None,
                   guess=None, stop_mode=CG_STOP_PER_ELEMENT):
    """Returns (pressure, iterations[B] int32, residual[B]); only the pressure is differentiable (w.r.t. div)."""
    ga = accuracy if gradient_accuracy is None else gradient_accuracy
    gi = max_iterations if max_gradient_iterations is None else max_gradient_iterations
    return _PressureSolve.apply(div, closed_code(closed), accuracy, int(max_iterations), ga, int(gi), guess, int(stop_mode))


class SlabSolver(object):
    """Row-slab decomposition of the pressure solve over the ranks of a process group (BASELINE config 5): every rank holds the
    whole (replicated) fields -- the stencil operators around the solve are < 1 % of a 2048^2 step -- and solves its slab of rows
    with the tiled on-chip CG whose per-iteration exchange runs over NVLink peer memory (pdec_cg_poisson_2d_slab); the slabs
    of the pressure are then all-gathered (NCCL).  The adjoint is the same solve applied to the incoming gradient."""

    def __init__(self, B, ny, nx, group=None):
        import torch.distributed as dist
        self.group = group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        assert ny % self.world == 0 and (ny // self.world) % 8 == 0, "slabs must be equal and a multiple of 8 rows"
        self.B, self.ny, self.nx, self.rows = B, ny, nx, ny // self.world
        self.comm = SlabComm(B, nx, group) if self.world > 1 else None

    def solve(self, rhs, closed, accuracy, max_iterations):
        """rhs [B, ny, nx] (replicated) -> (p [B, ny, nx] replicated, iterations, residual)"""
        import torch.distributed as dist
        B, ny, nx = rhs.shape
        assert (B, ny, nx) == (self.B, self.ny, self.nx)
        lo = self.rank * self.rows
        mine = rhs[:, lo:lo + self.rows].contiguous()
        out = torch.empty_like(mine)
        iters = torch.zeros(B, device=rhs.device, dtype=torch.int32)
        resid = torch.zeros(B, device=rhs.device, dtype=torch.float32)
        if self.world == 1:
            cg_poisson_raw(mine, out, closed, accuracy, max_iterations, None, iters, resid)
            return out, iters, resid
        cg_poisson_slab_raw(mine, out, closed, accuracy, max_iterations, ny_global=ny, row0=lo, comm=self.comm, iterations=iters, residual=resid)
        parts = [torch.empty_like(out) for _ in range(self.world)]
        dist.all_gather(parts, out, group=self.group)
        return torch.cat(parts, 1), iters, resid

    def close(self):
        if self.comm is not None:
            self.comm.close()
            self.comm = None


class _PressureSolveSlab(torch.autograd.Function):
    @staticmethod
    def forward(ctx, div, solver, closed, accuracy, max_iterations):
        _req_cuda(div)
        p, iters, resid = solver.solve(_c(div), closed, accuracy, max_iterations)
        ctx.cfg = (solver, closed, accuracy, max_iterations)
        ctx.mark_non_differentiable(iters, resid)
        return p, iters, resid

    @staticmethod
    def backward(ctx, g_p, _gi, _gr):
        solver, closed, accuracy, max_it = ctx.cfg
        g_div, _, _ = solver.solve(_c(g_p), closed, accuracy, max_it)
        return g_div, None, None, None, None


class _SubtractPressureGradient(torch.autograd.Function):
    """divergence_free tail (PF/physics/fluid.py:237-242)."""

    @staticmethod
    def forward(ctx, vy, vx, p, closed, size_y, size_x):
        _req_cuda(vy, vx, p)
        vy, vx, p = _c(vy), _c(vx), _c(p)
        B, ny, nx = p.shape
        oy, ox = torch.empty_like(vy), torch.empty_like(vx)
        with torch.cuda.device(p.device):
            _call("pdec_subtract_pressure_gradient_2d_fwd", _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(p), _lib.ptr(oy), _lib.ptr(ox), B, ny, nx,
                  int(closed), size_y, size_x, _lib.stream_ptr())
        ctx.cfg = (B, ny, nx, int(closed), size_y, size_x)
        return oy, ox

    @staticmethod
    def backward(ctx, g_oy, g_ox):
        B, ny, nx, closed, size_y, size_x = ctx.cfg
        g_oy, g_ox = _c(g_oy), _c(g_ox)
        g_vy, g_vx = torch.empty_like(g_oy), torch.empty_like(g_ox)
        g_p = torch.empty((B, ny, nx), device=g_oy.device, dtype=torch.float32)
        with torch.cuda.device(g_oy.device):
            _call("pdec_subtract_pressure_gradient_2d_bwd", _lib.ptr(g_oy), _lib.ptr(g_ox), _lib.ptr(g_vy), _lib.ptr(g_vx), _lib.ptr(g_p), 0,
                  B, ny, nx, closed, size_y, size_x, _lib.stream_ptr())
        return g_vy, g_vx, g_p, None, None, None


def subtract_pressure_gradient(vy, vx, p, closed, box=None):
    b = _box(box, p.shape[1], p.shape[2])
    return _SubtractPressureGradient.apply(vy, vx, p, closed_code(closed), b[2], b[3])


class _Normalize(torch.autograd.Function):
    """CenteredGrid.normalized -> math.normalize_to (grid.py:205-209, nd.py:28-40)."""

    @staticmethod
    def forward(ctx, target, source, epsilon):
        _req_cuda(target, source)
        target, source = _c(target), _c(source)
        B, ny, nx = target.shape
        out = torch.empty_like(target)
        nbytes = _lib.lib().pdec_normalize_2d_workspace_bytes(B, ny, nx)
        ws = _lib.workspace(target.device, nbytes)
        with torch.cuda.device(target.device):
            _call("pdec_normalize_2d_fwd", _lib.ptr(target), _lib.ptr(source), _lib.ptr(out), B, ny, nx, epsilon, _lib.ptr(ws), nbytes,
                  _lib.stream_ptr())
        ctx.save_for_backward(target, source)
        ctx.eps = epsilon
        return out

    @staticmethod
    def backward(ctx, g_out):
        target, source = ctx.saved_tensors
        B, ny, nx = target.shape
        g_out = _c(g_out)
        g_t, g_s = torch.empty_like(target), torch.empty_like(source)
        nbytes = _lib.lib().pdec_normalize_2d_workspace_bytes(B, ny, nx)
        ws = _lib.workspace(target.device, nbytes)
        with torch.cuda.device(target.device):
            _call("pdec_normalize_2d_bwd", _lib.ptr(target), _lib.ptr(source), _lib.ptr(g_out), _lib.ptr(g_t), _lib.ptr(g_s), B, ny, nx,
                  ctx.eps, _lib.ptr(ws), nbytes, _lib.stream_ptr())
        return g_t, g_s, None


def normalize_to(target, source, epsilon=1e-5):
    return _Normalize.apply(target, source, float(epsilon))


_FUSED_PROJECTION = bool(os.environ.get("PDEC_FUSED_PROJECTION"))      # default of divergence_free(fused=None); A/B switch (scripts/diag)


class _Project(torch.autograd.Function):
    """divergence_free (PF/physics/fluid.py:214-243) as ONE launch per direction: mask + divergence in the prologue of the on-chip CG,
    pressure-gradient subtraction in its epilogue (pdec_project_2d_fwd / _bwd).  Outputs (v'_y, v'_x, p, div, iterations); the
    backward launch is the adjoint of the whole triple with the second CG solve inside (sparse.py:93-94)."""

    @staticmethod
    def forward(ctx, vy, vx, closed, size_y, size_x, accuracy, max_iterations, gradient_accuracy, max_gradient_iterations):
        _req_cuda(vy, vx)
        vy, vx = _c(vy), _c(vx)
        B, ny, nx = vx.shape[0], vx.shape[1], vy.shape[2]
        oy, ox = torch.empty_like(vy), torch.empty_like(vx)
        p = torch.empty((B, ny, nx), device=vy.device, dtype=torch.float32)
        div = torch.empty_like(p)
        iters = torch.zeros(B, device=vy.device, dtype=torch.int32)
        with torch.cuda.device(vy.device):
            _call("pdec_project_2d_fwd", _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(oy), _lib.ptr(ox), _lib.ptr(p), _lib.ptr(div), _lib.ptr(iters),
                  None, B, ny, nx, int(closed), size_y, size_x, float(accuracy if accuracy is not None else 0.0), int(max_iterations),
                  _lib.stream_ptr())
        ctx.cfg = (B, ny, nx, int(closed), size_y, size_x, float(gradient_accuracy if gradient_accuracy is not None else 0.0),
                   int(max_gradient_iterations))
        ctx.mark_non_differentiable(iters)
        # p and div are returned for inspection: when nothing downstream uses them backward gets None for their gradients
        # (no zero tensors are materialised), which the kernel takes as "no upstream term"
        ctx.set_materialize_grads(False)
        return oy, ox, p, div, iters

    @staticmethod
    def backward(ctx, g_oy, g_ox, g_p, g_div, _gi):
        B, ny, nx, closed, size_y, size_x, accuracy, max_it = ctx.cfg
        dev = next(t.device for t in (g_oy, g_ox, g_p, g_div) if t is not None)
        g_oy = torch.zeros((B, ny + 1, nx), device=dev) if g_oy is None else _c(g_oy)
        g_ox = torch.zeros((B, ny, nx + 1), device=dev) if g_ox is None else _c(g_ox)
        g_p = None if g_p is None else _c(g_p)
        g_div = None if g_div is None else _c(g_div)
        g_vy, g_vx = torch.empty_like(g_oy), torch.empty_like(g_ox)
        with torch.cuda.device(g_oy.device):
            _call("pdec_project_2d_bwd", _lib.ptr(g_oy), _lib.ptr(g_ox), _lib.ptr(g_p), _lib.ptr(g_div), _lib.ptr(g_vy), _lib.ptr(g_vx),
                  None, None, B, ny, nx, closed, size_y, size_x, accuracy, max_it, _lib.stream_ptr())
        return g_vy, g_vx, None, None, None, None, None, None, None



def _project(vy, vx, closed, box, accuracy, max_iterations, gradient_accuracy, max_gradient_iterations):
    b = _box(box, vx.shape[1], vy.shape[2])
    ga = accuracy if gradient_accuracy is None else gradient_accuracy
    gi = max_iterations if max_gradient_iterations is None else max_gradient_iterations
    return _Project.apply(vy, vx, closed_code(closed), b[2], b[3], accuracy, int(max_iterations), ga, int(gi))


def _cg_from_direct(rhs, out, closed, accuracy, max_iterations, iters, stop_mode):
    """out <- CG(rhs) started from the exact solve (the chain SparseSciPy() & SparseCG()).  OPEN sides somewhere: the CG kernel takes the
    guess.  All sides closed: the matrix is singular (constants), and rounding leaves rhs - A p0 with a small mean no CG step can remove
    (the stopping rule would never be met); so the CORRECTION system A d = (rhs - A p0) - mean is solved from zero and out = p0 + d --
    the constant it drops is annihilated by the gradient that follows."""
    p0 = torch.empty_like(rhs)
    poisson_direct_raw(rhs, p0, closed)
    if closed_code(closed) != 1:
        cg_poisson_raw(rhs, out, closed, accuracy, max_iterations, p0, iters, None, stop_mode)
        return
    from .slab import laplace_rows
    r = rhs - laplace_rows(p0, 0, rhs.shape[1], True)
    r -= r.mean((1, 2), keepdim=True)
    cg_poisson_raw(r, out, closed, accuracy, max_iterations, None, iters, None, stop_mode)
    out += p0


class _ProjectComposed(torch.autograd.Function):
    """divergence_free (PF/physics/fluid.py:214-243) as ONE autograd node around the three launches (divergence, solve, gradient
    subtraction): the backward hands the velocity gradient of the gradient subtraction to the divergence adjoint to accumulate into
    (the kernels' `accumulate` flag) instead of leaving two sums per step to autograd.  Same kernels, same values as the three-node
    composition (the order of the one float addition per face is the same: subtraction term first, divergence term added).
    solver = ("direct",) or ("cg" | "cg_from_direct", accuracy, max_iterations, gradient_accuracy, max_gradient_iterations, stop_mode)."""

    @staticmethod
    def forward(ctx, vy, vx, closed, size_y, size_x, solver):
        _req_cuda(vy, vx)
        vy, vx = _c(vy), _c(vx)
        B, ny1, nx = vy.shape
        ny = ny1 - 1
        dev = vy.device
        div = torch.empty((B, ny, nx), device=dev, dtype=torch.float32)
        p = torch.empty_like(div)
        oy, ox = torch.empty_like(vy), torch.empty_like(vx)
        iters = torch.zeros(B, device=dev, dtype=torch.int32)
        with torch.cuda.device(dev):
            _call("pdec_mask_divergence_2d_fwd", _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(div), B, ny, nx, closed, _lib.stream_ptr())
            if solver[0] == "direct":
                poisson_direct_raw(div, p, closed)
            elif solver[0] == "cg_from_direct":                      # SparseSciPy() & SparseCG(...): the exact solve is the CG's initial guess
                _cg_from_direct(div, p, closed, solver[1], solver[2], iters, solver[5])
            else:
                cg_poisson_raw(div, p, closed, solver[1], solver[2], None, iters, None, solver[5])
            _call("pdec_subtract_pressure_gradient_2d_fwd", _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(p), _lib.ptr(oy), _lib.ptr(ox), B, ny, nx,
                  closed, size_y, size_x, _lib.stream_ptr())
        ctx.cfg = (B, ny, nx, closed, size_y, size_x, solver)
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(iters)
        return oy, ox, p, div, iters

    @staticmethod
    def backward(ctx, g_oy, g_ox, g_p, g_div, _gi):
        B, ny, nx, closed, size_y, size_x, solver = ctx.cfg
        if g_oy is None and g_ox is None and g_p is None and g_div is None:
            return None, None, None, None, None, None
        ref = next(t for t in (g_oy, g_ox, g_p, g_div) if t is not None)
        dev = ref.device
        g_vy = torch.empty((B, ny + 1, nx), device=dev, dtype=torch.float32)
        g_vx = torch.empty((B, ny, nx + 1), device=dev, dtype=torch.float32)
        with torch.cuda.device(dev):
            if g_oy is None and g_ox is None:
                g_vy.zero_(); g_vx.zero_()
                g_pt = _c(g_p) if g_p is not None else None
            else:
                g_oy = _c(g_oy) if g_oy is not None else torch.zeros_like(g_vy)
                g_ox = _c(g_ox) if g_ox is not None else torch.zeros_like(g_vx)
                g_pt = torch.empty((B, ny, nx), device=dev, dtype=torch.float32)
                _call("pdec_subtract_pressure_gradient_2d_bwd", _lib.ptr(g_oy), _lib.ptr(g_ox), _lib.ptr(g_vy), _lib.ptr(g_vx), _lib.ptr(g_pt), 0,
                      B, ny, nx, closed, size_y, size_x, _lib.stream_ptr())
                if g_p is not None:
                    g_pt = g_pt + g_p
            g_d = None
            if g_pt is not None:
                g_d = torch.empty_like(g_pt)
                if solver[0] == "direct":
                    poisson_direct_raw(g_pt, g_d, closed)
                elif solver[0] == "cg_from_direct":                  # the adjoint system is the same matrix: same chain
                    _cg_from_direct(g_pt, g_d, closed, solver[3], solver[4], None, solver[5])
                else:
                    cg_poisson_raw(g_pt, g_d, closed, solver[3], solver[4], None, None, None, solver[5])
            if g_div is not None:
                g_d = _c(g_div) if g_d is None else g_d + g_div
            if g_d is not None:
                _call("pdec_mask_divergence_2d_bwd", _lib.ptr(g_d), _lib.ptr(g_vy), _lib.ptr(g_vx), 1, B, ny, nx, closed, _lib.stream_ptr())
        return g_vy, g_vx, None, None, None, None


class _AdvectPair(torch.autograd.Function):
    """The two advections of IncompressibleFlow.step (PF/physics/fluid.py:102-103) -- density by the velocity, velocity by itself -- as
    one autograd node: the centred adjoint accumulates its velocity gradient into the staggered adjoint's (accumulate_v) instead of two
    separate sums.  Same kernels, same values as advect_centered + advect_staggered."""

    @staticmethod
    def forward(ctx, rho, vy, vx, dt, mode, box):
        _req_cuda(rho, vy, vx)
        rho, vy, vx = _c(rho), _c(vy), _c(vx)
        B, ny, nx = rho.shape
        assert vy.shape == (B, ny + 1, nx) and vx.shape == (B, ny, nx + 1), (rho.shape, vy.shape, vx.shape)
        out, ay, ax = torch.empty_like(rho), torch.empty_like(vy), torch.empty_like(vx)
        with torch.cuda.device(rho.device):
            _call("pdec_advect_centered_2d_fwd", _lib.ptr(rho), _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(out), B, ny, nx, mode, dt, *box, _lib.stream_ptr())
            _call("pdec_advect_staggered_2d_fwd", _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(ay), _lib.ptr(ax), B, ny, nx, mode, dt, *box, _lib.stream_ptr())
        ctx.save_for_backward(rho, vy, vx)
        ctx.cfg = (dt, mode, box)
        ctx.set_materialize_grads(False)
        return out, ay, ax

    @staticmethod
    def backward(ctx, g_out, g_ay, g_ax):
        rho, vy, vx = ctx.saved_tensors
        dt, mode, box = ctx.cfg
        B, ny, nx = rho.shape
        if g_out is None and g_ay is None and g_ax is None:
            return None, None, None, None, None, None
        g_vy, g_vx = torch.empty_like(vy), torch.empty_like(vx)
        g_rho = None
        nbytes = _lib.lib().pdec_advect_2d_workspace_bytes(B, ny, nx)
        ws = _lib.workspace(rho.device, nbytes)
        with torch.cuda.device(rho.device):
            if g_ay is None and g_ax is None:
                have_v = False
            else:
                g_ay = _c(g_ay) if g_ay is not None else torch.zeros_like(vy)
                g_ax = _c(g_ax) if g_ax is not None else torch.zeros_like(vx)
                _call("pdec_advect_staggered_2d_bwd", _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(g_ay), _lib.ptr(g_ax), _lib.ptr(g_vy), _lib.ptr(g_vx), 0,
                      B, ny, nx, mode, dt, *box, _lib.ptr(ws), nbytes, _lib.stream_ptr())
                have_v = True
            if g_out is not None:
                g_rho = torch.empty_like(rho)
                _call("pdec_advect_centered_2d_bwd", _lib.ptr(rho), _lib.ptr(vy), _lib.ptr(vx), _lib.ptr(_c(g_out)), _lib.ptr(g_rho), _lib.ptr(g_vy),
                      _lib.ptr(g_vx), 1 if have_v else 0, B, ny, nx, mode, dt, *box, _lib.ptr(ws), nbytes, _lib.stream_ptr())
            elif not have_v:
                g_vy.zero_(); g_vx.zero_()
        return g_rho, g_vy, g_vx, None, None, None


def advect_pair(rho, vy, vx, dt, mode=MODE_REPLICATE, box=None):
    """(advect_centered(rho, vy, vx), *advect_staggered(vy, vx)) as one autograd node."""
    return _AdvectPair.apply(rho, vy, vx, float(dt), int(mode), _box(box, rho.shape[1], rho.shape[2]))


def divergence_free(vy, vx, closed, box=None, accuracy=1e-5, max_iterations=2000, gradient_accuracy=None,
                    max_gradient_iterations=None, guess=None, slab=None, fused=None, direct=False, stop_mode=CG_STOP_PER_ELEMENT):
    """divergence_free (PF/physics/fluid.py:214-243) without obstacles -> (vy', vx', pressure, iterations, divergence).
    Default: divergence, solve and gradient subtraction are three launches.  fused=True (grids up to 128 x 128, zero guess) takes ONE
    launch per direction (the projection fused around the on-chip CG, _Project): bit-identical, two launches and two 4 MiB round trips
    fewer -- and MEASURED SLOWER at config 3 (14.19 against 13.37 ms per 16-step trajectory, profiles/r2f_fused_projection.md: the
    prologue / epilogue perturb the register allocation of the iteration loop by 9 instructions in 482, which costs more than the two
    stencil launches did inside a CUDA graph), so it is an option (PDEC_FUSED_PROJECTION=1 makes it the default), not the default.
    slab: a SlabSolver -> the pressure solve is decomposed over its process group.
    direct=True: the exact fast-diagonalisation solve instead of CG (pressure_solve_direct; iterations are reported as 0).
    guess="direct": the reference's solver chain `SparseSciPy() & SparseCG(...)` (solver_api.py:161-174): the exact solve is the initial
    guess of a CG that runs to the usual stopping rule max|r| < accuracy -- forward and adjoint (same matrix).  At 1e-5 the guess already
    satisfies the rule (0-1 iterations), at 1e-7 40-150 iterations remain instead of 320-480 (scripts/diag/cg_from_direct.py)."""
    ny, nx = vx.shape[1], vy.shape[2]
    b4 = _box(box, ny, nx)
    if direct:
        oy, ox, p, div, iters = _ProjectComposed.apply(vy, vx, closed_code(closed), b4[2], b4[3], ("direct",))
        return oy, ox, p, iters, div
    if fused is None:
        fused = _FUSED_PROJECTION
    if fused and slab is None and guess is None and stop_mode == CG_STOP_PER_ELEMENT and _lib.lib().pdec_project_2d_supported(ny, nx):
        oy, ox, p, div, iters = _project(vy, vx, closed, box, accuracy, max_iterations, gradient_accuracy, max_gradient_iterations)
        return oy, ox, p, iters, div
    from_direct = isinstance(guess, str) and guess == "direct"
    if from_direct:
        assert slab is None and direct_solver_supported(ny, nx), "guess='direct' needs a grid the direct solver takes (<= 128 x 128, one GPU)"
    if slab is None and (guess is None or from_direct):
        ga = accuracy if gradient_accuracy is None else gradient_accuracy
        gi = max_iterations if max_gradient_iterations is None else max_gradient_iterations
        oy, ox, p, div, iters = _ProjectComposed.apply(vy, vx, closed_code(closed), b4[2], b4[3],
                                                       ("cg_from_direct" if from_direct else "cg", accuracy, int(max_iterations), ga, int(gi),
                                                        int(stop_mode)))
        return oy, ox, p, iters, div
    div = mask_divergence(vy, vx, closed)
    if slab is not None:
        p, iters, _ = _PressureSolveSlab.apply(div, slab, closed_code(closed), accuracy, int(max_iterations))
    else:
        p, iters, _ = pressure_solve(div, closed, accuracy, max_iterations, gradient_accuracy, max_gradient_iterations, guess, stop_mode)
    oy, ox = subtract_pressure_gradient(vy, vx, p, closed, box)
    return oy, ox, p, iters, div


class ObstacleMasks(object):
    """Per-scene masks of a domain with obstacles (PoissonDomain, solver_api.py:51-121), built once from the cell mask
    `active` [ny, nx] (1 = fluid, 0 = obstacle; the reference samples the obstacle geometries at the cell centres,
    fluid.py:228-233):
      fy [ny+1, nx], fx [ny, nx+1]  face masks = min of the two adjacent `accessible` cells, the accessible grid extended with its
                                    own extrapolation (replicate on open, 0 on closed domains: _frictionless_velocity_mask :111-121)
      code [ny, nx] uint8           stencil of pdec_cg_poisson_2d_masked: off-diagonals active_self * active_neighbour (0 outside),
                                    diagonal -(accessible neighbours), outside = 1 on open / 0 on closed domains, clamped to <= -1
                                    (sparse.py:186-213 with active_tensor / accessible_tensor, solver_api.py:88-109)."""

    def __init__(self, active, closed):
        a = active.to(torch.float32).contiguous()
        assert a.dim() == 2 and a.is_cuda
        ny, nx = a.shape
        self.closed = bool(closed)
        self.active = a
        self.code = torch.empty((ny, nx), dtype=torch.uint8, device=a.device)
        self.fy = torch.empty((ny + 1, nx), dtype=torch.float32, device=a.device)
        self.fx = torch.empty((ny, nx + 1), dtype=torch.float32, device=a.device)
        with torch.cuda.device(a.device):
            _call("pdec_laplace_matrix_2d", _lib.ptr(a), _lib.ptr(self.code), _lib.ptr(self.fy), _lib.ptr(self.fx), ny, nx, int(self.closed),
                  _lib.stream_ptr())


def cg_poisson_masked_raw(rhs, out, code, accuracy, max_iterations, guess=None, iterations=None, residual=None, stop_mode=0):
    B, ny, nx = rhs.shape
    assert code.dtype == torch.uint8 and tuple(code.shape) == (ny, nx) and code.is_contiguous()
    L = _lib.lib()
    nbytes = L.pdec_cg_poisson_2d_workspace_bytes(B, ny, nx)
    ws = _lib.workspace(rhs.device, nbytes)
    with torch.cuda.device(rhs.device):
        _call("pdec_cg_poisson_2d_masked", _lib.ptr(rhs), _lib.ptr(guess), _lib.ptr(out), _lib.ptr(iterations), _lib.ptr(residual),
              _lib.ptr(code), B, ny, nx, float(accuracy if accuracy is not None else 0.0), int(max_iterations), int(stop_mode),
              _lib.ptr(ws), nbytes, _lib.stream_ptr())


class _PressureSolveMasked(torch.autograd.Function):
    """SparseCG on a domain with obstacles; backward = the same masked solve of the incoming gradient (sparse.py:93-94)."""

    @staticmethod
    def forward(ctx, div, code, accuracy, max_iterations, gradient_accuracy, max_gradient_iterations):
        _req_cuda(div)
        assert code.is_cuda and code.dtype == torch.uint8
        div = _c(div)
        B = div.shape[0]
        p = torch.empty_like(div)
        iters = torch.zeros(B, device=div.device, dtype=torch.int32)
        resid = torch.zeros(B, device=div.device, dtype=torch.float32)
        cg_poisson_masked_raw(div, p, code, accuracy, max_iterations, None, iters, resid)
        ctx.code = code
        ctx.cfg = (gradient_accuracy, max_gradient_iterations)
        ctx.mark_non_differentiable(iters, resid)
        return p, iters, resid

    @staticmethod
    def backward(ctx, g_p, _gi, _gr):
        accuracy, max_it = ctx.cfg
        g_p = _c(g_p)
        g_div = torch.empty_like(g_p)
        cg_poisson_masked_raw(g_p, g_div, ctx.code, accuracy, max_it)
        return g_div, None, None, None, None, None


def divergence_free_masked(vy, vx, masks, box=None, accuracy=1e-5, max_iterations=2000, gradient_accuracy=None,
                           max_gradient_iterations=None):
    """divergence_free (PF/physics/fluid.py:214-243) WITH obstacles -> (vy', vx', pressure, iterations, divergence).
    The face masks are applied with plain multiplies (autograd), the divergence / gradient kernels run unmasked on the masked
    velocity, the solve is the masked streaming CG."""
    ga = accuracy if gradient_accuracy is None else gradient_accuracy
    gi = max_iterations if max_gradient_iterations is None else max_gradient_iterations
    my, mx = vy * masks.fy, vx * masks.fx                      # with_hard_boundary_conditions(velocity)
    div = mask_divergence(my, mx, False)
    p, iters, _ = _PressureSolveMasked.apply(div, masks.code, accuracy, int(max_iterations), ga, int(gi))
    oy, ox = subtract_pressure_gradient(my, mx, p, False, box)  # my - grad(p dx)/dx, boundary-face gradient 0
    return oy * masks.fy, ox * masks.fx, p, iters, div         # = my - mask * grad p for binary masks


def incompressible_step(rho, vy, vx, dt, closed, box=None, accuracy=1e-5, max_iterations=2000, gradient_accuracy=None,
                        max_gradient_iterations=None, conserve_density=True, slab=None, direct=False, fused=None, guess=None):
    """IncompressibleFlow.step (PF/physics/fluid.py:90-114), default flags, no obstacles / effects / buoyancy.
    Field modes follow Material.extrapolation_mode / vector_extrapolation_mode: open -> constant, closed -> replicate.
    direct=True: exact fast-diagonalisation pressure solve instead of CG (see divergence_free)."""
    mode = MODE_REPLICATE if closed else MODE_CONSTANT
    rho2, ay, ax = advect_pair(rho, vy, vx, dt, mode, box)
    if closed and conserve_density:
        rho2 = normalize_to(rho2, rho)
    oy, ox, p, iters, _ = divergence_free(ay, ax, closed, box, accuracy, max_iterations, gradient_accuracy, max_gradient_iterations,
                                         slab=slab, direct=direct, fused=fused, guess=guess)
    return rho2, oy, ox, p, iters
