"""GPU parity: CUDA Burgers kernels (through the C-ABI) vs the oracle and the reference golden vectors.

Tolerances (BASELINE.json north_star): relative 1e-5 per step on states, 1e-4 over a full trajectory,
1e-3 on control-loss gradients; "relative" = max-norm error / max-norm of the reference field.
"""
import os

import numpy as np
import pytest
import torch

from oracle import restate as R
from conftest import GOLDEN, rel_err

pytestmark = pytest.mark.gpu


def _ops():
    from pde_control_b200 import ops
    return ops


def _kw(g):
    return dict(lower=float(g["lower"]), size=float(g["size"]), substeps=int(g["substeps"]), resubstep=bool(g["resubstep"]))


def test_config1_trajectory_vs_reference_golden():
    ops = _ops()
    g = np.load(os.path.join(GOLDEN, "burgers_config1.npz"))
    u0 = torch.tensor(g["u0"]).cuda()
    f = torch.tensor(g["force"]).cuda()
    fr = ops.burgers_trajectory(u0, 32, float(g["dt"]), float(g["viscosity"]), force=f, **_kw(g)).cpu().numpy()
    assert rel_err(fr[0], g["frames"][1]) < 1e-5
    assert max(rel_err(fr[t], g["frames"][t + 1]) for t in range(32)) < 1e-4


def test_single_steps_vs_reference_golden():
    ops = _ops()
    g = np.load(os.path.join(GOLDEN, "burgers_single_step.npz"))
    out = ops.burgers_step(torch.tensor(g["u"]).cuda(), float(g["dt"]), float(g["viscosity"]), **_kw(g)).cpu().numpy()
    assert rel_err(out, g["out"]) < 1e-5


@pytest.mark.parametrize("mode", [R.CONSTANT, R.REPLICATE, R.CIRCULAR])
@pytest.mark.parametrize("n", [128, 100, 37])
@pytest.mark.parametrize("resubstep,substeps", [(False, 1), (False, 4), (True, 4)])
def test_step_forward_and_adjoint_vs_oracle(mode, n, resubstep, substeps):
    ops = _ops()
    rng = np.random.RandomState(n + mode)
    B = 5
    u = (R.gaussian_clash(B, n, rng) + 0.05 * rng.randn(B, n)).astype(np.float32)
    f = R.gaussian_force(B, n, rng)
    w = rng.randn(B, n).astype(np.float32)
    kw = dict(lower=-0.5, size=2.0, mode=mode, substeps=substeps, resubstep=resubstep)
    # oracle (fp32 restatement + autograd)
    uo = torch.tensor(u, requires_grad=True)
    fo = torch.tensor(f, requires_grad=True)
    out_o = R.burgers_step(uo, 0.03, 0.003, force=fo, **kw)
    (out_o * torch.tensor(w)).sum().backward()
    # CUDA
    uc = torch.tensor(u, device="cuda", requires_grad=True)
    fc = torch.tensor(f, device="cuda", requires_grad=True)
    out_c = ops.burgers_step(uc, 0.03, 0.003, force=fc, **kw)
    (out_c * torch.tensor(w, device="cuda")).sum().backward()
    assert rel_err(out_c.detach().cpu(), out_o.detach()) < 1e-5
    assert rel_err(uc.grad.cpu(), uo.grad) < 1e-3
    assert rel_err(fc.grad.cpu(), fo.grad) < 1e-3


def test_trajectory_gradient_vs_oracle_config1():
    """Config 1 gradient half: loss = l2_loss(u_32 - target) w.r.t. u0 and the force, 32 steps, B=8."""
    ops = _ops()
    g = np.load(os.path.join(GOLDEN, "burgers_config1.npz"))
    kw = _kw(g)
    target = torch.tensor(g["frames"][32]) * 0.9
    uo = torch.tensor(g["u0"], requires_grad=True)
    fo = torch.tensor(g["force"], requires_grad=True)
    lo = R.l2_loss(R.burgers_trajectory(uo, 32, 0.03, 0.003, force=fo, **kw)[-1] - target)
    lo.backward()
    uc = torch.tensor(g["u0"], device="cuda", requires_grad=True)
    fc = torch.tensor(g["force"], device="cuda", requires_grad=True)
    fr = ops.burgers_trajectory(uc, 32, 0.03, 0.003, force=fc, **kw)
    d = fr[-1] - target.cuda()
    lc = (d ** 2).sum() / 2 / d.shape[0]
    lc.backward()
    assert abs(lc.item() - lo.item()) < 1e-4 * abs(lo.item())
    assert rel_err(uc.grad.cpu(), uo.grad) < 1e-3
    assert rel_err(fc.grad.cpu(), fo.grad) < 1e-3


def test_trajectory_gradient_with_intermediate_frame_losses():
    ops = _ops()
    rng = np.random.RandomState(3)
    u = R.gaussian_clash(4, 64, rng)
    w = rng.randn(6, 4, 64).astype(np.float32)
    kw = dict(substeps=4, resubstep=True)
    uo = torch.tensor(u, requires_grad=True)
    (R.burgers_trajectory(uo, 6, 0.03, 0.003, **kw)[1:] * torch.tensor(w)).sum().backward()
    uc = torch.tensor(u, device="cuda", requires_grad=True)
    (ops.burgers_trajectory(uc, 6, 0.03, 0.003, **kw) * torch.tensor(w, device="cuda")).sum().backward()
    assert rel_err(uc.grad.cpu(), uo.grad) < 1e-3


@pytest.mark.parametrize("mode", [R.CONSTANT, R.REPLICATE, R.CIRCULAR])
@pytest.mark.parametrize("B,n,T", [(8, 128, 32), (3, 50, 4), (1, 128, 1), (37, 128, 5), (700, 128, 3)])
def test_pair_loss_and_gradient_vs_oracle(mode, B, n, T):
    ops = _ops()
    rng = np.random.RandomState(B * n + T)
    g = np.load(os.path.join(GOLDEN, "burgers_config1.npz"))
    if n == 128 and B == 8:
        s = g["frames"][:T + 1] + 0.02 * rng.randn(T + 1, B, n).astype(np.float32)   # perturbed real trajectory
    else:
        s = np.stack([R.gaussian_clash(B, n, rng) for _ in range(T + 1)])
    s = s.astype(np.float32)
    so = torch.tensor(s, requires_grad=True)
    l2o, l1o = R.burgers_pair_loss(so, 0.03, 0.003, mode=mode)
    l2o.backward()
    sc = torch.tensor(s, device="cuda", requires_grad=True)
    l2c, l1c = ops.burgers_pair_loss(sc, 0.03, 0.003, mode=mode)
    l2c.backward()
    assert abs(l2c.item() - l2o.item()) < 1e-4 * abs(l2o.item())
    assert abs(l1c.item() - l1o.item()) < 1e-4 * abs(l1o.item())
    assert rel_err(sc.grad.cpu(), so.grad) < 1e-3


def test_pair_loss_is_bit_reproducible_and_batch_independent():
    """Determinism (no float atomics) and the reference's batch-independence invariant
    (PhiFlow/tests/test_fluid.py:80-100 states it for the fluid; Burgers rows never interact)."""
    ops = _ops()
    rng = np.random.RandomState(0)
    s = np.stack([R.gaussian_clash(64, 128, rng) for _ in range(9)]).astype(np.float32)
    sc = torch.tensor(s, device="cuda", requires_grad=True)
    grads = []
    for _ in range(3):
        sc.grad = None
        l2, _ = ops.burgers_pair_loss(sc, 0.03, 0.003)
        l2.backward()
        grads.append((l2.item(), sc.grad.clone()))
    assert grads[0][0] == grads[1][0] == grads[2][0]
    assert torch.equal(grads[0][1], grads[1][1]) and torch.equal(grads[0][1], grads[2][1])
    # element 0's gradient does not depend on the other batch elements (up to the 1/B scale)
    s1 = torch.tensor(s[:, :1].copy(), device="cuda", requires_grad=True)
    l2, _ = ops.burgers_pair_loss(s1, 0.03, 0.003)
    l2.backward()
    assert torch.equal(s1.grad[:, 0] , grads[0][1][:, 0] * 64) or rel_err((s1.grad[:, 0] / 64).cpu(), grads[0][1][:, 0].cpu()) < 1e-6


def test_full_size_config2_properties():
    """Config 2 size (s[33, 4096, 128]): size-independent properties instead of a CPU oracle run:
    (1) the loss of an exact solver trajectory is ~0 and its gradient ~0; (2) linearity of the L2 gradient
    in the residual for a fixed input frame pair: grad wrt s[T] equals d_T/B exactly."""
    ops = _ops()
    rng = np.random.RandomState(2)
    B, n, T = 4096, 128, 32
    u0 = torch.tensor(R.gaussian_clash(B, n, rng), device="cuda")
    frames = ops.burgers_trajectory(u0, T, 0.03, 0.003)            # 1 substep: same solver as the loss
    s = torch.cat([u0[None], frames]).detach().requires_grad_(True)
    l2, l1 = ops.burgers_pair_loss(s, 0.03, 0.003)
    l2.backward()
    assert l2.item() == 0.0 and l1.item() == 0.0
    assert float(s.grad.abs().max()) == 0.0
    s2 = (s.detach() + 0.01 * torch.randn_like(s)).requires_grad_(True)
    l2, l1 = ops.burgers_pair_loss(s2, 0.03, 0.003)
    l2.backward()
    d_last = s2.detach()[T] - ops.burgers_step(s2.detach()[T - 1], 0.03, 0.003)
    assert rel_err(s2.grad[T].cpu(), (d_last / B).cpu()) < 1e-6
    # loss value from independently computed residuals
    tot = 0.0
    for t in range(T):
        d = s2.detach()[t + 1] - ops.burgers_step(s2.detach()[t], 0.03, 0.003)
        tot += float((d.double() ** 2).sum()) / 2 / B
    assert abs(l2.item() - tot) < 1e-4 * tot


@pytest.mark.parametrize("frames_per_slice", [0, 1, 3, -1])
@pytest.mark.parametrize("B,chunk_rows,n,T", [(512, 128, 128, 8), (300, 128, 128, 5), (40, 16, 50, 3), (64, 256, 128, 4), (700, 128, 128, 3)])
def test_host_pipeline_equals_resident_call(B, chunk_rows, n, T, frames_per_slice, monkeypatch):
    """pdec_burgers_pair_loss_host (chunked H2D / kernel / D2H over three streams) == the resident single launch:
    gradient bit for bit when the chunk scale is a power of two, else to rounding of the 1/B scale; losses to 1e-6;
    ragged last chunk, a batch smaller than one chunk, the general-n kernel and two calls in a row (slot re-use).
    frames_per_slice > 0: the frame-sliced mode (windows of pairs through pdec_burgers_pair_loss_window; n = 128 only, other
    sizes fall back to row chunks): gradient bit for bit.
    frames_per_slice == 0 takes the direct-gradient path where it applies (the kernel stores into the page-locked host array itself);
    -1 = row chunks with the staged device-to-host copies (PDEC_PIPE_STAGED_D2H)."""
    ops = _ops()
    if frames_per_slice < 0:
        monkeypatch.setenv("PDEC_PIPE_STAGED_D2H", "1")
        frames_per_slice = 0
    rng = np.random.RandomState(B + T)
    s = np.stack([R.gaussian_clash(B, n, rng) for _ in range(T + 1)]).astype(np.float32)
    s += 0.02 * rng.randn(*s.shape).astype(np.float32)
    sc = torch.tensor(s, device="cuda", requires_grad=True)
    l2, l1 = ops.burgers_pair_loss(sc, 0.03, 0.003)
    l2.backward()
    pipe = ops.HostPipeline(n, T, chunk_rows=chunk_rows, max_batch=1024, frames_per_slice=frames_per_slice)
    hs = torch.tensor(s).pin_memory()
    hg = torch.full(s.shape, float("nan")).pin_memory()
    hl = torch.zeros(2).pin_memory()
    for _ in range(2):
        hg.fill_(float("nan"))
        pipe.pair_loss(hs, hg, hl, 0.03, 0.003)
        torch.cuda.current_stream().synchronize()
        pow2 = B % chunk_rows == 0 and ((B // chunk_rows) & (B // chunk_rows - 1)) == 0 or B <= chunk_rows or (frames_per_slice > 0 and n == 128)
        if pow2:
            assert torch.equal(hg, sc.grad.cpu())
        else:
            assert rel_err(hg, sc.grad.cpu()) < 1e-6
        assert abs(hl[0].item() - l2.item()) <= 1e-6 * abs(l2.item())
        assert abs(hl[1].item() - l1.item()) <= 1e-6 * abs(l1.item())
    # losses only
    hl.zero_()
    pipe.pair_loss(hs, None, hl, 0.03, 0.003)
    torch.cuda.current_stream().synchronize()
    assert abs(hl[0].item() - l2.item()) <= 1e-6 * abs(l2.item())
    with pytest.raises(ValueError):
        pipe.pair_loss(torch.tensor(s), hg, hl, 0.03, 0.003)        # not pinned
    pipe.close()


def test_edge_cases_empty_batch_and_bad_arguments():
    ops = _ops()
    from pde_control_b200._lib import PdecError
    out = ops.burgers_trajectory(torch.zeros(0, 16, device="cuda"), 2, 0.1, 0.1)
    assert out.shape == (2, 0, 16)
    with pytest.raises(PdecError):
        ops.burgers_step(torch.zeros(2, 1, device="cuda"), 0.1, 0.1)          # n < 2
    with pytest.raises(PdecError):
        ops.burgers_step(torch.zeros(1, 100000, device="cuda"), 0.1, 0.1)     # row does not fit shared memory


@pytest.mark.parametrize("mode_name", ["circular", "replicate"])
def test_periodic_unit_kernel_forward_is_bit_exact_and_matches_v1(monkeypatch, mode_name):
    """The periodic ('circular') and 'replicate' twins of the fused unit kernel: (1) the force loss of a trajectory produced by the general step kernel
    with the same mode is EXACTLY zero (the two forward steps agree bit for bit: wrapped taps, w1 = 1 - w0, wrapped stencil);
    (2) losses and gradients agree to rounding with the v1 kernel (PDEC_BURGERS_FAST_V1), on rows that cross the
    seam in both directions and on rows with shocks (folded back-trace: ranked scatter)."""
    ops = _ops()
    from pde_control_b200 import _lib
    MODE_CIRCULAR = _lib.MODE_CIRCULAR if mode_name == "circular" else _lib.MODE_REPLICATE      # the mode under test
    rng = np.random.RandomState(5)
    B, n, T = 600, 128, 6
    u0 = np.stack([R.gaussian_clash(1, n, rng)[0] for _ in range(B)]).astype(np.float32)
    u0[::3] += 1.5                                    # drift to the right across the seam
    u0[1::3] -= 1.5                                   # and to the left
    u0[2::7] *= 6.0                                   # steep: several cells per cell, folds
    u0c = torch.tensor(u0, device="cuda")
    frames = ops.burgers_trajectory(u0c, T, 0.03, 0.003, mode=MODE_CIRCULAR, substeps=1)
    s = torch.cat([u0c[None], frames]).contiguous()
    l2, l1 = ops.burgers_pair_loss(s, 0.03, 0.003, mode=MODE_CIRCULAR)
    assert l2.item() == 0.0 and l1.item() == 0.0
    sp = (s + 0.05 * torch.randn(s.shape, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))).requires_grad_(True)
    l2a, l1a = ops.burgers_pair_loss(sp, 0.03, 0.003, mode=MODE_CIRCULAR)
    l2a.backward()
    ga = sp.grad.clone()
    sp.grad = None
    monkeypatch.setenv("PDEC_BURGERS_FAST_V1", "1")
    l2b, l1b = ops.burgers_pair_loss(sp, 0.03, 0.003, mode=MODE_CIRCULAR)
    l2b.backward()
    # the two kernels cut the (row, pair) space differently, so the loss sums are grouped differently
    assert abs(l2a.item() - l2b.item()) <= 2e-6 * abs(l2b.item()) and abs(l1a.item() - l1b.item()) <= 2e-6 * abs(l1b.item())
    assert rel_err(ga.cpu(), sp.grad.cpu()) < 2e-6
