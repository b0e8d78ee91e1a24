"""GPU: the two benchmarked workloads compared with the oracle AT THEIR OWN SIZE, on the bench's own inputs (benchlib/inputs.py).

VERDICT round 1, "What's weak" item 1: config 2 (s[33, 4096, 128]) and config 3 (128 x 128, 16 projected steps) were only
property-tested at size.  Tolerances = the north star's: losses / states over a trajectory 1e-4, control-loss gradients 1e-3."""
import numpy as np
import pytest
import torch

from conftest import rel_err
from benchlib import inputs
from oracle import restate as R
from oracle import restate_fluid as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode", [R.CONSTANT, R.REPLICATE, R.CIRCULAR])
def test_config2_full_size_vs_oracle(mode):
    """s[33, 4096, 128] from the bench's generator, all three boundary modes: ops.burgers_pair_loss vs oracle.restate.burgers_pair_loss."""
    from pde_control_b200 import ops
    B, n, T = 4096, 128, 32
    s = inputs.burgers_config2_sets(B, n, T, 0, torch.device("cuda"), 1)[0]
    sc = s.clone().requires_grad_(True)
    l2c, l1c = ops.burgers_pair_loss(sc, inputs.DT, inputs.VISC, mode=mode)
    l2c.backward()
    torch.set_num_threads(max(1, torch.get_num_threads()))
    so = s.cpu().requires_grad_(True)
    l2o, l1o = R.burgers_pair_loss(so, inputs.DT, inputs.VISC, mode=mode)
    l2o.backward()
    assert abs(l2c.item() - l2o.item()) < 1e-4 * abs(l2o.item())
    assert abs(l1c.item() - l1o.item()) < 1e-4 * abs(l1o.item())
    assert rel_err(sc.grad.cpu(), so.grad) < 1e-3
    # per frame as well: no frame's gradient may hide behind a larger one
    per_frame = [rel_err(sc.grad[t].cpu(), so.grad[t]) for t in range(T + 1)]
    assert max(per_frame) < 1e-3, per_frame


def _fluid_at_size(bd, accuracy, elements):
    from pde_control_b200 import ops
    n, steps = 128, 16
    rho0, vy0, vx0, target = inputs.fluid_config3_inputs(64, n, 0)
    sel = list(elements)
    arrs = [a[sel] for a in (rho0, vy0, vx0)]
    tgt = target[sel]
    B = len(sel)
    geo = F.Geometry(n, n)
    to = [torch.tensor(a, requires_grad=True) for a in arrs]
    tc = [torch.tensor(a, device="cuda", requires_grad=True) for a in arrs]
    so, sc = tuple(to), tuple(tc)
    state_err, iters = 0.0, []
    for _ in range(steps):
        so = F.incompressible_step(*so, 1.0, geo, bd)[:3]
        out = ops.incompressible_step(*sc, 1.0, bd == "closed", accuracy=accuracy, max_iterations=2000)
        sc = out[:3]
        iters.append(out[4])
        state_err = max(state_err, max(rel_err(a.detach().cpu(), b.detach()) for a, b in zip(sc, so)))
    R.l2_loss(so[0] - torch.tensor(tgt)).backward()
    d = sc[0] - torch.tensor(tgt, device="cuda")
    ((d * d).sum() / (2 * B)).backward()
    grad_err = max(rel_err(a.grad.cpu(), b.grad) for a, b in zip(tc, to))
    return state_err, grad_err, int(torch.stack(iters).max())


@pytest.mark.parametrize("bd", ["open", "closed"])
@pytest.mark.parametrize("accuracy", [1e-5, 1e-6])
def test_config3_trajectory_vs_oracle_at_size(bd, accuracy):
    """128 x 128, 16 projected steps (IncompressibleFlow.step, PF/physics/fluid.py:90-114), 4 batch elements of the bench's own inputs,
    CG at the accuracy the bench times (1e-5) and at 1e-6; oracle pressure = fp64 sparse direct solve, gradient = autograd with the
    registered second solve (sparse.py:93-94).  States 1e-4 over the trajectory, gradients w.r.t. rho0, vy0, vx0 1e-3."""
    state_err, grad_err, max_it = _fluid_at_size(bd, accuracy, (0, 1, 31, 63))
    assert max_it < 2000
    assert state_err < 1e-4, (state_err, grad_err)
    assert grad_err < 1e-3, (state_err, grad_err)
