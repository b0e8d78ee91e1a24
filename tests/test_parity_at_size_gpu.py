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


def _cg_solver(accuracy):
    """The reference's own algorithm at its own accuracy: restated fp32 conjugate_gradient (PF/math/blas.py:5-64) with the registered
    gradient = the same CG on the incoming gradient (sparse.py:93-94)."""
    def solver(div, boundary):
        class S(torch.autograd.Function):
            @staticmethod
            def forward(ctx, d):
                return F.conjugate_gradient(d, boundary, accuracy, 2000)[0]

            @staticmethod
            def backward(ctx, g):
                return F.conjugate_gradient(g, boundary, accuracy, 2000)[0]
        return S.apply(div)
    return solver


def _oracle_run(arrs, tgt, bd, solver=None):
    geo = F.Geometry(128, 128)
    t = [torch.tensor(a, requires_grad=True) for a in arrs]
    s, traj = tuple(t), []
    for _ in range(16):
        s = (F.incompressible_step(*s, 1.0, geo, bd) if solver is None else F.incompressible_step(*s, 1.0, geo, bd, solver=solver))[:3]
        traj.append([x.detach() for x in s])
    R.l2_loss(s[0] - torch.tensor(tgt)).backward()
    return traj, [x.grad for x in t]


def _cuda_run(arrs, tgt, bd, accuracy, direct=False):
    from pde_control_b200 import ops
    t = [torch.tensor(a, device="cuda", requires_grad=True) for a in arrs]
    s, traj, iters = tuple(t), [], []
    for _ in range(16):
        out = ops.incompressible_step(*s, 1.0, bd == "closed", accuracy=accuracy, max_iterations=2000, direct=direct)
        s = out[:3]
        iters.append(out[4])
        traj.append([x.detach().cpu() for x in s])
    d = s[0] - torch.tensor(tgt, device="cuda")
    ((d * d).sum() / (2 * len(tgt))).backward()
    return traj, [x.grad.cpu() for x in t], int(torch.stack(iters).max())


def _dist(traj_a, grads_a, traj_b, grads_b):
    state = max(rel_err(a, b) for fa, fb in zip(traj_a, traj_b) for a, b in zip(fa, fb))
    grad = max(rel_err(a, b) for a, b in zip(grads_a, grads_b))
    return state, grad


def _inputs_at_size(elements=(0, 1, 31, 63)):
    rho0, vy0, vx0, target = inputs.fluid_config3_inputs(64, 128, 0)
    sel = list(elements)
    return [a[sel] for a in (rho0, vy0, vx0)], target[sel]


@pytest.mark.parametrize("bd", ["open", "closed"])
def test_config3_trajectory_vs_oracle_at_size(bd):
    """128 x 128, 16 projected steps (IncompressibleFlow.step, PF/physics/fluid.py:90-114), 4 batch elements of the bench's own inputs
    against the exact oracle (pressure = fp64 sparse direct solve, gradient = autograd with the registered second solve): states 1e-4
    over the trajectory, gradients w.r.t. rho0, vy0, vx0 1e-3 -- at CG accuracy 1e-7, the accuracy that HOLDS these tolerances
    (measured, scripts/diag/fluid_parity_at_size.py: 1e-6 holds them on OPEN domains, not on CLOSED ones)."""
    arrs, tgt = _inputs_at_size()
    traj_o, grads_o = _oracle_run(arrs, tgt, bd)
    traj_c, grads_c, max_it = _cuda_run(arrs, tgt, bd, 1e-7)
    state, grad = _dist(traj_c, grads_c, traj_o, grads_o)
    assert max_it < 2000
    assert state < 1e-4, (state, grad)
    assert grad < 1e-3, (state, grad)


@pytest.mark.parametrize("bd", ["open", "closed"])
def test_config3_trajectory_with_the_direct_solver_vs_oracle_at_size(bd):
    """The same trajectory with the exact fast-diagonalisation solver (physics.SparseSciPy / direct=True): no accuracy knob, the
    tolerances hold with an order of magnitude to spare."""
    arrs, tgt = _inputs_at_size()
    traj_o, grads_o = _oracle_run(arrs, tgt, bd)
    traj_c, grads_c, _ = _cuda_run(arrs, tgt, bd, None, direct=True)
    state, grad = _dist(traj_c, grads_c, traj_o, grads_o)
    assert state < 1e-4 and grad < 1e-3, (state, grad)
    assert state < 3e-5 and grad < 1e-4, (state, grad)


@pytest.mark.parametrize("bd", ["open", "closed"])
def test_config3_at_the_reference_default_accuracy_is_as_close_as_the_reference_algorithm(bd):
    """At the reference's DEFAULT accuracy (SparseCG(accuracy=1e-5), what the headline bench times) the reference's own algorithm does
    not pin the result to 1e-4 / 1e-3: the restated fp32 CG at 1e-5 is 9e-2 (states: one boundary sample of the 'constant' resample
    rule flips) / 3e-2 (gradients) away from the exact solve on OPEN, 1e-4 / 1e-3 on CLOSED.  What can be asserted there: the CUDA path
    is no further from the exact result than the reference algorithm itself (factor 3 + the tolerance floor), field by field."""
    arrs, tgt = _inputs_at_size()
    traj_o, grads_o = _oracle_run(arrs, tgt, bd)
    traj_r, grads_r = _oracle_run(arrs, tgt, bd, solver=_cg_solver(1e-5))
    traj_c, grads_c, max_it = _cuda_run(arrs, tgt, bd, 1e-5)
    ref_state, ref_grad = _dist(traj_r, grads_r, traj_o, grads_o)
    state, grad = _dist(traj_c, grads_c, traj_o, grads_o)
    assert max_it < 2000
    assert state <= 3 * ref_state + 1e-4, (state, ref_state)
    assert grad <= 3 * ref_grad + 1e-3, (grad, ref_grad)
