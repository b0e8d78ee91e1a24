"""CPU: the restatement oracle/restate.py against (a) golden vectors produced by the real reference and
(b) the real reference itself when /root/reference is present; plus gradient self-consistency."""
import os

import numpy as np
import pytest
import torch

from oracle import restate as R
from conftest import GOLDEN, rel_err


def _kw(g):
    return dict(lower=float(g["lower"]), size=float(g["size"]), substeps=int(g["substeps"]), resubstep=bool(g["resubstep"]))


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_config1_trajectory_matches_reference_golden(dtype):
    g = np.load(os.path.join(GOLDEN, "burgers_config1.npz"))
    tr = R.burgers_trajectory(torch.tensor(g["u0"], dtype=dtype), 32, float(g["dt"]), float(g["viscosity"]),
                              force=torch.tensor(g["force"], dtype=dtype), **_kw(g)).numpy()
    per_step = [rel_err(tr[t], g["frames"][t]) for t in range(1, 33)]
    assert rel_err(tr[1], g["frames"][1]) < 1e-5          # per-step tolerance (north_star)
    assert max(per_step) < 1e-4                            # full-trajectory tolerance


def test_single_steps_match_reference_golden():
    g = np.load(os.path.join(GOLDEN, "burgers_single_step.npz"))
    out = R.burgers_step(torch.tensor(g["u"]), float(g["dt"]), float(g["viscosity"]), **_kw(g)).numpy()
    # rows 32.. are rough random rows whose back-trace leaves the domain: the 'constant' rule must agree
    assert rel_err(out, g["out"]) < 1e-5


def test_tf_substep_collapse_differs_from_numpy_substepping():
    # on TF 4 substeps == 1 explicit step; on NumPy they are true substeps
    g = np.load(os.path.join(GOLDEN, "burgers_config1.npz"))
    u = torch.tensor(g["u0"])
    one = R.burgers_step(u, 0.03, 0.003, substeps=1)
    collapsed = R.burgers_step(u, 0.03, 0.003, substeps=4, resubstep=False)
    true_sub = R.burgers_step(u, 0.03, 0.003, substeps=4, resubstep=True)
    assert rel_err(collapsed, one) < 1e-6
    assert rel_err(true_sub, one) > 1e-4


@pytest.mark.parametrize("mode", [R.CONSTANT, R.REPLICATE, R.CIRCULAR])
def test_gradient_oracle_matches_finite_differences(mode):
    rng = np.random.RandomState(5)
    u = torch.tensor(R.gaussian_clash(2, 32, rng), dtype=torch.float64, requires_grad=True)
    f = torch.tensor(R.gaussian_force(2, 32, rng), dtype=torch.float64, requires_grad=True)
    w = torch.tensor(rng.randn(2, 32))

    def loss(u_, f_):
        return (R.burgers_trajectory(u_, 3, 0.03, 0.003, force=f_, mode=mode, substeps=2, resubstep=True)[-1] * w).sum()

    gu, gf = torch.autograd.grad(loss(u, f), [u, f])
    eps = 1e-6
    for (b, i) in [(0, 3), (1, 17), (0, 30)]:
        du = torch.zeros_like(u)
        du[b, i] = eps
        fd = (loss(u + du, f) - loss(u - du, f)) / (2 * eps)
        assert abs(fd.item() - gu[b, i].item()) < 1e-5 * max(1.0, abs(fd.item()))
        fd = (loss(u, f + du) - loss(u, f - du)) / (2 * eps)
        assert abs(fd.item() - gf[b, i].item()) < 1e-5 * max(1.0, abs(fd.item()))


def test_resample_known_answers():
    # PhiFlow/tests/test_fields.py:92-101 restated for the single-mode cases (2x2 grid [[1,2],[3,4]])
    data = torch.tensor([[[1., 2.], [3., 4.]]])
    # sample_at(points) with box = [0,2]^2: index = point - 0.5
    pts = torch.tensor([[[0.5, 0.5], [1.0, 1.0], [-0.5, 0.5], [1.5, 2.5]]]) - 0.5
    const = R.resample(data, pts, R.CONSTANT)[0]
    repl = R.resample(data, pts, R.REPLICATE)[0]
    assert const[0].item() == 1.0 and repl[0].item() == 1.0
    assert abs(const[1].item() - 2.5) < 1e-6 and abs(repl[1].item() - 2.5) < 1e-6
    assert const[2].item() == 0.0 and repl[2].item() == 1.0     # outside: zero vs. clamp
    assert const[3].item() == 0.0 and repl[3].item() == 4.0


@pytest.mark.reference
def test_restatement_vs_live_reference_random_modes():
    import warnings
    warnings.simplefilter("ignore")
    from oracle import ref_shim
    pf = ref_shim.flow()
    rng = np.random.RandomState(11)
    u = (rng.randn(4, 64) * 0.7).astype(np.float32)
    for extr, mode in (("constant", R.CONSTANT), ("boundary", R.REPLICATE)):
        grid = pf.CenteredGrid(u[..., None].copy(), box=pf.box[0:2], extrapolation=extr)
        ref = pf.Burgers().step_velocity(grid, 0.01, 0.05, (), 1).data[..., 0]
        got = R.burgers_step(torch.tensor(u), 0.05, 0.01, lower=0.0, size=2.0, mode=mode).numpy()
        assert rel_err(got, ref) < 1e-5, extr


def test_periodic_fft_diffusion_matches_reference_golden():
    """PERIODIC Burgers (circular advection + FFT diffusion + GROW force), 4 steps vs the live reference."""
    g = np.load(os.path.join(GOLDEN, "burgers_periodic.npz"))
    assert str(g["extrapolation"]) == "periodic"
    u = torch.tensor(g["frames"][0])
    f = torch.tensor(g["force"])
    for k in range(1, 5):
        u = R.burgers_step_periodic(u, float(g["dt"]), float(g["viscosity"]), force=f)
        assert rel_err(u, g["frames"][k]) < (1e-5 if k == 1 else 1e-4)
