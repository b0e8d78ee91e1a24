"""CPU: the fluid restatement oracle/restate_fluid.py against golden vectors produced by the real reference
(NumPy backend, oracle/make_golden_fluid.py), the reference's own known-answer tests restated, and gradient
self-consistency (finite differences)."""
import os

import numpy as np
import pytest
import torch

from oracle import restate as R
from oracle import restate_fluid as F
from conftest import GOLDEN, rel_err


def _load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


@pytest.fixture
def numpy_edge_rule():
    R.NUMPY_EDGE_RULE = True
    yield
    R.NUMPY_EDGE_RULE = False


@pytest.mark.parametrize("name", ["fluid_open_32", "fluid_closed_32x24", "fluid_open_128"])
def test_projected_step_trajectory_matches_reference_golden(name, numpy_edge_rule):
    """IncompressibleFlow.step x steps, fp32 restatement vs the reference (NumPy edge rule: the golden vectors
    come from the NumPy backend, which differs from the TF rule at xi == N-1 exactly)."""
    g = _load(name)
    B, ny, nx = g["rho0"].shape
    geo = F.Geometry(ny, nx)
    bd, dt = str(g["boundary"]), float(g["dt"])
    rho, vy, vx = [torch.tensor(g[k]) for k in ("rho0", "vy0", "vx0")]
    for k in range(1, int(g["steps"]) + 1):
        rho, vy, vx, p = F.incompressible_step(rho, vy, vx, dt, geo, bd)
        tol = 1e-5 if k == 1 else 1e-4
        assert rel_err(rho, g["rho%d" % k]) < tol
        assert rel_err(vy, g["vy%d" % k]) < tol and rel_err(vx, g["vx%d" % k]) < tol


def test_tf_edge_rule_differs_only_on_exact_edge_cells():
    """Per-step with the TF rule: every cell agrees with the golden step except samples with xi == N-1 exactly
    (boundary faces whose velocity is exactly 0 after the first projection)."""
    g = _load("fluid_open_32")
    geo = F.Geometry(32, 32)
    vy, vx = torch.tensor(g["vy1"]), torch.tensor(g["vx1"])
    ay, ax = F.advect_staggered(vy, vx, 1.0, geo, R.CONSTANT)
    bad = np.argwhere(np.abs(ay.numpy() - g["avy2"]) > 1e-5 * np.abs(g["avy2"]).max())
    assert len(bad) > 0 and all(i == 32 and g["vy1"][b, i, j] == 0.0 for b, i, j in bad)


def test_controlled_variant_matches_reference_golden():
    g = _load("fluid_controlled_64")
    assert str(g["density_extrapolation"]) == "constant" and str(g["velocity_extrapolation"]) == "boundary"
    geo = F.Geometry(64, 64)
    r2, vy, vx = F.controlled_step(torch.tensor(g["rho"]), torch.tensor(g["phi"]), float(g["dt"]), geo)
    assert rel_err(vy, g["vy"]) == 0.0 and rel_err(vx, g["vx"]) == 0.0
    assert rel_err(r2, g["rho_next"]) < 1e-5
    ay, ax = F.advect_staggered(vy, vx, float(g["dt"]), geo, R.REPLICATE)
    assert rel_err(ay, g["avy"]) < 1e-5 and rel_err(ax, g["avx"]) < 1e-5


@pytest.mark.parametrize("bd", ["open", "closed"])
def test_poisson_solution_matches_reference_cg(bd):
    g = _load("poisson_40x32")
    rhs = torch.tensor(g["rhs_" + bd])
    ref = g["p_" + bd]
    p = F.pressure_solve(rhs, bd).numpy()
    pc, it = F.conjugate_gradient(rhs, bd, 1e-6, 5000)
    pc = pc.numpy()
    if bd == "closed":      # singular Neumann matrix: pressure defined up to a constant per batch element
        ref = ref - ref.mean((1, 2), keepdims=True)
        p = p - p.mean((1, 2), keepdims=True)
        pc = pc - pc.mean((1, 2), keepdims=True)
    assert rel_err(p, ref) < 1e-5 and rel_err(pc, ref) < 1e-5 and 50 < it < 1000
    # PhiFlow/tests/test_poisson_solve.py:26-35: laplace(solve(rhs)) == rhs
    assert rel_err(F.laplace_apply(torch.tensor(p), bd), rhs) < 1e-4


def test_curl_is_divergence_free():
    """PhiFlow/tests/test_fields.py:114-119."""
    phi = torch.randn(2, 32, 32)
    vy, vx = F.curl_2d(phi, R.CONSTANT)
    assert float(F.divergence(vy, vx).abs().max()) < 1e-5


def test_open_projection_keeps_boundary_ring_divergence():
    """on OPEN domains boundary-face pressure gradients are 0, so the outermost ring of cells is
    not divergence-free while the interior is."""
    torch.manual_seed(0)
    vy, vx = torch.randn(1, 33, 32) * 0.5, torch.randn(1, 32, 33) * 0.5
    geo = F.Geometry(32, 32)
    vy2, vx2, p, div = F.divergence_free(vy, vx, geo, "open")
    d2 = F.divergence(vy2, vx2)[0]
    assert float(d2[1:-1, 1:-1].abs().max()) < 1e-4 and float(d2.abs().max()) > 1e-2
    assert torch.equal(vy2[:, 0], vy[:, 0]) and torch.equal(vx2[:, :, -1], vx[:, :, -1])
    vy3, vx3, _, _ = F.divergence_free(vy, vx, geo, "closed")
    assert float(F.divergence(vy3, vx3).abs().max()) < 1e-4


def test_batch_independence():
    """PhiFlow/tests/test_fluid.py:80-100."""
    torch.manual_seed(1)
    geo = F.Geometry(5, 4)
    rho, vy, vx = torch.rand(2, 5, 4), torch.randn(2, 6, 4), torch.randn(2, 5, 5)
    a = F.incompressible_step(rho, vy, vx, 1.0, geo, "closed")
    rho2, vy2, vx2 = rho.clone(), vy.clone(), vx.clone()
    rho2[1], vy2[1], vx2[1] = 0, 0, 0
    b = F.incompressible_step(rho2, vy2, vx2, 1.0, geo, "closed")
    for x, y in zip(a[:3], b[:3]):
        assert torch.equal(x[0], y[0])


@pytest.mark.parametrize("bd", ["open", "closed"])
def test_step_gradient_matches_finite_differences(bd):
    rng = np.random.RandomState(2)
    ny, nx = 10, 12
    geo = F.Geometry(ny, nx)
    rho = torch.tensor(rng.rand(1, ny, nx), dtype=torch.float64, requires_grad=True)
    vy = torch.tensor(rng.randn(1, ny + 1, nx) * 0.4, dtype=torch.float64, requires_grad=True)
    vx = torch.tensor(rng.randn(1, ny, nx + 1) * 0.4, dtype=torch.float64, requires_grad=True)
    w = [torch.tensor(rng.randn(*s)) for s in ((1, ny, nx), (1, ny + 1, nx), (1, ny, nx + 1))]

    def loss(r, y, x):
        r2, y2, x2, _ = F.incompressible_step(r, y, x, 0.7, geo, bd)
        return (r2 * w[0]).sum() + (y2 * w[1]).sum() + (x2 * w[2]).sum()

    gr, gy, gx = torch.autograd.grad(loss(rho, vy, vx), [rho, vy, vx])
    eps = 1e-6
    for t, gt, idx in ((rho, gr, (0, 4, 5)), (vy, gy, (0, 5, 6)), (vx, gx, (0, 3, 7)), (vy, gy, (0, 0, 3)), (vx, gx, (0, 2, 12))):
        d = torch.zeros_like(t)
        d[idx] = eps
        args_p = [a + d if a is t else a for a in (rho, vy, vx)]
        args_m = [a - d if a is t else a for a in (rho, vy, vx)]
        fd = (loss(*args_p) - loss(*args_m)) / (2 * eps)
        assert abs(fd.item() - gt[idx].item()) < 1e-5 * max(1.0, abs(fd.item())), (idx, fd.item(), gt[idx].item())


@pytest.mark.reference
def test_restatement_vs_live_reference_random_box():
    """Non-default box (dx != 1) and both boundary materials against the live reference."""
    import warnings
    warnings.simplefilter("ignore")
    from oracle import ref_shim, make_golden_fluid as M
    pf = ref_shim.flow()
    rng = np.random.RandomState(21)
    ny, nx, B = 16, 20, 2
    rho = np.abs(M._smooth(rng, (B, ny, nx), 1.0))
    vy, vx = M.random_staggered(rng, B, ny, nx, 0.2)
    R.NUMPY_EDGE_RULE = True
    try:
        for bd, material in (("open", pf.OPEN), ("closed", pf.CLOSED)):
            dom = pf.Domain([ny, nx], boundaries=material, box=pf.box[0:4, 0:5])
            fluid = pf.Fluid(dom, density=rho[..., None].copy(), velocity=M.staggered_tensor(vy, vx), batch_size=B)
            out = pf.IncompressibleFlow(pressure_solver=pf.SparseSciPy()).step(fluid, 0.5)
            geo = F.Geometry(ny, nx, lower=(0, 0), size=(4, 5))
            r2, y2, x2, p = F.incompressible_step(torch.tensor(rho), torch.tensor(vy), torch.tensor(vx), 0.5, geo, bd)
            assert rel_err(r2, out.density.data[..., 0]) < 1e-5
            assert rel_err(y2, out.velocity.data[0].data[..., 0]) < 2e-5
            assert rel_err(x2, out.velocity.data[1].data[..., 0]) < 2e-5
    finally:
        R.NUMPY_EDGE_RULE = False


@pytest.mark.parametrize("name", ["fluid_buoyancy_open_32", "fluid_buoyancy_closed_24x32", "fluid_obstacle_open_32",
                                  "fluid_obstacle_closed_24x32"])
def test_buoyancy_effects_and_obstacles_match_reference_golden(name, numpy_edge_rule):
    """SURVEY 8f N3 restatement (buoyancy through the face-sampled density, GROW density effect, obstacle masks in the
    divergence / matrix / gradient) vs the live reference's NumPy path."""
    g = _load(name)
    B, ny, nx = g["rho0"].shape
    geo = F.Geometry(ny, nx)
    bd, dt = str(g["boundary"]), float(g["dt"])
    src = torch.tensor(g["density_source"])[None] if "density_source" in g.files else None
    active = F.obstacle_active(ny, nx, g["obstacles"].tolist()) if "obstacles" in g.files else None
    rho, vy, vx = [torch.tensor(g[k]) for k in ("rho0", "vy0", "vx0")]
    for k in range(1, int(g["steps"]) + 1):
        rho, vy, vx, p = F.incompressible_step_general(rho, vy, vx, dt, geo, bd, buoyancy=float(g["buoyancy"]), density_source=src,
                                                       active=active)
        tol = 1e-5 if k == 1 else 1e-4
        assert rel_err(rho, g["rho%d" % k]) < tol
        assert rel_err(vy, g["vy%d" % k]) < 2 * tol and rel_err(vx, g["vx%d" % k]) < 2 * tol


@pytest.mark.parametrize("name", ["fluid_effects_open_32", "fluid_effects_closed_24x32"])
def test_geometry_effects_match_reference_golden(name, numpy_edge_rule):
    """SURVEY 8f N3 restatement of Inflow / Accelerator / FIX effects and of sphere / union obstacles (geometry masks at cell
    centres and face centres) vs the live reference's NumPy path."""
    from oracle.make_golden_fluid import EFFECT_SCENES
    g, s = _load(name), EFFECT_SCENES[name]
    geo = F.Geometry(s["ny"], s["nx"])
    rho, vy, vx = [torch.tensor(g[k]) for k in ("rho0", "vy0", "vx0")]
    for k in range(1, s["steps"] + 1):
        rho, vy, vx, p = F.incompressible_step_effects(rho, vy, vx, s["dt"], geo, s["boundary"], buoyancy=s["buoyancy"],
                                                       density_effects=s["density_effects"], velocity_effects=s["velocity_effects"],
                                                       obstacle=s["obstacle"])
        tol = 1e-5 if k == 1 else 1e-4
        assert rel_err(rho, g["rho%d" % k]) < tol
        assert rel_err(vy, g["vy%d" % k]) < 2 * tol and rel_err(vx, g["vx%d" % k]) < 2 * tol


@pytest.mark.parametrize("bd", ["open", "closed"])
def test_fourier_poisson_matches_reference(bd):
    """FourierSolver's guess (math.fourier_poisson) vs the live reference."""
    g = _load("poisson_40x32")
    assert rel_err(F.fourier_poisson(torch.tensor(g["rhs_" + bd])), g["fft_" + bd]) < 1e-5


_SIDE_CASES = {"fluid_noslip_24x20": ((1, 1, 1, 1), 1, True), "fluid_mixed_open_top_24x20": ((1, 0, 1, 1), 0, False),
               "fluid_mixed_half_open_x_20x24": ((0, 0, 1, 0), 0, False)}


@pytest.mark.parametrize("name", sorted(_SIDE_CASES))
def test_no_slip_and_mixed_boundaries_match_reference_golden(name):
    """SURVEY 8f N3 remainder: NO_SLIP (Material.vector_extrapolation_mode 'constant' on a solid domain, material.py:88-98) and per-side
    mixed boundaries Domain(..., [(CLOSED, OPEN), CLOSED]) (domain.py:13-35; sampling through _padded_resample, grid.py:211-216) against
    goldens from the live reference (conserve_density=False for the mixed cases: the reference's own flag test fails under NumPy 2)."""
    solid, friction, conserve = _SIDE_CASES[name]
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    B, ny, nx = g["rho0"].shape
    geo = F.Geometry(ny, nx)
    for k in range(1, int(g["steps"]) + 1):
        prev = [torch.tensor(g["%s%d" % (n, k - 1)]) for n in ("rho", "vy", "vx")]
        out = F.incompressible_step_sides(*prev, float(g["dt"]), geo, solid, friction=friction, buoyancy=float(g["buoyancy"]),
                                          conserve_density=conserve)
        for a, n in zip(out[:3], ("rho", "vy", "vx")):
            assert rel_err(a, g["%s%d" % (n, k)]) < 1e-5, (name, k, n)
        p, pref = out[3].numpy(), g["p%d" % k]
        if all(solid):
            p, pref = p - p.mean((1, 2), keepdims=True), pref - pref.mean((1, 2), keepdims=True)
        assert rel_err(p, pref) < 1e-4
