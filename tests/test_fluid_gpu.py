"""GPU parity: CUDA 2-D fluid kernels (through the C-ABI) vs the oracle (oracle/restate_fluid.py) and the golden
vectors produced by the real reference.

Tolerances (BASELINE.json north_star): relative 1e-5 per step on states, 1e-4 over a trajectory, 1e-3 on gradients;
"relative" = max-norm error / max-norm of the reference field.  Steps that contain the CG solve are compared at
the tolerance the solve was run to (stated per test).
"""
import os

import numpy as np
import pytest
import torch

from oracle import restate as R
from oracle import restate_fluid as F
from conftest import GOLDEN, rel_err

pytestmark = pytest.mark.gpu
MODES = [R.CONSTANT, R.REPLICATE, R.CIRCULAR]


def _ops():
    from pde_control_b200 import ops
    return ops


def _smooth(rng, shape, scale, passes=4):
    a = rng.randn(*shape)
    for _ in range(passes):
        a = (a + np.roll(a, 1, -1) + np.roll(a, -1, -1) + np.roll(a, 1, -2) + np.roll(a, -1, -2)) / 5
    return (a / np.abs(a).max() * scale).astype(np.float32)


def _fields(rng, B, ny, nx, vscale):
    return (np.abs(_smooth(rng, (B, ny, nx), 1.0)), _smooth(rng, (B, ny + 1, nx), vscale), _smooth(rng, (B, ny, nx + 1), vscale))


def _pair(arrs, grad=True):
    cpu = [torch.tensor(a, requires_grad=grad) for a in arrs]
    gpu = [torch.tensor(a, device="cuda", requires_grad=grad) for a in arrs]
    return cpu, gpu


def _backward_weighted(outs, ws):
    sum((o * w).sum() for o, w in zip(outs, ws)).backward()


@pytest.mark.parametrize("mode", MODES)
def test_curl_forward_adjoint_and_divergence_free(mode):
    ops = _ops()
    rng = np.random.RandomState(mode)
    phi = rng.randn(3, 20, 37).astype(np.float32)
    (po,), (pc,) = _pair([phi])
    vyo, vxo = F.curl_2d(po, mode)
    vyc, vxc = ops.curl_2d(pc, mode)
    assert torch.equal(vyc.detach().cpu(), vyo.detach()) and torch.equal(vxc.detach().cpu(), vxo.detach())     # exact
    wy, wx = rng.randn(*vyo.shape).astype(np.float32), rng.randn(*vxo.shape).astype(np.float32)
    _backward_weighted([vyo, vxo], [torch.tensor(wy), torch.tensor(wx)])
    _backward_weighted([vyc, vxc], [torch.tensor(wy).cuda(), torch.tensor(wx).cuda()])
    assert rel_err(pc.grad.cpu(), po.grad) < 1e-6
    # PhiFlow/tests/test_fields.py:114-119: div(curl) == 0
    assert float(ops.mask_divergence(vyc.detach(), vxc.detach(), False).abs().max()) < 1e-5


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("shape,box,vscale", [((2, 33, 40), None, 2.5), ((1, 128, 128), None, 1.5), ((2, 24, 18), (0.5, -1.0, 3.0, 4.5), 0.4)])
def test_advect_centered_forward_and_adjoint(mode, shape, box, vscale):
    ops = _ops()
    B, ny, nx = shape
    rng = np.random.RandomState(ny + mode)
    rho, vy, vx = _fields(rng, B, ny, nx, vscale)
    geo = F.Geometry(ny, nx) if box is None else F.Geometry(ny, nx, box[:2], box[2:])
    (ro, yo, xo), (rc, yc, xc) = _pair([rho, vy, vx])
    oo = F.advect_centered(ro, yo, xo, 0.8, geo, mode)
    oc = ops.advect_centered(rc, yc, xc, 0.8, mode, box)
    assert rel_err(oc.detach().cpu(), oo.detach()) < 1e-5
    w = rng.randn(B, ny, nx).astype(np.float32)
    (oo * torch.tensor(w)).sum().backward()
    (oc * torch.tensor(w).cuda()).sum().backward()
    assert rel_err(rc.grad.cpu(), ro.grad) < 1e-3
    assert rel_err(yc.grad.cpu(), yo.grad) < 1e-3 and rel_err(xc.grad.cpu(), xo.grad) < 1e-3


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("shape,box,vscale", [((2, 33, 40), None, 2.5), ((1, 128, 128), None, 1.5), ((2, 24, 18), (0.5, -1.0, 3.0, 4.5), 0.4)])
def test_advect_staggered_forward_and_adjoint(mode, shape, box, vscale):
    ops = _ops()
    B, ny, nx = shape
    rng = np.random.RandomState(nx + mode)
    _, vy, vx = _fields(rng, B, ny, nx, vscale)
    geo = F.Geometry(ny, nx) if box is None else F.Geometry(ny, nx, box[:2], box[2:])
    (yo, xo), (yc, xc) = _pair([vy, vx])
    ayo, axo = F.advect_staggered(yo, xo, 0.8, geo, mode)
    ayc, axc = ops.advect_staggered(yc, xc, 0.8, mode, box)
    assert rel_err(ayc.detach().cpu(), ayo.detach()) < 1e-5 and rel_err(axc.detach().cpu(), axo.detach()) < 1e-5
    wy, wx = rng.randn(*vy.shape).astype(np.float32), rng.randn(*vx.shape).astype(np.float32)
    _backward_weighted([ayo, axo], [torch.tensor(wy), torch.tensor(wx)])
    _backward_weighted([ayc, axc], [torch.tensor(wy).cuda(), torch.tensor(wx).cuda()])
    assert rel_err(yc.grad.cpu(), yo.grad) < 1e-3 and rel_err(xc.grad.cpu(), xo.grad) < 1e-3


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("shape,box,vscale", [((2, 33, 40), None, 2.5), ((3, 128, 128), None, 1.5), ((2, 24, 18), (0.5, -1.0, 3.0, 4.5), 0.4)])
def test_fused_staggered_force_loss_vs_oracle_and_vs_the_operator_composition(mode, shape, box, vscale):
    """pdec_staggered_force_loss_fwd / _bwd (one term of total_force_loss, src/control/pde/incompressible_flow.py:31-43): loss and the
    gradients w.r.t. BOTH velocities against oracle.restate_fluid.force_loss + autograd; against advect_staggered + torch arithmetic the
    gradient of the advected velocity is the same kernel on the same seed up to one rounding of the seed's scale."""
    ops = _ops()
    B, ny, nx = shape
    rng = np.random.RandomState(7 * nx + mode)
    _, vy, vx = _fields(rng, B, ny, nx, vscale)
    _, ty, tx = _fields(rng, B, ny, nx, vscale)
    geo = F.Geometry(ny, nx) if box is None else F.Geometry(ny, nx, box[:2], box[2:])
    (yo, xo, tyo, txo), (yc, xc, tyc, txc) = _pair([vy, vx, ty, tx])
    ay, ax = F.advect_staggered(yo, xo, 0.8, geo, mode)
    lo = ((ay - tyo) ** 2).sum() / 2 / B + ((ax - txo) ** 2).sum() / 2 / B
    (lo * 0.37).backward()
    lc = ops.staggered_force_loss(yc, xc, tyc, txc, 0.8, mode, box)
    (lc * 0.37).backward()
    assert abs(lc.item() - lo.item()) < 2e-6 * abs(lo.item())
    for a, b in ((yc, yo), (xc, xo), (tyc, tyo), (txc, txo)):
        assert rel_err(a.grad.cpu(), b.grad) < 1e-3
    # the unfused composition on the same inputs
    (y2, x2, ty2, tx2), = [[torch.tensor(a).cuda().requires_grad_() for a in (vy, vx, ty, tx)]]
    ay2, ax2 = ops.advect_staggered(y2, x2, 0.8, mode, box)
    l2 = ((ay2 - ty2) ** 2).sum() / 2 / B + ((ax2 - tx2) ** 2).sum() / 2 / B
    (l2 * 0.37).backward()
    assert abs(lc.item() - l2.item()) < 2e-6 * abs(l2.item())
    for a, b in ((yc, y2), (xc, x2), (tyc, ty2), (txc, tx2)):
        assert rel_err(a.grad.cpu(), b.grad.cpu()) < 2e-6
    # twice in a row: same bits (no atomics anywhere on the path)
    y3, x3 = yc.detach().clone().requires_grad_(), xc.detach().clone().requires_grad_()
    l3 = ops.staggered_force_loss(y3, x3, tyc.detach(), txc.detach(), 0.8, mode, box)
    (l3 * 0.37).backward()
    assert l3.item() == lc.item() and torch.equal(y3.grad, yc.grad) and torch.equal(x3.grad, xc.grad)


@pytest.mark.parametrize("pmode", MODES)
@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("shape,box", [((2, 33, 40), None), ((3, 128, 128), None), ((2, 24, 18), (0.5, -1.0, 3.0, 4.5))])
def test_fused_curl_advect_is_the_two_operator_composition(pmode, mode, shape, box):
    """pdec_curl_advect_2d_fwd / _bwd (ControlledIncompressibleFlow.step, src/control/pde/incompressible_flow.py:53-59): forward bit-identical
    to pdec_curl_2d_fwd + pdec_advect_centered_2d_fwd; gradients (density loss + terms on the velocity, which enter through add_vy /
    add_vx) equal the composition's up to the order of one float addition."""
    ops = _ops()
    B, ny, nx = shape
    rng = np.random.RandomState(3 * ny + 7 * mode + pmode)
    rho = rng.rand(B, ny, nx).astype(np.float32)
    phi = (rng.randn(B, ny, nx) * 1.5).astype(np.float32)
    wr = rng.randn(B, ny, nx).astype(np.float32)
    wy, wx = rng.randn(B, ny + 1, nx).astype(np.float32), rng.randn(B, ny, nx + 1).astype(np.float32)
    (r1, p1), (r2, p2) = [[torch.tensor(a).cuda().requires_grad_() for a in (rho, phi)] for _ in range(2)]
    W = [torch.tensor(a).cuda() for a in (wr, wy, wx)]
    o1, y1, x1 = ops.curl_advect(r1, p1, 0.8, pmode, mode, box)
    y2, x2 = ops.curl_2d(p2, pmode)
    o2 = ops.advect_centered(r2, y2, x2, 0.8, mode, box)
    assert torch.equal(o1, o2) and torch.equal(y1, y2) and torch.equal(x1, x2)
    ((o1 * W[0]).sum() + (y1 * W[1]).sum() + 0.5 * (x1 * W[2]).sum()).backward()
    ((o2 * W[0]).sum() + (y2 * W[1]).sum() + 0.5 * (x2 * W[2]).sum()).backward()
    assert torch.equal(r1.grad, r2.grad)
    assert rel_err(p1.grad.cpu(), p2.grad.cpu()) < 2e-6
    # only the density output used / only a velocity output used
    r3, p3 = [torch.tensor(a).cuda().requires_grad_() for a in (rho, phi)]
    o3, y3, x3 = ops.curl_advect(r3, p3, 0.8, pmode, mode, box)
    (o3 * W[0]).sum().backward()
    r4, p4 = [torch.tensor(a).cuda().requires_grad_() for a in (rho, phi)]
    y4, x4 = ops.curl_2d(p4, pmode)
    (ops.advect_centered(r4, y4, x4, 0.8, mode, box) * W[0]).sum().backward()
    assert torch.equal(r3.grad, r4.grad) and torch.equal(p3.grad, p4.grad)
    r5, p5 = [torch.tensor(a).cuda().requires_grad_() for a in (rho, phi)]
    o5, y5, x5 = ops.curl_advect(r5, p5, 0.8, pmode, mode, box)
    (y5 * W[1]).sum().backward()
    p6 = torch.tensor(phi).cuda().requires_grad_()
    y6, x6 = ops.curl_2d(p6, pmode)
    (y6 * W[1]).sum().backward()
    assert rel_err(p5.grad.cpu(), p6.grad.cpu()) < 2e-6 and float(r5.grad.abs().max()) == 0.0


def test_total_force_loss_of_the_control_module_takes_the_fused_kernel():
    """IncompFlowControl.total_force_loss equivalent (control/incompressible_flow.py) on StaggeredGrids from the curl: same value as the
    oracle's force_loss on the golden notebook frame."""
    from pde_control_b200.physics import Domain, CLOSED, StaggeredGrid, staggered_force_loss
    g = np.load(os.path.join(GOLDEN, "fluid_controlled_64.npz"))
    dt = float(g["dt"])
    dom = Domain([64, 64], boundaries=CLOSED)
    vy, vx = torch.tensor(g["vy"]).cuda(), torch.tensor(g["vx"]).cuda()
    v1 = StaggeredGrid([vy.unsqueeze(-1), vx.unsqueeze(-1)], box=dom.box, extrapolation="boundary")
    v2 = v1.with_tensors([vy * 0.9, vx * 1.1])
    lc = staggered_force_loss(v1, v2, dt)
    geo = F.Geometry(64, 64)
    lo = F.force_loss([torch.tensor(g["vy"]), torch.tensor(g["vy"]) * 0.9], [torch.tensor(g["vx"]), torch.tensor(g["vx"]) * 1.1], dt, geo)
    assert abs(lc.item() - lo.item()) < 1e-5 * abs(lo.item())


def test_circular_adjoint_with_back_traces_beyond_half_the_ring():
    """on an EVEN periodic grid a gather-transpose window clamped to half the ring must visit every source once (offsets
    -R and +R are the same cell).  Velocities large enough that the window covers the whole ring, against the oracle's autograd."""
    ops = _ops()
    rng = np.random.RandomState(77)
    B, ny, nx = 2, 16, 12
    rho, vy, vx = _fields(rng, B, ny, nx, 14.0)            # |v| dt up to ~11 cells: beyond ny / 2 and nx / 2
    geo = F.Geometry(ny, nx)
    (ro, yo, xo), (rc, yc, xc) = _pair([rho, vy, vx])
    oo = F.advect_centered(ro, yo, xo, 0.8, geo, R.CIRCULAR)
    oc = ops.advect_centered(rc, yc, xc, 0.8, R.CIRCULAR, None)
    assert rel_err(oc.detach().cpu(), oo.detach()) < 1e-4
    w = rng.randn(B, ny, nx).astype(np.float32)
    (oo * torch.tensor(w)).sum().backward()
    (oc * torch.tensor(w).cuda()).sum().backward()
    assert rel_err(rc.grad.cpu(), ro.grad) < 1e-3
    (yo2, xo2), (yc2, xc2) = _pair([vy, vx])
    ayo, axo = F.advect_staggered(yo2, xo2, 0.8, geo, R.CIRCULAR)
    ayc, axc = ops.advect_staggered(yc2, xc2, 0.8, R.CIRCULAR, None)
    wy, wx = rng.randn(*vy.shape).astype(np.float32), rng.randn(*vx.shape).astype(np.float32)
    _backward_weighted([ayo, axo], [torch.tensor(wy), torch.tensor(wx)])
    _backward_weighted([ayc, axc], [torch.tensor(wy).cuda(), torch.tensor(wx).cuda()])
    assert rel_err(yc2.grad.cpu(), yo2.grad) < 1e-3 and rel_err(xc2.grad.cpu(), xo2.grad) < 1e-3


@pytest.mark.parametrize("bd", ["open", "closed"])
def test_divergence_and_gradient_subtraction_forward_and_adjoint(bd):
    ops = _ops()
    rng = np.random.RandomState(8)
    B, ny, nx = 2, 19, 30
    _, vy, vx = _fields(rng, B, ny, nx, 1.0)
    p = rng.randn(B, ny, nx).astype(np.float32)
    box = (0.0, 0.0, 9.5, 15.0)           # dx = 0.5
    geo = F.Geometry(ny, nx, box[:2], box[2:])
    closed = bd == "closed"
    (yo, xo, po), (yc, xc, pc) = _pair([vy, vx, p])
    my, mx = F.hard_boundary_mask(yo, xo, bd)
    do = F.divergence(my, mx)
    oyo, oxo = F.subtract_pressure_gradient(my, mx, po, geo, bd)
    dc = ops.mask_divergence(yc, xc, closed)
    oyc, oxc = ops.subtract_pressure_gradient(yc, xc, pc, closed, box)
    assert rel_err(dc.detach().cpu(), do.detach()) < 1e-6
    assert rel_err(oyc.detach().cpu(), oyo.detach()) < 1e-6 and rel_err(oxc.detach().cpu(), oxo.detach()) < 1e-6
    ws = [rng.randn(*t.shape).astype(np.float32) for t in (do, oyo, oxo)]
    _backward_weighted([do, oyo, oxo], [torch.tensor(w) for w in ws])
    _backward_weighted([dc, oyc, oxc], [torch.tensor(w).cuda() for w in ws])
    for a, b in ((yc, yo), (xc, xo), (pc, po)):
        assert rel_err(a.grad.cpu(), b.grad) < 1e-5


def test_normalize_forward_and_adjoint():
    ops = _ops()
    rng = np.random.RandomState(9)
    t, s = rng.rand(3, 40, 50).astype(np.float32), rng.rand(3, 40, 50).astype(np.float32)
    t[2] = 0                                  # sum(target) < epsilon branch
    (to, so), (tc, sc) = _pair([t, s])
    oo, oc = F.normalize_to(to, so), ops.normalize_to(tc, sc)
    assert rel_err(oc.detach().cpu(), oo.detach()) < 1e-5
    w = rng.randn(3, 40, 50).astype(np.float32)
    (oo * torch.tensor(w)).sum().backward()
    (oc * torch.tensor(w).cuda()).sum().backward()
    assert rel_err(tc.grad.cpu(), to.grad) < 1e-3 and rel_err(sc.grad.cpu(), so.grad) < 1e-3


def _demean(p):
    return p - p.mean((1, 2), keepdims=True)


@pytest.mark.parametrize("bd", ["open", "closed"])
def test_cg_matches_reference_golden_and_direct_solve(bd):
    """poisson_40x32.npz: pressure from the reference's SparseCG(accuracy=1e-6); and the fp64 direct solve."""
    ops = _ops()
    g = np.load(os.path.join(GOLDEN, "poisson_40x32.npz"))
    rhs = torch.tensor(g["rhs_" + bd]).cuda()
    p, iters, resid = ops.pressure_solve(rhs, bd == "closed", accuracy=1e-6, max_iterations=5000)
    ref = g["p_" + bd]
    direct = F.pressure_solve(torch.tensor(g["rhs_" + bd]), bd).numpy()
    got = p.cpu().numpy()
    if bd == "closed":
        ref, direct, got = _demean(ref), _demean(direct), _demean(got)
    assert rel_err(got, ref) < 1e-5 and rel_err(got, direct) < 1e-5
    assert 50 < int(iters.max()) < 1000 and float(resid.max()) < 1e-6
    # PhiFlow/tests/test_poisson_solve.py:26-35: laplace(solve(rhs)) == rhs
    assert rel_err(F.laplace_apply(p.cpu().double(), bd), g["rhs_" + bd]) < 1e-4


@pytest.mark.parametrize("bd", ["open", "closed"])
@pytest.mark.parametrize("ny,nx", [(128, 128), (64, 64), (17, 23), (100, 90), (160, 144), (256, 256), (264, 200), (150, 140)])
def test_cg_solves_the_restated_operator(bd, ny, nx):
    """All kernel shapes: the one-CTA on-chip kernel at several block sizes; the grid of on-chip tiles with one grid-wide barrier
    per iteration (> 16384 cells, rows a multiple of 8: 160x144, 256x256, 264x200 with a ragged last tile column); and the
    cooperative streaming kernel (150x140).  Property: A p == rhs to the accuracy of the solve; p equals the fp64 direct solve."""
    ops = _ops()
    rng = np.random.RandomState(ny * nx)
    rhs = _smooth(rng, (3, ny, nx), 1.0, passes=2)
    if bd == "closed":
        rhs -= rhs.mean((1, 2), keepdims=True)       # compatibility condition of the Neumann problem
    p, iters, resid = ops.pressure_solve(torch.tensor(rhs).cuda(), bd == "closed", accuracy=2e-6, max_iterations=4000)
    assert float(resid.max()) < 2e-6 and int(iters.max()) < 4000
    # true residual (fp64 evaluation of A p): the recursive fp32 residual drifts by ~eps * |A| |p| * sqrt(iterations);
    # the oracle's own fp32 CG shows the same drift (scripts/diag_cg.py), the solution error below is what matters
    assert float((F.laplace_apply(p.cpu().double(), bd) - torch.tensor(rhs).double()).abs().max()) < 2e-5 * float(p.abs().max())
    direct = F.pressure_solve(torch.tensor(rhs), bd).numpy()
    got = p.cpu().numpy()
    if bd == "closed":
        direct, got = _demean(direct), _demean(got)
    assert rel_err(got, direct) < 2e-4


@pytest.mark.parametrize("bd", ["open", "closed"])
@pytest.mark.parametrize("ny,nx", [(128, 128), (64, 64), (17, 23), (100, 90), (40, 32), (2, 2), (128, 6)])
def test_direct_solver_vs_fp64_sparse_solve_and_reference_pressures(bd, ny, nx):
    """pdec_poisson_direct_2d (fast diagonalisation; the reference's SparseSciPy, sparse.py:14-37) against the fp64 sparse LU of the
    restated matrix: solution 1e-5, A p == rhs, rhs without the compatibility condition on CLOSED domains included (the singular mode is
    dropped, like the bordered system of the oracle); gradient = the same solve (symmetric); at 40 x 32 also the reference's own pressures."""
    ops = _ops()
    rng = np.random.RandomState(ny * 131 + nx)
    B = 3
    rhs = _smooth(rng, (B, ny, nx), 1.0, passes=2) if min(ny, nx) > 4 else rng.randn(B, ny, nx).astype(np.float32)
    if (ny, nx) == (40, 32):
        rhs = np.load(os.path.join(GOLDEN, "poisson_40x32.npz"))["rhs_" + bd].astype(np.float32)
    t = torch.tensor(rhs, device="cuda", requires_grad=True)
    p = ops.pressure_solve_direct(t, bd == "closed")
    w = torch.tensor(rng.randn(B, ny, nx).astype(np.float32), device="cuda")
    (p * w).sum().backward()
    direct = F.pressure_solve(torch.tensor(rhs), bd).numpy()
    got = p.detach().cpu().numpy()
    if bd == "closed":
        assert float(np.abs(got.mean((1, 2))).max()) < 1e-4 * float(np.abs(got).max())        # the mean-free solution
        direct, got = _demean(direct), _demean(got)
    assert rel_err(got, direct) < 1e-5
    # A p == rhs (minus its mean on closed domains: that component is outside the range of A)
    target = rhs - rhs.mean((1, 2), keepdims=True) if bd == "closed" else rhs
    assert float((F.laplace_apply(p.detach().cpu().double(), bd) - torch.tensor(target).double()).abs().max()) < 2e-5 * max(1.0, float(p.abs().max()))
    gdirect = F.pressure_solve(w.cpu(), bd).numpy()
    ggot = t.grad.cpu().numpy()
    if bd == "closed":
        gdirect, ggot = _demean(gdirect), _demean(ggot)
    assert rel_err(ggot, gdirect) < 1e-5
    if (ny, nx) == (40, 32):
        ref = np.load(os.path.join(GOLDEN, "poisson_40x32.npz"))["p_" + bd]                # reference SparseCG(1e-6) pressures
        mine = p.detach().cpu().numpy()
        if bd == "closed":
            ref, mine = _demean(ref), _demean(mine)
        assert rel_err(mine, ref) < 1e-5


def test_cg_iteration_semantics_guess_and_batch_independence():
    ops = _ops()
    rng = np.random.RandomState(1)
    rhs = _smooth(rng, (4, 32, 32), 1.0, passes=1)
    rhs_c = torch.tensor(rhs).cuda()
    # fixed iteration count (accuracy=None: "the algorithm runs until max_iterations", blas.py:13) == oracle CG iterates
    p5, it5, _ = ops.pressure_solve(rhs_c, False, accuracy=None, max_iterations=5)
    o5, _ = F.conjugate_gradient(torch.tensor(rhs), "open", accuracy=None, max_iterations=5)
    assert int(it5.min()) == 5 and rel_err(p5.cpu(), o5) < 1e-5
    # already-converged input: zero iterations, zero pressure
    p0, it0, _ = ops.pressure_solve(rhs_c * 1e-9, False, accuracy=1e-5)
    assert int(it0.max()) == 0 and float(p0.abs().max()) == 0.0
    # a good initial guess needs fewer iterations and gives the same solution
    p, it, _ = ops.pressure_solve(rhs_c, False, accuracy=1e-5)
    pg, itg, _ = ops.pressure_solve(rhs_c, False, accuracy=1e-5, guess=p)
    assert int(itg.max()) <= 2 and rel_err(pg.cpu(), p.cpu()) < 1e-4
    # per-element stop: iteration counts differ per element; element 0 is bit-identical whatever else is in the batch
    # (PhiFlow/tests/test_fluid.py:80-100)
    rhs2 = rhs.copy()
    rhs2[1:] = 0
    pb, itb, _ = ops.pressure_solve(torch.tensor(rhs2).cuda(), False, accuracy=1e-5)
    assert torch.equal(pb[0], p[0]) and int(itb[1]) == 0
    # the reference's whole-batch stop rule (blas.py:33-37): every element iterates until the batch-wide max|r| < accuracy;
    # iteration count equals the restated whole-batch CG's (fp32 summation order may move it by a step or two)
    scaled = rhs * np.array([1.0, 0.3, 0.05, 1e-3], np.float32)[:, None, None]
    pw, itw, rw = ops.pressure_solve(torch.tensor(scaled).cuda(), False, accuracy=1e-5, stop_mode=ops.CG_STOP_WHOLE_BATCH)
    pe, ite, _ = ops.pressure_solve(torch.tensor(scaled).cuda(), False, accuracy=1e-5)
    ow, k = F.conjugate_gradient(torch.tensor(scaled), "open", accuracy=1e-5)
    assert int(itw.min()) == int(itw.max()) and abs(int(itw[0]) - k) <= 2 and float(rw.max()) < 1e-5
    assert int(ite.min()) < int(ite.max()) <= int(itw[0])
    assert rel_err(pw.cpu(), ow) < 1e-4


def test_pressure_solve_gradient_is_a_second_solve():
    ops = _ops()
    rng = np.random.RandomState(2)
    rhs = _smooth(rng, (2, 24, 24), 1.0, passes=1)
    w = _smooth(rng, (2, 24, 24), 1.0, passes=1)
    (ro,), (rc,) = _pair([rhs])
    (F.pressure_solve(ro, "open") * torch.tensor(w)).sum().backward()
    p, _, _ = ops.pressure_solve(rc, False, accuracy=1e-6, max_iterations=3000)
    (p * torch.tensor(w).cuda()).sum().backward()
    assert rel_err(rc.grad.cpu(), ro.grad) < 1e-3


@pytest.mark.parametrize("name", ["fluid_open_32", "fluid_closed_32x24", "fluid_open_128"])
def test_projected_step_vs_reference_golden(name):
    """One IncompressibleFlow.step from every golden frame.  The kernels follow the TF edge rule; the golden vectors come
    from the NumPy backend, so samples with xi == N-1 exactly (boundary faces with velocity exactly 0) are excluded --
    they are identified with the oracle under both rules and reported."""
    ops = _ops()
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    B, ny, nx = g["rho0"].shape
    bd, dt = str(g["boundary"]), float(g["dt"])
    geo = F.Geometry(ny, nx)
    for k in range(int(g["steps"])):
        rho, vy, vx = [g["%s%d" % (n, k)] for n in ("rho", "vy", "vx")]
        rc, yc, xc, pc, iters = ops.incompressible_step(*[torch.tensor(a).cuda() for a in (rho, vy, vx)], dt, bd == "closed",
                                                        accuracy=1e-6, max_iterations=3000)
        # advected (pre-projection) velocity: per-step tolerance 1e-5 away from exact-edge samples
        ay, ax = ops.advect_staggered(torch.tensor(vy).cuda(), torch.tensor(vx).cuda(), dt, F.modes_of(bd)[1])
        t = [torch.tensor(a) for a in (vy, vx)]
        tf_rule = F.advect_staggered(t[0], t[1], dt, geo, F.modes_of(bd)[1])
        R.NUMPY_EDGE_RULE = True
        try:
            np_rule = F.advect_staggered(t[0], t[1], dt, geo, F.modes_of(bd)[1])
        finally:
            R.NUMPY_EDGE_RULE = False
        flips = sum(int((a != b).sum()) for a, b in zip(tf_rule, np_rule))
        for got, gold, a, b in ((ay, g["avy%d" % (k + 1)], tf_rule[0], np_rule[0]), (ax, g["avx%d" % (k + 1)], tf_rule[1], np_rule[1])):
            keep = (a == b).numpy()
            assert rel_err(got.cpu().numpy() * keep, gold * keep) < 1e-5
        assert rel_err(rc.cpu(), g["rho%d" % (k + 1)]) < 1e-5
        if flips == 0:      # the projection is global: compare projected velocities only when no edge sample flipped
            assert rel_err(yc.cpu(), g["vy%d" % (k + 1)]) < 2e-5 and rel_err(xc.cpu(), g["vx%d" % (k + 1)]) < 2e-5


@pytest.mark.parametrize("bd", ["open", "closed"])
def test_trajectory_and_gradient_vs_oracle(bd):
    """4 projected steps: states within 1e-4 over the trajectory, control-loss gradients within 1e-3.
    Oracle pressure = fp64 direct solve; CUDA CG runs to 1e-6."""
    ops = _ops()
    rng = np.random.RandomState(12)
    B, ny, nx = 2, 32, 32
    rho, vy, vx = _fields(rng, B, ny, nx, 1.2)
    geo = F.Geometry(ny, nx)
    (ro, yo, xo), (rc, yc, xc) = _pair([rho, vy, vx])
    so, sc = (ro, yo, xo), (rc, yc, xc)
    for _ in range(4):
        so = F.incompressible_step(*so, 0.9, geo, bd)[:3]
        sc = ops.incompressible_step(*sc, 0.9, bd == "closed", accuracy=1e-6, max_iterations=3000)[:3]
    for a, b in zip(sc, so):
        assert rel_err(a.detach().cpu(), b.detach()) < 1e-4
    target = _fields(rng, B, ny, nx, 1.0)[0]
    (R.l2_loss(so[0] - torch.tensor(target)) + 0.1 * R.l2_loss(so[1]) + 0.1 * R.l2_loss(so[2])).backward()
    d = sc[0] - torch.tensor(target).cuda()
    ((d ** 2).sum() / 2 / B + 0.1 * (sc[1] ** 2).sum() / 2 / B + 0.1 * (sc[2] ** 2).sum() / 2 / B).backward()
    for a, b in ((rc, ro), (yc, yo), (xc, xo)):
        assert rel_err(a.grad.cpu(), b.grad) < 1e-3


def test_controlled_variant_and_force_loss_vs_golden_and_oracle():
    """Notebook variant (src/control/pde/incompressible_flow.py:53-59, 31-43): curl -> advect density; force loss =
    l2(semi_lagrangian(v, v) - v_next)."""
    ops = _ops()
    g = np.load(os.path.join(GOLDEN, "fluid_controlled_64.npz"))
    dt = float(g["dt"])
    (ro, po), (rc, pc) = _pair([g["rho"], g["phi"]])
    vyc, vxc = ops.curl_2d(pc, R.CONSTANT)
    nxt = ops.advect_centered(rc, vyc, vxc, dt, R.CONSTANT)
    ay, ax = ops.advect_staggered(vyc, vxc, dt, R.REPLICATE)
    assert rel_err(vyc.detach().cpu(), g["vy"]) == 0.0 and rel_err(vxc.detach().cpu(), g["vx"]) == 0.0
    assert rel_err(nxt.detach().cpu(), g["rho_next"]) < 1e-5
    assert rel_err(ay.detach().cpu(), g["avy"]) < 1e-5 and rel_err(ax.detach().cpu(), g["avx"]) < 1e-5
    # gradient of (density loss + force loss) w.r.t. the potential and the density
    geo = F.Geometry(64, 64)
    rng = np.random.RandomState(4)
    tgt = rng.rand(*g["rho"].shape).astype(np.float32)
    v2y, v2x = g["avy"] * 0.9, g["avx"] * 1.1
    n2, vyo, vxo = F.controlled_step(ro, po, dt, geo)
    lo = R.l2_loss(n2 - torch.tensor(tgt)) + F.force_loss([vyo, torch.tensor(v2y)], [vxo, torch.tensor(v2x)], dt, geo)
    lo.backward()
    B = tgt.shape[0]
    lc = ((nxt - torch.tensor(tgt).cuda()) ** 2).sum() / 2 / B + ((ay - torch.tensor(v2y).cuda()) ** 2).sum() / 2 / B \
        + ((ax - torch.tensor(v2x).cuda()) ** 2).sum() / 2 / B
    lc.backward()
    assert abs(lc.item() - lo.item()) < 1e-4 * abs(lo.item())
    assert rel_err(pc.grad.cpu(), po.grad) < 1e-3 and rel_err(rc.grad.cpu(), ro.grad) < 1e-3


@pytest.mark.parametrize("closed", [False, True, (True, False, False, True), (False, True, True, True)])
@pytest.mark.parametrize("B,ny,nx", [(5, 128, 128), (2, 64, 128), (3, 100, 90), (1, 37, 50), (2, 2, 3), (150, 16, 24)])
def test_tensor_core_direct_solver_against_the_fp32_kernel_and_the_residual(closed, B, ny, nx):
    """pdec_poisson_direct_tc_2d (tcgen05.mma kind::tf32, three products on hi / lo splits, accumulators in tensor memory) against
    pdec_poisson_direct_2d (fp32 FFMA) on the same right-hand sides -- uniform and per-side boundaries, ragged sizes (zero padding), more
    elements than SMs -- and, independently, A p == rhs through the oracle's stencil on the uniform domains.  Same bits twice in a row."""
    ops = _ops()
    rng = np.random.RandomState(ny * 7 + nx + B)
    rhs = _smooth(rng, (B, ny, nx), 1.0, passes=2) if min(ny, nx) > 4 else rng.randn(B, ny, nx).astype(np.float32)
    t = torch.tensor(rhs, device="cuda")
    p_tc, p_tc2, p_ff = torch.empty_like(t), torch.empty_like(t), torch.empty_like(t)
    ops.poisson_direct_raw(t, p_tc, closed, tensor_cores=True)
    ops.poisson_direct_raw(t, p_ff, closed, tensor_cores=False)
    ops.poisson_direct_raw(t, p_tc2, closed, tensor_cores=True)
    assert torch.equal(p_tc, p_tc2)
    scale = float(p_ff.abs().max())
    assert float((p_tc - p_ff).abs().max()) < 2e-5 * scale
    if isinstance(closed, bool):
        bd = "closed" if closed else "open"
        target = rhs - rhs.mean((1, 2), keepdims=True) if closed else rhs
        resid = (F.laplace_apply(p_tc.cpu().double(), bd) - torch.tensor(target).double()).abs().max()
        assert float(resid) < 4e-5 * max(1.0, scale)


@pytest.mark.parametrize("closed", [False, True, (True, False, False, True)])
@pytest.mark.parametrize("B,ny,nx", [(2, 200, 160), (1, 512, 384)])
def test_direct_solver_beyond_128_uses_library_gemms_and_matches_the_fp64_formula(closed, B, ny, nx):
    """ops.poisson_direct_raw on grids the on-chip kernels do not take: p = Ey [(Ey^T rhs Ex) / (l_y + l_x)] Ex^T with fp32 library GEMMs,
    against the same formula in fp64 (whose stencil residual is at 1e-11) and, on uniform domains, against A p == rhs; the differentiable
    wrapper's gradient is the same solve."""
    ops = _ops()
    rng = np.random.RandomState(ny + nx)
    rhs = _smooth(rng, (B, ny, nx), 1.0, passes=3)
    t = torch.tensor(rhs, device="cuda", requires_grad=True)
    p = ops.pressure_solve_direct(t, closed)
    Ey, EyT, Ex, ExT, inv = ops._spectral_large(t.device, ny, nx, closed)
    code = ops.closed_code(closed)
    sides = (0xF if code == 1 else 0) if code < 0x10 else (code & 0xF)
    Ey64, ly = ops._eig_1d(ny, sides & 3, t.device)
    Ex64, lx = ops._eig_1d(nx, (sides >> 2) & 3, t.device)
    assert float((Ey64.T @ Ey64 - torch.eye(ny, device="cuda", dtype=torch.float64)).abs().max()) < 1e-10      # orthonormal bases
    lam = ly[:, None] + lx[None, :]
    inv64 = torch.where(lam.abs() < 1e-9, torch.zeros_like(lam), 1.0 / lam)
    p64 = Ey64 @ ((Ey64.T @ t.detach().double() @ Ex64) * inv64) @ Ex64.T
    assert rel_err(p.detach().cpu(), p64.cpu()) < 2e-5
    if isinstance(closed, bool):
        bd = "closed" if closed else "open"
        target = rhs - rhs.mean((1, 2), keepdims=True) if closed else rhs
        assert float((F.laplace_apply(p64.cpu(), bd) - torch.tensor(target).double()).abs().max()) < 1e-6
    w = torch.tensor(rng.randn(B, ny, nx).astype(np.float32), device="cuda")
    (p * w).sum().backward()
    g64 = Ey64 @ ((Ey64.T @ w.double() @ Ex64) * inv64) @ Ex64.T
    assert rel_err(t.grad.cpu(), g64.cpu()) < 2e-5


def test_tensor_core_direct_solver_c_abi_codes():
    """Return codes of the tensor-core entry points: sizes beyond 128, too small or misaligned spectral block, null pointers, B = 0."""
    ops = _ops()
    from pde_control_b200 import _lib
    L = _lib.lib()
    n = L.pdec_poisson_direct_tc_2d_spectral_bytes()
    assert n % 16 == 0 and n > 4 * 128 * 128 * 4
    blk = torch.empty(n // 4 + 4, device="cuda")
    st = _lib.stream_ptr()
    assert L.pdec_poisson_direct_tc_2d_setup(_lib.ptr(blk), n, 129, 64, 0, st) == -2
    assert L.pdec_poisson_direct_tc_2d_setup(_lib.ptr(blk), n - 16, 64, 64, 0, st) == -4
    assert L.pdec_poisson_direct_tc_2d_setup(blk.data_ptr() + 4, n, 64, 64, 0, st) == -1
    assert L.pdec_poisson_direct_tc_2d_setup(0, n, 64, 64, 0, st) == -1
    assert L.pdec_poisson_direct_tc_2d_setup(_lib.ptr(blk), n, 64, 64, 0, st) == 0
    x = torch.zeros(2, 64, 64, device="cuda")
    assert L.pdec_poisson_direct_tc_2d(_lib.ptr(x), _lib.ptr(x), 0, 64, 64, _lib.ptr(blk), st) == 0
    assert L.pdec_poisson_direct_tc_2d(0, _lib.ptr(x), 2, 64, 64, _lib.ptr(blk), st) == -1
    assert L.pdec_poisson_direct_tc_2d(_lib.ptr(x), _lib.ptr(x), 2, 200, 64, _lib.ptr(blk), st) == -2
    torch.cuda.synchronize()


@pytest.mark.parametrize("closed", [False, True])
@pytest.mark.parametrize("B,ny,nx", [(3, 128, 128), (2, 64, 128), (2, 100, 90), (1, 37, 50), (2, 16, 24)])
def test_fused_projection_is_bit_identical_to_the_three_operator_composition(closed, B, ny, nx):
    """pdec_project_2d_fwd / _bwd (divergence in the prologue of the on-chip CG, gradient subtraction in its epilogue) against
    pdec_mask_divergence_2d -> pdec_cg_poisson_2d -> pdec_subtract_pressure_gradient_2d and their adjoints: same bits, forward and
    reverse, including upstream gradients on the pressure and the divergence outputs, on a non-square box."""
    ops = _ops()
    rng = np.random.RandomState(ny * 7 + nx + int(closed))
    vy = (rng.randn(B, ny + 1, nx) * 0.5).astype(np.float32)
    vx = (rng.randn(B, ny, nx + 1) * 0.5).astype(np.float32)
    box = (0.0, 0.0, float(ny) * 1.5, float(nx) * 0.75)
    w = [torch.tensor(rng.randn(*s).astype(np.float32), device="cuda") for s in ((B, ny + 1, nx), (B, ny, nx + 1), (B, ny, nx), (B, ny, nx))]
    res = []
    for fused, use_p in ((True, True), (False, True), (True, False), (False, False)):      # fused=False is the default path
        t = [torch.tensor(a, device="cuda", requires_grad=True) for a in (vy, vx)]
        oy, ox, p, iters, div = ops.divergence_free(t[0], t[1], closed, box, accuracy=1e-5, max_iterations=1500, fused=fused)
        loss = (oy * w[0]).sum() + (ox * w[1]).sum()
        if use_p:
            loss = loss + 0.3 * (p * w[2]).sum() + 0.7 * (div * w[3]).sum()
        loss.backward()
        res.append([oy.detach(), ox.detach(), p.detach(), div.detach(), iters, t[0].grad, t[1].grad])
    for a, b in ((res[0], res[1]), (res[2], res[3])):
        for x, y in zip(a, b):
            assert torch.equal(x, y)
    assert int(res[0][4].max()) < 1500


def test_controlled_variant_at_config3_size_vs_reference_golden():
    """The notebook variant at 128 x 128 (the size bench.py's fluid_notebook workload runs) against the live reference's golden:
    curl exact, advected density and self-advected velocity 1e-5."""
    ops = _ops()
    g = np.load(os.path.join(GOLDEN, "fluid_controlled_128.npz"))
    rho, phi = torch.tensor(g["rho"]).cuda(), torch.tensor(g["phi"]).cuda()
    vy, vx = ops.curl_2d(phi, R.CONSTANT)
    assert rel_err(vy.cpu(), g["vy"]) == 0.0 and rel_err(vx.cpu(), g["vx"]) == 0.0
    nxt = ops.advect_centered(rho, vy, vx, float(g["dt"]), R.CONSTANT)
    ay, ax = ops.advect_staggered(vy, vx, float(g["dt"]), R.REPLICATE)
    # the NumPy reference interpolates at a sample coordinate of exactly N - 1 where the TF rule (which the kernels follow) returns 0:
    # compare where the oracle agrees under both rules
    geo = F.Geometry(128, 128)
    R.NUMPY_EDGE_RULE = True
    try:
        np_rule = F.controlled_step(torch.tensor(g["rho"]), torch.tensor(g["phi"]), float(g["dt"]), geo)[0]
    finally:
        R.NUMPY_EDGE_RULE = False
    tf_rule = F.controlled_step(torch.tensor(g["rho"]), torch.tensor(g["phi"]), float(g["dt"]), geo)[0]
    keep = (np_rule == tf_rule).numpy()
    assert keep.mean() > 0.99
    assert rel_err(nxt.cpu().numpy() * keep, g["rho_next"] * keep) < 1e-5
    assert rel_err(ay.cpu(), g["avy"]) < 1e-5 and rel_err(ax.cpu(), g["avx"]) < 1e-5


def test_adjoint_is_bit_reproducible_and_batch_independent():
    ops = _ops()
    rng = np.random.RandomState(6)
    rho, vy, vx = _fields(rng, 4, 48, 48, 2.0)

    def run(r, y, x):
        t = [torch.tensor(a, device="cuda", requires_grad=True) for a in (r, y, x)]
        out = ops.incompressible_step(*t, 1.0, False, accuracy=1e-5)
        (out[0].sum() + (out[1] ** 2).sum() + (out[2] ** 2).sum()).backward()
        return [a.grad.clone() for a in t] + [o.detach() for o in out[:3]]

    a, b = run(rho, vy, vx), run(rho, vy, vx)
    assert all(torch.equal(p, q) for p, q in zip(a, b))
    rho2, vy2, vx2 = rho.copy(), vy.copy(), vx.copy()
    rho2[1:], vy2[1:], vx2[1:] = 0, 0, 0
    c = run(rho2, vy2, vx2)
    assert all(torch.equal(p[0], q[0]) for p, q in zip(a, c))      # PhiFlow/tests/test_fluid.py:80-100


def test_full_size_config3_properties():
    """BASELINE config 3 size (128x128, batch 64): size-independent properties.
    (1) CLOSED projection is divergence-free everywhere, OPEN only in the interior;
    (2) the projection is linear and its adjoint kernels are its transpose: <P u, w> == <u, P^T w>;
    (3) projecting twice changes nothing (idempotence)."""
    ops = _ops()
    rng = np.random.RandomState(3)
    B, n = 64, 128
    vy = torch.tensor(_smooth(rng, (B, n + 1, n), 0.5), device="cuda")
    vx = torch.tensor(_smooth(rng, (B, n, n + 1), 0.5), device="cuda")
    for closed in (False, True):
        u = [vy.clone().requires_grad_(True), vx.clone().requires_grad_(True)]
        py, px, p, iters, div = ops.divergence_free(u[0], u[1], closed, accuracy=1e-5)
        d2 = ops.mask_divergence(py.detach(), px.detach(), closed)
        inner = d2[:, 1:-1, 1:-1] if not closed else d2
        assert float(inner.abs().max()) < 5e-5 and int(iters.max()) < 2000
        if not closed:
            assert float(d2.abs().max()) > 1e-3
        wy, wx = torch.randn_like(py), torch.randn_like(px)
        ((py * wy).sum() + (px * wx).sum()).backward()
        lhs = float((py.detach().double() * wy.double()).sum() + (px.detach().double() * wx.double()).sum())
        rhs = float((u[0].grad.double() * vy.double()).sum() + (u[1].grad.double() * vx.double()).sum())
        assert abs(lhs - rhs) < 1e-3 * max(abs(lhs), 1.0)
        qy, qx, _, _, _ = ops.divergence_free(py.detach(), px.detach(), closed, accuracy=1e-5)
        if closed:
            assert rel_err(qy.cpu(), py.detach().cpu()) < 1e-4 and rel_err(qx.cpu(), px.detach().cpu()) < 1e-4


def test_edge_cases():
    ops = _ops()
    from pde_control_b200._lib import PdecError
    z = torch.zeros(0, 8, 8, device="cuda")
    assert ops.advect_centered(z, torch.zeros(0, 9, 8, device="cuda"), torch.zeros(0, 8, 9, device="cuda"), 1.0).shape == (0, 8, 8)
    with pytest.raises(PdecError):
        ops.curl_2d(torch.zeros(1, 1, 8, device="cuda"))                       # ny < 2
    with pytest.raises(RuntimeError):
        ops.mask_divergence(torch.zeros(1, 9, 8), torch.zeros(1, 8, 9), False)  # CPU tensors: no fallback
    # huge velocities: every back-trace leaves the domain -> zeros ('constant') and finite gradients
    rho = torch.rand(1, 16, 16, device="cuda", requires_grad=True)
    vy = torch.full((1, 17, 16), 1e4, device="cuda", requires_grad=True)
    vx = torch.full((1, 16, 17), -1e4, device="cuda", requires_grad=True)
    out = ops.advect_centered(rho, vy, vx, 1.0, R.CONSTANT)
    out.sum().backward()
    assert float(out.abs().max()) == 0.0 and bool(torch.isfinite(rho.grad).all()) and float(rho.grad.abs().max()) == 0.0


def test_edge_cases_of_the_round_2_entry_points():
    """C-ABI return codes and empty / ragged inputs of pdec_project_2d_*, pdec_poisson_direct_2d*, pdec_halo_* and the per-side closed codes:
    empty batches are no-ops, sizes beyond the on-chip kernels are refused with PDEC_ERR_UNSUPPORTED_SIZE (-2), bad arguments with -1,
    too-small buffers with -4 -- never a crash, never a silent fallback."""
    from pde_control_b200 import _lib, ops
    L = _lib.lib()
    st = _lib.stream_ptr()
    dev = torch.device("cuda")
    vy, vx = torch.zeros(1, 9, 8, device=dev), torch.zeros(1, 8, 9, device=dev)
    oy, ox = torch.empty_like(vy), torch.empty_like(vx)
    P = _lib.ptr
    # fused projection
    assert L.pdec_project_2d_supported(128, 128) == 1 and L.pdec_project_2d_supported(129, 64) == 0 and L.pdec_project_2d_supported(1, 8) == 0
    assert L.pdec_project_2d_fwd(P(vy), P(vx), P(oy), P(ox), None, None, None, None, 0, 8, 8, 0, 8.0, 8.0, 1e-5, 100, st) == 0        # B = 0
    assert L.pdec_project_2d_fwd(P(vy), P(vx), P(oy), P(ox), None, None, None, None, 1, 200, 8, 0, 200.0, 8.0, 1e-5, 100, st) == -2
    assert L.pdec_project_2d_fwd(None, P(vx), P(oy), P(ox), None, None, None, None, 1, 8, 8, 0, 8.0, 8.0, 1e-5, 100, st) == -1
    assert L.pdec_project_2d_bwd(P(vy), P(vx), None, None, P(oy), P(ox), None, None, 1, 8, 8, 1, 8.0, -1.0, 1e-5, 100, st) == -1   # size <= 0
    # a zero field projects to zero in zero iterations, with every per-side code
    it = torch.full((1,), -1, dtype=torch.int32, device=dev)
    for code in (0, 1, 0x10 | 0b0101, 0x10 | 0b1010):
        assert L.pdec_project_2d_fwd(P(vy), P(vx), P(oy), P(ox), None, None, P(it), None, 1, 8, 8, code, 8.0, 8.0, 1e-5, 100, st) == 0
        torch.cuda.synchronize()
        assert int(it[0]) == 0 and float(oy.abs().max()) == 0.0 and float(ox.abs().max()) == 0.0
    # direct solver
    nbytes = L.pdec_poisson_direct_2d_spectral_bytes()
    spec = torch.empty(nbytes // 4, device=dev)
    assert L.pdec_poisson_direct_2d_supported(128, 128) == 1 and L.pdec_poisson_direct_2d_supported(130, 8) == 0
    assert L.pdec_poisson_direct_2d_setup(P(spec), nbytes - 4, 8, 8, 0, st) == -4
    assert L.pdec_poisson_direct_2d_setup(P(spec), nbytes, 300, 8, 0, st) == -2
    assert L.pdec_poisson_direct_2d_setup(P(spec), nbytes, 8, 8, 0x10 | 0b0110, st) == 0
    rhs, p = torch.zeros(0, 8, 8, device=dev), torch.zeros(0, 8, 8, device=dev)
    assert L.pdec_poisson_direct_2d(P(rhs), P(p), 0, 8, 8, P(spec), st) == 0                                              # empty batch
    assert L.pdec_poisson_direct_2d(None, None, 1, 8, 8, P(spec), st) == -1
    assert ops.pressure_solve_direct(torch.zeros(0, 8, 8, device=dev), False).shape == (0, 8, 8)
    # a ragged shape (rows and columns not multiples of the thread tiles) against the fp64 solve, mixed sides
    rng = np.random.RandomState(9)
    r = rng.randn(2, 13, 27).astype(np.float32)
    got = ops.pressure_solve_direct(torch.tensor(r, device=dev), (True, False, False, True)).cpu().numpy()
    want = F._solve_sides_np(r, (True, False, False, True))
    assert rel_err(got, want) < 1e-5
    # halo exchange: argument checks (the transport itself needs two GPUs: tests/test_slab_multi_gpu.py)
    assert L.pdec_halo_inbox_bytes(2, 16, 4) == 4 * 2 * 4 * 16 * 8 and L.pdec_halo_inbox_bytes(0, 16, 4) == 0
    f = torch.zeros(2, 8, 16, device=dev)
    box = torch.zeros(L.pdec_halo_inbox_bytes(2, 16, 4) // 4, device=dev)
    assert L.pdec_halo_inbox_init(P(box), 2, 16, 4, st) == 0
    assert L.pdec_halo_push_rows(P(f), 2, 8, 16, 0, 0, P(box), 0, 1, 4, st) == 0              # zero rows: nothing to do
    assert L.pdec_halo_push_rows(P(f), 2, 8, 16, 6, 4, P(box), 0, 1, 4, st) == -1             # rows beyond the field
    assert L.pdec_halo_push_rows(P(f), 2, 8, 16, 0, 5, P(box), 0, 1, 4, st) == -1             # more rows than the inbox holds
    assert L.pdec_halo_push_rows(P(f), 2, 8, 16, 0, 2, P(box), 0, 0, 4, st) == -1             # sequence number 0 is "never written"
    # a rank may push to ITSELF (one process): rows arrive, tag-checked, and accumulate in reverse mode
    f[:, :2] = torch.arange(32, device=dev, dtype=torch.float32).reshape(1, 2, 16) + 1.0
    assert L.pdec_halo_push_rows(P(f), 2, 8, 16, 0, 2, P(box), 1, 7, 4, st) == 0
    g = torch.ones(2, 8, 16, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    assert L.pdec_halo_pull_rows(P(g), 2, 8, 16, 6, 2, P(box), 1, 7, 1, 4, P(status), st) == 0
    torch.cuda.synchronize()
    assert int(status[0]) == 0 and torch.equal(g[:, 6:], f[:, :2] + 1.0) and float(g[:, :6].min()) == 1.0
