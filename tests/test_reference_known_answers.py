"""The known-answer cases the reference's own test-suite holds for this path, against the oracle on CPU and against the CUDA
path on the GPU.  Values are the ones written in the reference's tests; nothing here reads /root/reference."""
import numpy as np
import pytest
import torch

from oracle import restate as R
from oracle import restate_fluid as F
from conftest import rel_err


def _poisson_examples():
    """PhiFlow/tests/test_poisson_solve.py:14-23: a linear ramp and a single column, mean removed, on a 4 x 5 grid."""
    ex1 = np.tile(np.linspace(1, 0, 5), [4, 1])
    ex1 = ex1 - ex1.mean()
    ex2 = np.zeros([4, 5])
    ex2[:, 2] = 1
    ex2 -= ex2.mean()
    return np.stack([ex1, ex2]).astype(np.float32)


def test_inner_interpolation_known_answer():
    """PhiFlow/tests/test_fields.py:23-35: a 2 x 3 grid on box[0:2, 0:3] resampled at the centres of a 2 x 2 grid on
    box[0:2, 0.5:2.5] gives [[1.5, 2.5], [4.5, 5.5]] (fast path and generic path alike)."""
    data = torch.tensor([[[1., 2., 3.], [4., 5., 6.]]])
    ys = (np.array([0.5, 1.5]) - 0.0) / 2.0 * 2 - 0.5            # sample_at: local * resolution - 0.5 (grid.py:104-105)
    xs = (np.array([1.0, 2.0]) - 0.0) / 3.0 * 3 - 0.5
    coords = torch.tensor(np.stack(np.meshgrid(ys, xs, indexing="ij"), -1)[None], dtype=torch.float32)
    # the grid's extrapolation there is the default 'boundary' (replicate)
    np.testing.assert_array_equal(R.resample(data, coords, R.REPLICATE)[0].numpy(), [[1.5, 2.5], [4.5, 5.5]])
    # 'constant': the reference test runs the NumPy backend, which keeps a sample at index N - 1 exactly (interpn, scipy_backend.py:
    # 149-158); TensorFlow's resample -- the rule the kernels follow -- zeroes it (tf_backend.py:466-472)
    R.NUMPY_EDGE_RULE = True
    try:
        np.testing.assert_array_equal(R.resample(data, coords, R.CONSTANT)[0].numpy(), [[1.5, 2.5], [4.5, 5.5]])
    finally:
        R.NUMPY_EDGE_RULE = False
    np.testing.assert_array_equal(R.resample(data, coords, R.CONSTANT)[0].numpy(), [[1.5, 2.5], [0.0, 0.0]])


def test_gradient_and_laplace_padding_known_answers():
    """PhiFlow/tests/test_math.py:118-126 (forward difference of arange(5): replicate [1,1,1,1,0], circular [1,1,1,1,-4], dx = 0.1 ->
    [10,10,10,10,0]) and :35-52 (laplace of a constant field is 0 for every padding) with the oracle's pad_axis / laplace_1d."""
    t = torch.arange(5, dtype=torch.float32)[None]
    rep = R.pad_axis(t, 1, 0, 1, R.REPLICATE)
    cir = R.pad_axis(t, 1, 0, 1, R.CIRCULAR)
    np.testing.assert_array_equal((rep[:, 1:] - rep[:, :-1])[0].numpy(), [1, 1, 1, 1, 0])
    np.testing.assert_array_equal((cir[:, 1:] - cir[:, :-1])[0].numpy(), [1, 1, 1, 1, -4])
    np.testing.assert_allclose(((rep[:, 1:] - rep[:, :-1]) / 0.1)[0].numpy(), [10, 10, 10, 10, 0], rtol=1e-6)
    const = torch.full((2, 4), 3.0)
    for mode in (R.REPLICATE, R.CIRCULAR):
        assert float(R.laplace_1d(const, 1.0, mode).abs().max()) == 0.0


def test_divide_no_nan_known_answer():
    """PhiFlow/tests/test_math.py:94-103: [1,-1,0,1,-1] / [1,2,0,0,0] = [1,-0.5,0,0,0].  The oracle's CG uses the same rule: a zero
    right-hand side gives a zero, finite pressure with no iteration diverging."""
    p, it = F.conjugate_gradient(torch.zeros(2, 4, 5), F.OPEN, 1e-5, 10)
    assert bool(torch.isfinite(p).all()) and float(p.abs().max()) == 0.0


@pytest.mark.parametrize("bd", [F.CLOSED, F.OPEN])
def test_fixed_poisson_examples_oracle(bd):
    """PhiFlow/tests/test_poisson_solve.py:26-35 on the reference's two fixed right-hand sides (4 x 5, CLOSED and OPEN, the two domain
    types on this path): laplace(solve(rhs)) == rhs to 5 decimals; :67-73 two solvers agree to 4 decimals (here CG and the direct
    solve)."""
    rhs = torch.tensor(_poisson_examples())
    p, it = F.conjugate_gradient(rhs, bd, 1e-6, 200)
    np.testing.assert_almost_equal(F.laplace_apply(p, bd).numpy(), rhs.numpy(), decimal=5)
    direct = F.pressure_solve(rhs, bd)
    a, b = p.numpy(), direct.numpy()
    if bd == F.CLOSED:
        a, b = a - a.mean((1, 2), keepdims=True), b - b.mean((1, 2), keepdims=True)
    np.testing.assert_almost_equal(a, b, decimal=4)


@pytest.mark.gpu
@pytest.mark.parametrize("bd", ["closed", "open"])
def test_fixed_poisson_examples_cuda(bd):
    """The same two right-hand sides through poisson_solve / SparseCG on the GPU (one CTA per element on a 4 x 5 grid)."""
    from pde_control_b200.physics import CLOSED, OPEN, CenteredGrid, Domain, SparseCG, poisson_solve
    rhs = _poisson_examples()
    domain = Domain([4, 5], boundaries=CLOSED if bd == "closed" else OPEN)
    field = CenteredGrid(torch.tensor(rhs[..., None], device="cuda"), box=domain.box)
    p, it = poisson_solve(field, domain, solver=SparseCG(accuracy=1e-6, max_iterations=200))
    got = p.data[..., 0].cpu()
    np.testing.assert_almost_equal(F.laplace_apply(got, bd).numpy(), rhs, decimal=5)
    want, _ = F.conjugate_gradient(torch.tensor(rhs), bd, 1e-6, 200)
    a, b = got.numpy(), want.numpy()
    if bd == "closed":
        a, b = a - a.mean((1, 2), keepdims=True), b - b.mean((1, 2), keepdims=True)
    np.testing.assert_almost_equal(a, b, decimal=4)
    assert 0 < int(torch.as_tensor(it).max()) <= 200


@pytest.mark.gpu
def test_zero_right_hand_side_cuda():
    """divide_no_nan in the CUDA CG kernels: a zero right-hand side returns a zero pressure, zero iterations, no NaN, on the on-chip,
    masked and streaming kernels."""
    from pde_control_b200 import ops
    for ny, nx in ((4, 5), (128, 128), (150, 140)):
        p, iters, resid = ops.pressure_solve(torch.zeros(2, ny, nx, device="cuda"), False, accuracy=1e-5, max_iterations=50)
        assert bool(torch.isfinite(p).all()) and float(p.abs().max()) == 0.0 and int(iters.max()) == 0
    masks = ops.ObstacleMasks(torch.ones(24, 32, device="cuda"), True)
    out = torch.full((2, 24, 32), float("nan"), device="cuda")
    ops.cg_poisson_masked_raw(torch.zeros(2, 24, 32, device="cuda"), out, masks.code, 1e-5, 50)
    assert float(out.abs().max()) == 0.0


@pytest.mark.gpu
def test_inner_interpolation_cuda():
    """test_fields.py:23-35 through CenteredGrid.at of another grid (generic path: sample_at -> pdec_resample_fwd)."""
    from pde_control_b200.physics import CenteredGrid, box
    data = torch.zeros(1, 2, 3, 1, device="cuda")
    data[0, :, :, 0] = torch.tensor([[1., 2., 3.], [4., 5., 6.]])
    f = CenteredGrid(data, box[0:2, 0:3])
    g = CenteredGrid(torch.zeros(1, 2, 2, 1, device="cuda"), box[0:2, 0.5:2.5])
    resampled = f.at(g)
    assert resampled.compatible(g)
    np.testing.assert_array_equal(resampled.data[0, ..., 0].cpu().numpy(), [[1.5, 2.5], [4.5, 5.5]])


@pytest.mark.gpu
@pytest.mark.parametrize("mode", [R.CONSTANT, R.REPLICATE, R.CIRCULAR])
@pytest.mark.parametrize("rank,shared", [(1, True), (1, False), (2, True), (2, False)])
def test_resample_operator_vs_oracle(mode, rank, shared):
    """pdec_resample_fwd / _bwd (the backend operator, tf_backend.py:145-154, 451-502) vs the oracle's restatement: values to 1e-6,
    gradients w.r.t. data and coordinates to 1e-5; points inside, on the last cell (index N - 1 exactly) and outside the grid."""
    from pde_control_b200 import ops
    from pde_control_b200._lib import MODE_CONSTANT, MODE_REPLICATE, MODE_CIRCULAR
    cmode = {R.CONSTANT: MODE_CONSTANT, R.REPLICATE: MODE_REPLICATE, R.CIRCULAR: MODE_CIRCULAR}[mode]
    rng = np.random.RandomState(7 + rank)
    B, S = 3, ((19,) if rank == 1 else (11, 14))
    data = rng.randn(B, *S).astype(np.float32)
    P = 400
    coords = np.stack([rng.uniform(-3.0, s + 2.0, size=(1 if shared else B, P)) for s in S], -1).astype(np.float32)
    coords[:, :4] = np.array([[0.0] * rank, [s - 1.0 for s in S], [s - 1.5 for s in S], [-0.5] * rank], np.float32)
    w = rng.randn(B, P).astype(np.float32)
    dc, cc = torch.tensor(data, requires_grad=True), torch.tensor(coords, requires_grad=True)
    want = R.resample(dc, cc, mode)
    (want * torch.tensor(w)).sum().backward()
    dg, cg = torch.tensor(data, device="cuda", requires_grad=True), torch.tensor(coords, device="cuda", requires_grad=True)
    got = ops.resample(dg, cg, cmode)
    (got * torch.tensor(w, device="cuda")).sum().backward()
    assert rel_err(got.detach().cpu(), want.detach()) < 1e-6
    assert rel_err(dg.grad.cpu(), dc.grad) < 1e-5
    assert rel_err(cg.grad.cpu(), cc.grad) < 1e-5
    # the data gradient is a scatter at arbitrary points: fixed-point integer atomics, so it is bit-reproducible (no float atomics);
    # many points on one cell, gradients spanning 12 orders of magnitude
    big = torch.tensor(np.tile(coords[:, :8], (1, 500, 1)), device="cuda")
    wbig = torch.tensor((rng.randn(B, big.shape[1]) * 10.0 ** rng.uniform(-6, 6, size=(B, big.shape[1]))).astype(np.float32), device="cuda")
    grads = []
    for _ in range(3):
        d2 = torch.tensor(data, device="cuda", requires_grad=True)
        (ops.resample(d2, big, cmode) * wbig).sum().backward()
        grads.append(d2.grad.clone())
    assert torch.equal(grads[0], grads[1]) and torch.equal(grads[0], grads[2])
    d3 = torch.tensor(data, requires_grad=True)
    (R.resample(d3.double(), big.cpu().double(), mode) * wbig.cpu().double()).sum().backward()
    assert rel_err(grads[0].cpu(), d3.grad) < 1e-5


@pytest.mark.gpu
def test_resample_boundary_table_cuda():
    """PhiFlow/tests/test_fields.py:92-101 restated for the single-mode cases on the 2 x 2 grid [[1, 2], [3, 4]] (box [0, 2]^2):
    centre of cell (0, 0) -> 1; the grid centre -> 2.5; a point outside -> 0 ('constant') or the edge value ('boundary')."""
    from pde_control_b200.physics import CenteredGrid, box
    data = torch.tensor([[1., 2.], [3., 4.]], device="cuda").reshape(1, 2, 2, 1)
    pts = torch.tensor([[[0.5, 0.5], [1.0, 1.0], [-0.5, 0.5], [1.5, 2.5]]], device="cuda")
    const = CenteredGrid(data, box[0:2, 0:2], extrapolation="constant").sample_at(pts)[0, :, 0].cpu().numpy()
    repl = CenteredGrid(data, box[0:2, 0:2], extrapolation="boundary").sample_at(pts)[0, :, 0].cpu().numpy()
    assert const[0] == 1.0 and repl[0] == 1.0
    assert abs(const[1] - 2.5) < 1e-6 and abs(repl[1] - 2.5) < 1e-6
    assert const[2] == 0.0 and repl[2] == 1.0
    assert const[3] == 0.0 and repl[3] == 4.0
