"""CPU checks of the closed-form eigenpairs the direct Poisson solvers are built on (csrc/poisson_direct.cu eig_vec / eig_val and their
torch twin ops._eig_1d): for every combination of open / solid ends, E is orthonormal and E^T T E = diag(lambda) for the 1-D operator
T = tridiag(1, d, 1), d = -2 with -1 at a solid end (PF/physics/pressuresolver/sparse.py:151-221 restricted to one axis), and the 2-D
fast-diagonalisation formula solves the restated 5-point system of oracle.restate_fluid."""
import numpy as np
import pytest
import torch

from oracle import restate_fluid as F


def _T(n, ends):
    T = np.zeros((n, n))
    for i in range(n):
        T[i, i] = -2.0
        if i > 0:
            T[i, i - 1] = 1.0
        if i < n - 1:
            T[i, i + 1] = 1.0
    if ends & 1:
        T[0, 0] = -1.0
    if ends & 2:
        T[n - 1, n - 1] = -1.0
    return T


@pytest.mark.parametrize("ends", [0, 1, 2, 3])
@pytest.mark.parametrize("n", [2, 5, 37, 128, 300])
def test_closed_form_eigenpairs_diagonalise_the_1d_operator(n, ends):
    from pde_control_b200 import ops
    E, lam = ops._eig_1d(n, ends, torch.device("cpu"))
    E, lam = E.numpy(), lam.numpy()
    assert np.abs(E.T @ E - np.eye(n)).max() < 1e-12
    assert np.abs(E.T @ _T(n, ends) @ E - np.diag(lam)).max() < 1e-11


@pytest.mark.parametrize("bd,ends", [("open", 0), ("closed", 3)])
def test_fast_diagonalisation_solves_the_restated_poisson_system(bd, ends):
    from pde_control_b200 import ops
    ny, nx = 24, 17
    rng = np.random.RandomState(1)
    rhs = rng.randn(2, ny, nx)
    if bd == "closed":
        rhs -= rhs.mean((1, 2), keepdims=True)
    Ey, ly = ops._eig_1d(ny, ends, torch.device("cpu"))
    Ex, lx = ops._eig_1d(nx, ends, torch.device("cpu"))
    lam = ly[:, None] + lx[None, :]
    inv = torch.where(lam.abs() < 1e-9, torch.zeros_like(lam), 1.0 / lam)
    p = Ey @ ((Ey.T @ torch.tensor(rhs) @ Ex) * inv) @ Ex.T
    assert float((F.laplace_apply(p, bd) - torch.tensor(rhs)).abs().max()) < 1e-10
