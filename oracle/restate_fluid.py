"""TEST INFRASTRUCTURE ONLY -- CPU restatement (PyTorch-CPU) of the reference's 2-D incompressible-flow hot
path (PhiFlow 1.4.0, TF-1.14 semantics): SURVEY rows F1-F8.  Autograd over this op graph is the gradient
oracle; the pressure solve carries the reference's registered gradient (a second solve, A symmetric:
PF/physics/pressuresolver/sparse.py:93-94).

Parity pinning: forwards validated against the UNMODIFIED reference NumPy backend (oracle/ref_shim.py) by
tests/test_oracle_fluid.py (build container) and against tests/golden/fluid_*.npz (generated from the real
reference by oracle/make_golden_fluid.py).  The pressure is pinned as the *solution* of A p = div (direct
sparse solve in fp64), not as a CG iterate sequence.

Layouts (channel dropped): rho[B,ny,nx], vy[B,ny+1,nx], vx[B,ny,nx+1]; axis 0 = y, axis 1 = x
(PF/physics/field/staggered_grid.py:33-39).  PF/ = /root/reference/PhiFlow/phi/.
"""
import numpy as np
import torch

from .restate import CONSTANT, REPLICATE, CIRCULAR, resample, pad_axis, l2_loss  # noqa: F401

OPEN, CLOSED = "open", "closed"


class Geometry:
    """Domain box + derived staggered component boxes (PF/physics/field/staggered_grid.py:49-53)."""

    def __init__(self, ny, nx, lower=(0.0, 0.0), size=None):
        self.res = (ny, nx)
        size = (float(ny), float(nx)) if size is None else size
        self.lower = np.asarray(lower, np.float32)
        self.size = np.asarray(size, np.float32)
        self.dx = self.size.astype(np.float64) / np.asarray(self.res, np.float64)     # grid.py:77-79

    def component(self, axis):
        """(resolution, lower, size) of staggered component `axis` as a CenteredGrid."""
        res = list(self.res)
        res[axis] += 1
        unit = np.zeros(2)
        unit[axis] = self.dx[axis]
        lower = (self.lower.astype(np.float64) - unit / 2).astype(np.float32)
        upper = (self.lower.astype(np.float64) + self.size.astype(np.float64) + unit / 2).astype(np.float32)
        return tuple(res), lower, (upper - lower).astype(np.float32)


def _points(res, lower, size, dtype):
    """CenteredGrid.getpoints (grid.py:176-181): float64 linspace cast to float32, then local_to_global."""
    out = []
    for d in range(2):
        loc = torch.from_numpy(np.linspace(0.5 / res[d], 1 - 0.5 / res[d], res[d]).astype(np.float32)).to(dtype)
        out.append(loc * torch.tensor(float(size[d]), dtype=dtype) + torch.tensor(float(lower[d]), dtype=dtype))
    return out  # [y coords (res0), x coords (res1)]


def _index_coords(x, lower, size, res, axis, dtype):
    """sample_at: local = (x - lower)/size (geom/_box.py:50-52); idx = local*res - 0.5 (grid.py:104-105)."""
    local = (x - torch.tensor(float(lower[axis]), dtype=dtype)) / torch.tensor(float(size[axis]), dtype=dtype)
    return local * torch.tensor(float(res[axis]), dtype=dtype) - 0.5


def _half_weights(self_lower, self_size, self_res, other_lower, axis, dtype):
    """CenteredGrid.at fast path (grid.py:112-120): upper_weight = origin_in_local % 1.0 where
    origin_in_local = global_to_local(other.lower) * resolution; lower_weight = 1 - upper_weight
    (math/nd.py:386).  Computed like the reference (float32 box arithmetic, float64 product)."""
    origin = (np.float32(other_lower[axis]) - np.float32(self_lower[axis])) / np.float32(self_size[axis])
    origin = np.float64(origin) * self_res[axis]
    wu = origin % 1.0
    return (torch.tensor(float(np.float32(wu)), dtype=dtype), torch.tensor(float(np.float32(1 - wu)), dtype=dtype))


def curl_2d(phi, mode=CONSTANT):
    """F3: staggered_curl_2d (PF/physics/field/util.py:60-67): phi padded [1,2]x[1,2] with its own mode, 3x3
    'valid' correlation -> vy[i,j] = phi[i,j]-phi[i,j+1] ([ny+1,nx]), vx[i,j] = phi[i+1,j]-phi[i,j] ([ny,nx+1])."""
    p = pad_axis(pad_axis(phi, 1, 1, 2, mode), 2, 1, 2, mode)       # p[a,b] = phi[a-1,b-1]
    ny, nx = phi.shape[1:]
    vy = p[:, 1:ny + 2, 1:nx + 1] - p[:, 1:ny + 2, 2:nx + 2]
    vx = p[:, 2:ny + 2, 1:nx + 2] - p[:, 1:ny + 1, 1:nx + 2]
    return vy, vx


def velocity_at_centers(vy, vx, geo):
    """StaggeredGrid.at(centered points) (staggered_grid.py:126-137) -> per component CenteredGrid.at fast path:
    crop + interpolate_linear along the staggered axis (math/nd.py:378-392): upper*wu + lower*wl."""
    dt_ = vy.dtype
    (_, ly, sy), (_, lx, sx) = geo.component(0), geo.component(1)
    ry, rx = geo.component(0)[0], geo.component(1)[0]
    wu, wl = _half_weights(ly, sy, ry, geo.lower, 0, dt_)
    cy = vy[:, 1:, :] * wu + vy[:, :-1, :] * wl
    wu, wl = _half_weights(lx, sx, rx, geo.lower, 1, dt_)
    cx = vx[:, :, 1:] * wu + vx[:, :, :-1] * wl
    return cy, cx


def advect_centered(rho, vy, vx, dt, geo, mode):
    """F1: semi_lagrangian(density, velocity, dt) (PF/physics/field/advect.py:24-37)."""
    dtp = rho.dtype
    ny, nx = geo.res
    py, px = _points(geo.res, geo.lower, geo.size, dtp)
    cy, cx = velocity_at_centers(vy, vx, geo)
    dt_ = torch.tensor(dt, dtype=dtp)
    y = py[None, :, None] - cy * dt_
    x = px[None, None, :] - cx * dt_
    iy = _index_coords(y, geo.lower, geo.size, geo.res, 0, dtp)
    ix = _index_coords(x, geo.lower, geo.size, geo.res, 1, dtp)
    return resample(rho, torch.stack([iy, ix], -1), mode)


def _other_component_at(v_other, other_axis, geo, mode, dtp):
    """Component `other_axis` brought to the face centres of the OTHER staggered component
    (grid.py:112-123): pad 1 cell along the target's axis with the field's mode, then crop +
    interpolate_linear over both axes (axis order y then x)."""
    target_axis = 1 - other_axis
    res_o, lo_o, sz_o = geo.component(other_axis)
    res_t, lo_t, sz_t = geo.component(target_axis)
    padded = pad_axis(v_other, 1 + target_axis, 1, 1, mode)
    # padded box: lower - dx along target axis
    lo_p = lo_o.astype(np.float64).copy()
    lo_p[target_axis] -= geo.dx[target_axis]
    sz_p = sz_o.astype(np.float64).copy()
    sz_p[target_axis] += 2 * geo.dx[target_axis]
    res_p = list(res_o)
    res_p[target_axis] += 2
    lo_p = lo_p.astype(np.float32)
    sz_p = sz_p.astype(np.float32)
    t = padded
    for axis in (0, 1):
        wu, wl = _half_weights(lo_p, sz_p, res_p, lo_t, axis, dtp)
        if axis == 0:
            t = t[:, 1:, :] * wu + t[:, :-1, :] * wl
        else:
            t = t[:, :, 1:] * wu + t[:, :, :-1] * wl
    return t


def advect_staggered(vy, vx, dt, geo, mode):
    """F2: semi_lagrangian(v, v, dt) on a StaggeredGrid (advect.py:38-40 recursion per component)."""
    dtp = vy.dtype
    dt_ = torch.tensor(dt, dtype=dtp)
    out = []
    for c, vc in ((0, vy), (1, vx)):
        res_c, lo_c, sz_c = geo.component(c)
        py, px = _points(res_c, lo_c, sz_c, dtp)
        other = _other_component_at(vx if c == 0 else vy, 1 - c, geo, mode, dtp)
        uy, ux = (vc, other) if c == 0 else (other, vc)
        y = py[None, :, None] - uy * dt_
        x = px[None, None, :] - ux * dt_
        iy = _index_coords(y, lo_c, sz_c, res_c, 0, dtp)
        ix = _index_coords(x, lo_c, sz_c, res_c, 1, dtp)
        out.append(resample(vc, torch.stack([iy, ix], -1), mode))
    return out[0], out[1]


def hard_boundary_mask(vy, vx, boundary):
    """F4: PoissonDomain.with_hard_boundary_conditions (solver_api.py:111-121), no obstacles: faces on a solid
    domain boundary are multiplied by 0, all others by 1."""
    if boundary == OPEN:
        return vy, vx
    my = torch.ones_like(vy)
    my[:, 0, :] = 0
    my[:, -1, :] = 0
    mx = torch.ones_like(vx)
    mx[:, :, 0] = 0
    mx[:, :, -1] = 0
    return vy * my, vx * mx


def divergence(vy, vx):
    """F4: StaggeredGrid.divergence(physical_units=False) (staggered_grid.py:201-209)."""
    return (vy[:, 1:, :] - vy[:, :-1, :]) + (vx[:, :, 1:] - vx[:, :, :-1])


def laplace_apply(p, boundary):
    """F5 matrix (sparse.py:114-154 / 157-221), no obstacles: off-diagonal 1 towards in-domain neighbours,
    diagonal -(number of accessible neighbours) clamped <= -1: OPEN -4; CLOSED -(in-domain neighbours)."""
    z = pad_axis(pad_axis(p, 1, 1, 1, CONSTANT), 2, 1, 1, CONSTANT)
    nb = z[:, 2:, 1:-1] + z[:, :-2, 1:-1] + z[:, 1:-1, 2:] + z[:, 1:-1, :-2]
    if boundary == OPEN:
        diag = torch.full_like(p[0], -4.0)
    else:
        o = pad_axis(pad_axis(torch.ones_like(p[:1]), 1, 1, 1, CONSTANT), 2, 1, 1, CONSTANT)
        diag = -(o[:, 2:, 1:-1] + o[:, :-2, 1:-1] + o[:, 1:-1, 2:] + o[:, 1:-1, :-2])[0]
        diag = torch.minimum(diag, torch.tensor(-1.0, dtype=p.dtype))
    return nb + diag[None] * p


def conjugate_gradient(div, boundary, accuracy=1e-5, max_iterations=2000, guess=None):
    """F5: conjugate_gradient with TF (immutable tensor) semantics, PF/math/blas.py:5-64.
    Returns (x, iterations).  Stop rule: max|r| over the WHOLE batch < accuracy."""
    k = div
    if guess is None:
        x = torch.zeros_like(k)
        m = k
    else:
        x = guess
        m = k - laplace_apply(x, boundary)
    r = m
    Am = laplace_apply(m, boundary)
    it = 0

    def nn(a, b):   # divide_no_nan
        return torch.where(b == 0, torch.zeros_like(a), a / torch.where(b == 0, torch.ones_like(b), b))

    while it < max_iterations and (accuracy is None or float(r.abs().max()) >= accuracy):
        tmp = (m * Am).sum((1, 2), keepdim=True)
        a = nn((m * r).sum((1, 2), keepdim=True), tmp)
        x = x + a * m
        r = r - a * Am
        m = r - nn((r * Am).sum((1, 2), keepdim=True) * m, tmp)
        Am = laplace_apply(m, boundary)
        it += 1
    return x, it


_LU = {}


def _direct_factor(ny, nx, boundary):
    """fp64 sparse LU of the restated matrix.  CLOSED is singular (constants): bordered with the mean-zero
    constraint, which is the solution CG converges to from a zero guess."""
    key = (ny, nx, boundary)
    if key not in _LU:
        import scipy.sparse as sp
        import scipy.sparse.linalg as spl
        N = ny * nx
        idx = np.arange(N).reshape(ny, nx)
        rows, cols, vals = [], [], []
        deg = np.zeros((ny, nx))
        for sl_a, sl_b in ((np.s_[1:, :], np.s_[:-1, :]), (np.s_[:, 1:], np.s_[:, :-1])):
            a, b = idx[sl_a].ravel(), idx[sl_b].ravel()
            rows += [a, b]
            cols += [b, a]
            vals += [np.ones(a.size), np.ones(a.size)]
            deg[sl_a] += 1
            deg[sl_b] += 1
        diag = -4.0 * np.ones(N) if boundary == OPEN else np.minimum(-deg.ravel(), -1.0)
        rows.append(idx.ravel())
        cols.append(idx.ravel())
        vals.append(diag)
        A = sp.csc_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(N, N))
        if boundary == CLOSED:
            one = sp.csc_matrix(np.ones((N, 1)))
            A = sp.bmat([[A, one], [one.T, None]], format="csc")
        _LU[key] = spl.splu(A)
    return _LU[key]


def _direct_solve_np(rhs, boundary):
    B, ny, nx = rhs.shape
    lu = _direct_factor(ny, nx, boundary)
    out = np.empty((B, ny, nx))
    for b in range(B):
        v = rhs[b].astype(np.float64).ravel()
        if boundary == CLOSED:
            v = np.concatenate([v, [0.0]])
        out[b] = lu.solve(v)[:ny * nx].reshape(ny, nx)
    return out


class _PressureSolve(torch.autograd.Function):
    """p = A^-1 div (converged CG), gradient dL/ddiv = A^-1 dL/dp (sparse.py:93-94, A symmetric)."""

    @staticmethod
    def forward(ctx, div, boundary):
        ctx.boundary = boundary
        return torch.from_numpy(_direct_solve_np(div.detach().numpy(), boundary)).to(div.dtype)

    @staticmethod
    def backward(ctx, g):
        return torch.from_numpy(_direct_solve_np(g.detach().numpy(), ctx.boundary)).to(g.dtype), None


def pressure_solve(div, boundary):
    return _PressureSolve.apply(div, boundary)


def subtract_pressure_gradient(vy, vx, p, geo, boundary):
    """F6: divergence_free tail (PF/physics/fluid.py:240-242): p *= dx[0]; StaggeredGrid.gradient with
    replicate padding (staggered_grid.py:219-231) -> boundary-face gradients are 0;
    v -= mask * grad."""
    dtp = p.dtype
    p = p * torch.tensor(float(geo.dx[0]), dtype=dtp)
    py = pad_axis(p, 1, 1, 1, REPLICATE)
    px = pad_axis(p, 2, 1, 1, REPLICATE)
    gy = (py[:, 1:, :] - py[:, :-1, :]) / torch.tensor(float(geo.dx[0]), dtype=dtp)
    gx = (px[:, :, 1:] - px[:, :, :-1]) / torch.tensor(float(geo.dx[1]), dtype=dtp)
    gy, gx = hard_boundary_mask(gy, gx, boundary)
    return vy - gy, vx - gx


def divergence_free(vy, vx, geo, boundary, solver=pressure_solve):
    """F4-F6: divergence_free (PF/physics/fluid.py:214-243), no obstacles."""
    vy, vx = hard_boundary_mask(vy, vx, boundary)
    div = divergence(vy, vx)
    p = solver(div, boundary)
    vy2, vx2 = subtract_pressure_gradient(vy, vx, p, geo, boundary)
    return vy2, vx2, p, div


def normalize_to(target, source, epsilon=1e-5):
    """CenteredGrid.normalized -> math.normalize_to (PF/math/nd.py:28-40)."""
    tt = target.sum((1, 2), keepdim=True)
    den = torch.maximum(tt, torch.tensor(epsilon, dtype=target.dtype))
    st = source.sum((1, 2), keepdim=True)
    return target * (st / den)


def modes_of(boundary):
    """(density mode, velocity mode): Material.extrapolation_mode / vector_extrapolation_mode
    (PF/physics/material.py:66-98): OPEN -> constant/constant; CLOSED (friction 0) -> boundary/boundary."""
    return (CONSTANT, CONSTANT) if boundary == OPEN else (REPLICATE, REPLICATE)


def incompressible_step(rho, vy, vx, dt, geo, boundary, solver=pressure_solve):
    """F8: IncompressibleFlow.step (PF/physics/fluid.py:90-114) with default flags, no obstacles, no effects,
    buoyancy_factor 0."""
    rmode, vmode = modes_of(boundary)
    rho2 = advect_centered(rho, vy, vx, dt, geo, rmode)
    ay, ax = advect_staggered(vy, vx, dt, geo, vmode)
    if boundary == CLOSED:                       # conserve_density and all boundaries solid (fluid.py:100-101)
        rho2 = normalize_to(rho2, rho)
    vy2, vx2, p, div = divergence_free(ay, ax, geo, boundary, solver)
    return rho2, vy2, vx2, p


def controlled_step(rho, phi, dt, geo, rho_mode=CONSTANT, phi_mode=CONSTANT):
    """Notebook variant: ControlledIncompressibleFlow.step (src/control/pde/incompressible_flow.py:53-59):
    v = staggered_curl_2d(potential); density advected by v.  Returns (rho', vy, vx)."""
    vy, vx = curl_2d(phi, phi_mode)
    return advect_centered(rho, vy, vx, dt, geo, rho_mode), vy, vx


def force_loss(vys, vxs, dt, geo, mode=REPLICATE):
    """IncompressibleFluidPDE.total_force_loss (src/control/pde/incompressible_flow.py:31-43): mean over
    consecutive pairs of l2_loss(semi_lagrangian(v1, v1) - v2).  l2_loss of a StaggeredGrid difference =
    l2 over the staggered tensor (sum over both components)."""
    loss = 0
    n = len(vys) - 1
    for t in range(n):
        ay, ax = advect_staggered(vys[t], vxs[t], dt, geo, mode)
        loss = loss + l2_loss(ay - vys[t + 1]) + l2_loss(ax - vxs[t + 1])
    return loss / n


# ---- obstacles and buoyancy (restated for the CPU checker; test infrastructure only) -------------------------

def obstacle_active(ny, nx, boxes, lower=(0.0, 0.0), size=None):
    """active = 1 - union of the obstacle boxes sampled at the cell centres (fluid.py:228-233; _box.py:58-61 bounds included)."""
    size = size or (float(ny), float(nx))
    cy = (np.arange(ny) + 0.5) / ny * size[0] + lower[0]
    cx = (np.arange(nx) + 0.5) / nx * size[1] + lower[1]
    solid = np.zeros((ny, nx), bool)
    for (y0, y1, x0, x1) in boxes:
        solid |= ((cy >= y0) & (cy <= y1))[:, None] & ((cx >= x0) & (cx <= x1))[None, :]
    return (~solid).astype(np.float64)


def masked_operator(active, boundary):
    """Dense matrix of sparse.py:157-221 for a [ny, nx] active mask: off-diagonal active_self * active_nb (0 outside the domain),
    diagonal -(sum of accessible neighbours; outside: 1 on open, 0 on closed domains: solver_api.py:98-109) clamped to <= -1;
    and the face masks of _frictionless_velocity_mask (:111-121: accessible grid extended by its own extrapolation)."""
    ny, nx = active.shape
    out = 0.0 if boundary == CLOSED else 1.0
    eact = np.pad(active, 1, constant_values=0.0)
    eacc = np.pad(active, 1, constant_values=out)
    N = ny * nx
    A = np.zeros((N, N))
    for i in range(ny):
        for j in range(nx):
            k = i * nx + j
            d = 0.0
            for di, dj in ((-1, 0), (1, 0), (0, -1), (0, 1)):
                d -= eacc[i + 1 + di, j + 1 + dj]
                if 0 <= i + di < ny and 0 <= j + dj < nx:
                    A[k, (i + di) * nx + (j + dj)] = eact[i + 1 + di, j + 1 + dj] * active[i, j]
            A[k, k] = min(d, -1.0)
    mode = "constant" if boundary == CLOSED else "edge"
    py = np.pad(active, ((1, 1), (0, 0)), mode=mode)
    px = np.pad(active, ((0, 0), (1, 1)), mode=mode)
    return A, np.minimum(py[:-1], py[1:]), np.minimum(px[:, :-1], px[:, 1:])


def divergence_free_obstacles(vy, vx, geo, boundary, active):
    """divergence_free (fluid.py:214-243) with obstacles; direct dense solve in float64 (least squares: the closed matrix is singular)."""
    A, fy, fx = masked_operator(active, boundary)
    dtp = vy.dtype
    fy_t, fx_t = torch.tensor(fy, dtype=dtp), torch.tensor(fx, dtype=dtp)
    my, mx = vy * fy_t, vx * fx_t
    div = divergence(my, mx)
    B, ny, nx = div.shape
    if boundary == OPEN:        # non-singular: differentiable dense solve (gradient oracle for the masked projection)
        p = torch.linalg.solve(torch.tensor(A), div.double().reshape(B, -1, 1)).reshape(B, ny, nx).to(dtp)
    else:
        sol = np.linalg.lstsq(A, div.detach().double().numpy().reshape(B, -1).T, rcond=None)[0].T.reshape(B, ny, nx)
        p = torch.tensor(sol, dtype=dtp)
    p_s = p * torch.tensor(float(geo.dx[0]), dtype=dtp)
    py = pad_axis(p_s, 1, 1, 1, REPLICATE)
    px = pad_axis(p_s, 2, 1, 1, REPLICATE)
    gy = (py[:, 1:, :] - py[:, :-1, :]) / torch.tensor(float(geo.dx[0]), dtype=dtp)
    gx = (px[:, :, 1:] - px[:, :, :-1]) / torch.tensor(float(geo.dx[1]), dtype=dtp)
    return my - gy * fy_t, mx - gx * fx_t, p, div


def centered_to_faces(rho, mode):
    """CenteredGrid.at(StaggeredGrid) on the fast path (grid.py:112-120): pad one cell with the field's extrapolation, average
    adjacent cells -> ([B, ny+1, nx], [B, ny, nx+1])."""
    py = pad_axis(rho, 1, 1, 1, mode)
    px = pad_axis(rho, 2, 1, 1, mode)
    return py[:, :-1, :] * 0.5 + py[:, 1:, :] * 0.5, px[:, :, :-1] * 0.5 + px[:, :, 1:] * 0.5


def incompressible_step_general(rho, vy, vx, dt, geo, boundary, buoyancy=0.0, gravity=-9.81, density_source=None, active=None,
                                solver=pressure_solve):
    """IncompressibleFlow.step (fluid.py:90-114) with a GROW density effect, buoyancy (gravity along -axis 0) and obstacles."""
    rmode, vmode = modes_of(boundary)
    rho2 = advect_centered(rho, vy, vx, dt, geo, rmode)
    ay, ax = advect_staggered(vy, vx, dt, geo, vmode)
    if boundary == CLOSED:
        rho2 = normalize_to(rho2, rho)
    if density_source is not None:
        rho2 = rho2 + density_source * dt                       # effect_applied GROW (effect.py:48-51)
    if buoyancy != 0.0:
        # fluid.py:107, up = axis 0.  The two-component lift grid is unstacked before resampling and CenteredGrid.unstack
        # (grid.py:126-129) drops the extrapolation: the components are padded with the default 'boundary' (replicate)
        ly, _lx = centered_to_faces(rho2 * (-gravity * buoyancy * dt), REPLICATE)
        ay = ay + ly
    if active is not None:
        vy2, vx2, p, div = divergence_free_obstacles(ay, ax, geo, boundary, active)
    else:
        vy2, vx2, p, div = divergence_free(ay, ax, geo, boundary, solver)
    return rho2, vy2, vx2, p


def fourier_poisson(rhs):
    """FourierSolver (fourier.py:17-18) = math.fourier_poisson (nd.py:305-314): spectrum / (-(2 pi)^2 |k|^2), k = fftfreq in cycles
    per cell, the k = 0 mode set to 0.  rhs [B, ny, nx]."""
    ny, nx = rhs.shape[1:]
    ky = torch.fft.fftfreq(ny, dtype=torch.float64)[:, None]
    kx = torch.fft.fftfreq(nx, dtype=torch.float64)[None, :]
    lap = -(2 * np.pi) ** 2 * (ky ** 2 + kx ** 2).to(torch.float32).to(torch.float64)       # fftfreq(..., dtype=np.float32)
    inv = torch.where(lap == 0, torch.zeros_like(lap), 1.0 / torch.where(lap == 0, torch.ones_like(lap), lap))
    return torch.fft.ifft2(torch.fft.fft2(rhs.to(torch.complex128)) * inv).real.to(rhs.dtype)


# ---- geometry masks and effects (Inflow, Accelerator / Fan, FIX effects, sphere / union obstacles) ----
def lies_inside(points, geometry):
    """Geometry.lies_inside for a spec tuple: ("box", (y0, x0), (y1, x1)) bounds included in float32 (_box.py:58-61);
    ("sphere", (cy, cx), r) |x - c|^2 <= r^2 in float64 (_sphere.py:24-28); ("union", spec, ...) (_union.py:24-25).
    points [ny, nx, 2] float32 -> bool [ny, nx]."""
    kind = geometry[0]
    if kind == "box":
        lo, hi = np.asarray(geometry[1], np.float32), np.asarray(geometry[2], np.float32)
        return np.all((points >= lo) & (points <= hi), axis=-1)
    if kind == "sphere":
        return np.sum((points - np.asarray(geometry[1])) ** 2, axis=-1) <= geometry[2] ** 2
    if kind == "union":
        out = lies_inside(points, geometry[1])
        for g in geometry[2:]:
            out = out | lies_inside(points, g)
        return out
    raise ValueError(kind)


def sample_points(res, lower, size):
    """float32 [res0, res1, 2] sample points of a centred grid (grid.py:176-181)."""
    ys, xs = _points(res, lower, size, torch.float32)
    return np.stack(np.meshgrid(ys.numpy(), xs.numpy(), indexing="ij"), -1)


def geometry_mask(geometry, geo, axis=None):
    """GeometryMask.at(grid) (mask.py:45-55): 1 inside, 0 outside, at the cell centres (axis None) or at the faces of staggered
    component `axis`."""
    res, lower, size = (geo.res, geo.lower, geo.size) if axis is None else geo.component(axis)
    return lies_inside(sample_points(res, lower, size), geometry).astype(np.float64)


def apply_effect(arrays, effect, dt, geo, staggered):
    """effect_applied (effect.py:48-59).  effect = dict(mode 'grow' | 'add' | 'fix', geometry spec or None (constant field),
    value: number or one per component, bounds: spec for 'fix').  arrays: [rho] or [vy, vx]."""
    out = []
    for d, a in enumerate(arrays):
        axis = d if staggered else None
        value = effect["value"]
        value = float(value[d]) if np.ndim(value) > 0 else float(value)
        field = value if effect.get("geometry") is None else torch.tensor(geometry_mask(effect["geometry"], geo, axis), dtype=a.dtype) * value
        if effect["mode"] == "grow":
            out.append(a + field * dt)
        elif effect["mode"] == "add":
            out.append(a + field)
        else:
            inside = torch.tensor(geometry_mask(effect["bounds"], geo, axis) > 0)
            out.append(torch.where(inside, torch.as_tensor(field, dtype=a.dtype).expand_as(a) if not torch.is_tensor(field)
                                   else field.expand_as(a), a))
    return out


def incompressible_step_effects(rho, vy, vx, dt, geo, boundary, buoyancy=0.0, gravity=-9.81, density_effects=(), velocity_effects=(),
                                obstacle=None):
    """IncompressibleFlow.step (fluid.py:90-114): advect, normalise (closed), density effects, velocity effects, buoyancy, projection
    with the obstacle geometry's active mask."""
    rmode, vmode = modes_of(boundary)
    rho2 = advect_centered(rho, vy, vx, dt, geo, rmode)
    ay, ax = advect_staggered(vy, vx, dt, geo, vmode)
    if boundary == CLOSED:
        rho2 = normalize_to(rho2, rho)
    for e in density_effects:
        rho2, = apply_effect([rho2], e, dt, geo, False)
    for e in velocity_effects:
        ay, ax = apply_effect([ay, ax], e, dt, geo, True)
    if buoyancy != 0.0:
        ly, _lx = centered_to_faces(rho2 * (-gravity * buoyancy * dt), REPLICATE)
        ay = ay + ly
    if obstacle is not None:
        active = 1.0 - geometry_mask(obstacle, geo)
        vy2, vx2, p, div = divergence_free_obstacles(ay, ax, geo, boundary, active)
    else:
        vy2, vx2, p, div = divergence_free(ay, ax, geo, boundary)
    return rho2, vy2, vx2, p


# ---- SURVEY 8f N3 remainder: NO_SLIP and per-side mixed boundaries (test infrastructure only) --------------------------------
# A Domain's boundaries may differ per side (PF/physics/domain.py:13-35).  Fields then carry a per-side extrapolation and
#   * sampling goes through CenteredGrid._padded_resample (grid.py:211-216): pad ONE cell on every side with the side's mode, resample
#     the padded array with 'replicate' at coordinates shifted by one cell;
#   * paddings run in passes (backend_helper.py:22-36): every 'replicate' side first, then every 'constant' side;
#   * solid sides zero their boundary faces (solver_api.py:111-121) and do not count as accessible neighbours (sparse.py:186-213).
# `sides` below = [(lower, upper) per axis] of CONSTANT / REPLICATE; `solid` = (y-lower, y-upper, x-lower, x-upper) booleans.

def sides_of(solid, friction=0):
    """(density sides, velocity sides) from the per-side solid flags: Material.extrapolation_mode / vector_extrapolation_mode."""
    rho = [(REPLICATE if solid[0] else CONSTANT, REPLICATE if solid[1] else CONSTANT),
           (REPLICATE if solid[2] else CONSTANT, REPLICATE if solid[3] else CONSTANT)]
    vel = [tuple(REPLICATE if (m == REPLICATE and friction == 0) else CONSTANT for m in pair) for pair in rho]
    return rho, vel


def _uniform(sides):
    modes = {m for pair in sides for m in pair}
    return modes.pop() if len(modes) == 1 else None


def pad_sides(t, widths, sides):
    """math.pad of a [B, ny, nx] array with per-side modes, in the reference's pass order."""
    for single in (REPLICATE, CONSTANT):
        for axis in (0, 1):
            lo = widths[axis][0] if sides[axis][0] == single else 0
            hi = widths[axis][1] if sides[axis][1] == single else 0
            if lo or hi:
                t = pad_axis(t, 1 + axis, lo, hi, single)
    return t


def sample_sides(field, iy, ix, sides):
    """CenteredGrid.sample_at at index coordinates (iy, ix): one mode -> the backend's resample; per-side -> _padded_resample."""
    mode = _uniform(sides)
    if mode is not None:
        return resample(field, torch.stack([iy, ix], -1), mode)
    padded = pad_sides(field, [[1, 1], [1, 1]], sides)
    return resample(padded, torch.stack([iy + 1.0, ix + 1.0], -1), REPLICATE)


def _other_component_at_sides(v_other, other_axis, geo, sides, dtp):
    """_other_component_at with the component padded per side along the target's axis."""
    target_axis = 1 - other_axis
    res_o, lo_o, sz_o = geo.component(other_axis)
    _res_t, lo_t, _sz_t = geo.component(target_axis)
    widths = [[0, 0], [0, 0]]
    widths[target_axis] = [1, 1]
    padded = pad_sides(v_other, widths, sides)
    lo_p = lo_o.astype(np.float64).copy()
    lo_p[target_axis] -= geo.dx[target_axis]
    sz_p = sz_o.astype(np.float64).copy()
    sz_p[target_axis] += 2 * geo.dx[target_axis]
    res_p = list(res_o)
    res_p[target_axis] += 2
    lo_p, sz_p = lo_p.astype(np.float32), sz_p.astype(np.float32)
    t = padded
    for axis in (0, 1):
        wu, wl = _half_weights(lo_p, sz_p, res_p, lo_t, axis, dtp)
        t = (t[:, 1:, :] * wu + t[:, :-1, :] * wl) if axis == 0 else (t[:, :, 1:] * wu + t[:, :, :-1] * wl)
    return t


def advect_sides(rho, vy, vx, dt, geo, rho_sides, v_sides):
    """F1 + F2 with per-side extrapolations -> (rho', advected vy, advected vx)."""
    dtp = rho.dtype
    dt_ = torch.tensor(dt, dtype=dtp)
    py, px = _points(geo.res, geo.lower, geo.size, dtp)
    cy, cx = velocity_at_centers(vy, vx, geo)
    iy = _index_coords(py[None, :, None] - cy * dt_, geo.lower, geo.size, geo.res, 0, dtp)
    ix = _index_coords(px[None, None, :] - cx * dt_, geo.lower, geo.size, geo.res, 1, dtp)
    rho2 = sample_sides(rho, iy, ix, rho_sides)
    out = []
    for c, vc in ((0, vy), (1, vx)):
        res_c, lo_c, sz_c = geo.component(c)
        qy, qx = _points(res_c, lo_c, sz_c, dtp)
        other = _other_component_at_sides(vx if c == 0 else vy, 1 - c, geo, v_sides, dtp)
        uy, ux = (vc, other) if c == 0 else (other, vc)
        iy = _index_coords(qy[None, :, None] - uy * dt_, lo_c, sz_c, res_c, 0, dtp)
        ix = _index_coords(qx[None, None, :] - ux * dt_, lo_c, sz_c, res_c, 1, dtp)
        out.append(sample_sides(vc, iy, ix, v_sides))
    return rho2, out[0], out[1]


_LU_SIDES = {}


def _solve_sides_np(rhs, solid):
    """fp64 sparse LU of the matrix with per-side boundary rules; singular (constants) only if every side is solid."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    B, ny, nx = rhs.shape
    key = (ny, nx, tuple(bool(s) for s in solid))
    singular = all(solid)
    if key not in _LU_SIDES:
        N = ny * nx
        idx = np.arange(N).reshape(ny, nx)
        rows, cols, vals = [], [], []
        acc = np.zeros((ny, nx))
        for sl_a, sl_b in ((np.s_[1:, :], np.s_[:-1, :]), (np.s_[:, 1:], np.s_[:, :-1])):
            a, b = idx[sl_a].ravel(), idx[sl_b].ravel()
            rows += [a, b]
            cols += [b, a]
            vals += [np.ones(a.size), np.ones(a.size)]
            acc[sl_a] += 1
            acc[sl_b] += 1
        # the outside counts as an accessible neighbour on OPEN sides
        acc[0, :] += 0 if solid[0] else 1
        acc[-1, :] += 0 if solid[1] else 1
        acc[:, 0] += 0 if solid[2] else 1
        acc[:, -1] += 0 if solid[3] else 1
        rows.append(idx.ravel())
        cols.append(idx.ravel())
        vals.append(np.minimum(-acc.ravel(), -1.0))
        A = sp.csc_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(N, N))
        if singular:
            one = sp.csc_matrix(np.ones((N, 1)))
            A = sp.bmat([[A, one], [one.T, None]], format="csc")
        _LU_SIDES[key] = spl.splu(A)
    lu = _LU_SIDES[key]
    out = np.empty((B, ny, nx))
    for b in range(B):
        v = rhs[b].astype(np.float64).ravel()
        if singular:
            v = np.concatenate([v, [0.0]])
        out[b] = lu.solve(v)[:ny * nx].reshape(ny, nx)
    return out


class _PressureSolveSides(torch.autograd.Function):
    @staticmethod
    def forward(ctx, div, solid):
        ctx.solid = solid
        return torch.from_numpy(_solve_sides_np(div.detach().numpy(), solid)).to(div.dtype)

    @staticmethod
    def backward(ctx, g):
        return torch.from_numpy(_solve_sides_np(g.detach().numpy(), ctx.solid)).to(g.dtype), None


def divergence_free_sides(vy, vx, geo, solid):
    my, mx = torch.ones_like(vy), torch.ones_like(vx)
    if solid[0]:
        my[:, 0, :] = 0
    if solid[1]:
        my[:, -1, :] = 0
    if solid[2]:
        mx[:, :, 0] = 0
    if solid[3]:
        mx[:, :, -1] = 0
    vy, vx = vy * my, vx * mx
    div = divergence(vy, vx)
    p = _PressureSolveSides.apply(div, tuple(bool(s) for s in solid))
    dtp = p.dtype
    q = p * torch.tensor(float(geo.dx[0]), dtype=dtp)
    qy, qx = pad_axis(q, 1, 1, 1, REPLICATE), pad_axis(q, 2, 1, 1, REPLICATE)
    gy = (qy[:, 1:, :] - qy[:, :-1, :]) / torch.tensor(float(geo.dx[0]), dtype=dtp)
    gx = (qx[:, :, 1:] - qx[:, :, :-1]) / torch.tensor(float(geo.dx[1]), dtype=dtp)
    return vy - gy * my, vx - gx * mx, p, div


def incompressible_step_sides(rho, vy, vx, dt, geo, solid, friction=0, buoyancy=0.0, gravity=-9.81, conserve_density=True):
    """IncompressibleFlow.step (fluid.py:90-114) on a domain whose sides are solid per `solid` (NO_SLIP: all solid, friction 1)."""
    rho_sides, v_sides = sides_of(solid, friction)
    rho2, ay, ax = advect_sides(rho, vy, vx, dt, geo, rho_sides, v_sides)
    if conserve_density and all(solid):
        rho2 = normalize_to(rho2, rho)
    if buoyancy != 0.0:
        ly, _lx = centered_to_faces(rho2 * (-gravity * buoyancy * dt), REPLICATE)
        ay = ay + ly
    vy2, vx2, p, _div = divergence_free_sides(ay, ax, geo, solid)
    return rho2, vy2, vx2, p
