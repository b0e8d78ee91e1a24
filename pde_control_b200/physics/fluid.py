"""Fluid state, IncompressibleFlow physics, divergence_free and the pressure-solver objects over the CUDA kernels.

Mirrors  Fluid               PF/physics/fluid.py:21-60
         IncompressibleFlow  PF/physics/fluid.py:73-114   (no obstacles; buoyancy through Gravity, PF/physics/field/effect.py:77-112)
         divergence_free     PF/physics/fluid.py:214-243
         SparseCG            PF/physics/pressuresolver/sparse.py:40-102  (parameters only; the solve is pdec_cg_poisson_2d)
         poisson_solve       PF/physics/pressuresolver/solver_api.py:132-158
         advect.semi_lagrangian, staggered_curl_2d  PF/physics/field/advect.py:24-40, util.py:60-67
"""
import numpy as np
import torch

from .. import ops
from .._lib import MODE_OF_EXTRAPOLATION
from .burgers import DomainState
from .domain import Domain, OPEN, MixedBoundaries
from .field import CenteredGrid, StaggeredGrid, effect_applied
from .world import Physics, State, StateDependency, Gravity


class _Chainable(object):
    def __and__(self, other):
        return _SolverChain((self, other))


class SparseCG(_Chainable):
    """Conjugate-gradient pressure solver parameters (sparse.py:42-71).  'mirror' gradient iterations are treated as
    'same' (the device decides the iteration count)."""

    def __init__(self, accuracy=1e-5, gradient_accuracy="same", max_iterations=2000, max_gradient_iterations="same", autodiff=False,
                 stop_mode="per_element"):
        """stop_mode: 'per_element' (default here: every batch element stops on ITS max|r| < accuracy -- on chip, and batch shards on
        different GPUs then behave like the whole batch) or 'whole_batch' (the reference's rule, PF/math/blas.py:33-37: all elements
        iterate until the batch-wide max|r| < accuracy; runs on the streaming cooperative kernel for B > 1).  The SOLUTION is the parity
        oracle either way (converged elements only move below the tolerance); iteration counts differ."""
        assert not autodiff, "autodiff through the CG loop is not provided: the gradient is a second solve (sparse.py:93-94)"
        assert stop_mode in ("per_element", "whole_batch")
        self.stop_mode = ops.CG_STOP_WHOLE_BATCH if stop_mode == "whole_batch" else ops.CG_STOP_PER_ELEMENT
        self.accuracy = accuracy
        self.gradient_accuracy = accuracy if gradient_accuracy == "same" else gradient_accuracy
        self.max_iterations = max_iterations
        self.max_gradient_iterations = max_iterations if max_gradient_iterations in ("same", "mirror") else max_gradient_iterations

    def solve(self, field, domain, guess=None):
        """field [B, ny, nx, 1] or [B, ny, nx] -> (pressure of the same shape, iterations[B])."""
        t = field[..., 0] if field.dim() == 4 else field
        g = None if guess is None else (guess[..., 0] if guess.dim() == 4 else guess)
        p, iters, _ = ops.pressure_solve(t, _closed(domain), self.accuracy, self.max_iterations, self.gradient_accuracy,
                                         self.max_gradient_iterations, g, self.stop_mode)
        return (p.unsqueeze(-1) if field.dim() == 4 else p), iters


class SparseSciPy(_Chainable):
    """The reference's DIRECT solver class (pressuresolver/sparse.py:14-37: scipy spsolve of the sparse matrix; its default on the NumPy
    backend, solver_api.py:144-148).  Here: the exact fast-diagonalisation solve on the GPU for rectangular domains without obstacles
    -- on the tensor cores on chip up to 128 x 128 (pdec_poisson_direct_tc_2d), four library GEMMs beyond (ops.poisson_direct_large_raw;
    2048 x 2048 in 1.3 ms) -- no iteration count, no accuracy knob; the gradient is a second exact solve.  Obstacles: use SparseCG."""

    accuracy = 0.0
    max_iterations = 0

    def solve(self, field, domain, guess=None):
        t = field[..., 0] if field.dim() == 4 else field
        if not ops.direct_solver_supported(t.shape[1], t.shape[2]):
            raise NotImplementedError("SparseSciPy (direct solve): grid beyond ops.DIRECT_LARGE_MAX; use SparseCG")
        p = ops.pressure_solve_direct(t, _closed(domain))
        return (p.unsqueeze(-1) if field.dim() == 4 else p), None


class _SolverChain(object):
    """solver1 & solver2 (solver_api.py:161-174): each solver's result is the next one's initial guess.  The reference's default
    on the TensorFlow backend is FourierSolver() & SparseCG() (solver_api.py:149-155)."""

    def __init__(self, solvers):
        self.solvers = tuple(solvers)

    def __and__(self, other):
        return _SolverChain(self.solvers + (other,))

    def solve(self, field, domain, guess=None):
        iterations = None
        for solver in self.solvers:
            guess, iterations = solver.solve(field, domain, guess)
        return guess, iterations


class FourierSolver(object):
    """Inverse Laplace in Fourier space (pressuresolver/fourier.py:5-18 -> math.fourier_poisson, nd.py:305-314): exact for periodic
    domains, an initial guess for the others.  cuFFT through torch.fft, as the reference goes through its backend's fft."""

    def __and__(self, other):
        return _SolverChain((self, other))

    def solve(self, field, domain, guess=None):
        t = field[..., 0] if field.dim() == 4 else field
        ny, nx = t.shape[1:]
        ky = torch.fft.fftfreq(ny, device=t.device, dtype=torch.float32)[:, None]
        kx = torch.fft.fftfreq(nx, device=t.device, dtype=torch.float32)[None, :]
        lap = -(2.0 * np.pi) ** 2 * (ky * ky + kx * kx)
        inv = torch.where(lap == 0, torch.zeros_like(lap), 1.0 / torch.where(lap == 0, torch.ones_like(lap), lap))
        p = torch.fft.ifft2(torch.fft.fft2(t.to(torch.complex64)) * inv).real.contiguous()
        return (p.unsqueeze(-1) if field.dim() == 4 else p), None


def _closed(domain):
    domain = domain.domain if hasattr(domain, "domain") else domain
    m = domain.boundaries
    if m.periodic:
        raise NotImplementedError("periodic pressure solves are not on the B200 path (SURVEY 8f N4)")
    if isinstance(m, MixedBoundaries):
        return tuple(domain.solid_sides)          # per-side solid flags -> PDEC_CLOSED_SIDES (ops.closed_code)
    return bool(m.solid)


def poisson_solve(input_field, poisson_domain, solver=None, guess=None):
    """Solves laplace(p) = input_field (solver_api.py:132-158) -> (pressure CenteredGrid, iterations)."""
    assert isinstance(input_field, CenteredGrid)
    solver = solver or SparseCG()
    p, it = solver.solve(input_field.data, poisson_domain, None if guess is None else guess.data)
    return CenteredGrid(p, input_field.box, extrapolation=input_field.extrapolation, name="pressure"), it


def semi_lagrangian(field, velocity_field, dt):
    """advect.semi_lagrangian (advect.py:24-40) for the two cases on the path: a centred scalar field advected by a
    staggered velocity, and a staggered velocity advected by itself.  Uniform extrapolations take the fused kernels; per-side (mixed
    boundary) extrapolations take the reference's generic route, operator by operator (_semi_lagrangian_generic)."""
    mixed = not isinstance(field.extrapolation, str) or not isinstance(velocity_field.extrapolation, str)
    if isinstance(field, StaggeredGrid):
        assert velocity_field is field or velocity_field.compatible(field), "staggered fields are self-advected on the B200 path"
        if mixed:
            return field.with_data([_semi_lagrangian_generic(c, field, dt) for c in field.data])
        vy, vx = field.tensors
        ay, ax = ops.advect_staggered(vy, vx, dt, MODE_OF_EXTRAPOLATION[field.extrapolation], _box4(field.box))
        return field.with_tensors([ay, ax])
    assert isinstance(field, CenteredGrid) and isinstance(velocity_field, StaggeredGrid) and field.rank == 2
    if mixed:
        return field.copied_with(data=_semi_lagrangian_generic(field, velocity_field, dt))
    vy, vx = velocity_field.tensors
    out = ops.advect_centered(field.tensor, vy, vx, dt, MODE_OF_EXTRAPOLATION[field.extrapolation], _box4(field.box))
    return field.with_tensor(out)


def staggered_force_loss(v1, v2, dt):
    """math.l2_loss(advect.semi_lagrangian(v1, v1, dt) - v2) for two StaggeredGrids of one lattice (one term of total_force_loss,
    src/control/pde/incompressible_flow.py:37-38): the fused kernel where v1's extrapolation is uniform, else the operator route."""
    if isinstance(v1.extrapolation, str) and v1.rank == 2 and v1.compatible(v2):
        vy, vx = v1.tensors
        ty, tx = v2.tensors
        return ops.staggered_force_loss(vy, vx, ty, tx, dt, MODE_OF_EXTRAPOLATION[v1.extrapolation], _box4(v1.box))
    adv, loss = semi_lagrangian(v1, v1, dt), 0
    for a, b in zip(adv.data, v2.data):
        d = a.data - b.data
        loss = loss + (d ** 2).sum() / 2 / d.shape[0]
    return loss


def _staggered_at_grid(velocity, grid):
    """StaggeredGrid.at(points of a CenteredGrid on the same lattice) (staggered_grid.py:126-137 -> CenteredGrid.at fast path,
    grid.py:112-123): every component is padded (its own extrapolation, per side) by what the target box needs, cropped and linearly
    interpolated at the fractional offset -- 0.5 along every axis where the two lattices are staggered against each other.
    Returns [B, *grid.resolution, rank]."""
    comps = []
    for comp in velocity.data:
        lower = np.ceil(np.maximum(0, comp.box.lower - grid.box.lower) / comp.dx - 1e-9).astype(int)
        upper = np.ceil(np.maximum(0, grid.box.upper - comp.box.upper) / comp.dx - 1e-9).astype(int)
        src = comp.padded([[int(a), int(b)] for a, b in zip(lower, upper)]) if (lower.sum() + upper.sum()) > 0 else comp
        origin = (grid.box.lower - src.box.lower) / src.dx                    # in cells of the padded component
        data = src.data
        for ax in range(grid.rank):
            o = int(np.floor(origin[ax] + 1e-9))
            frac = float(origin[ax] - o)
            n = int(grid.resolution[ax])
            if abs(frac) < 1e-6:
                data = data.narrow(1 + ax, o, n)
            else:
                assert abs(frac - 0.5) < 1e-6, "grids of one lattice are staggered by half a cell"
                lo, hi = data.narrow(1 + ax, o, n), data.narrow(1 + ax, o + 1, n)
                data = lo * 0.5 + hi * 0.5                                   # interpolate_linear: x0 (1 - w) + x1 w (nd.py:378-392)
        comps.append(data)
    return torch.cat(comps, -1)


def _semi_lagrangian_generic(grid, velocity, dt):
    """x0 = grid.points; v = velocity.at(x0); data = grid.sample_at(x0 - v dt) (advect.py:33-37) for one CenteredGrid (a scalar field or one
    staggered component); plain torch + the resample operator.  Returns the new data [B, ..., 1]."""
    v = _staggered_at_grid(velocity, grid)
    x = grid.points.to(v.device) - v * float(dt)
    B = v.shape[0]
    sampled = grid.sample_at(x.reshape(B, -1, grid.rank))
    return sampled.reshape(grid.data.shape)


def staggered_curl_2d(grid):
    """staggered_curl_2d (util.py:60-67): result has the StaggeredGrid default extrapolation 'boundary'."""
    assert isinstance(grid, CenteredGrid) and grid.rank == 2
    vy, vx = ops.curl_2d(grid.tensor, MODE_OF_EXTRAPOLATION[grid.extrapolation])
    return StaggeredGrid([vy.unsqueeze(-1), vx.unsqueeze(-1)], box=grid.box, batch_size=grid._batch_size)


def curl_advect(density, potential, dt):
    """(semi_lagrangian(density, staggered_curl_2d(potential), dt), staggered_curl_2d(potential)) -- the controlled step
    (src/control/pde/incompressible_flow.py:56-58) -- through the fused kernel where both extrapolations are uniform."""
    if isinstance(density.extrapolation, str) and isinstance(potential.extrapolation, str) and density.rank == 2 and potential.compatible(density):
        out, vy, vx = ops.curl_advect(density.tensor, potential.tensor, dt, MODE_OF_EXTRAPOLATION[potential.extrapolation],
                                      MODE_OF_EXTRAPOLATION[density.extrapolation], _box4(density.box))
        velocity = StaggeredGrid([vy.unsqueeze(-1), vx.unsqueeze(-1)], box=potential.box, batch_size=potential._batch_size)
        return density.with_tensor(out), velocity
    velocity = staggered_curl_2d(potential)
    return semi_lagrangian(density, velocity, dt), velocity


def _box4(box):
    return tuple(float(v) for v in box.lower) + tuple(float(v) for v in box.size)


class Obstacle(State):
    """Solid body inside the domain (PF/physics/obstacle.py:10-27): `geometry` is an AABox (or a list of them) in physical units;
    cells whose centre lies inside (bounds included, geom/_box.py:58-61) are inactive and inaccessible."""

    def __init__(self, geometry, material=None, velocity=0, tags=("obstacle",), name=None, **kwargs):
        State.__init__(self, name=name, tags=tags, **kwargs)
        self.geometry = geometry
        self.material = material
        self.velocity = velocity


_MASK_CACHE = {}


def obstacle_masks(obstacles, velocity, domain):
    """active = 1 - union of the obstacle geometries sampled at the cell centres (fluid.py:228-233) -> ops.ObstacleMasks (cached
    per geometry set, resolution, box and boundary type: obstacles are static on this path).  Geometries: AABox, Sphere, unions."""
    from .geom import geometry_mask_tensor, union
    geos = []
    for o in obstacles:
        geo = o.geometry if hasattr(o, "geometry") else o
        geos += list(geo) if isinstance(geo, (list, tuple)) else [geo]
    geometry = union(geos)
    device = velocity.data[0].data.device
    key = (geometry.key(), tuple(int(r) for r in velocity.resolution), _box4(velocity.box), _closed(domain), str(device))
    if key not in _MASK_CACHE:
        ny, nx = (int(r) for r in velocity.resolution)
        centers = CenteredGrid(torch.zeros(1, ny, nx, 1, device=device), box=velocity.box)         # velocity.center_points
        active = 1.0 - geometry_mask_tensor(geometry, centers)[0, ..., 0]
        _MASK_CACHE[key] = ops.ObstacleMasks(active.contiguous(), _closed(domain))
    return _MASK_CACHE[key]


def divergence_free(velocity, domain=None, obstacles=(), pressure_solver=None, return_info=False):
    """Pressure projection (fluid.py:214-243); with obstacles the masked operators of ops.divergence_free_masked."""
    assert isinstance(velocity, StaggeredGrid)
    if domain is None:
        domain = Domain(velocity.resolution, OPEN, box=velocity.box)
    solver = pressure_solver or SparseCG()
    vy, vx = velocity.tensors
    guess = None
    if isinstance(solver, _SolverChain):
        # solver1 & solver2 (solver_api.py:161-174; the reference's TF default is FourierSolver() & SparseCG()): the leading solvers turn
        # the divergence into an initial guess for the last one
        ny_, nx_ = int(vx.shape[1]), int(vy.shape[2])
        if (len(solver.solvers) == 2 and isinstance(solver.solvers[0], SparseSciPy) and isinstance(solver.solvers[1], SparseCG)
                and len(obstacles) == 0 and ops.direct_solver_supported(ny_, nx_)):
            # SparseSciPy() & SparseCG(...): the exact solve as the CG's initial guess, in ONE autograd node -- the adjoint solve takes
            # the same chain (same matrix) instead of a CG from zero
            guess = "direct"
        else:
            div0 = ops.mask_divergence(vy, vx, _closed(domain)).detach()
            for lead in solver.solvers[:-1]:
                guess, _ = lead.solve(div0, domain, guess)
        solver = solver.solvers[-1]
    if isinstance(solver, FourierSolver):
        raise NotImplementedError("FourierSolver alone is exact on periodic domains only: chain it, FourierSolver() & SparseCG()")
    if len(obstacles) > 0:
        assert velocity.rank == 2
        if isinstance(domain.boundaries, MixedBoundaries):
            raise NotImplementedError("obstacles inside a domain with per-side mixed boundaries")
        if not isinstance(solver, SparseCG):
            raise NotImplementedError("domains with obstacles are solved with SparseCG")
        masks = obstacle_masks(obstacles, velocity, domain)
        oy, ox, p, iters, div = ops.divergence_free_masked(vy, vx, masks, _box4(velocity.box), solver.accuracy, solver.max_iterations,
                                                           solver.gradient_accuracy, solver.max_gradient_iterations)
    elif isinstance(solver, SparseSciPy):
        oy, ox, p, iters, div = ops.divergence_free(vy, vx, _closed(domain), _box4(velocity.box), direct=True)
    else:
        oy, ox, p, iters, div = ops.divergence_free(vy, vx, _closed(domain), _box4(velocity.box), solver.accuracy, solver.max_iterations,
                                                    solver.gradient_accuracy, solver.max_gradient_iterations, guess=guess,
                                                    stop_mode=getattr(solver, "stop_mode", ops.CG_STOP_PER_ELEMENT))
    result = velocity.with_tensors([oy, ox])
    if not return_info:
        return result
    pressure = CenteredGrid(p.unsqueeze(-1), velocity.box, extrapolation="boundary", name="pressure") * 1.0
    pressure = pressure.copied_with(data=pressure.data * float(velocity.dx[0]))       # fluid.py:240
    divergence = CenteredGrid(div.unsqueeze(-1), velocity.box, name="div")
    return result, {"pressure": pressure, "iterations": iters, "divergence": divergence}


class Fluid(DomainState):

    def __init__(self, domain, density=0.0, velocity=0.0, buoyancy_factor=0.0, tags=("fluid", "velocityfield", "velocity"),
                 name="fluid", age=0.0, batch_size=None, device="cuda"):
        DomainState.__init__(self, domain, tags=tags, name=name, age=age, batch_size=batch_size)
        self.buoyancy_factor = buoyancy_factor
        self.density = self.centered_grid("density", density, device=device)
        self.velocity = self._staggered("velocity", velocity, device)
        self.solve_info = {}

    def _staggered(self, name, value, device):
        mode = self.domain.boundaries.vector_extrapolation_mode
        if isinstance(value, StaggeredGrid):
            return value if value.extrapolation == mode else value.copied_with(extrapolation=mode)
        if isinstance(value, (int, float)):
            return StaggeredGrid.sample(value, self.domain, batch_size=self._batch_size, name=name, device=device)
        t = value if isinstance(value, torch.Tensor) else torch.as_tensor(value, dtype=torch.float32)
        return StaggeredGrid(t.to(device), box=self.domain.box, name=name, extrapolation=mode, batch_size=self._batch_size)

    def default_physics(self):
        return INCOMPRESSIBLE_FLOW


def gravity_tensor(gravity, rank):
    """effect.py:106-112: [rank] vector; scalar g -> g * up_vector = (g, 0, ...)."""
    if isinstance(gravity, Gravity):
        gravity = gravity.gravity
    if isinstance(gravity, (tuple, list)) and len(gravity) > 0 and isinstance(gravity[0], Gravity):
        gravity = sum(np.asarray(gravity_tensor(g, rank)) for g in gravity)      # several gravity states add up (effect.py:88-96)
    g = np.asarray(gravity, np.float32)
    if g.ndim == 0:
        g = g * np.asarray([1.0] + [0.0] * (rank - 1), np.float32)
    assert g.shape[-1] == rank
    return g.reshape(-1)[-rank:]


class IncompressibleFlow(Physics):

    def __init__(self, pressure_solver=None, make_input_divfree=False, make_output_divfree=True, conserve_density=True):
        Physics.__init__(self, [StateDependency("obstacles", "obstacle"),
                                StateDependency("gravity", "gravity", single_state=True),
                                StateDependency("density_effects", "density_effect", blocking=True),
                                StateDependency("velocity_effects", "velocity_effect", blocking=True)])
        self.pressure_solver = pressure_solver
        self.make_input_divfree = make_input_divfree
        self.make_output_divfree = make_output_divfree
        self.conserve_density = conserve_density

    def step(self, fluid, dt=1.0, obstacles=(), gravity=None, density_effects=(), velocity_effects=()):
        gravity = gravity_tensor(Gravity() if gravity is None or (isinstance(gravity, (tuple, list)) and len(gravity) == 0) else gravity,
                                 fluid.rank)
        velocity, density = fluid.velocity, fluid.density
        solve_info = {}
        if self.make_input_divfree:
            velocity, solve_info = divergence_free(velocity, fluid.domain, obstacles, self.pressure_solver, return_info=True)
        new_density = semi_lagrangian(density, velocity, dt)
        velocity = advected_velocity = semi_lagrangian(velocity, velocity, dt)
        if self.conserve_density and fluid.domain.boundaries.solid:       # every side solid (np.all(Material.solid(boundaries)), fluid.py:100)
            new_density = new_density.with_tensor(ops.normalize_to(new_density.tensor, density.tensor))
        density = new_density
        for effect in density_effects:
            density = effect_applied(effect, density, dt)
        for effect in velocity_effects:
            velocity = effect_applied(effect, velocity, dt)
        if fluid.buoyancy_factor != 0:
            # velocity += (density * -gravity * buoyancy_factor * dt).at(velocity)   (fluid.py:107)
            factors = [float(-g) * float(fluid.buoyancy_factor) * float(dt) for g in gravity]
            lift = CenteredGrid(torch.cat([density.data * f for f in factors], -1), density.box, extrapolation=density.extrapolation)
            velocity = velocity + lift.at(velocity)
        divergent_velocity = velocity
        if self.make_output_divfree:
            velocity, solve_info = divergence_free(velocity, fluid.domain, obstacles, self.pressure_solver, return_info=True)
        solve_info = dict(solve_info, advected_velocity=advected_velocity, divergent_velocity=divergent_velocity)
        return fluid.copied_with(density=density, velocity=velocity, age=fluid.age + dt, solve_info=solve_info)


INCOMPRESSIBLE_FLOW = IncompressibleFlow()
