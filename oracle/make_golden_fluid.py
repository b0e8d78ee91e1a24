"""TEST INFRASTRUCTURE ONLY -- fluid golden vectors from the UNMODIFIED reference NumPy backend
(oracle/ref_shim.py).  Called by oracle/make_golden.py.  Build container only."""
import os

import numpy as np


def _smooth(rng, shape, scale, passes=6):
    a = rng.randn(*shape)
    for _ in range(passes):
        a = (a + np.roll(a, 1, -1) + np.roll(a, -1, -1) + np.roll(a, 1, -2) + np.roll(a, -1, -2)) / 5
    return (a / np.abs(a).max() * scale).astype(np.float32)


def random_staggered(rng, B, ny, nx, scale):
    return _smooth(rng, (B, ny + 1, nx), scale), _smooth(rng, (B, ny, nx + 1), scale)


def staggered_tensor(vy, vx):
    """[B, ny+1, nx+1, 2] staggered tensor (staggered_grid.py:41-46)."""
    B, ny1, nx = vy.shape
    t = np.zeros((B, ny1, nx + 1, 2), np.float32)
    t[:, :, :nx, 0] = vy
    t[:, :ny1 - 1, :, 1] = vx
    return t


BOUNDARY_SPECS = {
    # name -> per axis a material name or a (lower, upper) pair (Domain boundaries, PF/physics/domain.py:13-35)
    "open": "OPEN", "closed": "CLOSED", "no_slip": "NO_SLIP",
    "open_top": [("CLOSED", "OPEN"), "CLOSED"],                 # the reference's own example: a box open at the top
    "open_y_half_open_x": ["OPEN", ("CLOSED", "OPEN")],
}


def ref_boundaries(pf, name):
    from phi.physics import material as M
    spec = BOUNDARY_SPECS[name]
    get = lambda n: getattr(M, n)
    if isinstance(spec, str):
        return get(spec)
    return [get(e) if isinstance(e, str) else (get(e[0]), get(e[1])) for e in spec]


def fluid_step_case(pf, ny, nx, B, boundary, seed, steps, dt, solver="direct", buoyancy=0.0, density_effect=False, obstacles=(),
                    conserve_density=True):
    rng = np.random.RandomState(seed)
    rho = np.abs(_smooth(rng, (B, ny, nx), 1.0))
    vy, vx = random_staggered(rng, B, ny, nx, 1.5)
    dom = pf.Domain([ny, nx], boundaries=ref_boundaries(pf, boundary))
    fluid = pf.Fluid(dom, density=rho[..., None].copy(), velocity=staggered_tensor(vy, vx), batch_size=B, buoyancy_factor=buoyancy)
    effects = []
    if density_effect:          # a GROW effect on the density (what Inflow is, effect.py:62) given as a plain grid
        src = np.zeros((1, ny, nx, 1), np.float32)
        src[:, ny // 4:ny // 2, nx // 3:2 * nx // 3] = 0.25
        effects = [pf.FieldEffect(pf.CenteredGrid(src, box=dom.box), ["density"], pf.GROW)]
    psolver = pf.SparseSciPy() if solver == "direct" else pf.SparseCG(accuracy=1e-7, max_iterations=5000)
    phys = pf.IncompressibleFlow(pressure_solver=psolver, conserve_density=conserve_density)
    out = dict(rho0=rho, vy0=vy, vx0=vx, dt=dt, boundary=boundary, steps=steps, buoyancy=buoyancy)
    obs = [pf.Obstacle(pf.box[y0:y1, x0:x1]) for (y0, y1, x0, x1) in obstacles]
    if obstacles:
        out["obstacles"] = np.asarray(obstacles, np.float32)
    if density_effect:
        out["density_source"] = src[0, ..., 0]
    for s in range(steps):
        fluid = phys.step(fluid, dt, obstacles=obs, density_effects=effects)
        out["rho%d" % (s + 1)] = np.asarray(fluid.density.data[..., 0], np.float32)
        out["vy%d" % (s + 1)] = np.asarray(fluid.velocity.data[0].data[..., 0], np.float32)
        out["vx%d" % (s + 1)] = np.asarray(fluid.velocity.data[1].data[..., 0], np.float32)
        out["p%d" % (s + 1)] = np.asarray(fluid.solve_info["pressure"].data[..., 0], np.float32)
        out["div%d" % (s + 1)] = np.asarray(fluid.solve_info["divergence"].data[..., 0], np.float32)
        out["avy%d" % (s + 1)] = np.asarray(fluid.solve_info["advected_velocity"].data[0].data[..., 0], np.float32)
        out["avx%d" % (s + 1)] = np.asarray(fluid.solve_info["advected_velocity"].data[1].data[..., 0], np.float32)
    return out


def _ref_geometry(pf, spec):
    """Geometry spec tuple of oracle/restate_fluid.lies_inside -> the reference's geometry object."""
    if spec[0] == "box":
        return pf.AABox(np.asarray(spec[1], np.float32), np.asarray(spec[2], np.float32))
    if spec[0] == "sphere":
        return pf.Sphere(tuple(spec[1]), spec[2])
    return pf.union([_ref_geometry(pf, s) for s in spec[1:]])


def _ref_effect(pf, e, target):
    """Effect dict of oracle/restate_fluid.apply_effect -> the reference's FieldEffect (through its own shorthands where one exists)."""
    if e["mode"] == "fix":
        if target == "velocity":
            return pf.ConstantVelocity(_ref_geometry(pf, e["bounds"]), e["value"])
        return pf.FieldEffect(pf.ConstantField(e["value"]), [target], pf.FIX, bounds=_ref_geometry(pf, e["bounds"]))
    assert e["mode"] == "grow"
    if target == "velocity":
        return pf.Accelerator(_ref_geometry(pf, e["geometry"]), e["value"])
    return pf.Inflow(_ref_geometry(pf, e["geometry"]), rate=e["value"])


EFFECT_SCENES = {
    # Inflow through a sphere, a fan with a per-component acceleration, a FIX velocity region, a sphere obstacle, buoyancy
    "fluid_effects_open_32": dict(ny=32, nx=32, B=2, boundary="open", seed=23, steps=3, dt=0.5, buoyancy=0.1,
                                  density_effects=[dict(mode="grow", geometry=("sphere", (8, 16), 4), value=0.3)],
                                  velocity_effects=[dict(mode="grow", geometry=("box", (16, 8), (24, 20)), value=[0.2, -0.1]),
                                                    dict(mode="fix", bounds=("box", (26, 20), (30, 28)), value=[0.3, 0.0], geometry=None)],
                                  obstacle=("sphere", (16, 26), 3)),
    # closed box: FIX density region, scalar fan, obstacle = union of a sphere and a box in the corner
    "fluid_effects_closed_24x32": dict(ny=24, nx=32, B=2, boundary="closed", seed=29, steps=2, dt=0.5, buoyancy=0.0,
                                       density_effects=[dict(mode="fix", bounds=("sphere", (6, 10), 3), value=0.7, geometry=None),
                                                        dict(mode="grow", geometry=("box", (14, 4), (18, 9)), value=0.25)],
                                       velocity_effects=[dict(mode="grow", geometry=("box", (2, 16), (8, 28)), value=0.2)],
                                       obstacle=("union", ("sphere", (14, 20), 3.5), ("box", (20, 0), (24, 4)))),
}


def effects_case(pf, name):
    s = EFFECT_SCENES[name]
    ny, nx, B, dt = s["ny"], s["nx"], s["B"], s["dt"]
    rng = np.random.RandomState(s["seed"])
    rho = np.abs(_smooth(rng, (B, ny, nx), 1.0))
    vy, vx = random_staggered(rng, B, ny, nx, 1.5)
    dom = pf.Domain([ny, nx], boundaries=pf.OPEN if s["boundary"] == "open" else pf.CLOSED)
    fluid = pf.Fluid(dom, density=rho[..., None].copy(), velocity=staggered_tensor(vy, vx), batch_size=B, buoyancy_factor=s["buoyancy"])
    phys = pf.IncompressibleFlow(pressure_solver=pf.SparseSciPy())
    de = [_ref_effect(pf, e, "density") for e in s["density_effects"]]
    ve = [_ref_effect(pf, e, "velocity") for e in s["velocity_effects"]]
    obs = [pf.Obstacle(_ref_geometry(pf, s["obstacle"]))]
    out = dict(rho0=rho, vy0=vy, vx0=vx)
    for k in range(s["steps"]):
        fluid = phys.step(fluid, dt, obstacles=obs, density_effects=de, velocity_effects=ve)
        out["rho%d" % (k + 1)] = np.asarray(fluid.density.data[..., 0], np.float32)
        out["vy%d" % (k + 1)] = np.asarray(fluid.velocity.data[0].data[..., 0], np.float32)
        out["vx%d" % (k + 1)] = np.asarray(fluid.velocity.data[1].data[..., 0], np.float32)
        out["p%d" % (k + 1)] = np.asarray(fluid.solve_info["pressure"].data[..., 0], np.float32)
        out["dvy%d" % (k + 1)] = np.asarray(fluid.solve_info["divergent_velocity"].data[0].data[..., 0], np.float32)
        out["dvx%d" % (k + 1)] = np.asarray(fluid.solve_info["divergent_velocity"].data[1].data[..., 0], np.float32)
    return out


def controlled_case(pf, ny, nx, B, seed, dt):
    """Notebook variant: curl of a potential ('constant' extrapolation, like the U-Net output) advects the
    density; staggered self-advection of that velocity (force loss)."""
    rng = np.random.RandomState(seed)
    rho = np.abs(_smooth(rng, (B, ny, nx), 1.0))
    phi = _smooth(rng, (B, ny, nx), 6.0)
    dom = pf.Domain([ny, nx])
    density = pf.CenteredGrid.sample(0, dom, batch_size=B).copied_with(data=rho[..., None].copy())
    pot = pf.CenteredGrid(phi[..., None].copy(), box=dom.box, extrapolation="constant")
    vel = pf.staggered_curl_2d(pot)
    nxt = pf.advect.semi_lagrangian(density, vel, dt)
    adv = pf.advect.semi_lagrangian(vel, vel, dt)
    return dict(rho=rho, phi=phi, dt=dt, vy=np.asarray(vel.data[0].data[..., 0], np.float32),
                vx=np.asarray(vel.data[1].data[..., 0], np.float32), rho_next=np.asarray(nxt.data[..., 0], np.float32),
                avy=np.asarray(adv.data[0].data[..., 0], np.float32), avx=np.asarray(adv.data[1].data[..., 0], np.float32),
                density_extrapolation=str(density.extrapolation), velocity_extrapolation=str(vel.extrapolation))


def poisson_cases(pf):
    """PhiFlow/tests/test_poisson_solve.py:14-23 fixed right-hand sides + random 40x32 on OPEN/CLOSED."""
    out = {}
    rng = np.random.RandomState(4)
    for name, material in (("open", pf.OPEN), ("closed", pf.CLOSED)):
        dom = pf.Domain([40, 32], boundaries=material)
        rhs = rng.randn(3, 40, 32).astype(np.float32)
        if name == "closed":
            rhs -= rhs.mean((1, 2), keepdims=True)
        field = pf.CenteredGrid(rhs[..., None].copy(), box=dom.box)
        p, _ = pf.poisson_solve(field, dom, solver=pf.SparseCG(accuracy=1e-6, max_iterations=5000))
        out["rhs_" + name] = rhs
        out["p_" + name] = np.asarray(p.data[..., 0], np.float32)
        # FourierSolver (fourier.py:17-18 -> math.fourier_poisson, nd.py:305-314): the first link of the TF-default solver chain
        out["fft_" + name] = np.asarray(pf.math.fourier_poisson(rhs[..., None].copy())[..., 0], np.float32)
    return out


def main(pf, out_dir):
    np.savez_compressed(os.path.join(out_dir, "fluid_open_32.npz"), **fluid_step_case(pf, 32, 32, 2, "open", 3, 3, 1.0))
    np.savez_compressed(os.path.join(out_dir, "fluid_closed_32x24.npz"), **fluid_step_case(pf, 32, 24, 2, "closed", 5, 3, 0.7))
    np.savez_compressed(os.path.join(out_dir, "fluid_open_128.npz"), **fluid_step_case(pf, 128, 128, 1, "open", 7, 1, 1.0))
    np.savez_compressed(os.path.join(out_dir, "fluid_buoyancy_open_32.npz"),
                        **fluid_step_case(pf, 32, 32, 2, "open", 11, 3, 0.5, buoyancy=0.2, density_effect=True))
    np.savez_compressed(os.path.join(out_dir, "fluid_buoyancy_closed_24x32.npz"),
                        **fluid_step_case(pf, 24, 32, 2, "closed", 13, 3, 0.5, buoyancy=0.2, density_effect=True))
    # obstacles: one inside the domain, one touching the lower x boundary (accessible-mask extrapolation), SURVEY 8f N3
    np.savez_compressed(os.path.join(out_dir, "fluid_obstacle_open_32.npz"),
                        **fluid_step_case(pf, 32, 32, 2, "open", 17, 2, 0.5, obstacles=[(8, 14, 10, 20), (20, 26, 0, 5)]))
    np.savez_compressed(os.path.join(out_dir, "fluid_obstacle_closed_24x32.npz"),
                        **fluid_step_case(pf, 24, 32, 2, "closed", 19, 2, 0.5, obstacles=[(6, 12, 10, 20), (16, 24, 27, 32)]))
    # SURVEY 8f N3 remainder: NO_SLIP (velocity extrapolated with 0 outside) and per-side mixed boundaries.  The reference's own
    # conserve_density test (np.all over a ragged list, fluid.py:100) fails under NumPy 2 for mixed boundaries: conserve_density=False
    np.savez_compressed(os.path.join(out_dir, "fluid_noslip_24x20.npz"), **fluid_step_case(pf, 24, 20, 2, "no_slip", 31, 3, 0.6, buoyancy=0.1))
    np.savez_compressed(os.path.join(out_dir, "fluid_mixed_open_top_24x20.npz"),
                        **fluid_step_case(pf, 24, 20, 2, "open_top", 37, 3, 0.6, buoyancy=0.1, conserve_density=False))
    np.savez_compressed(os.path.join(out_dir, "fluid_mixed_half_open_x_20x24.npz"),
                        **fluid_step_case(pf, 20, 24, 2, "open_y_half_open_x", 41, 3, 0.6, conserve_density=False))
    for name in EFFECT_SCENES:          # Inflow / Accelerator / FIX effects, sphere and union obstacles
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **effects_case(pf, name))
    np.savez_compressed(os.path.join(out_dir, "fluid_controlled_64.npz"), **controlled_case(pf, 64, 64, 2, 9, 1.0))
    np.savez_compressed(os.path.join(out_dir, "fluid_controlled_128.npz"), **controlled_case(pf, 128, 128, 1, 15, 1.0))      # config-3 size
    np.savez_compressed(os.path.join(out_dir, "poisson_40x32.npz"), **poisson_cases(pf))
