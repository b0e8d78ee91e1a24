"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by running the UNMODIFIED reference
(PhiFlow 1.4.0 NumPy/SciPy backend through oracle/ref_shim.py).  Build container only.

    python -m oracle.make_golden            # regenerate everything

Every fixture stores its inputs, the reference outputs and the parameters, so tests on the GPU box need
neither /root/reference nor this script.
"""
import os
import warnings

import numpy as np

from . import ref_shim, restate as R

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def _burgers_world(pf, u0, force, viscosity, substeps, box01=True):
    B, N = u0.shape
    dom = pf.Domain([N], box=pf.box[0:1])
    world = pf.World(batch_size=B)
    st = world.add(pf.BurgersVelocity(dom, velocity=u0[..., None].copy(), viscosity=viscosity, batch_size=B),
                   physics=pf.Burgers(diffusion_substeps=substeps))
    if force is not None:
        world.add(pf.FieldEffect(pf.CenteredGrid(force[..., None].copy(), box=dom.box), ['velocity']))
    return world, st


def burgers_config1(pf):
    """BASELINE config 1: Burgers res 128, 32 steps, batch 8, GaussianClash + GaussianForce, seed 0,
    nu=0.003, dt=0.03, Burgers(diffusion_substeps=4) on NumPy (notebook cell 9)."""
    rng = np.random.RandomState(0)
    B, N, n = 8, 128, 32
    u0 = R.gaussian_clash(B, N, rng)
    force = R.gaussian_force(B, N, rng)
    world, st = _burgers_world(pf, u0, force, 0.003, 4)
    frames = [u0]
    for _ in range(n):
        world.step(dt=0.03)
        frames.append(st.velocity.data[..., 0].astype(np.float32).copy())
    np.savez_compressed(os.path.join(OUT, "burgers_config1.npz"), u0=u0, force=force, frames=np.stack(frames),
                        dt=0.03, viscosity=0.003, substeps=4, resubstep=1, lower=0.0, size=1.0)


def burgers_single_steps(pf):
    """Independent single steps (the training-loss shape): Burgers() = 1 substep, no force,
    on frames of the config-1 trajectory plus rough random rows that leave the domain ('constant' edge rule)."""
    rng = np.random.RandomState(1)
    g = np.load(os.path.join(OUT, "burgers_config1.npz"))
    rows = [g["frames"][t] for t in (0, 5, 17, 31)]
    rows.append((rng.randn(8, 128) * 1.5).astype(np.float32))
    u = np.concatenate(rows, 0)
    world, st = _burgers_world(pf, u, None, 0.003, 1)
    world.step(dt=0.03)
    out = st.velocity.data[..., 0].astype(np.float32).copy()
    np.savez_compressed(os.path.join(OUT, "burgers_single_step.npz"), u=u, out=out, dt=0.03, viscosity=0.003,
                        substeps=1, resubstep=0, lower=0.0, size=1.0)


def burgers_periodic(pf):
    """PERIODIC domain -> diffuse takes the FFT branch (PF/physics/field/util.py:30-37); advection wraps around
    ('circular' resample); with a GROW force.  4 steps, batch 3, N = 128, box [0, 1]."""
    rng = np.random.RandomState(21)
    B, N = 3, 128
    x = (np.arange(N) + 0.5) / N
    u0 = np.stack([a * np.sin(2 * np.pi * (k * x + ph)) for a, k, ph in zip(rng.uniform(0.5, 2.5, B), (1, 2, 3), rng.rand(B))]).astype(np.float32)
    force = (0.5 * np.cos(2 * np.pi * x))[None].repeat(B, 0).astype(np.float32)
    dom = pf.Domain([N], boundaries=pf.PERIODIC, box=pf.box[0:1])
    world = pf.World(batch_size=B)
    st = world.add(pf.BurgersVelocity(dom, velocity=u0[..., None].copy(), viscosity=0.01, batch_size=B), physics=pf.Burgers())
    world.add(pf.FieldEffect(pf.CenteredGrid(force[..., None].copy(), box=dom.box, extrapolation="periodic"), ['velocity']))
    frames = [u0]
    for _ in range(4):
        world.step(dt=0.03)
        frames.append(np.asarray(st.velocity.data[..., 0], np.float32).copy())
    np.savez_compressed(os.path.join(OUT, "burgers_periodic.npz"), frames=np.stack(frames), force=force, dt=0.03, viscosity=0.01,
                        extrapolation=str(st.velocity.extrapolation))


def main():
    os.makedirs(OUT, exist_ok=True)
    warnings.simplefilter("ignore")
    pf = ref_shim.flow()
    burgers_config1(pf)
    burgers_single_steps(pf)
    burgers_periodic(pf)
    try:
        from . import make_golden_fluid
        make_golden_fluid.main(pf, OUT)
    except ImportError:
        pass
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
