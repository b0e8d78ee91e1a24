"""GPU: the reference-facing API (World.step / Physics / PDE / ControlTraining) driving the CUDA kernels.
Golden vectors come from the real reference run through the SAME API calls (oracle/make_golden*.py)."""
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, rel_err

pytestmark = pytest.mark.gpu


def test_world_step_burgers_with_force_matches_reference_golden():
    """Notebook cell 9: world.add(BurgersVelocity, physics=Burgers(diffusion_substeps=4)); world.add(FieldEffect(force));
    32 x world.step(dt=0.03) -- NumPy substep semantics (tf_substeps=False), SURVEY fact 5."""
    from pde_control_b200.physics import World, Domain, box, BurgersVelocity, Burgers, FieldEffect, CenteredGrid
    g = np.load(os.path.join(GOLDEN, "burgers_config1.npz"))
    B, N = g["u0"].shape
    dom = Domain([N], box=box[0:1])
    world = World(batch_size=B)
    u = world.add(BurgersVelocity(dom, velocity=g["u0"][..., None], viscosity=0.003, batch_size=B),
                  physics=Burgers(diffusion_substeps=4, tf_substeps=False))
    world.add(FieldEffect(CenteredGrid(torch.tensor(g["force"][..., None]).cuda(), box=dom.box, extrapolation="constant"), ["velocity"]))
    errs = []
    for t in range(1, 33):
        world.step(dt=0.03)
        errs.append(rel_err(u.velocity.data[..., 0].cpu(), g["frames"][t]))
    assert errs[0] < 1e-5 and max(errs) < 1e-4
    assert abs(u.age - 32 * 0.03) < 1e-9 and u.velocity.extrapolation == "constant"


@pytest.mark.parametrize("name", ["fluid_open_32", "fluid_closed_32x24"])
def test_world_step_fluid_matches_reference_golden(name):
    from pde_control_b200.physics import World, Domain, Fluid, IncompressibleFlow, SparseCG, OPEN, CLOSED
    from oracle.make_golden_fluid import staggered_tensor
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    B, ny, nx = g["rho0"].shape
    dom = Domain([ny, nx], boundaries=OPEN if str(g["boundary"]) == "open" else CLOSED)
    world = World(batch_size=B)
    fluid = world.add(Fluid(dom, density=g["rho0"][..., None], velocity=staggered_tensor(g["vy0"], g["vx0"]), batch_size=B),
                      physics=IncompressibleFlow(pressure_solver=SparseCG(accuracy=1e-6, max_iterations=3000)))
    world.step(dt=float(g["dt"]))
    assert rel_err(fluid.density.data[..., 0].cpu(), g["rho1"]) < 1e-5
    assert rel_err(fluid.velocity.data[0].data[..., 0].cpu(), g["vy1"]) < 2e-5
    assert rel_err(fluid.velocity.data[1].data[..., 0].cpu(), g["vx1"]) < 2e-5
    info = fluid.solve_info
    assert rel_err(info["divergence"].data[..., 0].cpu(), g["div1"]) < 1e-5 and int(info["iterations"].max()) > 10
    assert rel_err(info["advected_velocity"].data[0].data[..., 0].cpu(), g["avy1"]) < 1e-5


def test_control_training_burgers_runs_the_fused_force_loss():
    from pde_control_b200 import ops, datagen
    from pde_control_b200.physics import Domain, box
    from pde_control_b200.control import ControlTraining, BurgersPDE, StaggeredSequence
    rng = np.random.RandomState(0)
    n, sims, N = 8, 32, 128
    u0 = torch.tensor(datagen.gaussian_clash(sims, N, rng)).cuda()
    f = torch.tensor(datagen.gaussian_force(sims, N, rng)).cuda()
    frames = torch.cat([u0[None], ops.burgers_trajectory(u0, n, 0.03, 0.003, force=f, substeps=4, resubstep=True)])
    data = {"u": frames.permute(1, 0, 2)[..., None].cpu().numpy()}            # [sims, frames, N, 1]
    torch.manual_seed(0)
    pde = BurgersPDE(Domain([N], box=box[0:1]), 0.003, 0.03)
    app = ControlTraining(n, pde, data, [0, 1], range(sims), trace_to_channel=lambda _: "u", obs_loss_frames=[],
                          trainable_networks=["OP2", "OP4", "OP8"], sequence_class=StaggeredSequence, batch_size=16,
                          learning_rate=1e-3, dt=0.03)
    losses = [app.step()["Loss"] for _ in range(12)]
    assert all(np.isfinite(losses)) and min(losses[6:]) < losses[0]
    assert "Total Force" in app.scalars and set(pde.networks.keys()) == {"OP2", "OP4", "OP8"}
    states = app.infer_all_frames([0, 1])
    assert len(states) == n + 1 and states[3].burgers.velocity.data.shape == (2, N, 1)
    assert rel_err(states[0].burgers.velocity.data.cpu(), data["u"][:2, 0]) == 0.0


def test_control_training_fluid_cfe_and_op():
    from pde_control_b200 import datagen
    from pde_control_b200.physics import Domain
    from pde_control_b200.control import ControlTraining, IncompressibleFluidPDE, StaggeredSequence
    import random
    n, sims, res = 4, 8, 32
    shapes = datagen.procedural_shapes(6, 8, 0)
    r = random.Random(0)
    start = datagen.distribute_random_shape([res, res], sims, shapes, rng=r)
    end = datagen.distribute_random_shape([res, res], sims, shapes, rng=r)
    frames = np.stack([start * (1 - t / n) + end * (t / n) for t in range(n + 1)], 1)[..., None]
    torch.manual_seed(0)
    pde = IncompressibleFluidPDE(Domain([res, res]), dt=1.0)
    app = ControlTraining(n, pde, {"density": frames}, None, range(sims), obs_loss_frames=[], trainable_networks=["CFE", "OP2", "OP4"],
                          sequence_class=StaggeredSequence, batch_size=4, learning_rate=2e-4, dt=1.0)
    losses = [app.step()["Loss"] for _ in range(4)]
    assert all(np.isfinite(losses))
    assert set(pde.networks.keys()) == {"CFE", "OP2", "OP4"}
    grads = [p.grad for p in pde.networks["CFE"].parameters()]
    assert all(g is not None and bool(torch.isfinite(g).all()) for g in grads) and any(float(g.abs().max()) > 0 for g in grads)
    states = app.infer_all_frames(range(2))
    assert hasattr(states[1].density, "velocity") and states[1].density.velocity.data[0].data.shape == (2, res + 1, res, 1)
