"""The README's quick start, runnable (B200)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from pde_control_b200.physics import *  # noqa: F401,F403

rng = np.random.RandomState(0)
rho0 = np.abs(rng.randn(64, 128, 128, 1)).astype(np.float32)
v0 = (0.3 * rng.randn(64, 129, 129, 2)).astype(np.float32)
world = World(batch_size=64)
fluid = world.add(Fluid(Domain([128, 128], boundaries=CLOSED), density=rho0, velocity=v0, batch_size=64, buoyancy_factor=0.1),
                  physics=IncompressibleFlow(pressure_solver=SparseCG(accuracy=1e-5, max_iterations=2000)))
world.add(Inflow(Sphere((16, 64), 8), rate=0.2))
world.add(Obstacle(box[60:70, 40:90]))
world.step(dt=1.0)
print("fluid:", fluid.density.data.shape, fluid.density.data.device, float(fluid.density.data.sum()), int(fluid.solve_info["iterations"].max()))

u0 = np.sin(2 * np.pi * np.linspace(0, 1, 128, endpoint=False))[None, :, None].astype(np.float32).repeat(8, 0)
w2 = World(batch_size=8)
u = w2.add(BurgersVelocity(Domain([128], box=box[0:1]), velocity=u0, viscosity=0.003, batch_size=8), physics=Burgers())
w2.step(dt=0.03)
print("burgers:", u.velocity.data.shape, float(u.velocity.data.abs().max()))

from pde_control_b200 import ops
frames = torch.randn(5, 16, 128, device="cuda", requires_grad=True)
l2, l1 = ops.burgers_pair_loss(frames, dt=0.03, viscosity=0.003)
l2.backward()
print("pair loss:", float(l2), float(l1), frames.grad.shape)
