"""Obstacle scene at the config-3 size (128x128, B=64): masked Poisson solve on chip against the streaming kernel, and the full
projected step (advection, masks, divergence, masked CG, gradient subtraction) through IncompressibleFlow.step."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from pde_control_b200 import ops
from pde_control_b200.physics import (CLOSED, OPEN, Domain, Fluid, IncompressibleFlow, Obstacle, SparseCG, Sphere, box)


def timed(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def smooth(rng, shape, scale, passes=4):
    a = rng.randn(*shape)
    for _ in range(passes):
        a = (a + np.roll(a, 1, -1) + np.roll(a, -1, -1) + np.roll(a, 1, -2) + np.roll(a, -1, -2)) / 5
    return (a / np.abs(a).max() * scale).astype(np.float32)


out = {}
B, n = 64, 128
rng = np.random.RandomState(0)
for boundary in (OPEN, CLOSED):
    name = "open" if boundary is OPEN else "closed"
    dom = Domain([n, n], boundaries=boundary)
    vel = np.zeros((B, n + 1, n + 1, 2), np.float32)
    vel[:, :, :n, 0] = smooth(rng, (B, n + 1, n), 1.0)
    vel[:, :n, :, 1] = smooth(rng, (B, n, n + 1), 1.0)
    fluid = Fluid(dom, density=np.abs(smooth(rng, (B, n, n, 1), 1.0)), velocity=vel, batch_size=B)
    obstacles = [Obstacle(box[40:60, 30:90]), Obstacle(Sphere((96, 64), 12))]
    phys = IncompressibleFlow(pressure_solver=SparseCG(accuracy=1e-5, max_iterations=2000))
    state = {"f": fluid}

    def step():
        state["f"] = phys.step(state["f"], 1.0, obstacles=obstacles)

    ms_step = timed(step, reps=8, warm=2)
    its = state["f"].solve_info["iterations"]
    # the solve alone, both kernels, on this scene's divergence
    from pde_control_b200.physics.fluid import obstacle_masks
    masks = obstacle_masks(obstacles, fluid.velocity, dom)
    div = state["f"].solve_info["divergence"].tensor.contiguous()
    div = torch.randn_like(div) * masks.active
    if boundary is CLOSED:
        div = (div - div.sum((1, 2), keepdim=True) / masks.active.sum()) * masks.active
    res = {}
    for kname, flag in (("on_chip", 0), ("streaming", 1 << 8)):
        p, it = torch.empty_like(div), torch.zeros(B, dtype=torch.int32, device="cuda")
        ms = timed(lambda: ops.cg_poisson_masked_raw(div, p, masks.code, 1e-5, 2000, None, it, None, flag), reps=5, warm=2)
        res[kname] = dict(ms_per_solve=ms, iterations_mean=float(it.float().mean()), iterations_max=int(it.max()),
                          us_per_iteration=ms * 1e3 / max(int(it.max()), 1))
    out[name] = dict(ms_per_step=ms_step, step_iterations_mean=float(its.float().mean()), solve=res)
    print(name, json.dumps(out[name]), flush=True)
print(json.dumps(dict(config="128x128 B=64, box + sphere obstacles, accuracy 1e-5", **out)))
