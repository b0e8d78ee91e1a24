"""us per launch group of the fluid operators at config-3 size (B = 64, 128 x 128), forward and adjoint, CUDA events."""
import os, sys, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
from benchlib import inputs
from pde_control_b200 import ops
rho0, vy0, vx0, target = inputs.fluid_config3_inputs(64, 128, 0)
scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
rho, vy, vx = [torch.tensor(a, device="cuda") for a in (rho0, vy0 * scale, vx0 * scale)]

def timeit(fn, reps=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3

out = {"velocity scale": scale}
out["advect_centered fwd"] = timeit(lambda: ops.advect_centered(rho, vy, vx, 1.0, 0))
out["advect_staggered fwd"] = timeit(lambda: ops.advect_staggered(vy, vx, 1.0, 0))
r, y, x = [t.clone().requires_grad_(True) for t in (rho, vy, vx)]
o = ops.advect_centered(r, y, x, 1.0, 0)
g = torch.randn_like(o)
out["advect_centered bwd"] = timeit(lambda: torch.autograd.grad(o, (r, y, x), g, retain_graph=True))
ay, ax = ops.advect_staggered(y, x, 1.0, 0)
gy, gx = torch.randn_like(ay), torch.randn_like(ax)
out["advect_staggered bwd"] = timeit(lambda: torch.autograd.grad((ay, ax), (y, x), (gy, gx), retain_graph=True))
d = ops.mask_divergence(y, x, False)
out["mask_divergence fwd"] = timeit(lambda: ops.mask_divergence(vy, vx, False))
print(json.dumps(out))
