"""Direct Poisson solver: tcgen05 3 x TF32 kernel against the fp32 FFMA kernel -- error of both against the fp64 sparse solve of the
oracle, difference between the two, microseconds per batch-64 solve (CUDA events, 200 back-to-back solves)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from pde_control_b200 import ops
from oracle import restate_fluid as F

def rel(a, b): return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))

for (ny, nx, bd) in [(128, 128, "open"), (128, 128, "closed"), (100, 90, "open"), (37, 50, "closed")]:
    rng = np.random.RandomState(ny + nx)
    B = 64
    rhs = rng.randn(B, ny, nx).astype(np.float32)
    for _ in range(2):
        rhs = (rhs + np.roll(rhs, 1, 1) + np.roll(rhs, 1, 2) + np.roll(rhs, -1, 1) + np.roll(rhs, -1, 2)) / 5
    t = torch.tensor(rhs, device="cuda")
    outs = {}
    for tc in (False, True):
        out = torch.empty_like(t)
        ops.poisson_direct_raw(t, out, bd == "closed", tensor_cores=tc)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(200): ops.poisson_direct_raw(t, out, bd == "closed", tensor_cores=tc)
        b.record(); torch.cuda.synchronize()
        outs[tc] = (out.cpu().numpy(), a.elapsed_time(b) / 200 * 1e3)
    exact = F.pressure_solve(torch.tensor(rhs[:4]), bd).numpy()
    dm = (lambda x: x - x.mean((1, 2), keepdims=True)) if bd == "closed" else (lambda x: x)
    print("%3dx%3d %-6s  ffma: %.2e vs fp64, %6.1f us | tcgen05: %.2e vs fp64, %6.1f us | tc vs ffma %.2e"
          % (ny, nx, bd, rel(dm(outs[False][0][:4]), dm(exact)), outs[False][1], rel(dm(outs[True][0][:4]), dm(exact)), outs[True][1],
             rel(outs[True][0], outs[False][0])), flush=True)
