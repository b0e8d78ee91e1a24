"""How fast is a bilinear scatter with 64-bit fixed-point global atomics (pdec_resample_bwd) at the fluid's size (B = 64, 128 x 128)?
Decides whether the advection adjoints should scatter instead of gather-transposing.  Uses torch.profiler kernel times."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from pde_control_b200 import ops
B, n = 64, 128
rng = np.random.RandomState(0)
yy, xx = np.meshgrid(np.arange(n, dtype=np.float32), np.arange(n, dtype=np.float32), indexing="ij")
disp = rng.randn(B, n, n, 2).astype(np.float32)
for _ in range(4): disp = (disp + np.roll(disp, 1, 1) + np.roll(disp, 1, 2) + np.roll(disp, -1, 1) + np.roll(disp, -1, 2)) / 5
coords = torch.tensor(np.stack([yy, xx], -1)[None] + 3 * disp / np.abs(disp).max(), device="cuda").requires_grad_()
data = torch.rand(B, n, n, device="cuda", requires_grad=True)
w = torch.randn(B, n, n, device="cuda")
for _ in range(3):
    out = ops.resample(data, coords, 0); (out * w).sum().backward()
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(10):
        out = ops.resample(data, coords, 0); (out * w).sum().backward()
    torch.cuda.synchronize()
for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:8]:
    print("%-70s %8.1f us x %d" % (e.key[:70], e.device_time_total / e.count, e.count))
