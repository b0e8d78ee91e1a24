"""Micro-benchmark of the Poisson CG kernels (128x128, batch 64 and 148): time per solve and per iteration."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pde_control_b200 import ops

def smooth(rng, shape, scale, passes=2):
    a = rng.randn(*shape)
    for _ in range(passes):
        a = (a + np.roll(a, 1, -1) + np.roll(a, -1, -1) + np.roll(a, 1, -2) + np.roll(a, -1, -2)) / 5
    return (a / np.abs(a).max() * scale).astype(np.float32)

rng = np.random.RandomState(0)
for B, n in ((64, 128), (148, 128), (296, 128), (64, 64)):
    rhs = torch.tensor(smooth(rng, (B, n, n), 1.0)).cuda()
    out = torch.empty_like(rhs)
    it = torch.zeros(B, dtype=torch.int32, device="cuda")
    res = torch.zeros(B, device="cuda")
    for variant, name in ((0, "block"), (5, "block-fused"), (4, "block-2red"), (3, "tile")):
        if B > 148 and variant: continue
        for closed in (False, True):
            for acc, mi in ((1e-5, 2000), (0.0, 200)):
                args = (rhs, out, closed, acc, mi, None, it, res, variant << 8)
                for _ in range(3): ops.cg_poisson_raw(*args)
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(10): ops.cg_poisson_raw(*args)
                e1.record(); torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / 10
                mean_it = float(it.float().mean()); max_it = int(it.max())
                print("B=%d n=%d %s closed=%d acc=%g: %.3f ms/solve, iters mean %.0f max %d -> %.2f us/iter (max), %.3g cell-iter/s" % (
                    B, n, name, closed, acc, ms, mean_it, max_it, ms * 1e3 / max(max_it, 1), B * n * n * mean_it / (ms * 1e-3)), flush=True)
