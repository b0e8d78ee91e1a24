"""Config-5 building block on ONE GPU: the cooperative CG kernel on large grids (time per iteration, iterations to 1e-5)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pde_control_b200 import ops

def smooth(rng, shape, scale, passes=2):
    a = rng.randn(*shape).astype(np.float32)
    for _ in range(passes):
        a = (a + np.roll(a, 1, -1) + np.roll(a, -1, -1) + np.roll(a, 1, -2) + np.roll(a, -1, -2)) / 5
    return (a / np.abs(a).max() * scale).astype(np.float32)

rng = np.random.RandomState(0)
for B, n in ((1, 512), (1, 1024), (1, 2048), (4, 2048)):
    rhs = torch.tensor(smooth(rng, (B, n, n), 1.0)).cuda()
    out = torch.empty_like(rhs)
    it = torch.zeros(B, dtype=torch.int32, device="cuda")
    res = torch.zeros(B, device="cuda")
    for closed in (False, True):
        if closed:
            rhs = rhs - rhs.mean((1, 2), keepdim=True)
        for acc, mi in ((0.0, 200), (1e-5, 20000)):
            args = (rhs, out, closed, acc, mi, None, it, res, 0)
            ops.cg_poisson_raw(*args)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ops.cg_poisson_raw(*args)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            k = int(it.max())
            print("B=%d n=%d closed=%d acc=%g: %.2f ms/solve, iterations %d, residual %.2e -> %.2f us/iter, %.0f GB/s algorithmic (36 B/cell/iter)" % (
                B, n, closed, acc, ms, k, float(res.max()), ms * 1e3 / max(k, 1), 36.0 * B * n * n * k / (ms * 1e-3) / 1e9), flush=True)
