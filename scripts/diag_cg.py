"""Diagnostic: CUDA CG vs oracle fp32 CG vs fp64 direct solve (true residual drift on the singular closed operator)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import restate_fluid as F
from pde_control_b200 import ops

def smooth(rng, shape, scale, passes=2):
    a = rng.randn(*shape)
    for _ in range(passes):
        a = (a + np.roll(a, 1, -1) + np.roll(a, -1, -1) + np.roll(a, 1, -2) + np.roll(a, -1, -2)) / 5
    return (a / np.abs(a).max() * scale).astype(np.float32)

def dm(p): return p - p.mean((1, 2), keepdims=True)
for bd in ("open", "closed"):
    for ny, nx in ((40, 32), (64, 64), (128, 128), (160, 144)):
        rng = np.random.RandomState(ny * nx)
        rhs = smooth(rng, (3, ny, nx), 1.0)
        if bd == "closed": rhs -= rhs.mean((1, 2), keepdims=True)
        for acc in (1e-5, 2e-6):
            p, it, res = ops.pressure_solve(torch.tensor(rhs).cuda(), bd == "closed", accuracy=acc, max_iterations=4000)
            pc = p.cpu()
            true_res = float((F.laplace_apply(pc.double(), bd) - torch.tensor(rhs).double()).abs().max())
            direct = F.pressure_solve(torch.tensor(rhs), bd).numpy()
            got = pc.numpy()
            if bd == "closed": direct, got = dm(direct), dm(got)
            err = np.abs(got - direct).max() / np.abs(direct).max()
            line = "%s %dx%d acc %.0e: it %s res %.2e true_res %.2e |p| %.1f err_vs_direct %.2e" % (bd, ny, nx, acc, it.tolist(), float(res.max()), true_res, np.abs(direct).max(), err)
            if ny <= 64:
                po, ito = F.conjugate_gradient(torch.tensor(rhs), bd, acc, 4000)
                tr = float((F.laplace_apply(po.double(), bd) - torch.tensor(rhs).double()).abs().max())
                o = po.numpy()
                if bd == "closed": o = dm(o)
                line += " | oracle fp32 CG: it %d true_res %.2e err %.2e" % (ito, tr, np.abs(o - direct).max() / np.abs(direct).max())
            print(line, flush=True)
