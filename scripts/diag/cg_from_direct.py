"""How many CG iterations does the reference's stopping rule (max|r| <= accuracy) still need when the solve starts from the
tensor-core direct solution instead of zero?  Right-hand sides = divergences of the config-3 bench inputs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from pde_control_b200 import ops
from benchlib import inputs
rho0, vy0, vx0, target = inputs.fluid_config3_inputs(64, 128, 0)
vy, vx = torch.tensor(vy0).cuda(), torch.tensor(vx0).cuda()
for closed in (False, True):
    div = ops.mask_divergence(vy, vx, closed)
    for acc in (1e-5, 1e-6, 1e-7):
        B = div.shape[0]
        it0 = torch.zeros(B, dtype=torch.int32, device="cuda"); it1 = torch.zeros_like(it0)
        p0, p1, pd = torch.empty_like(div), torch.empty_like(div), torch.empty_like(div)
        ops.cg_poisson_raw(div, p0, closed, acc, 2000, None, it0, None)
        ops.poisson_direct_raw(div, pd, closed)
        ops.cg_poisson_raw(div, p1, closed, acc, 2000, pd, it1, None)
        torch.cuda.synchronize()
        def t(fn):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(20): fn()
            b.record(); torch.cuda.synchronize(); return a.elapsed_time(b) / 20 * 1e3
        t0 = t(lambda: ops.cg_poisson_raw(div, p0, closed, acc, 2000, None, it0, None))
        t1 = t(lambda: (ops.poisson_direct_raw(div, pd, closed), ops.cg_poisson_raw(div, p1, closed, acc, 2000, pd, it1, None)))
        print("closed=%d acc=%.0e: from zero %d..%d iterations %.0f us | from direct %d..%d iterations %.0f us | |p1-p0|/|p0| = %.2e"
              % (closed, acc, int(it0.min()), int(it0.max()), t0, int(it1.min()), int(it1.max()), t1, float((p1 - p0).abs().max() / p0.abs().max())), flush=True)
