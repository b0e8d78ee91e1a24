"""A/B of the 128 x 128 Poisson CG: one CTA per element (cg_block_kernel) against a cluster of two CTAs per element (cg_pair_kernel)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch

from pde_control_b200 import ops

rng = np.random.RandomState(0)
for B in (1, 16, 64, 74, 148):
    rhs = torch.tensor(rng.randn(B, 128, 128).astype(np.float32)).cuda()
    for closed in (False, True):
        if closed:
            rhs = rhs - rhs.mean(dim=(1, 2), keepdim=True)
        line = "B=%d closed=%d:" % (B, closed)
        for name, variant in (("block", 4), ("pair", 6), ("p2p", 7)):
            out, it = torch.empty_like(rhs), torch.zeros(B, dtype=torch.int32, device="cuda")
            args = (rhs, out, closed, 1e-5, 2000, None, it, None, variant << 8)
            for _ in range(3):
                ops.cg_poisson_raw(*args)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(10):
                ops.cg_poisson_raw(*args)
            b.record()
            torch.cuda.synchronize()
            ms = a.elapsed_time(b) / 10
            line += "  %s %.3f ms, %d it, %.2f us/it" % (name, ms, int(it.max()), ms * 1e3 / int(it.max()))
        print(line, flush=True)
