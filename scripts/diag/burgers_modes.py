"""Pair-loss kernel time per boundary mode at BASELINE config 2's size (which kernel serves which mode: DESIGN 4.1)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

import bench
from pde_control_b200 import ops, _lib

wl = bench.BurgersWorkload(torch.device("cuda", 0), 0)
T1, B, n = wl.s[0].shape
for name, mode in (("constant", _lib.MODE_CONSTANT), ("replicate", _lib.MODE_REPLICATE), ("circular", _lib.MODE_CIRCULAR)):
    def step(i=[0]):
        k = i[0] % wl.nsets
        i[0] += 1
        ops.burgers_pair_loss_raw(wl.s[k], wl.grad[k], wl.losses, wl.ws, bench.DT, bench.VISC, mode=mode)
    for _ in range(10):
        step()
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(200):
            step()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b) / 200 * 1e3)
    gbs = 8.0 * (T1 - 1) * B * n / (best * 1e-6) / 1e9
    print("%-9s %.2f us/step  %.0f GB/s algorithmic  losses %s" % (name, best, gbs, wl.losses.tolist()), flush=True)
