"""Diagnostic: Burgers pair-loss step time -- single launches (events around each), back-to-back loop, CUDA-graph replay."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import bench
wl = bench.BurgersWorkload(torch.device("cuda", 0), 0)
for _ in range(5): wl.step()
torch.cuda.synchronize()
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(50)]
for a, b in ev:
    a.record(); wl.step(); b.record(); torch.cuda.synchronize()
t = sorted(a.elapsed_time(b) for a, b in ev)
print("single launches (sync between): median %.1f us, min %.1f us" % (t[25] * 1e3, t[0] * 1e3))
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(200): wl.step()
b.record(); torch.cuda.synchronize()
print("back-to-back loop: %.1f us per step" % (a.elapsed_time(b) / 200 * 1e3))
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    wl._args = None
    for _ in range(4): wl.step()
    side.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        for _ in range(wl.nsets * 4): wl.step()
    g.replay(); side.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(side)
    for _ in range(25): g.replay()
    b.record(side); side.synchronize()
print("graph replay (%d steps per graph): %.1f us per step" % (wl.nsets * 4, a.elapsed_time(b) / 25 / (wl.nsets * 4) * 1e3))
