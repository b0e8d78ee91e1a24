"""PCIe legs of the host pipeline: contiguous H2D / D2H alone and together, then the pipeline (losses only, with gradient) for chunk sizes."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from pde_control_b200 import ops
T, B, n = 32, 4096, 128
hs = torch.randn(T + 1, B, n).pin_memory(); hg = torch.empty(T + 1, B, n).pin_memory(); hl = torch.zeros(2).pin_memory()
ds = torch.empty(T + 1, B, n, device="cuda"); dg = torch.randn(T + 1, B, n, device="cuda")
mb = hs.numel() * 4 / 1e6
def timeit(fn, reps=20):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
t = timeit(lambda: ds.copy_(hs, non_blocking=True)); print("H2D contiguous %.0f MB: %.3f ms = %.1f GB/s" % (mb, t, mb / t))
t = timeit(lambda: hg.copy_(dg, non_blocking=True)); print("D2H contiguous: %.3f ms = %.1f GB/s" % (t, mb / t))
def both():
    cur = torch.cuda.current_stream()
    s1.wait_stream(cur); s2.wait_stream(cur)
    with torch.cuda.stream(s1): ds.copy_(hs, non_blocking=True)
    with torch.cuda.stream(s2): hg.copy_(dg, non_blocking=True)
    cur.wait_stream(s1); cur.wait_stream(s2)
t = timeit(both); print("H2D + D2H concurrently: %.3f ms = %.1f GB/s each way" % (t, mb / t))
for fps in (1, 2, 4, 8):
    pipe = ops.HostPipeline(n, T, chunk_rows=512, max_batch=B, frames_per_slice=fps)
    t0 = timeit(lambda: pipe.pair_loss(hs, None, hl, 0.03, 0.003))
    t1 = timeit(lambda: pipe.pair_loss(hs, hg, hl, 0.03, 0.003))
    print("pipeline frames_per_slice=%d: losses only %.3f ms, with gradient %.3f ms" % (fps, t0, t1))
    pipe.close()
for rows in (512,):
    pipe = ops.HostPipeline(n, T, chunk_rows=rows, max_batch=B, frames_per_slice=0)
    t0 = timeit(lambda: pipe.pair_loss(hs, None, hl, 0.03, 0.003))
    t1 = timeit(lambda: pipe.pair_loss(hs, hg, hl, 0.03, 0.003))
    print("pipeline chunk_rows=%d: losses only %.3f ms, with gradient %.3f ms" % (rows, t0, t1))
    pipe.close()
