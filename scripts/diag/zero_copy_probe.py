"""Does the fused pair-loss kernel run well straight out of / into page-locked HOST memory (UVA zero copy, no copy engines)?
Variants: s from host + grad to host; one direction only; a copy engine doing the other direction at the same time (the steady state of
a chunked pipeline); host-pipeline chunk sizes.  Checks the zero-copy result against the resident call."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch

from pde_control_b200 import ops, _lib

T, B, n = 32, 4096, 128
rng = np.random.RandomState(0)
s = (0.5 * rng.randn(T + 1, B, n)).astype(np.float32)
for _ in range(4):
    s = (s + np.roll(s, 1, -1) + np.roll(s, -1, -1)) / 3
hs = torch.tensor(s).pin_memory()
hg = torch.zeros_like(hs).pin_memory()
ds = hs.cuda()
dg = torch.zeros_like(ds)
losses = torch.zeros(2, device="cuda")
L = _lib.lib()
ws_bytes = L.pdec_burgers_pair_loss_workspace_bytes(B, n, T)
ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")


def call(sp, gp):
    rc = L.pdec_burgers_pair_loss(sp, gp, _lib.ptr(losses), T, B, n, 0, 0.03, 0.003, 0.0, 1.0, 1, 0, 1.0, 0.0, _lib.ptr(ws), ws_bytes,
                                  _lib.stream_ptr())
    assert rc == 0, rc


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


hp, gp = ctypes.c_void_p(hs.data_ptr()), ctypes.c_void_p(hg.data_ptr())
print("resident          %.3f ms" % timed(lambda: call(_lib.ptr(ds), _lib.ptr(dg))))
ref_l = losses.cpu().clone()
print("s host, g device  %.3f ms" % timed(lambda: call(hp, _lib.ptr(dg))))
print("s device, g host  %.3f ms" % timed(lambda: call(_lib.ptr(ds), gp)))
print("s host, g host    %.3f ms" % timed(lambda: call(hp, gp)))
torch.cuda.synchronize()
print("losses equal", torch.equal(losses.cpu(), ref_l), "grad equal", torch.equal(hg, dg.cpu()), "(%.1f MB each way)" % (hs.numel() * 4e-6))

side = torch.cuda.Stream()
other_h = torch.empty_like(hs).pin_memory()
other_d = torch.empty_like(ds)


def hybrid(kernel_call, copy_call):
    def fn():
        ev = torch.cuda.Event()
        ev.record()
        side.wait_event(ev)
        with torch.cuda.stream(side):
            copy_call()
        kernel_call()
        torch.cuda.current_stream().wait_stream(side)
    return fn


print("kernel reads host + CE D2H   %.3f ms" % timed(hybrid(lambda: call(hp, _lib.ptr(dg)), lambda: other_h.copy_(other_d, non_blocking=True))))
print("CE H2D + kernel writes host  %.3f ms" % timed(hybrid(lambda: call(_lib.ptr(ds), gp), lambda: other_d.copy_(other_h, non_blocking=True))))
print("CE both ways                 %.3f ms" % timed(hybrid(lambda: ds.copy_(hs, non_blocking=True), lambda: other_h.copy_(other_d, non_blocking=True))))

hl = torch.zeros(2).pin_memory()
for staged in (True, False):
    if staged:
        os.environ["PDEC_PIPE_STAGED_D2H"] = "1"
    else:
        os.environ.pop("PDEC_PIPE_STAGED_D2H", None)
    for rows in (1024, 512, 256, 128):
        pipe = ops.HostPipeline(n, T, chunk_rows=rows, max_batch=B)
        t = timed(lambda: pipe.pair_loss(hs, hg, hl, 0.03, 0.003))
        torch.cuda.synchronize()
        ok = torch.equal(hg, dg.cpu())
        print("host pipeline %s chunk_rows=%4d: %.3f ms  (grad equal: %s)" % ("staged D2H" if staged else "direct    ", rows, t, ok))
        pipe.close()
