"""Timeline of chunked H2D on one stream and D2H on another (D2H chunk c waits for H2D chunk c): start/end of every copy."""
import sys, torch
nch = int(sys.argv[1]) if len(sys.argv) > 1 else 8
N = 33 * 4096 * 128
hs = torch.randn(N).pin_memory(); hg = torch.empty(N).pin_memory()
ds = torch.empty(N, device="cuda"); dg = torch.randn(N, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
per = N // nch
def run(record):
    ev = []
    t0 = torch.cuda.Event(enable_timing=True); t0.record()
    s1.wait_event(t0); s2.wait_event(t0)
    for c in range(nch):
        a, b, d, e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        with torch.cuda.stream(s1):
            a.record(); ds[c * per:(c + 1) * per].copy_(hs[c * per:(c + 1) * per], non_blocking=True); b.record()
        with torch.cuda.stream(s2):
            s2.wait_event(b)
            d.record(); hg[c * per:(c + 1) * per].copy_(dg[c * per:(c + 1) * per], non_blocking=True); e.record()
        ev.append((a, b, d, e))
    torch.cuda.synchronize()
    if record:
        for c, (a, b, d, e) in enumerate(ev):
            print("chunk %2d  H2D %.3f-%.3f (%.1f GB/s)   D2H %.3f-%.3f (%.1f GB/s)" % (
                c, t0.elapsed_time(a), t0.elapsed_time(b), per * 4e-6 / a.elapsed_time(b), t0.elapsed_time(d), t0.elapsed_time(e), per * 4e-6 / d.elapsed_time(e)))
run(False); run(True)
