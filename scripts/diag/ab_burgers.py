"""A/B of Burgers pair-loss builds on one box: for every library given (PDEC_LIB_PATH variants) a subprocess runs the bench
workload (B=4096, T=32, n=128, rotating input sets): 3 x 300 back-to-back launches timed with CUDA events (min reported), and the
loss + gradient of set 0 are compared with the first library's.
usage: python scripts/diag/ab_burgers.py [name=path|name=path,ENV=1 ...]   (first entry = baseline)"""
import os, subprocess, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

WORKER = r'''
import sys, os, json
sys.path.insert(0, %r)
import torch
from benchlib.burgers import BurgersConfig2
BurgersConfig2.mode = int(os.environ.get("AB_MODE", "0"))      # 2 = periodic rows
wl = BurgersConfig2(torch.device("cuda", 0), 0, 1)
for _ in range(10): wl.step()
torch.cuda.synchronize()
best = 1e9
for rep in range(3):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(300): wl.step()
    b.record(); torch.cuda.synchronize()
    best = min(best, a.elapsed_time(b) / 300 * 1e3)
wl.i = 0
wl.step(); torch.cuda.synchronize()
out = sys.argv[1]
torch.save({"grad": wl.grad[0].cpu(), "losses": wl.losses.cpu()}, out)
print(json.dumps({"us_per_step": best, "losses": wl.losses.cpu().tolist()}))
''' % ROOT

def main():
    specs = sys.argv[1:]
    base = None
    import torch
    for spec in specs:
        name, rest = spec.split("=", 1)
        parts = rest.split(",")
        env = dict(os.environ)
        if parts[0]:
            env["PDEC_LIB_PATH"] = os.path.join(ROOT, parts[0])
        for kv in parts[1:]:
            k, v = kv.split("=")
            env[k] = v
        out = "/tmp/ab_%s.pt" % name
        r = subprocess.run([sys.executable, "-c", WORKER, out], env=env, capture_output=True, text=True)
        if r.returncode != 0:
            print(name, "FAILED", r.stderr[-2000:])
            continue
        res = json.loads(r.stdout.strip().splitlines()[-1])
        cur = torch.load(out)
        if base is None:
            base = cur
            cmp = "baseline"
        else:
            g0, g1 = base["grad"].double(), cur["grad"].double()
            cmp = "grad rel diff vs baseline %.3e, loss rel diff %.3e" % (
                float((g0 - g1).abs().max() / g0.abs().max()), abs(cur["losses"][0] - base["losses"][0]).item() / abs(base["losses"][0]).item())
        print("%-12s %7.2f us/step  losses %s  %s" % (name, res["us_per_step"], res["losses"], cmp), flush=True)

if __name__ == "__main__":
    main()
