"""A/B of two libpdec builds on the config-3 trajectory: bit-identity of every output and gradient, and the replay time.
usage: python scripts/diag/ab_fluid_bits.py name=path [name=path ...]   (path relative to the repo root; empty = the built library)"""
import os, subprocess, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
WORKER = r'''
import sys, json
sys.path.insert(0, %r)
import torch
from benchlib import inputs
from pde_control_b200 import graphs
rho0, vy0, vx0, target = inputs.fluid_config3_inputs(64, 128, 0)
args = [torch.tensor(a, device="cuda") for a in (rho0, vy0, vx0, target)]
res = {}
saved = {}
for closed in (False, True):
    g = graphs.GraphedFunction(graphs.fluid_trajectory(16, 1.0, closed, direct=True), args, grad_inputs=(0, 1, 2))
    for _ in range(3): g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(9):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    res["closed" if closed else "open"] = sorted(ts)[4]
    saved["closed" if closed else "open"] = [t.cpu() for t in list(g.outputs[:4]) + list(g.grads)]
torch.save(saved, sys.argv[1])
print(json.dumps(res))
''' % ROOT
import torch
base = None
for spec in sys.argv[1:]:
    name, path = spec.split("=", 1)
    env = dict(os.environ)
    if path:
        env["PDEC_LIB_PATH"] = os.path.join(ROOT, path)
    out = "/tmp/abf_%s.pt" % name
    r = subprocess.run([sys.executable, "-c", WORKER, out], env=env, capture_output=True, text=True)
    if r.returncode != 0:
        print(name, "FAILED", r.stderr[-1500:]); continue
    cur = torch.load(out)
    same = "baseline" if base is None else "bit-identical to baseline: %s" % all(torch.equal(a, b) for k in cur for a, b in zip(cur[k], base[k]))
    base = base or cur
    print(name, r.stdout.strip().splitlines()[-1], same, flush=True)
