"""A/B on one box: config-3 trajectory (128^2, 16 steps, B=64, fwd+adjoint, CUDA graph) with the fused projection launch and with
the three-operator composition; several alternating repetitions."""
import os, sys, subprocess, json
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..")
code = r'''
import sys, torch, json
sys.path.insert(0, %r)
from benchlib import inputs
from pde_control_b200 import graphs
rho0, vy0, vx0, target = inputs.fluid_config3_inputs(64, 128, 0)
args = [torch.tensor(a, device="cuda") for a in (rho0, vy0, vx0, target)]
out = {}
for acc in (1e-5, 1e-7):
    g = graphs.GraphedFunction(graphs.fluid_trajectory(16, 1.0, False, accuracy=acc), args, grad_inputs=(0, 1, 2))
    for _ in range(3): g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(7):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    ts.sort()
    out[str(acc)] = {"ms_median": ts[3], "ms_min": ts[0], "mean_fwd_iterations": float(g.outputs[4].float().mean()), "max_fwd_iterations": int(g.outputs[4].max())}
print(json.dumps(out))
''' % os.path.abspath(ROOT)
for rep in range(2):
    for env, name in (({"PDEC_FUSED_PROJECTION": "1"}, "fused"), ({}, "three operators")):
        e = dict(os.environ); e.update(env)
        r = subprocess.run([sys.executable, "-c", code], env=e, capture_output=True, text=True)
        print(name, r.stdout.strip() or r.stderr[-800:], flush=True)
