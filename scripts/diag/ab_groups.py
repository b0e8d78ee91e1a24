"""Config-3 trajectory (128^2, 16 steps, B=64, fwd+adjoint, one CUDA graph) with the batch cut into sub-batches on parallel graph branches."""
import os, sys, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
from benchlib import inputs
from pde_control_b200 import graphs
rho0, vy0, vx0, target = inputs.fluid_config3_inputs(64, 128, 0)
args = [torch.tensor(a, device="cuda") for a in (rho0, vy0, vx0, target)]

def timeit(fn, reps=7):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]

out = {}
ref = None
for name, kw in (("cg 1e-5", dict(accuracy=1e-5)), ("cg 1e-7", dict(accuracy=1e-7)), ("direct", dict(direct=True))):
    for groups in (1, 2, 4, 8, 16):
        g = graphs.GraphedFunction(graphs.fluid_trajectory(16, 1.0, False, groups=groups, **kw), args, grad_inputs=(0, 1, 2))
        ms = timeit(g.replay)
        grads = [x.clone() for x in g.grads]
        if groups == 1:
            ref = grads
        same = all(torch.equal(a, b) for a, b in zip(grads, ref))
        out["%s groups=%d" % (name, groups)] = {"ms": ms, "loss": float(g.outputs[0]), "grads_bit_identical_to_groups_1": same}
        print(name, groups, ms, same, flush=True)
        del g
        torch.cuda.empty_cache()
print(json.dumps(out, indent=1))
