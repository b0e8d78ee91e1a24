"""Config-3 trajectory (128^2, 16 steps, B=64, fwd+adjoint, one CUDA graph): CG at 1e-5 / 1e-7 against the direct solver; and the
solver kernels alone."""
import os, sys, json
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
from benchlib import inputs
from pde_control_b200 import graphs, ops
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
for closed in (False, True):
    for name, kw in (("cg 1e-5", dict(accuracy=1e-5)), ("cg 1e-7", dict(accuracy=1e-7)), ("direct", dict(direct=True))):
        g = graphs.GraphedFunction(graphs.fluid_trajectory(16, 1.0, closed, **kw), args, grad_inputs=(0, 1, 2))
        out["%s %s" % ("closed" if closed else "open", name)] = {"ms_trajectory": timeit(g.replay), "loss": float(g.outputs[0]),
                                                                  "mean_fwd_iterations": float(g.outputs[4].float().mean())}
rhs = ops.mask_divergence(args[1], args[2], False)
p = torch.empty_like(rhs)
out["solver kernel alone: direct, us"] = 1e3 * timeit(lambda: ops.poisson_direct_raw(rhs, p, False), 21)
it = torch.zeros(64, dtype=torch.int32, device="cuda")
out["solver kernel alone: cg 1e-5, us"] = 1e3 * timeit(lambda: ops.cg_poisson_raw(rhs, p, False, 1e-5, 2000, None, it, None), 21)
out["cg iterations mean/max"] = [float(it.float().mean()), int(it.max())]
print(json.dumps(out, indent=1))
