"""One eager config-3 trajectory (128^2, 16 steps, B=64, fwd + adjoint) after one warm-up, for an ncu launch list:
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python scripts/diag/fluid_one_trajectory.py [accuracy]"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
from benchlib import inputs
from pde_control_b200 import graphs
acc = float(sys.argv[1]) if len(sys.argv) > 1 else 1e-5
rho0, vy0, vx0, target = inputs.fluid_config3_inputs(64, 128, 0)
args = [torch.tensor(a, device="cuda") for a in (rho0, vy0, vx0, target)]
fn = graphs.fluid_trajectory(16, 1.0, False, accuracy=acc)
for rep in range(2):
    leaves = [a.clone().requires_grad_(True) for a in args[:3]]
    torch.cuda.synchronize()
    if rep == 1:
        torch.cuda.nvtx.range_push("trajectory")
    out = fn(*leaves, args[3])
    grads = torch.autograd.grad(out[0], leaves)
    torch.cuda.synchronize()
    if rep == 1:
        torch.cuda.nvtx.range_pop()
print("loss", float(out[0]), "mean fwd iterations", float(out[4].float().mean()))
