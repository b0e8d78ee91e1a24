"""BASELINE config 5: 2-D incompressible flow 2048 x 2048 (B=1), STEPS projected steps forward + adjoint, the pressure solve
slab-decomposed over the GPUs (ops.SlabSolver).  CG accuracy 1e-5; the reference's max_iterations=2000 (sparse.py:42) would cap
the solve at this size (4500+ iterations needed), so max_iterations=20000 and the iteration counts are reported.
Run: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_fluid_2048.py [steps] [n]"""
import os, sys, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import torch.distributed as dist
from pde_control_b200 import ops

world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29590")
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
STEPS = int(sys.argv[1]) if len(sys.argv) > 1 else 4
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
closed = False
g = torch.Generator(device="cuda"); g.manual_seed(3)                   # same fields on every rank (replicated)
def smooth(x, passes=6):
    for _ in range(passes):
        x = (x + torch.roll(x, 1, -1) + torch.roll(x, -1, -1) + torch.roll(x, 1, -2) + torch.roll(x, -1, -2)) / 5
    return x / x.abs().max()
rho0 = (smooth(torch.randn(1, n, n, device="cuda", generator=g)) > 0.2).float()
vy0 = smooth(torch.randn(1, n + 1, n, device="cuda", generator=g)); vx0 = smooth(torch.randn(1, n, n + 1, device="cuda", generator=g))
target = (smooth(torch.randn(1, n, n, device="cuda", generator=g)) > 0.2).float()
solver = ops.SlabSolver(1, n, n)

def trajectory():
    rho, vy, vx = rho0.clone().requires_grad_(True), vy0.clone().requires_grad_(True), vx0.clone().requires_grad_(True)
    s = (rho, vy, vx); its = []
    for _ in range(STEPS):
        out = ops.incompressible_step(*s, 0.8, closed, accuracy=1e-5, max_iterations=20000, slab=solver)
        s = out[:3]; its.append(out[4])
    loss = ((s[0] - target) ** 2).sum() / 2
    grads = torch.autograd.grad(loss, (rho, vy, vx))
    return loss, grads, its

loss, grads, its = trajectory()                      # warm-up
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); loss, grads, its = trajectory(); e1.record(); torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    ms = float(ms)
    print(json.dumps({"workload": "incompressible_flow %dx%d B=1 steps=%d OPEN fwd+adjoint, slab CG" % (n, n, STEPS), "n_gpus": world,
                      "ms_per_trajectory": ms, "ms_per_step": ms / STEPS, "cell_updates_per_s": n * n * STEPS / (ms * 1e-3),
                      "cg_iterations_forward": [int(i.max()) for i in its], "loss": float(loss),
                      "grad_norms": [float(x.norm()) for x in grads]}), flush=True)
solver.close()
dist.destroy_process_group()
