"""The two config-3 solver kernels alone (B = 64, 128 x 128, right-hand side = divergence of the bench's velocity), a few launches each:
ncu --set full -k regex:'poisson_direct_kernel|cg_block_kernel' target."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
from benchlib import inputs
from pde_control_b200 import ops
rho0, vy0, vx0, target = inputs.fluid_config3_inputs(64, 128, 0)
vy, vx = torch.tensor(vy0, device="cuda"), torch.tensor(vx0, device="cuda")
rhs = ops.mask_divergence(vy, vx, False)
p = torch.empty_like(rhs)
it = torch.zeros(64, dtype=torch.int32, device="cuda")
for _ in range(4):
    ops.poisson_direct_raw(rhs, p, False)
    ops.cg_poisson_raw(rhs, p, False, 1e-5, 2000, None, it, None)
torch.cuda.synchronize()
print("cg iterations mean/max", float(it.float().mean()), int(it.max()))
