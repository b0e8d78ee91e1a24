"""Where does the pair-loss forward differ from burgers_step bitwise?  (exact solver trajectory -> residual must be exactly 0)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from pde_control_b200 import ops, datagen
rng = np.random.RandomState(2)
B, n, T = 4096, 128, 32
u0 = torch.tensor(datagen.gaussian_clash(B, n, rng), device="cuda")
frames = ops.burgers_trajectory(u0, T, 0.03, 0.003)
s = torch.cat([u0[None], frames]).detach().requires_grad_(True)
l2, l1 = ops.burgers_pair_loss(s, 0.03, 0.003)
l2.backward()
print("l2", l2.item(), "l1", l1.item())
g = s.grad[T] * B      # = d_T exactly
idx = torch.nonzero(g)
print("nonzero residual cells in last frame:", idx.shape[0])
for b, j in idx[:8].tolist():
    lo, hi = max(0, j - 2), min(n, j + 3)
    print("row", b, "cell", j, "d", g[b, j].item(), "s[T-1]", s[T - 1, b, lo:hi].tolist(), "s[T]", s[T, b, lo:hi].tolist())
