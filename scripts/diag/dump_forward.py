"""Extract the bit-exact forward step of the pair-loss kernel (s = [u, 0] -> grad[1] * B = -step(u)) next to burgers_step."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from pde_control_b200 import ops
rng = np.random.RandomState(5)
B, n = 64, 128
u = torch.tensor((rng.randn(B, n) * 6).astype(np.float32), device="cuda")
s = torch.stack([u, torch.zeros_like(u)]).requires_grad_(True)
l2, l1 = ops.burgers_pair_loss(s, 0.03, 0.003)
l2.backward()
v_pl = (-s.grad[1] * B).cpu().numpy()
v_st = ops.burgers_step(u, 0.03, 0.003).cpu().numpy()
np.savez(sys.argv[1], u=u.cpu().numpy(), v_pl=v_pl, v_st=v_st)
print("mismatching cells:", int((v_pl != v_st).sum()), "of", v_pl.size)
