"""2048 x 2048: how many CG iterations does the CORRECTION system A d = rhs - A p0 need (p0 = the fp32 library-GEMM direct solve) to
reach the reference's stopping rule max|r| < 1e-5, against a CG from zero on the original system?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from pde_control_b200 import ops
n = 2048
g = torch.Generator(device="cuda"); g.manual_seed(3)
def smooth(x, passes=6):
    for _ in range(passes): x = (x + torch.roll(x, 1, -1) + torch.roll(x, -1, -1) + torch.roll(x, 1, -2) + torch.roll(x, -1, -2)) / 5
    return x / x.abs().max()
vy = smooth(torch.randn(1, n + 1, n, device="cuda", generator=g)); vx = smooth(torch.randn(1, n, n + 1, device="cuda", generator=g))
div = ops.mask_divergence(vy, vx, False)
p0 = torch.empty_like(div); ops.poisson_direct_raw(div, p0, False)
def A(p):
    q = torch.nn.functional.pad(p, (1, 1, 1, 1))
    return q[:, :-2, 1:-1] + q[:, 2:, 1:-1] + q[:, 1:-1, :-2] + q[:, 1:-1, 2:] - 4 * p
r0 = div - A(p0)
print("|div| %.3f  |p0| %.1f  |r0| fp32 %.3e  |r0| fp64 %.3e" % (div.abs().max(), p0.abs().max(), r0.abs().max(), (div.double() - A(p0.double())).abs().max()))
for acc in (1e-4, 1e-5):
    it = torch.zeros(1, dtype=torch.int32, device="cuda"); d = torch.empty_like(div)
    ops.cg_poisson_raw(r0, d, False, acc, 20000, None, it, None)
    p = p0 + d
    it0 = torch.zeros(1, dtype=torch.int32, device="cuda"); pz = torch.empty_like(div)
    ops.cg_poisson_raw(div, pz, False, acc, 20000, None, it0, None)
    res = lambda x: float((div.double() - A(x.double())).abs().max())
    print("accuracy %.0e: correction solve %d iterations (true residual %.2e) | from zero %d iterations (true residual %.2e) | |p - pz|/|pz| %.2e"
          % (acc, int(it), res(p), int(it0), res(pz), float((p - pz).abs().max() / pz.abs().max())), flush=True)
