"""torchrun --nproc-per-node 2: chained vs plain slab solve on smooth and rough right-hand sides, against the fp64 fast-diagonalisation."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.distributed as dist
rank = int(os.environ["RANK"]); torch.cuda.set_device(rank)
dist.init_process_group("cpu:gloo,cuda:nccl")
from pde_control_b200 import ops, slab
B, ny, nx = 2, 256, 192
st = slab.SlabStepper(B, ny, nx, halo=4); lay = st.layout
g = torch.Generator(device="cuda"); g.manual_seed(1)
for name in ("smooth", "rough", "rough_again"):
    full = torch.randn(B, ny, nx, device="cuda", generator=g)
    if name == "smooth":
        for _ in range(4): full = (full + full.roll(1, 1) + full.roll(-1, 1) + full.roll(1, 2) + full.roll(-1, 2)) / 5
    mine = full[:, lay.r0:lay.r1].contiguous()
    Ey, ly = ops._eig_1d(ny, 0, full.device); Ex, lx = ops._eig_1d(nx, 0, full.device)
    p64 = Ey @ ((Ey.T @ full.double() @ Ex) / (ly[:, None] + lx[None, :])) @ Ex.T
    pa, ia = slab._SlabSolve.apply(mine, st, 0, 1e-7, 8000)
    pb, ib = slab._SlabSolveChained.apply(mine, st, 0, 1e-7, 8000)
    pc = slab._SlabSolveChained._solve(mine, st, 0, 1e-7, 8000, None)
    ref = p64[:, lay.r0:lay.r1]
    rel = lambda a: float((a.double() - ref).abs().max() / ref.abs().max())
    print(rank, name, "plain %.2e (%d it) chained %.2e (%d it) chained-no-iters %.2e" % (rel(pa), int(ia.max()), rel(pb), int(ib.max()), rel(pc)), flush=True)
# closed domain: mean-free right-hand sides
for name in ("closed_smooth", "closed_rough"):
    full = torch.randn(B, ny, nx, device="cuda", generator=g)
    if name == "closed_smooth":
        for _ in range(4): full = (full + full.roll(1, 1) + full.roll(-1, 1) + full.roll(1, 2) + full.roll(-1, 2)) / 5
    full -= full.mean((1, 2), keepdim=True)
    mine = full[:, lay.r0:lay.r1].contiguous()
    Ey, ly = ops._eig_1d(ny, 3, full.device); Ex, lx = ops._eig_1d(nx, 3, full.device)
    lam = ly[:, None] + lx[None, :]; inv = torch.where(lam.abs() < 1e-9, torch.zeros_like(lam), 1 / lam)
    p64 = Ey @ ((Ey.T @ full.double() @ Ex) * inv) @ Ex.T
    pa, ia = slab._SlabSolve.apply(mine, st, 1, 1e-7, 8000)
    pb, ib = slab._SlabSolveChained.apply(mine, st, 1, 1e-7, 8000)
    ref = p64[:, lay.r0:lay.r1]
    dm = lambda a: a.double() - st.gather(a).double().mean((1, 2), keepdim=True)
    rel = lambda a: float((dm(a) - ref).abs().max() / ref.abs().max())
    print(rank, name, "plain %.2e (%d it) chained %.2e (%d it)" % (rel(pa), int(ia.max()), rel(pb), int(ib.max())), flush=True)
st.close(); dist.destroy_process_group()
