"""Small cases of the kernels added late in round 1, for compute-sanitizer (memcheck / racecheck): masked on-chip CG, the CTA-pair CG
(both synchronisation variants), the strided-gradient unit kernel storing into page-locked host memory through the host pipeline."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch

from pde_control_b200 import ops

rng = np.random.RandomState(0)
# masked CG, ragged and full sizes
for ny, nx in ((24, 32), (97, 128), (128, 128)):
    active = np.ones((ny, nx), np.float32)
    active[ny // 4:ny // 2, nx // 3:2 * nx // 3] = 0
    masks = ops.ObstacleMasks(torch.tensor(active, device="cuda"), False)
    rhs = torch.tensor(rng.randn(2, ny, nx).astype(np.float32) * active, device="cuda")
    out, it = torch.empty_like(rhs), torch.zeros(2, dtype=torch.int32, device="cuda")
    ops.cg_poisson_masked_raw(rhs, out, masks.code, 1e-4, 60, None, it, None, 0)
    torch.cuda.synchronize()
    print("masked", ny, nx, it.tolist(), bool(torch.isfinite(out).all()))
# CTA-pair CG
rhs = torch.tensor(rng.randn(3, 128, 128).astype(np.float32), device="cuda")
for variant in (4, 6, 7):
    out, it = torch.empty_like(rhs), torch.zeros(3, dtype=torch.int32, device="cuda")
    ops.cg_poisson_raw(rhs, out, False, 1e-3, 40, None, it, None, variant << 8)
    torch.cuda.synchronize()
    print("cg variant", variant, it.tolist(), float(out.abs().max()))
# host pipeline, direct gradient
T, B, n = 4, 300, 128
s = (0.3 * rng.randn(T + 1, B, n)).astype(np.float32)
hs, hg, hl = torch.tensor(s).pin_memory(), torch.zeros(T + 1, B, n).pin_memory(), torch.zeros(2).pin_memory()
pipe = ops.HostPipeline(n, T, chunk_rows=128, max_batch=512)
pipe.pair_loss(hs, hg, hl, 0.03, 0.003)
torch.cuda.current_stream().synchronize()
sc = torch.tensor(s, device="cuda", requires_grad=True)
l2, l1 = ops.burgers_pair_loss(sc, 0.03, 0.003)
l2.backward()
print("pipeline", float(hl[0]), float(l2), float((hg - sc.grad.cpu()).abs().max()))
pipe.close()
