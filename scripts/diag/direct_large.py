"""Fast-diagonalisation Poisson solve at 2048 x 2048 with library GEMMs: accuracy of fp32 (plain SGEMM, and three TF32 GEMMs on hi / lo
splits) against the same formula in fp64 and against the stencil residual, and time per solve."""
import math, torch
torch.backends.cuda.matmul.allow_tf32 = False
dev = "cuda"

def eig(n, ends, dtype=torch.float64):
    i = torch.arange(n, device=dev, dtype=torch.float64)[:, None]; k = torch.arange(n, device=dev, dtype=torch.float64)[None, :]
    if ends == 3:
        c = torch.where(k == 0, math.sqrt(1.0 / n), math.sqrt(2.0 / n)); E = c * torch.cos(math.pi * (i + 0.5) * k / n)
        lam = -2.0 + 2.0 * torch.cos(math.pi * k[0] / n)
    elif ends == 0:
        E = math.sqrt(2.0 / (n + 1)) * torch.sin(math.pi * (i + 1) * (k + 1) / (n + 1)); lam = -2.0 + 2.0 * torch.cos(math.pi * (k[0] + 1) / (n + 1))
    else:
        raise NotImplementedError
    return E.to(dtype), lam

def apply_A(p, closed):           # 5-point Laplacian, zero outside (open) or one neighbour fewer (closed)
    q = torch.nn.functional.pad(p, (1, 1, 1, 1), mode="replicate" if closed else "constant")
    return q[:, :-2, 1:-1] + q[:, 2:, 1:-1] + q[:, 1:-1, :-2] + q[:, 1:-1, 2:] - 4 * p

def split(x):
    hi = (x.view(torch.int32) & -8192).view(torch.float32); return hi, x - hi

def mm3(a, b):                   # three TF32 tensor-core GEMMs on exact splits
    torch.backends.cuda.matmul.allow_tf32 = True
    ah, al = split(a); bh, bl = split(b)
    r = ah @ bh + (al @ bh + ah @ bl)
    torch.backends.cuda.matmul.allow_tf32 = False
    return r

for n, closed in ((2048, False), (2048, True), (1024, False)):
    ends = 3 if closed else 0
    E64, lam = eig(n, ends); E32 = E64.float()
    s = lam[:, None] + lam[None, :]
    inv64 = torch.where(s.abs() < 1e-9, torch.zeros_like(s), 1.0 / s); inv32 = inv64.float()
    rhs = torch.randn(1, n, n, device=dev)
    for _ in range(3): rhs = (rhs + rhs.roll(1, 1) + rhs.roll(-1, 1) + rhs.roll(1, 2) + rhs.roll(-1, 2)) / 5
    if closed: rhs = rhs - rhs.mean()
    p64 = E64 @ ((E64.T @ rhs.double() @ E64) * inv64) @ E64.T
    res64 = (apply_A(p64, closed) - rhs.double()).abs().max().item()
    def solve32(mm):
        return mm(mm(E32, (mm(mm(E32.T.contiguous(), rhs[0]), E32) * inv32)), E32.T.contiguous())[None]
    for name, mm in (("sgemm", lambda a, b: a @ b), ("3xtf32", mm3)):
        p = solve32(mm); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10): p = solve32(mm)
        b.record(); torch.cuda.synchronize()
        err = ((p.double() - p64).abs().max() / p64.abs().max()).item()
        res = (apply_A(p.double(), closed) - rhs.double()).abs().max().item()
        print("n=%d closed=%d %-7s: %.3f ms per solve, rel err vs fp64 %.2e, max residual %.2e (fp64 formula: %.1e), |rhs|max %.2f |p|max %.1f"
              % (n, closed, name, a.elapsed_time(b) / 10, err, res, res64, rhs.abs().max().item(), p64.abs().max().item()), flush=True)
