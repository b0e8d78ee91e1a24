"""BASELINE config 5 building block: CG Poisson solve on 2048 x 2048 (B=1), row slabs over the GPUs of one box.
Run: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_slab_cg.py [n]
(N = 1: the single-GPU path of pdec_cg_poisson_2d).  Prints per-iteration time (max over ranks, CUDA events)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import torch.distributed as dist
from pde_control_b200 import ops

def smooth(rng, shape, scale, passes=2):
    a = rng.randn(*shape).astype(np.float32)
    for _ in range(passes):
        a = (a + np.roll(a, 1, -1) + np.roll(a, -1, -1) + np.roll(a, 1, -2) + np.roll(a, -1, -2)) / 5
    return (a / np.abs(a).max() * scale).astype(np.float32)

world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
B = 1
rng = np.random.RandomState(0)
full = smooth(rng, (B, n, n), 1.0)
rows = n // world
for closed in (False, True):
    rhs_full = full - full.mean((1, 2), keepdims=True) if closed else full
    mine = torch.tensor(rhs_full[:, rank * rows:(rank + 1) * rows]).cuda().contiguous()
    out = torch.empty_like(mine)
    it = torch.zeros(B, dtype=torch.int32, device="cuda"); res = torch.zeros(B, device="cuda")
    comm = ops.SlabComm(B, n) if world > 1 else None
    for acc, mi in ((0.0, 500), (1e-5, 20000)):
        def solve():
            if world > 1:
                ops.cg_poisson_slab_raw(mine, out, closed, acc, mi, ny_global=n, row0=rank * rows, comm=comm, iterations=it, residual=res)
            else:
                ops.cg_poisson_raw(mine, out, closed, acc, mi, None, it, res)
        solve(); torch.cuda.synchronize()
        if world > 1: dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); solve(); e1.record(); torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
        if world > 1: dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        k = int(it.max())
        if rank == 0:
            print(json.dumps({"workload": "cg_poisson %dx%d B=%d closed=%d" % (n, n, B, closed), "n_gpus": world, "accuracy": acc, "iterations": k,
                              "residual": float(res.max()), "ms_per_solve": float(ms), "us_per_iteration": float(ms) * 1e3 / max(k, 1),
                              "algorithmic_GBps_36B_per_cell_iter": 36.0 * B * n * n * k / (float(ms) * 1e-3) / 1e9}), flush=True)
    if comm is not None: comm.close()
if world > 1: dist.destroy_process_group()
