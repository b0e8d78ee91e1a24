"""BASELINE config 4: Burgers batch sweep (1e4 .. 1e6 trajectories in total), batch-sharded over the ranks (torchrun), one fused
force-loss fwd+adjoint per step plus ONE NCCL all-reduce of an OP-network-sized gradient bucket (32 641 floats + the two loss
scalars).  Time = max over ranks (CUDA events, barrier on both sides).  One JSON line per total batch on rank 0."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from pde_control_b200 import ops, datagen, _lib

world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
n, T, DT, NU = 128, 32, 0.03, 0.003
totals = [int(float(x)) for x in (sys.argv[1:] or ["1e4", "1e5", "1e6"])]
for total in totals:
    B = total // world
    rng = np.random.RandomState(1 + rank)
    base = torch.tensor(datagen.gaussian_clash(min(B, 4096), n, rng), device=dev)
    u0 = base.repeat((B + base.shape[0] - 1) // base.shape[0], 1)[:B].contiguous()
    s = torch.empty(T + 1, B, n, device=dev); s[0] = u0
    s[1:] = ops.burgers_trajectory(u0, T, DT, NU, substeps=4, resubstep=True)
    s += 0.02 * torch.randn_like(s)
    grad = torch.empty_like(s); losses = torch.zeros(2, device=dev)
    ws = torch.empty((_lib.lib().pdec_burgers_pair_loss_workspace_bytes(B, n, T) + 3) // 4, device=dev)
    bucket = torch.randn(32641 + 2, device=dev)
    def step():
        ops.burgers_pair_loss_raw(s, grad, losses, ws, DT, NU)
        bucket[-2:].copy_(losses)
        if world > 1:
            dist.all_reduce(bucket)
    def barrier():
        if world > 1: dist.barrier()
        torch.cuda.synchronize()
    for _ in range(3): step()
    barrier()
    reps = 20 if total <= 1e5 else 5
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): step()
    b.record(); barrier()
    ms = torch.tensor([a.elapsed_time(b) / reps], device=dev, dtype=torch.float64)
    if world > 1: dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        ms = float(ms)
        print(json.dumps({"workload": "config 4: burgers_pair_loss res=128 T=32, total batch %d over %d GPU(s) + all-reduce of 32643 floats" % (B * world, world),
                          "n_gpus": world, "batch_per_gpu": B, "ms_per_step": ms, "cell_updates_per_s": T * B * world * n / (ms * 1e-3),
                          "hbm_GBps_per_gpu_algorithmic": 8.0 * T * B * n / (ms * 1e-3) / 1e9, "bytes_resident_per_gpu": 2 * s.numel() * 4}), flush=True)
    del s, grad, ws
    torch.cuda.empty_cache()
if world > 1:
    dist.destroy_process_group()
