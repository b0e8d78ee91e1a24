#!/usr/bin/env python
"""bench.py -- headline measurement of the B200-native PDE-Control physics path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload burgers|burgers_periodic|fluid]

Metric (BASELINE.json): forward+adjoint cell-updates/s.  One "step" = one pass of the hot path over one batch:

  burgers (default; BASELINE config 2): the fused force-loss + adjoint (pdec_burgers_pair_loss) over
      s[33, 4096, 128] per GPU = 32*4096*128 = 1.68e7 cell-updates per step.
  fluid (BASELINE config 3): 16 projected IncompressibleFlow steps + adjoint at 128x128, batch 64 per GPU.

N > 1: launched by torchrun, one rank per GPU; the batch is sharded (every rank owns its own 4096 trajectories,
weak scaling); the physics has no data-path collective; time = max over ranks.

Lines printed: ONE JSON line on rank 0 (see the contract in the task statement).
`--impl reference`: the CPU oracle port (oracle/restate.py, PyTorch-CPU fp32 restatement of the reference's TF
semantics incl. autograd) on all host threads, bounded sample.  The reference itself is Python over
TensorFlow 1.14 (not installable) / NumPy (cannot travel to the GPU box, forward-only) -- see DESIGN.md.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

L2_BYTES = 126 * 1024 * 1024
DT, VISC = 0.03, 0.003


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        threading.Thread.__init__(self, daemon=True)
        self.index = index
        self.samples = []
        self.stop_flag = False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [x.strip() for x in out.strip().split(",")]
                if len(parts) >= 7:
                    self.samples.append(parts)
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(s[0]) for s in self.samples)
        reasons = []
        for k, name in ((3, "hw_slowdown"), (4, "hw_thermal_slowdown"), (5, "sw_thermal_slowdown"), (6, "sw_power_cap")):
            if any(s[k].lower().startswith("active") for s in self.samples):
                reasons.append(name)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.samples[0][1]), "reasons": reasons,
                "samples": len(self.samples)}


# ----------------------------------------------------------------------------------------------------
# workloads
# ----------------------------------------------------------------------------------------------------

def _burgers_frames(B, n, T, seed):
    """Seeded stand-in for the OP-network predictions: frames of a GaussianClash
    trajectory, perturbed so that every pair has a non-zero residual."""
    import numpy as np
    from pde_control_b200 import datagen
    rng = np.random.RandomState(seed)
    u0 = datagen.gaussian_clash(B, n, rng)
    return u0, rng


class BurgersWorkload:
    name = "burgers_pair_loss res=128 T=32 B=4096/gpu (BASELINE config 2: force loss fwd+adjoint)"
    algorithmic_bytes_per_cell_update = 8.0     # read each frame once + write its gradient once
    dtype = "f32"
    dominant_kernel = "burgers_pair_loss_unit_kernel"
    ncu_traffic_bytes = 129.0e6     # steady state: dram read 66.1 MB + write 62.9 MB per launch, profiles/r1k_burgers_summary.md
    mode = 0                        # PDEC_MODE_CONSTANT: the reference's Burgers domain (OPEN boundaries -> 'constant' extrapolation)

    def __init__(self, device, rank, B=4096, n=128, T=32):
        import torch
        from pde_control_b200 import ops, _lib
        self.torch, self.ops, self.dev = torch, ops, device
        self.B, self.n, self.T = B, n, T
        self.cell_updates = T * B * n
        u0, rng = _burgers_frames(B, n, T, 1 + rank)
        u0 = torch.tensor(u0, device=device)
        frames = ops.burgers_trajectory(u0, T, DT, VISC, substeps=4, resubstep=True)
        base = torch.cat([u0[None], frames]).contiguous()
        set_bytes = 2 * base.numel() * 4
        self.nsets = max(2, (2 * L2_BYTES + set_bytes - 1) // set_bytes)    # rotate over > 2x L2 of distinct data
        g = torch.Generator(device=device)
        g.manual_seed(7 + rank)
        self.s = [(base + 0.02 * torch.randn(base.shape, device=device, generator=g)).contiguous() for _ in range(self.nsets)]
        self.grad = [torch.empty_like(base) for _ in range(self.nsets)]
        self.losses = torch.zeros(2, device=device)
        ws_bytes = _lib.lib().pdec_burgers_pair_loss_workspace_bytes(B, n, T)
        self.ws = torch.empty((ws_bytes + 3) // 4, device=device)
        from pde_control_b200 import hostmem
        self.host_s = hostmem.pinned_empty(base.shape, device)             # page-locked, on the GPU's NUMA node
        self.host_s.copy_(self.s[0].cpu())
        self.host_grad = hostmem.pinned_empty(base.shape, device)
        self.host_losses = torch.empty(2, dtype=torch.float32).pin_memory()
        self.h2d = self.host_s.numel() * 4
        self.d2h = self.host_grad.numel() * 4 + 8
        self.pipe = ops.HostPipeline(n, T, chunk_rows=512, max_batch=B)     # the public host-buffer call (e2e)
        self.launches_per_step = 1      # fused pair-loss kernel (its last CTA does the fixed-order loss reduction)
        self.l2_policy = "inputs rotate over %d resident sets (%.0f MB > 2x 126 MB L2)" % (self.nsets, self.nsets * set_bytes / 1e6)
        self.i = 0
        self._args = None

    def step(self):
        """One fused kernel launch through the C-ABI.  The ctypes argument tuples are built once per input set (same stream,
        same buffers), so the host costs a few microseconds per step and the device, not Python, sets the pace."""
        k = self.i % self.nsets
        self.i += 1
        if self._args is None:
            import ctypes
            from pde_control_b200 import _lib
            T1, B, n = self.s[0].shape
            st = _lib.stream_ptr()
            self._fn = _lib.lib().pdec_burgers_pair_loss
            self._args = [(_lib.ptr(self.s[j]), _lib.ptr(self.grad[j]), _lib.ptr(self.losses), T1 - 1, B, n, self.mode, float(DT), float(VISC),
                           0.0, 1.0, 1, 0, 1.0, 0.0, _lib.ptr(self.ws), self.ws.numel() * 4, st) for j in range(self.nsets)]
        rc = self._fn(*self._args[k])
        if rc != 0:
            from pde_control_b200 import _lib
            _lib.check(rc, "pdec_burgers_pair_loss")

    def step_e2e(self):
        """Host buffers in, host buffers out, through the public host-buffer op (ops.HostPipeline.pair_loss ->
        pdec_burgers_pair_loss_host): per 512-row chunk H2D of the frames, the fused kernel, D2H of the gradient, the three
        legs overlapped on three streams; plus the two loss scalars.  The step ends when the host holds the results."""
        self.i += 1
        self.pipe.pair_loss(self.host_s, self.host_grad, self.host_losses, DT, VISC, mode=self.mode)
        self.torch.cuda.current_stream().synchronize()

    def config(self):
        return {"workload": self.name, "batch_per_gpu": self.B, "resolution": self.n, "frames": self.T + 1,
                "l2": self.l2_policy}

    # ---- CPU legs (oracle port) ---------------------------------------------------------------
    @classmethod
    def cpu_step(cls, Bs, n=128, T=32, seed=1):
        import torch
        from oracle import restate as R
        import numpy as np
        rng = np.random.RandomState(seed)
        s = np.stack([R.gaussian_clash(Bs, n, rng) for _ in range(T + 1)]).astype(np.float32)
        s = torch.tensor(s, requires_grad=True)
        t0 = time.perf_counter()
        l2, l1 = R.burgers_pair_loss(s, DT, VISC, mode=R.CIRCULAR if cls.mode == 2 else R.CONSTANT)
        l2.backward()
        return time.perf_counter() - t0, T * Bs * n


class PeriodicBurgersWorkload(BurgersWorkload):
    """The same step on a periodic grid (the north star's wording; the reference's own Burgers scenes are not periodic)."""
    name = "burgers_pair_loss PERIODIC res=128 T=32 B=4096/gpu (config 2 with 'circular' extrapolation)"
    ncu_traffic_bytes = 90.8e6      # as captured (read 70.9 MB + write 19.9 MB; the gradient's write-back outlives the launch): profiles/r1r_unit_periodic_summary.md
    mode = 2                        # PDEC_MODE_CIRCULAR


def _make_workload(name, device, rank):
    if name == "burgers":
        return BurgersWorkload(device, rank)
    if name == "burgers_periodic":
        return PeriodicBurgersWorkload(device, rank)
    if name == "fluid":
        from pde_control_b200 import bench_fluid
        return bench_fluid.FluidWorkload(device, rank)
    raise SystemExit("unknown workload %s" % name)


def _cpu_class(name):
    if name == "burgers":
        return BurgersWorkload
    if name == "burgers_periodic":
        return PeriodicBurgersWorkload
    from pde_control_b200 import bench_fluid
    return bench_fluid.FluidWorkload


def cpu_baseline(name, budget_s=12.0):
    """Oracle port timed on the host cores, bounded sample (about `budget_s` seconds of CPU work)."""
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cls = _cpu_class(name)
    Bs = cls.cpu_sample_batch if hasattr(cls, "cpu_sample_batch") else 64
    cls.cpu_step(max(1, Bs // 8))               # warm-up (thread pool, allocator)
    t_total, units, reps = 0.0, 0, 0
    while t_total < budget_s and reps < 20:
        t, u = cls.cpu_step(Bs, seed=reps + 1)
        t_total += t
        units += u
        reps += 1
    return {"value": units / t_total, "unit": "cell-updates/s", "cores": cores, "kind": "port",
            "sample": "%d x oracle.restate fwd+autograd backward on batch %d (same resolution/frames as the GPU workload), "
                      "torch CPU fp32, %d threads" % (reps, Bs, cores)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cls = _cpu_class(args.workload)
    Bs = cls.cpu_sample_batch if hasattr(cls, "cpu_sample_batch") else 64
    for i in range(args.warmup):
        cls.cpu_step(Bs, seed=100 + i)
    t_total, units = 0.0, 0
    for i in range(args.steps):
        t, u = cls.cpu_step(Bs, seed=i + 1)
        t_total += t
        units += u
    value = units / t_total
    name = cls.name
    line = {"impl": "reference", "metric": "fwd+adjoint cell-updates/s", "value": value, "unit": "cell-updates/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": name, "sample_batch": Bs},
            "cpu_baseline": {"value": value, "unit": "cell-updates/s", "cores": cores, "kind": "port",
                             "sample": "each step = oracle.restate fwd+autograd backward on batch %d, torch CPU fp32, %d threads; "
                                       "the reference's TF-1.14 CPU path is not installable here (DESIGN.md)" % (Bs, cores)},
            "e2e": {"value": value, "unit": "cell-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_b200(args):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU oracle")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    wl = _make_workload(args.workload, dev, rank)
    for _ in range(max(args.warmup, 3)):
        wl.step()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    # kernel-resident timing: CUDA events on the launching (current) stream
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        wl.step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    # end-to-end timing: pinned host buffers in/out every step
    for _ in range(2):
        wl.step_e2e()
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for _ in range(args.steps):
        wl.step_e2e()
    f1.record()
    barrier()
    ms_e2e = f0.elapsed_time(f1)
    clocks = sampler.summary() if sampler else None
    if world > 1:
        t = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e = float(t[0]), float(t[1])
    if rank == 0:
        peak, peak_src = _peaks()
        units_step = wl.cell_updates * world
        value = units_step * args.steps / (ms * 1e-3)
        e2e = units_step * args.steps / (ms_e2e * 1e-3)
        bytes_launch = wl.algorithmic_bytes_per_cell_update * wl.cell_updates
        achieved = bytes_launch / (ms * 1e-3 / args.steps) / 1e9            # per GPU
        line = {"metric": "fwd+adjoint cell-updates/s", "value": value, "unit": "cell-updates/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": wl.dtype,
                "data": "synthetic", "config": dict(wl.config(), parallelism="batch-sharded x%d" % world),
                "clocks": clocks,
                "e2e": {"value": e2e, "unit": "cell-updates/s", "h2d_bytes_per_step": wl.h2d, "d2h_bytes_per_step": wl.d2h,
                        "ms_per_step": ms_e2e / args.steps},
                "gpu_launches": wl.launches_per_step * args.steps,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": getattr(wl, "ncu_traffic_bytes", None), "peak_source": peak_src,
                             "algorithmic_bytes_per_cell_update": wl.algorithmic_bytes_per_cell_update,
                             "kernel": wl.dominant_kernel}}
        if hasattr(wl, "extra"):
            line["extra"] = wl.extra()
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(args.workload)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="burgers", choices=["burgers", "burgers_periodic", "fluid"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
