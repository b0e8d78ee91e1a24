#!/usr/bin/env python
"""bench.py -- headline measurement of the B200-native PDE-Control physics path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--only key,key] [--no-cpu] [--no-parity]

Metric (BASELINE.json): forward+adjoint cell-updates/s.  One "step" = one pass of the hot path over one batch.

ONE JSON line on rank 0.  Its top level is BASELINE config 2 (the configuration the metric is quoted on that fits one GPU: the fused
Burgers force loss + adjoint over s[33, 4096, 128] per GPU); `workloads` carries the rest of the target with the same keys each:
  fluid_config3               128 x 128, 16 projected steps + adjoint, B = 64: CG at the accuracy that HOLDS the parity tolerances (1e-7)
  fluid_config3_ref_accuracy  the same at the reference's default CG accuracy 1e-5
  fluid_config3_direct        the same with the opt-in exact solver (physics.SparseSciPy; tcgen05 tensor cores)
  fluid_config3_chained       SparseSciPy() & SparseCG(1e-7): the exact solve as the CG's initial guess (the reference's solver chain)
  fluid_notebook              the IncompressibleFluidPDE path: potential -> curl -> advection + staggered force loss
  burgers_config1             B = 8 trajectory + control-loss gradient (latency-bound)
  config4        (N >= 2)     config 2 + the OP-network gradient all-reduce inside the timed region
  config5_direct              2048 x 2048 projected steps + adjoint on one GPU with the exact solver (library GEMMs beyond 128 x 128)
  config5_slab   (N >= 2)     2048 x 2048 slab-decomposed projected steps + a solve checked against its fp64 residual
  config5_slab_chained (N >= 2)  the same with the solver chain: whole-domain exact solve as the guess, slab CG on the correction
Every workload: W untimed warm-up steps (at least 3), then the K-step loop repeated R times so that the timed region lasts >= 0.6 s
(`timing.repeats`), CUDA events on the launching stream, barrier + synchronize on both sides of every repetition, max over ranks,
median repetition; `e2e` = the same through the public API with pinned HOST buffers copied in and results copied out every step;
`parity` = the timed inputs (a slice, or all of them) checked against the oracle after the timed region.

`--impl reference`: the reference arm on the host cores (rank 0 only): the oracle PORT (PyTorch-CPU fp32 restatement incl. the adjoint --
the reference's reverse mode is TF-1.14 graph autodiff, not installable) at the SAME batch as the GPU arm, and, as
`cpu_baseline.reference_forward`, the reference's own unmodified NumPy code (forward only), staged under baseline/_ref.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC, UNIT = "fwd+adjoint cell-updates/s", "cell-updates/s"


def measure(wl, args, proto, sampler, peak, peak_src, world, rank, top=False):
    """warm-up, device-timed loop, end-to-end loop, roofline, parity, cpu baseline of one workload -> its result dict (rank 0)."""
    import torch
    W, K = max(args.warmup, 3), args.steps
    for _ in range(W):
        wl.step()
    proto.barrier()
    mark = sampler.mark() if sampler else 0
    ms, R = proto.timed(wl.step, K)
    res = {}
    has_e2e = hasattr(wl, "step_e2e")
    if has_e2e:
        for _ in range(2):
            wl.step_e2e()
        ms_e2e, R2 = proto.timed(wl.step_e2e, K)
    clocks = sampler.summary(since=mark) if sampler else None
    if rank != 0:
        if hasattr(wl, "solve_parity"):
            wl.solve_parity()           # collective: every rank takes part
        return None
    strong = getattr(wl, "key", "").startswith("config5_slab")
    units = wl.cell_updates * (1 if strong else world)
    res.update({"metric": METRIC, "value": units * K / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None,
                "dtype": wl.dtype, "data": "synthetic",
                "config": wl.config(),          # identical dict in the reference arm (nothing run-dependent in here)
                "timing": {"repeats": R, "timed_seconds": R * ms * 1e-3,
                           "protocol": "median of %d repetitions of the %d-step loop, each bracketed by barrier + synchronize, CUDA events" % (R, K)},
                "clocks": clocks})
    if has_e2e:
        res["e2e"] = {"value": units * K / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": wl.h2d, "d2h_bytes_per_step": wl.d2h,
                      "ms_per_step": ms_e2e / K, "repeats": R2}
    res["gpu_launches"] = int(getattr(wl, "launches_per_step", 1) * K)
    if hasattr(wl, "roofline"):
        res["roofline"] = wl.roofline(ms / K, peak, peak_src)
    if not args.no_parity:
        if hasattr(wl, "parity"):
            res["parity"] = wl.parity()
        elif hasattr(wl, "solve_parity"):
            res["parity"] = wl.solve_parity()
    if world == 1 and not args.no_cpu and hasattr(wl, "cpu_baseline"):
        res["cpu_baseline"] = wl.cpu_baseline()
    if hasattr(wl, "extra"):
        res["extra"] = wl.extra()
    return res


def run_b200(args):
    import torch
    import torch.distributed as dist
    from benchlib import timing
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    proto = timing.Protocol(torch, dist, world, dev, args.min_seconds)
    peak, peak_src = timing.peaks()
    sampler = timing.ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    from benchlib.burgers import BurgersConfig2, BurgersConfig1, BurgersConfig4
    from benchlib.fluid import FluidConfig3, FluidNotebook, Config5Slab, Config5Direct
    only = set(args.only.split(",")) if args.only else None
    want = lambda key: only is None or key in only

    base = BurgersConfig2(dev, rank, world)
    line = measure(base, args, proto, sampler, peak, peak_src, world, rank, top=True)
    if rank == 0 and world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline_config2()
    subs = {}
    if world > 1 and want("config4"):
        subs["config4"] = measure(BurgersConfig4(dev, rank, world, base), args, proto, sampler, peak, peak_src, world, rank)
    for key in ("fluid_config3", "fluid_config3_ref_accuracy", "fluid_config3_direct", "fluid_config3_chained"):
        if want(key):
            wl = FluidConfig3(dev, rank, world, key)
            subs[key] = measure(wl, args, proto, sampler, peak, peak_src, world, rank)
            del wl
            torch.cuda.empty_cache()
    if want("fluid_notebook"):
        subs["fluid_notebook"] = measure(FluidNotebook(dev, rank, world), args, proto, sampler, peak, peak_src, world, rank)
        torch.cuda.empty_cache()
    if want("burgers_config1"):
        subs["burgers_config1"] = measure(BurgersConfig1(dev, rank, world), args, proto, sampler, peak, peak_src, world, rank)
    if want("config5_direct"):
        sub_args = argparse.Namespace(**vars(args))
        sub_args.steps, sub_args.warmup = 3, 3
        subs["config5_direct"] = measure(Config5Direct(dev, rank, world), sub_args, proto, sampler, peak, peak_src, world, rank)
        torch.cuda.empty_cache()
    for key, chained in (("config5_slab", False), ("config5_slab_chained", True)):
        if world > 1 and want(key):
            sub_args = argparse.Namespace(**vars(args))
            sub_args.steps, sub_args.warmup = 1, 1          # one 2-step trajectory per "step"
            wl = Config5Slab(dev, rank, world, chained=chained)
            try:
                subs[key] = measure(wl, sub_args, proto, sampler, peak, peak_src, world, rank)
            finally:
                wl.close()
    if rank == 0:
        if sampler:
            sampler.summary(stop=True)
        line["workloads"] = subs
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def cpu_baseline_config2(budget_s=12.0, Bs=512):
    """The oracle port on the host cores on a bounded sample of config 2 (same frames statistics, batch 512) + the reference's own forward."""
    import torch
    from benchlib.burgers import BurgersConfig2 as W
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sets = W.cpu_inputs(Bs, 2)
    W.cpu_step(sets[0])
    elapsed, reps = 0.0, 0
    while elapsed < budget_s and reps < 40:
        elapsed += W.cpu_step(sets[reps % 2])
        reps += 1
    out = {"value": reps * W.T * Bs * W.n / elapsed, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": "%d x oracle.restate.burgers_pair_loss + autograd backward on batch %d (same resolution / frames / statistics as the GPU "
                     "workload), torch CPU fp32, %d threads" % (reps, Bs, cores)}
    try:
        rf = W.reference_forward()
        if rf:
            out["reference_forward"] = rf
    except Exception as e:          # the staged reference is optional on a box
        out["reference_forward"] = {"unavailable": repr(e)[:200]}
    return out


def run_reference(args):
    """Reference arm: rank 0 only, host cores, the GPU arm's config / metric / unit."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import torch
    from benchlib.burgers import BurgersConfig2 as W
    from benchlib.fluid import FluidConfig3
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    sets = W.cpu_inputs(W.B, 2)                              # the GPU arm's batch: 4096 rows, two rotating input sets
    for i in range(args.warmup):
        W.cpu_step(sets[i % 2])
    t_total = 0.0
    for i in range(args.steps):
        t_total += W.cpu_step(sets[i % 2])
    units = W.T * W.B * W.n
    value = units * args.steps / t_total
    cpu = {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": "each step = oracle.restate.burgers_pair_loss + autograd backward on the FULL batch %d (the GPU arm's config), torch CPU fp32, "
                     "%d threads; the reference's TF-1.14 CPU path (reverse mode) is not installable here (DESIGN.md)" % (W.B, cores)}
    try:
        rf = W.reference_forward()
        if rf:
            cpu["reference_forward"] = rf
    except Exception as e:
        cpu["reference_forward"] = {"unavailable": repr(e)[:200]}
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": W.describe(world), "cpu_baseline": cpu,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    # the fluid half of the target, bounded samples (the port with the reference's CG; the reference's own forward)
    subs = {}
    try:
        fl = _FluidCpuOnly()
        sub = fl.cpu_baseline(budget_s=20.0, Bs=2)
        entry = {"impl": "reference", "metric": METRIC, "value": sub["value"], "unit": UNIT, "cpu_baseline": sub}
        rf = FluidConfig3.reference_forward()
        if rf:
            entry["cpu_baseline"]["reference_forward"] = rf
        subs["fluid_config3"] = entry
    except Exception as e:
        subs["fluid_config3"] = {"unavailable": repr(e)[:300]}
    line["workloads"] = subs
    print(json.dumps(line), flush=True)


class _FluidCpuOnly(object):
    """The CPU legs of FluidConfig3 without a CUDA device (reference arm): same inputs, host tensors."""

    def __init__(self):
        import torch
        from benchlib import inputs
        from benchlib.fluid import FluidConfig3
        self.N, self.STEPS, self.B, self.closed = FluidConfig3.N, FluidConfig3.STEPS, FluidConfig3.B, False
        rho0, vy0, vx0, target = inputs.fluid_config3_inputs(8, self.N, 0)
        self.args = [torch.tensor(a) for a in (rho0, vy0, vx0, target)]
        self.cpu_baseline = lambda **kw: FluidConfig3.cpu_baseline(self, **kw)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--only", default="", help="comma-separated sub-workload keys (default: all that apply at this N)")
    ap.add_argument("--min-seconds", type=float, default=0.6, help="minimum timed region per workload")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs (profiling runs)")
    ap.add_argument("--no-parity", action="store_true", help="skip the in-run oracle checks")
    args = ap.parse_args()
    t0 = time.time()
    # stdout carries the ONE JSON line and nothing else: libraries that write to file descriptor 1 themselves (NCCL prints its version
    # there at the first communicator) are sent to stderr for the length of the run
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(json_fd, "w")
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
    sys.stdout.flush()
    if int(os.environ.get("RANK", "0")) == 0:
        print("bench.py wall %.1f s" % (time.time() - t0), file=sys.stderr)


if __name__ == "__main__":
    main()
