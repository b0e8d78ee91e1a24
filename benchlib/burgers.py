"""Burgers workloads of bench.py: BASELINE config 2 (the headline: fused force loss fwd+adjoint over s[33, 4096, 128]), config 1
(B = 8 trajectory + control-loss gradient, CUDA graph) and config 4 (config 2 + the OP-network gradient all-reduce, N >= 2)."""
import os
import time

import numpy as np
import torch

from . import inputs
from .inputs import DT, VISC

L2_BYTES = 126 * 1024 * 1024


def _rel(a, b):
    a = a.detach().cpu().double()
    b = b.detach().cpu().double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


class BurgersConfig2(object):
    key = "burgers_config2"
    algorithmic_bytes_per_cell_update = 8.0     # read each frame once + write its gradient once
    dtype = "f32"
    dominant_kernel = "burgers_pair_loss_unit_kernel"
    ncu_traffic_bytes = 130.4e6     # steady state: dram read 67.1 MB + write 63.3 MB per launch (ncu --set full, profiles/r2ac_burgers_summary.md)
    mode = 0                        # PDEC_MODE_CONSTANT: the reference's Burgers domain (OPEN boundaries -> 'constant' extrapolation)
    B, n, T = 4096, 128, 32

    @classmethod
    def describe(cls, world=1):
        """The workload's `config`, identical for the GPU arm and the reference arm."""
        set_mb = 2 * (cls.T + 1) * cls.B * cls.n * 4 / 1e6
        nsets = max(2, int(-(-2 * L2_BYTES // (set_mb * 1e6))))
        return {"workload": "burgers_pair_loss res=%d T=%d B=%d/gpu (BASELINE config 2: force loss fwd+adjoint)" % (cls.n, cls.T, cls.B),
                "batch_per_gpu": cls.B, "resolution": cls.n, "frames": cls.T + 1,
                "l2": "inputs rotate over %d distinct sets per GPU (%.0f MB > 2x 126 MB L2)" % (nsets, nsets * set_mb),
                "parallelism": "batch-sharded x%d" % world}

    def __init__(self, device, rank, world):
        from pde_control_b200 import ops, _lib, hostmem
        self.ops, self.dev, self.rank, self.world = ops, device, rank, world
        B, n, T = self.B, self.n, self.T
        self.cell_updates = T * B * n
        set_bytes = 2 * (T + 1) * B * n * 4
        self.nsets = max(2, (2 * L2_BYTES + set_bytes - 1) // set_bytes)    # rotate over > 2x L2 of distinct data
        self.s = inputs.burgers_config2_sets(B, n, T, rank, device, self.nsets)
        self.grad = [torch.empty_like(self.s[0]) for _ in range(self.nsets)]
        self.losses = torch.zeros(2, device=device)
        ws_bytes = _lib.lib().pdec_burgers_pair_loss_workspace_bytes(B, n, T)
        self.ws = torch.empty((ws_bytes + 3) // 4, device=device)
        self.host_s = hostmem.pinned_empty(self.s[0].shape, device)             # page-locked, on the GPU's NUMA node
        self.host_s.copy_(self.s[0].cpu())
        self.host_grad = hostmem.pinned_empty(self.s[0].shape, device)
        self.host_losses = torch.empty(2, dtype=torch.float32).pin_memory()
        self.h2d = self.host_s.numel() * 4
        self.d2h = self.host_grad.numel() * 4 + 8
        self.pipe = ops.HostPipeline(n, T, chunk_rows=512, max_batch=B)     # the public host-buffer call (e2e)
        self.launches_per_step = 1      # fused pair-loss kernel (its last CTA does the fixed-order loss reduction)
        self.i = 0
        self._args = None

    def step(self):
        """One fused kernel launch through the C-ABI.  The ctypes argument tuples are built once per input set (same stream,
        same buffers), so the host costs a few microseconds per step and the device, not Python, sets the pace."""
        from pde_control_b200 import _lib
        k = self.i % self.nsets
        self.i += 1
        if self._args is None:
            T1, B, n = self.s[0].shape
            st = _lib.stream_ptr()
            self._fn = _lib.lib().pdec_burgers_pair_loss
            self._args = [(_lib.ptr(self.s[j]), _lib.ptr(self.grad[j]), _lib.ptr(self.losses), T1 - 1, B, n, self.mode, float(DT), float(VISC),
                           0.0, 1.0, 1, 0, 1.0, 0.0, _lib.ptr(self.ws), self.ws.numel() * 4, st) for j in range(self.nsets)]
        rc = self._fn(*self._args[k])
        if rc != 0:
            _lib.check(rc, "pdec_burgers_pair_loss")

    def step_e2e(self):
        """Host buffers in, host buffers out, through the public host-buffer op (ops.HostPipeline.pair_loss ->
        pdec_burgers_pair_loss_host): per 512-row chunk H2D of the frames, the fused kernel, the gradient stored into the pinned host
        array, the legs overlapped on the pipeline's streams; plus the two loss scalars.  The step ends when the host holds the results."""
        self.pipe.pair_loss(self.host_s, self.host_grad, self.host_losses, DT, VISC, mode=self.mode)
        torch.cuda.current_stream().synchronize()

    def config(self):
        return self.describe(self.world)

    def roofline(self, ms_per_step, peak, peak_src):
        achieved = self.algorithmic_bytes_per_cell_update * self.cell_updates / (ms_per_step * 1e-3) / 1e9
        return {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": self.ncu_traffic_bytes, "peak_source": peak_src,
                "algorithmic_bytes_per_cell_update": self.algorithmic_bytes_per_cell_update, "kernel": self.dominant_kernel,
                "note": "the fused kernel is the whole step: its launch duration is the step time (CUDA events on the launching stream)"}

    def parity(self):
        """The timed launch's own outputs (input set 0: loss scalars and the whole gradient array) against the oracle on the same
        frames, whole batch (about 1.5 s of CPU)."""
        from oracle import restate as R
        torch.cuda.synchronize()
        self.i = 0
        self.step()                                     # set 0 -> grad[0], losses
        torch.cuda.synchronize()
        so = self.s[0].cpu().requires_grad_(True)
        l2o, l1o = R.burgers_pair_loss(so, DT, VISC, mode=R.CIRCULAR if self.mode == 2 else R.CONSTANT)
        l2o.backward()
        l2, l1 = float(self.losses[0]), float(self.losses[1])
        out = {"checked": True, "against": "oracle.restate.burgers_pair_loss + autograd on the timed input set 0, all %d rows" % self.B,
               "max_rel_state": max(abs(l2 - float(l2o)) / abs(float(l2o)), abs(l1 - float(l1o)) / abs(float(l1o))),
               "max_rel_grad": _rel(self.grad[0], so.grad), "tolerance": {"state": 1e-4, "grad": 1e-3}}
        out["ok"] = out["max_rel_state"] < 1e-4 and out["max_rel_grad"] < 1e-3
        # the host-buffer path returned the same numbers?
        self.step_e2e()
        out["e2e_grad_equals_device_path"] = bool(torch.equal(self.host_grad, self.grad[0].cpu()))
        return out

    # ---- CPU legs ---------------------------------------------------------------------------------
    @classmethod
    def cpu_inputs(cls, Bs, nsets=2, seed=1):
        """Frames for the CPU legs with the GPU arm's statistics: GaussianClash trajectories of the oracle's own solver + noise."""
        from oracle import restate as R
        rng = np.random.RandomState(seed)
        sets = []
        u0 = torch.tensor(R.gaussian_clash(Bs, cls.n, rng))
        with torch.no_grad():
            tr = R.burgers_trajectory(u0, cls.T, DT, VISC, substeps=4, resubstep=True)      # [T+1, B, n] (frame 0 = u0)
        for _ in range(nsets):
            sets.append((tr + 0.02 * torch.tensor(rng.randn(*tr.shape).astype(np.float32))).contiguous())
        return sets

    @classmethod
    def cpu_step(cls, s):
        """One step of the oracle PORT: loss + full gradient (PyTorch-CPU fp32 restatement of the reference's TF semantics)."""
        from oracle import restate as R
        s = s.detach().requires_grad_(True)
        t0 = time.perf_counter()
        l2, l1 = R.burgers_pair_loss(s, DT, VISC, mode=R.CIRCULAR if cls.mode == 2 else R.CONSTANT)
        l2.backward()
        return time.perf_counter() - t0

    @classmethod
    def reference_forward(cls, Bs=64, budget_s=8.0):
        """The reference's OWN code (unmodified PhiFlow, NumPy backend, staged under baseline/_ref on the GPU box): forward only -- it has
        no reverse mode without TensorFlow.  Per frame pair: Burgers().step on a BurgersVelocity state + the residual's l2_loss."""
        from oracle import ref_shim
        if not ref_shim.available():
            return None
        pf = ref_shim.flow()
        s = cls.cpu_inputs(Bs, 1, seed=9)[0].numpy()
        dom = pf.Domain([cls.n], box=pf.box[0:1])
        physics = pf.Burgers()

        def one_pass():
            t0 = time.perf_counter()
            total = 0.0
            for t in range(cls.T):
                state = pf.BurgersVelocity(dom, velocity=s[t][..., None], viscosity=VISC, batch_size=Bs)
                nxt = physics.step(state, dt=DT)
                total += float(pf.math.l2_loss(s[t + 1][..., None] - nxt.velocity.data))
            return time.perf_counter() - t0, total

        one_pass()
        elapsed, reps = 0.0, 0
        while elapsed < budget_s and reps < 10:
            dt_, _ = one_pass()
            elapsed += dt_
            reps += 1
        return {"value": reps * cls.T * Bs * cls.n / elapsed, "unit": "cell-updates/s (FORWARD ONLY)", "cores": 1, "kind": "reference-forward",
                "sample": "%d x %d frame pairs of batch %d through the unmodified reference: phi.flow Burgers().step + math.l2_loss, NumPy/SciPy "
                          "backend (single-threaded Python loops over the batch)" % (reps, cls.T, Bs)}


class BurgersConfig1(object):
    """BASELINE config 1: res 128, 32 steps, batch 8: forward simulation with a GaussianForce + gradient of l2_loss(u_32 - target) w.r.t.
    u0 and the force; the whole thing (2 physics launches + the loss arithmetic) replayed from one CUDA graph (graphs.GraphedFunction)."""
    key = "burgers_config1"
    dtype = "f32"
    dominant_kernel = "burgers_trajectory_bwd_cta_kernel"
    algorithmic_bytes_per_cell_update = 20.0    # fwd read u_t + write u_t+1, adjoint read u_t + lambda in/out
    B, n, T = 8, 128, 32

    def __init__(self, device, rank, world):
        from pde_control_b200 import graphs, datagen
        self.dev, self.world = device, world
        rng = np.random.RandomState(100 + rank)
        B, n = self.B, self.n
        host = [datagen.gaussian_clash(B, n, rng), datagen.gaussian_force(B, n, rng), datagen.gaussian_clash(B, n, rng)]
        self.host_in = [torch.tensor(a).pin_memory() for a in host]
        self.args = [t.to(device) for t in self.host_in]
        self.fn = graphs.burgers_trajectory_loss(self.T, DT, VISC, substeps=4, resubstep=True)
        self.graphed = graphs.GraphedFunction(self.fn, self.args, grad_inputs=(0, 1))
        self.host_out = [torch.empty(B, n).pin_memory() for _ in range(2)]
        self.host_loss = torch.empty(1).pin_memory()
        self.cell_updates = self.T * B * n
        self.h2d = sum(t.numel() * 4 for t in self.host_in)
        self.d2h = 2 * B * n * 4 + 4
        self.launches_per_step = 2      # trajectory fwd + bwd kernels (+ a handful of torch elementwise kernels for the loss)

    def step(self):
        self.graphed.replay()

    def step_e2e(self):
        outs, grads = self.graphed(*self.host_in)
        for dst, src in zip(self.host_out, grads):
            dst.copy_(src, non_blocking=True)
        self.host_loss.copy_(outs[0].reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def config(self):
        return {"workload": "burgers trajectory res=128 steps=32 B=8 + grad(u0, force) of l2_loss(u_32 - target), GaussianForce, 4 substeps "
                            "(BASELINE config 1)", "batch_per_gpu": self.B, "resolution": self.n, "steps": self.T, "cuda_graph": True,
                "l2": "135 KB of state: latency-bound by construction, everything is cache resident (SURVEY 8d)",
                "parallelism": "replicas x%d" % self.world}

    def roofline(self, ms_per_step, peak, peak_src):
        achieved = self.algorithmic_bytes_per_cell_update * self.cell_updates / (ms_per_step * 1e-3) / 1e9
        return {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "peak_source": peak_src, "algorithmic_bytes_per_cell_update": self.algorithmic_bytes_per_cell_update,
                "kernel": self.dominant_kernel,
                "note": "8 rows x 32 SEQUENTIAL steps: the bound is the dependent chain of one step (about 1 us), not bandwidth; frac ~ 0 by construction"}

    def parity(self):
        from oracle import restate as R
        outs, grads = self.graphed(*self.args)
        torch.cuda.synchronize()
        uo, fo = [a.cpu().requires_grad_(True) for a in self.args[:2]]
        tr = R.burgers_trajectory(uo, self.T, DT, VISC, force=fo, substeps=4, resubstep=True)
        lo = R.l2_loss(tr[-1] - self.args[2].cpu())
        lo.backward()
        out = {"checked": True, "against": "oracle.restate.burgers_trajectory + autograd, the timed inputs",
               "max_rel_state": max(_rel(outs[1][-1], tr[-1]), abs(float(outs[0]) - float(lo)) / abs(float(lo))),
               "max_rel_grad": max(_rel(grads[0], uo.grad), _rel(grads[1], fo.grad)), "tolerance": {"state": 1e-4, "grad": 1e-3}}
        out["ok"] = out["max_rel_state"] < 1e-4 and out["max_rel_grad"] < 1e-3
        return out

    def cpu_baseline(self, budget_s=6.0):
        from oracle import restate as R
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        u, f, tgt = [a.cpu() for a in self.args]

        def once():
            uo, fo = u.clone().requires_grad_(True), f.clone().requires_grad_(True)
            t0 = time.perf_counter()
            tr = R.burgers_trajectory(uo, self.T, DT, VISC, force=fo, substeps=4, resubstep=True)
            R.l2_loss(tr[-1] - tgt).backward()
            return time.perf_counter() - t0
        once()
        elapsed, reps = 0.0, 0
        while elapsed < budget_s and reps < 50:
            elapsed += once()
            reps += 1
        return {"value": reps * self.cell_updates / elapsed, "unit": "cell-updates/s", "cores": cores, "kind": "port",
                "sample": "%d x the same trajectory + gradient with oracle.restate (torch CPU fp32, %d threads)" % (reps, cores)}


class BurgersConfig4(object):
    """BASELINE config 4 (N >= 2): the config-2 step followed by ONE NCCL all-reduce of an OP-network-sized gradient bucket
    (32 641 parameters of OPResNet + the two loss scalars) INSIDE the timed region; total batch = N x 4096."""
    key = "config4"
    dtype = "f32"

    def __init__(self, device, rank, world, base):
        import torch.distributed as dist
        self.dist, self.base, self.world = dist, base, world
        self.bucket = torch.randn(32641 + 2, device=device)
        self.cell_updates = base.cell_updates
        self.launches_per_step = 1

    def step(self):
        self.base.step()
        self.bucket[-2:].copy_(self.base.losses)
        self.dist.all_reduce(self.bucket)

    def config(self):
        return {"workload": "config-2 step + NCCL all-reduce of 32643 floats (OP-network gradients + losses) per step, total batch %d"
                            % (self.base.B * self.world), "batch_per_gpu": self.base.B, "allreduce_bytes": self.bucket.numel() * 4,
                "parallelism": "batch-sharded x%d, one all-reduce per step" % self.world}
