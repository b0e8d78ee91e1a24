"""bench.py workload for BASELINE config 3: 2-D incompressible smoke, 128x128, 16 projected steps
(IncompressibleFlow.step, PF/physics/fluid.py:90-114) + the adjoint of a density-matching loss, batch 64 per GPU.

The whole trajectory (16 forward steps, loss, 16 adjoint steps = 32 CG solves) is captured in ONE CUDA graph;
the CG iteration counts are decided on the device.
"""
import time

import numpy as np
import torch


def _smooth(rng, shape, scale, passes=6):
    a = rng.randn(*shape)
    for _ in range(passes):
        a = (a + np.roll(a, 1, -1) + np.roll(a, -1, -1) + np.roll(a, 1, -2) + np.roll(a, -1, -2)) / 5
    return (a / np.abs(a).max() * scale).astype(np.float32)


class FluidWorkload:
    name = "incompressible_flow 128x128 steps=16 B=64/gpu OPEN (BASELINE config 3: projected step + CG + adjoint)"
    dtype = "f32"
    dominant_kernel = "cg_block_kernel"
    ncu_traffic_bytes = 4.23e6      # dram read per launch (rhs); the solution is still in L2 at kernel end, profiles/r1b_fluid_summary.md
    cpu_sample_batch = 2
    STEPS, N, B = 16, 128, 64
    ACCURACY, MAX_IT = 1e-5, 2000

    def __init__(self, device, rank, closed=False, use_graph=True):
        import os
        use_graph = use_graph and not os.environ.get("PDEC_BENCH_NO_GRAPH")
        from . import ops, datagen
        self.ops, self.dev, self.closed = ops, device, closed
        B, n = self.B, self.N
        rng = np.random.RandomState(3 + rank)
        import random
        shapes = datagen.procedural_shapes(10, 32, 0)
        rho0 = datagen.distribute_random_shape([n, n], B, shapes, rng=random.Random(rank))
        target = datagen.distribute_random_shape([n, n], B, shapes, rng=random.Random(1000 + rank))
        vy0, vx0 = _smooth(rng, (B, n + 1, n), 1.0), _smooth(rng, (B, n, n + 1), 1.0)
        self.host_in = [torch.tensor(a).pin_memory() for a in (rho0, vy0, vx0)]
        self.inputs = [t.to(device).requires_grad_(True) for t in self.host_in]
        self.target = torch.tensor(target, device=device)
        self.host_out = [torch.empty_like(t).pin_memory() for t in self.host_in]
        self.host_loss = torch.empty(1).pin_memory()
        self.cell_updates = B * n * n * self.STEPS
        self.h2d = sum(t.numel() * 4 for t in self.host_in)
        self.d2h = self.h2d + 4
        self.iters = []
        self.graph = None
        # Everything that touches autograd runs on ONE side stream (warm-up and capture): autograd binds a leaf's gradient
        # sink to the stream of its first use, and a sink bound to the legacy stream cannot be captured.
        side = torch.cuda.Stream(device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                self._trajectory()      # warm-up (also sizes the workspaces)
        side.synchronize()
        if use_graph:
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph, stream=side):
                self._trajectory()
        else:
            self.side = side
        torch.cuda.synchronize()
        self.step()                      # one replay so that the iteration counters hold real values
        torch.cuda.synchronize()
        its = torch.stack(self.iters).float()
        self.mean_iterations_fwd = float(its.mean())
        # adjoint solves run the same CG on the incoming gradient: measured separately below
        self.algorithmic_bytes_per_cell_update = 128.0 + 36.0 * 2 * self.mean_iterations_fwd
        self.launches_per_step = self._count_launches()
        self.l2_policy = "working set %.0f MB is L2 resident by design (all CG vectors stay on chip); inputs identical each step" % (
            (len(self.inputs) * 2 + 4) * B * n * n * 4 / 1e6)

    def _trajectory(self):
        ops = self.ops
        rho, vy, vx = self.inputs
        self.iters = []
        for _ in range(self.STEPS):
            rho, vy, vx, p, it = ops.incompressible_step(rho, vy, vx, 1.0, self.closed, accuracy=self.ACCURACY,
                                                         max_iterations=self.MAX_IT)
            self.iters.append(it)
        d = rho - self.target
        self.loss = (d * d).sum() / (2 * self.B)          # math.l2_loss (PF/math/nd.py:111-124)
        # autograd.grad (not .backward()): no AccumulateGrad nodes, which are bound to the stream the leaf was created on
        # and would break CUDA-graph capture on a side stream
        self.grads = torch.autograd.grad(self.loss, self.inputs)

    def _count_launches(self):
        # per projected step: fwd advect (1 + 2) + divergence + CG + subtract = 6; bwd: subtract^T + CG + divergence^T +
        # staggered advect^T (memset + 2 prep + 1) + centred advect^T (memset + prep + 1) = 10 (+ torch adds for the loss)
        return self.STEPS * 16 + 4

    def step(self):
        if self.graph is not None:
            self.graph.replay()
        else:
            self.side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self.side):
                self._trajectory()
            torch.cuda.current_stream().wait_stream(self.side)

    def step_e2e(self):
        with torch.no_grad():
            for dst, src in zip(self.inputs, self.host_in):
                dst.copy_(src, non_blocking=True)
        self.step()
        for dst, src in zip(self.host_out, self.grads):
            dst.copy_(src, non_blocking=True)
        self.host_loss.copy_(self.loss.detach().reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def config(self):
        return {"workload": self.name, "batch_per_gpu": self.B, "resolution": [self.N, self.N], "steps": self.STEPS,
                "cg_accuracy": self.ACCURACY, "cg_mean_iterations_forward": self.mean_iterations_fwd,
                "cuda_graph": self.graph is not None, "l2": self.l2_policy}

    def extra(self):
        return {"loss": float(self.loss.detach()), "cg_iterations_last_step": [int(v) for v in self.iters[-1][:4].tolist()]}

    # ---- CPU leg (oracle port) -------------------------------------------------------------------------------
    @staticmethod
    def cpu_step(Bs, seed=1, steps=2, n=128):
        from oracle import restate_fluid as F
        from oracle import restate as R
        rng = np.random.RandomState(seed)
        geo = F.Geometry(n, n)
        rho = torch.tensor(np.abs(_smooth(rng, (Bs, n, n), 1.0)), requires_grad=True)
        vy = torch.tensor(_smooth(rng, (Bs, n + 1, n), 1.0), requires_grad=True)
        vx = torch.tensor(_smooth(rng, (Bs, n, n + 1), 1.0), requires_grad=True)
        F._direct_factor(n, n, "open")          # factorisation is set-up, not part of a step
        t0 = time.perf_counter()
        s = (rho, vy, vx)
        for _ in range(steps):
            s = F.incompressible_step(*s, 1.0, geo, "open")[:3]
        R.l2_loss(s[0]).backward()
        return time.perf_counter() - t0, Bs * n * n * steps
