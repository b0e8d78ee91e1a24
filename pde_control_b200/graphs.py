"""CUDA-graph capture of whole trajectories as a product feature.

The reference builds ONE static TensorFlow graph per training app and replays it with session.run
(src/control/control_training.py:50-84); eager PyTorch re-launches every kernel from Python instead (a 16-step projected fluid
trajectory with its adjoint is a few hundred launches of a few microseconds each).  `GraphedFunction` is the B200 analogue of the static
graph: it runs a function of tensors once under stream capture -- forward, the scalar loss and, if asked, `torch.autograd.grad`
w.r.t. chosen inputs -- and afterwards replays the captured launches with one `cudaGraphLaunch`.  Every C-ABI call of this library only
enqueues work on the current stream, allocates nothing and decides iteration counts on the device (include/pdec.h), which is what
makes a trajectory with 32 conjugate-gradient solves capturable.

Rules (those of torch.cuda.graph): shapes are fixed at capture; inputs are copied into static buffers before each replay; returned
tensors are static buffers that the next replay overwrites (clone what must survive).
"""
import torch


class GraphedFunction(object):
    """graphed = GraphedFunction(fn, example_inputs, grad_inputs=(0, 2)); outputs, grads = graphed(*inputs)

    fn(*tensors) -> tensor or tuple of tensors.  With `grad_inputs` (indices into the inputs) the FIRST output must be a scalar loss and
    the graph also holds d loss / d inputs[k] for those k (torch.autograd.grad inside the capture).  `outputs` / `grads` are tuples of
    static tensors.  `graphed.replay()` re-runs on the current contents of `graphed.inputs` without any copy."""

    def __init__(self, fn, example_inputs, grad_inputs=(), warmup=2, device=None):
        self.fn = fn
        self.grad_inputs = tuple(grad_inputs)
        example_inputs = tuple(example_inputs)
        dev = device or next(t.device for t in example_inputs if isinstance(t, torch.Tensor))
        assert dev.type == "cuda", "CUDA graphs need CUDA tensors (there is no CPU fallback)"
        self.device = dev
        self.inputs = tuple(t.detach().clone().contiguous() for t in example_inputs)
        for k in self.grad_inputs:
            self.inputs[k].requires_grad_(True)
        # Everything that touches autograd runs on ONE side stream (warm-up and capture): autograd binds a leaf's gradient sink to the
        # stream of its first use, and a sink bound to the legacy default stream cannot be captured.
        self._stream = torch.cuda.Stream(dev)
        self._stream.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(self._stream):
            for _ in range(max(1, warmup)):         # sizes the grow-only workspaces, loads the kernels, warms the allocator
                self._run()
        self._stream.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=self._stream):
            self.outputs, self.grads = self._run()
        torch.cuda.current_stream(dev).wait_stream(self._stream)

    def _run(self):
        out = self.fn(*self.inputs)
        outs = out if isinstance(out, (tuple, list)) else (out,)
        grads = ()
        if self.grad_inputs:
            # autograd.grad, not .backward(): no AccumulateGrad nodes (they are stream-bound and would break the capture)
            grads = torch.autograd.grad(outs[0], [self.inputs[k] for k in self.grad_inputs])
        return tuple(outs), tuple(grads)

    def replay(self):
        self.graph.replay()
        return self.outputs, self.grads

    def __call__(self, *inputs):
        assert len(inputs) == len(self.inputs)
        with torch.no_grad():
            for dst, src in zip(self.inputs, inputs):
                if src is not dst:
                    dst.copy_(src, non_blocking=True)
        return self.replay()


def fluid_trajectory(steps, dt, closed, accuracy=1e-5, max_iterations=2000, box=None, loss=None, direct=False, groups=1, guess=None):
    """fn(rho, vy, vx[, target]) for GraphedFunction: `steps` projected IncompressibleFlow steps (PF/physics/fluid.py:90-114); returns
    (loss, rho, vy, vx, cg_iterations [steps, B]) with loss = l2_loss(rho_final - target) (PF/math/nd.py:111-124) unless `loss` is given.

    groups > 1: the batch is cut into that many sub-batches whose trajectories run on PARALLEL streams (parallel branches of the captured
    graph).  Batch elements never interact, but a CG launch lasts as long as its slowest element: with one launch per step for the whole
    batch every step waits for the slowest of all B elements, with sub-batches only for the slowest of its own group, and the groups drift
    apart instead of idling (64 one-CTA solves still fit the 148 SMs at once).  Per-element results are bit-identical; the loss is summed
    group by group.
    guess="direct": the solver chain SparseSciPy() & SparseCG(accuracy) (ops.divergence_free)."""
    from . import ops
    streams = []

    def run(rho, vy, vx, target):
        iters = []
        for _ in range(steps):
            rho, vy, vx, _p, it = ops.incompressible_step(rho, vy, vx, dt, closed, box=box, accuracy=accuracy, max_iterations=max_iterations,
                                                          direct=direct, guess=guess)
            iters.append(it)
        return rho, vy, vx, torch.stack(iters)

    def fn(rho, vy, vx, target=None):
        B = rho.shape[0]
        if groups <= 1 or B < 2 * groups:
            rho2, vy2, vx2, iters = run(rho, vy, vx, target)
        else:
            main = torch.cuda.current_stream(rho.device)
            while len(streams) < groups:
                streams.append(torch.cuda.Stream(rho.device))
            bounds = [B * g // groups for g in range(groups + 1)]
            parts = []
            for g in range(groups):
                lo, hi = bounds[g], bounds[g + 1]
                streams[g].wait_stream(main)
                with torch.cuda.stream(streams[g]):
                    parts.append(run(rho[lo:hi], vy[lo:hi], vx[lo:hi], target))
            for g in range(groups):
                main.wait_stream(streams[g])
            rho2, vy2, vx2 = (torch.cat([p[k] for p in parts]) for k in range(3))
            iters = torch.cat([p[3] for p in parts], 1)
        if loss is not None:
            value = loss(rho2, vy2, vx2, target)
        else:
            d = rho2 if target is None else rho2 - target
            value = (d * d).sum() / (2 * B)
        return value, rho2, vy2, vx2, iters

    return fn


def burgers_trajectory_loss(steps, dt, viscosity, substeps=1, resubstep=False, **kw):
    """fn(u0, force, target) for GraphedFunction: BASELINE config 1 -- `steps` Burgers steps with a GROW force (one trajectory kernel),
    loss = l2_loss(u_final - target); returns (loss, frames [steps, B, n])."""
    from . import ops

    def fn(u0, force, target):
        frames = ops.burgers_trajectory(u0, steps, dt, viscosity, force=force, substeps=substeps, resubstep=resubstep, **kw)
        d = frames[-1] - target
        return (d * d).sum() / (2 * u0.shape[0]), frames

    return fn
