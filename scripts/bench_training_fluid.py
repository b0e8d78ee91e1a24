"""BASELINE config 3 as the reference's notebook states it: one CFE / OP control-network TRAINING step on the shape-transition task
(2-D incompressible smoke, 128 x 128, n = 16 frames, batch 64) through the reference-facing API -- ControlTraining(n=16,
IncompressibleFluidPDE, StaggeredSequence) with the U-Nets as ordinary PyTorch modules and the physics (curl, advection, force loss and
their adjoints) on the CUDA kernels.  Prints one JSON line: seconds per optimisation step eager and as one CUDA graph, and the physics
share (the same operators timed alone on the same shapes).  Data: synthetic shape pairs (no dataset on the box)."""
import os, sys, json, time, random
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pde_control_b200 import datagen, graphs
from pde_control_b200.physics import Domain
from pde_control_b200.control import ControlTraining, IncompressibleFluidPDE, StaggeredSequence
from benchlib.fluid import FluidNotebook

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
n, res, steps = 16, 128, int(sys.argv[2]) if len(sys.argv) > 2 else 5
shapes = datagen.procedural_shapes(10, 32, 0)
r = random.Random(0)
start = datagen.distribute_random_shape([res, res], B, shapes, rng=r)
end = datagen.distribute_random_shape([res, res], B, shapes, rng=r)
frames = np.stack([start * (1 - t / n) + end * (t / n) for t in range(n + 1)], 1)[..., None].astype(np.float32)
out = {"workload": "ControlTraining IncompressibleFluidPDE n=%d res=%dx%d batch=%d, StaggeredSequence, CFE + OP2..OP16 trainable" % (n, res, res, B)}
for graphed in (False, True):
    torch.manual_seed(0)
    pde = IncompressibleFluidPDE(Domain([res, res]), dt=1.0)
    app = ControlTraining(n, pde, {"density": frames}, None, range(B), obs_loss_frames=[], trainable_networks=["CFE", "OP2", "OP4", "OP8", "OP16"],
                          sequence_class=StaggeredSequence, batch_size=B, learning_rate=2e-4, dt=1.0, cuda_graph=graphed)
    for _ in range(5):
        app.step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    losses = [app.step()["Loss"] for _ in range(steps)]
    torch.cuda.synchronize()
    out["s_per_training_step_%s" % ("cuda_graph" if graphed else "eager")] = (time.perf_counter() - t0) / steps
    out["loss_%s" % ("cuda_graph" if graphed else "eager")] = losses[-1]
    out["parameters"] = sum(p.numel() for p in pde.networks.parameters())
    del app, pde
    torch.cuda.empty_cache()
# the physics alone: the notebook-variant trajectory of bench.py (16 steps: curl, density advection, force loss; forward + adjoint)
wl = FluidNotebook(torch.device("cuda"), 0, 1)
wl.B = B
for _ in range(3):
    wl.step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(10):
    wl.step()
torch.cuda.synchronize()
out["physics_fwd_adjoint_s"] = (time.perf_counter() - t0) / 10
out["physics_share_of_graphed_step"] = out["physics_fwd_adjoint_s"] / out["s_per_training_step_cuda_graph"]
print(json.dumps(out), flush=True)
