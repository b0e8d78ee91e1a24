import os, sys, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from pde_control_b200 import ops, datagen
from pde_control_b200.physics import Domain, box
from pde_control_b200.control import ControlTraining, BurgersPDE, StaggeredSequence
B, n, N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096, 32, 128
rng = np.random.RandomState(0)
u0 = torch.tensor(datagen.gaussian_clash(B, N, rng)).cuda(); f = torch.tensor(datagen.gaussian_force(B, N, rng)).cuda()
frames = torch.cat([u0[None], ops.burgers_trajectory(u0, n, 0.03, 0.003, force=f, substeps=4, resubstep=True)])
data = {"u": frames.permute(1, 0, 2)[..., None].cpu().numpy()}
pde = BurgersPDE(Domain([N], box=box[0:1]), 0.003, 0.03)
app = ControlTraining(n, pde, data, None, range(B), trace_to_channel=lambda _: "u", obs_loss_frames=[],
                      trainable_networks=["OP2", "OP4", "OP8", "OP16", "OP32"], sequence_class=StaggeredSequence, batch_size=B, learning_rate=1e-3, dt=0.03)
app.step(); torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
app.step(); torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
