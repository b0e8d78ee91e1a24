"""BASELINE config 1: Burgers res 128, 32 steps, batch 8: forward simulation + control-loss gradient w.r.t. u0 and the force
(l2_loss(u_32 - target)), through the C-ABI trajectory kernels.  135 KB of state: launch-latency bound by construction,
so the figure of merit is time per trajectory-batch; reported eager (2 launches + the loss kernels) and as a CUDA-graph replay,
next to the oracle port on the host cores.  Prints one JSON line."""
import os, sys, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pde_control_b200 import ops, datagen, _lib

B, n, T, DT, NU = int(sys.argv[1]) if len(sys.argv) > 1 else 8, 128, 32, 0.03, 0.003
rng = np.random.RandomState(0)
u0 = torch.tensor(datagen.gaussian_clash(B, n, rng)).cuda()
force = torch.tensor(datagen.gaussian_force(B, n, rng)).cuda()
target = torch.tensor(datagen.gaussian_clash(B, n, rng)).cuda()
L = _lib.lib()
frames = torch.empty(T, B, n, device="cuda"); g_frames = torch.zeros(B, n, device="cuda")
g_u0 = torch.empty(B, n, device="cuda"); g_f = torch.empty(B, n, device="cuda"); loss = torch.zeros((), device="cuda")

def step():
    st = _lib.stream_ptr()
    _lib.check(L.pdec_burgers_trajectory_fwd(_lib.ptr(u0), _lib.ptr(force), _lib.ptr(frames), B, n, T, 1, 0, DT, NU, 0.0, 1.0, 4, 1, 1, st), "fwd")
    d = frames[T - 1] - target                                      # l2_loss(u_32 - target) = sum(d^2)/2/B, gradient d/B
    torch.mul(d, 1.0 / B, out=g_frames)
    loss.copy_((d * d).sum() * (0.5 / B))
    _lib.check(L.pdec_burgers_trajectory_bwd(_lib.ptr(u0), _lib.ptr(frames), _lib.ptr(g_frames), _lib.ptr(g_u0), _lib.ptr(g_f), B, n, T, 0, 0,
                                             DT, NU, 0.0, 1.0, 4, 1, 1, st), "bwd")

def timeit(fn, reps=200):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3

t_eager = timeit(step)
side = torch.cuda.Stream()
with torch.cuda.stream(side):
    step(); side.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        step()
    t_graph = timeit(g.replay)
# oracle port on the host cores
from oracle import restate as R
torch.set_num_threads(os.cpu_count() or 1)
uo = torch.tensor(u0.cpu().numpy(), requires_grad=True); fo = torch.tensor(force.cpu().numpy(), requires_grad=True)
def cpu():
    uo.grad = None; fo.grad = None
    tr = R.burgers_trajectory(uo, T, DT, NU, force=fo, substeps=4, resubstep=True)
    l = R.l2_loss(tr[-1] - target.cpu()); l.backward()
    return l
cpu(); t0 = time.perf_counter()
for _ in range(5): lo = cpu()
t_cpu = (time.perf_counter() - t0) / 5 * 1e6
step(); torch.cuda.synchronize()
rel = lambda a, b: float((a.cpu() - b).abs().max() / b.abs().max())
print(json.dumps({"workload": "config 1: Burgers res=128 steps=32 batch=%d fwd + grad(u0, force), substeps 4" % B,
                  "us_eager": t_eager, "us_cuda_graph": t_graph, "cell_updates_per_s_graph": T * B * n / (t_graph * 1e-6),
                  "us_cpu_oracle_port": t_cpu, "cpu_threads": os.cpu_count(), "loss": float(loss), "loss_oracle": float(lo),
                  "grad_u0_rel_err_vs_oracle": rel(g_u0, uo.grad), "grad_force_rel_err_vs_oracle": rel(g_f, fo.grad),
                  "note": "launch-latency bound: 135 KB of state, 2 physics launches + 3 small loss kernels"}), flush=True)
