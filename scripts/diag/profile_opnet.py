"""Where does an OPResNet forward+backward spend its time?  (diagnostic for the N1 training step)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from torch.profiler import profile, ProfilerActivity
from pde_control_b200.control.networks import OPResNet

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
net = OPResNet().cuda()
a = torch.randn(B, 128, 1, device="cuda"); b = torch.randn(B, 128, 1, device="cuda")


def run():
    out = net(a, b)
    out.square().sum().backward()


def timed(label):
    for _ in range(3):
        run()
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(5):
        run()
    torch.cuda.synchronize(); print(label, (time.perf_counter() - t) / 5 * 1e3, "ms per fwd+bwd", flush=True)


timed("default")
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    run(); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14, max_name_column_width=70))

from pde_control_b200.control.networks import UNet
net2 = UNet(2, 1, levels=3).cuda()                  # fluid OP network, control/incompressible_flow.py:28
x = torch.randn(64, 128, 128, 2, device="cuda")


def run():
    net2(x).square().sum().backward()


timed("UNet B=64 128^2")
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    run(); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=12, max_name_column_width=70))
