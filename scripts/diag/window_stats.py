"""How large are the gather-transpose windows along a config-3 trajectory? per-step max displacement (element-wide) and the share of 8 x 32
tiles whose own bound is smaller."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import torch
from benchlib import inputs
from pde_control_b200 import ops
rho0, vy0, vx0, target = inputs.fluid_config3_inputs(64, 128, 0)
s = [torch.tensor(a, device="cuda") for a in (rho0, vy0, vx0)]
for step in range(16):
    vy, vx = s[1], s[2]
    dy = (vy.abs() * 1.0).floor() + 1          # window half-width of a y-face sample along y: int(|v| dt / dx) + 1
    tiles = dy[:, :128].reshape(64, 16, 8, 4, 32).amax((2, 4))
    print(step, "max|vy| %.2f max|vx| %.2f" % (float(vy.abs().max()), float(vx.abs().max())), "element-wide R", int(dy.amax()),
          "share of tiles with R=1: %.2f" % float((tiles == 1).float().mean()), "mean per-element max R %.2f" % float(dy.amax((1, 2)).mean()))
    s = list(ops.incompressible_step(*s, 1.0, False, accuracy=1e-5)[:3])
