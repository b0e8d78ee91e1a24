"""Diagnostic: config 3 (128^2, 16 projected steps) vs the oracle at several CG accuracies: where the state error sits (boundary flips of
the 'constant' resample rule or everywhere), gradient errors per input, adjoint iteration counts, and time per trajectory."""
import sys, os, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import numpy as np
import torch
from benchlib import inputs
from oracle import restate as R
from oracle import restate_fluid as F
from pde_control_b200 import ops

n, steps = 128, 16
rho0, vy0, vx0, target = inputs.fluid_config3_inputs(64, n, 0)
sel = [0, 1, 31, 63]
arrs = [a[sel] for a in (rho0, vy0, vx0)]
tgt = target[sel]
B = len(sel)
geo = F.Geometry(n, n)


def rel(a, b):
    a = a.detach().cpu().double(); b = b.detach().cpu().double()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


for bd in ("open", "closed"):
    to = [torch.tensor(a, requires_grad=True) for a in arrs]
    so = tuple(to)
    traj_o = []
    for _ in range(steps):
        so = F.incompressible_step(*so, 1.0, geo, bd)[:3]
        traj_o.append([t.detach() for t in so])
    R.l2_loss(so[0] - torch.tensor(tgt)).backward()
    for acc in (1e-5, 1e-6, 1e-7, 3e-8):
        tc = [torch.tensor(a, device="cuda", requires_grad=True) for a in arrs]
        sc = tuple(tc)
        worst = []
        its = []
        for k in range(steps):
            out = ops.incompressible_step(*sc, 1.0, bd == "closed", accuracy=acc, max_iterations=2000)
            sc = out[:3]
            its.append(int(out[4].max()))
            errs = [rel(a, b) for a, b in zip(sc, traj_o[k])]
            worst.append(errs)
        w = np.array(worst)
        # where is the density error at the last step?
        e = (sc[0].detach().cpu() - traj_o[-1][0]).abs()
        big = (e > 1e-3 * traj_o[-1][0].abs().max()).nonzero()
        d = sc[0] - torch.tensor(tgt, device="cuda")
        ((d * d).sum() / (2 * B)).backward()
        gerr = [rel(a.grad, b.grad) for a, b in zip(tc, to)]
        print(bd, acc, "state err per field max over steps", w.max(0), "at step", w.argmax(0), "fwd its max", max(its),
              "cells>1e-3 at last step", len(big), big[:6].tolist(), "grad err", gerr, flush=True)
    # how far is the restated fp32 CG (the reference's algorithm at its own accuracy) from the exact solve?
    for acc in (1e-5, 1e-6):
        def cg_solver(div, boundary, acc=acc):
            class S(torch.autograd.Function):
                @staticmethod
                def forward(ctx, d):
                    return F.conjugate_gradient(d, boundary, acc, 2000)[0]
                @staticmethod
                def backward(ctx, g):
                    return F.conjugate_gradient(g, boundary, acc, 2000)[0]
            return S.apply(div)
        tr = [torch.tensor(a, requires_grad=True) for a in arrs]
        sr = tuple(tr)
        w = []
        for k in range(steps):
            sr = F.incompressible_step(*sr, 1.0, geo, bd, solver=cg_solver)[:3]
            w.append([rel(a, b) for a, b in zip(sr, traj_o[k])])
        R.l2_loss(sr[0] - torch.tensor(tgt)).backward()
        print(bd, "restated fp32 CG (reference algorithm) at", acc, "vs exact solve: state", np.array(w).max(0), "grad",
              [rel(a.grad, b.grad) for a, b in zip(tr, to)], flush=True)
