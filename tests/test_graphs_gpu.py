"""GPU: CUDA-graph capture of whole trajectories (pde_control_b200/graphs.py) replays exactly what the eager path computes."""
import numpy as np
import pytest
import torch

from benchlib import inputs

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("closed", [False, True])
def test_graphed_fluid_trajectory_is_bit_identical_to_eager(closed):
    from pde_control_b200 import graphs
    rho0, vy0, vx0, target = inputs.fluid_config3_inputs(4, 64, 0)
    dev = torch.device("cuda")
    args = [torch.tensor(a, device=dev) for a in (rho0, vy0, vx0, target)]
    fn = graphs.fluid_trajectory(5, 1.0, closed, accuracy=1e-5)
    # eager, with autograd
    leaves = [a.clone().requires_grad_(True) for a in args[:3]]
    out = fn(*leaves, args[3])
    grads = torch.autograd.grad(out[0], leaves)
    graphed = graphs.GraphedFunction(fn, args, grad_inputs=(0, 1, 2))
    outs, g = graphed(*args)
    torch.cuda.synchronize()
    for a, b in zip(outs, out):
        assert torch.equal(a, b.detach())
    for a, b in zip(g, grads):
        assert torch.equal(a, b)
    # a replay on different inputs follows them (the CG iteration counts are decided on the device)
    other = [torch.tensor(a, device=dev) for a in inputs.fluid_config3_inputs(4, 64, 1)]
    outs2, g2 = graphed(*other)
    leaves = [a.clone().requires_grad_(True) for a in other[:3]]
    out2 = fn(*leaves, other[3])
    grads2 = torch.autograd.grad(out2[0], leaves)
    torch.cuda.synchronize()
    assert not torch.equal(out2[0].detach(), out[0].detach())
    for a, b in zip(outs2, out2):
        assert torch.equal(a, b.detach())
    for a, b in zip(g2, grads2):
        assert torch.equal(a, b)


def test_graphed_burgers_config1_matches_eager():
    from pde_control_b200 import graphs, datagen
    rng = np.random.RandomState(0)
    dev = torch.device("cuda")
    u0 = torch.tensor(datagen.gaussian_clash(8, 128, rng), device=dev)
    force = torch.tensor(datagen.gaussian_force(8, 128, rng), device=dev)
    target = torch.tensor(datagen.gaussian_clash(8, 128, rng), device=dev)
    fn = graphs.burgers_trajectory_loss(32, 0.03, 0.003, substeps=4, resubstep=True)
    leaves = [u0.clone().requires_grad_(True), force.clone().requires_grad_(True)]
    out = fn(leaves[0], leaves[1], target)
    grads = torch.autograd.grad(out[0], leaves)
    graphed = graphs.GraphedFunction(fn, (u0, force, target), grad_inputs=(0, 1))
    outs, g = graphed(u0, force, target)
    torch.cuda.synchronize()
    assert torch.equal(outs[0], out[0].detach()) and torch.equal(outs[1], out[1].detach())
    assert torch.equal(g[0], grads[0]) and torch.equal(g[1], grads[1])
