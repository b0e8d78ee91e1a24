"""Seeded synthetic inputs of the benchmarked workloads, shared by bench.py and by the at-size parity tests so that
the tests check exactly what the bench times."""
import random

import numpy as np
import torch

DT, VISC = 0.03, 0.003


def burgers_config2_base(B, n, T, rank, device):
    """[T+1, B, n]: an exact GaussianClash solver trajectory (4 NumPy-style diffusion substeps, like the notebook's data generator);
    the OP-network stand-in is this plus seeded noise (burgers_config2_set)."""
    from pde_control_b200 import ops, datagen
    rng = np.random.RandomState(1 + rank)
    u0 = torch.tensor(datagen.gaussian_clash(B, n, rng), device=device)
    frames = ops.burgers_trajectory(u0, T, DT, VISC, substeps=4, resubstep=True)
    return torch.cat([u0[None], frames]).contiguous()


def burgers_config2_sets(B, n, T, rank, device, nsets):
    base = burgers_config2_base(B, n, T, rank, device)
    g = torch.Generator(device=device)
    g.manual_seed(7 + rank)
    return [(base + 0.02 * torch.randn(base.shape, device=device, generator=g)).contiguous() for _ in range(nsets)]


def smooth(rng, shape, scale, passes=6):
    a = rng.randn(*shape)
    for _ in range(passes):
        a = (a + np.roll(a, 1, -1) + np.roll(a, -1, -1) + np.roll(a, 1, -2) + np.roll(a, -1, -2)) / 5
    return (a / np.abs(a).max() * scale).astype(np.float32)


def fluid_config3_inputs(B, n, rank):
    """rho0, target = distribute_random_shape (src/shape_utils.py:14-22 restated in datagen), v0 = smooth staggered noise (numpy)."""
    from pde_control_b200 import datagen
    rng = np.random.RandomState(3 + rank)
    shapes = datagen.procedural_shapes(10, 32, 0)
    rho0 = datagen.distribute_random_shape([n, n], B, shapes, rng=random.Random(rank))
    target = datagen.distribute_random_shape([n, n], B, shapes, rng=random.Random(1000 + rank))
    vy0, vx0 = smooth(rng, (B, n + 1, n), 1.0), smooth(rng, (B, n, n + 1), 1.0)
    return rho0.astype(np.float32), vy0, vx0, target.astype(np.float32)


def fluid_notebook_inputs(B, n, T, rank):
    """rho0 = shapes, phi_t = smooth noise standing in for the CFE output at every step (notebook variant)."""
    from pde_control_b200 import datagen
    rng = np.random.RandomState(5 + rank)
    shapes = datagen.procedural_shapes(10, 32, 0)
    rho0 = datagen.distribute_random_shape([n, n], B, shapes, rng=random.Random(rank)).astype(np.float32)
    target = datagen.distribute_random_shape([n, n], B, shapes, rng=random.Random(1000 + rank)).astype(np.float32)
    phi = np.stack([smooth(rng, (B, n, n), 2.0) for _ in range(T)])
    return rho0, phi, target
