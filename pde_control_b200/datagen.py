"""Synthetic initial conditions / forcings of the reference experiments (host-side NumPy, no physics).

GaussianClash / GaussianForce follow src/control/pde/burgers.py:92-134 (random draws in the same order so a
seeded `np.random.RandomState` reproduces the reference's fields); random shapes follow
src/shape_utils.py:14-22 with a procedural shape library (the notebook's .npz shapes do not travel).
"""
import numpy as np


def cell_centers(n, lower=0.0, size=1.0):
    """CenteredGrid.getpoints (PF/physics/field/grid.py:176-181): float64 linspace cast to float32."""
    loc = np.linspace(0.5 / n, 1 - 0.5 / n, n).astype(np.float32)
    return loc * np.float32(size) + np.float32(lower)


def gaussian_clash(batch_size, n, rng=np.random):
    """Two opposing Gaussian bumps (src/control/pde/burgers.py:98-111), sampled at the cell centres of [0,1]."""
    x = cell_centers(n)[None, :]
    draw = lambda lo, hi: rng.uniform(lo, hi, batch_size)[:, None]
    leftloc, leftamp, leftsig = draw(0.2, 0.4), draw(0, 3), draw(0.05, 0.15)
    rightloc, rightamp, rightsig = draw(0.6, 0.8), draw(-3, 0), draw(0.05, 0.15)
    left = leftamp * np.exp(-0.5 * (x - leftloc) ** 2 / leftsig ** 2)
    right = rightamp * np.exp(-0.5 * (x - rightloc) ** 2 / rightsig ** 2)
    return (left + right).astype(np.float32)


def gaussian_force(batch_size, n, rng=np.random):
    """Stationary Gaussian forcing (src/control/pde/burgers.py:118-134)."""
    x = cell_centers(n)[None, :]
    loc = rng.uniform(0.4, 0.6, batch_size)[:, None]
    amp = (rng.uniform(-0.05, 0.05, batch_size) * 32)[:, None]
    sig = rng.uniform(0.1, 0.4, batch_size)[:, None]
    return (amp * np.exp(-0.5 * (x - loc) ** 2 / sig ** 2)).astype(np.float32)


def procedural_shapes(count=10, size=32, seed=0):
    """A small library of binary-ish shapes (discs, boxes, rings, crosses) standing in for
    notebooks/shapes/*.npz (10 hand-drawn shapes; data files are not part of the build)."""
    rng = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:size, 0:size].astype(np.float32)
    c = (size - 1) / 2.0
    out = []
    for k in range(count):
        kind = k % 4
        r = size * (0.25 + 0.15 * rng.rand())
        if kind == 0:
            m = ((yy - c) ** 2 + (xx - c) ** 2) <= r * r
        elif kind == 1:
            m = (np.abs(yy - c) <= r) & (np.abs(xx - c) <= r * (0.5 + 0.5 * rng.rand()))
        elif kind == 2:
            d2 = (yy - c) ** 2 + (xx - c) ** 2
            m = (d2 <= r * r) & (d2 >= (0.5 * r) ** 2)
        else:
            m = ((np.abs(yy - c) <= r) & (np.abs(xx - c) <= r / 3)) | ((np.abs(xx - c) <= r) & (np.abs(yy - c) <= r / 3))
        out.append(m.astype(np.float32))
    return out


def distribute_random_shape(resolution, batch_size, shape_library, margin=1, rng=None):
    """Place one random library shape per batch element at a random position (src/shape_utils.py:14-22:
    `random.choice`, then `random.randint(margin, res - margin - extent - 2)` (inclusive) for y, then x).
    `rng` is a `random.Random` (default: the global `random` module, as in the reference).
    Returns [B, ny, nx] float32."""
    import random
    rng = rng or random
    ny, nx = resolution
    arr = np.zeros((batch_size, ny, nx), np.float32)
    for b in range(batch_size):
        shape = rng.choice(shape_library)
        h, w = shape.shape
        y = rng.randint(margin, ny - margin - h - 2)
        x = rng.randint(margin, nx - margin - w - 2)
        arr[b, y:y + h, x:x + w] = shape
    return arr
