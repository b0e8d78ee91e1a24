"""AABox, Material and Domain: geometry / boundary metadata of the hot path (host side, NumPy only).

Mirrors  AABox + `box[...]`   PF/geom/_box.py:8-56, 120-135
         Material, OPEN ...   PF/physics/material.py:8-104
         Domain               PF/physics/domain.py:13-205
"""
import numpy as np


class AABox(object):
    """Axis-aligned box with float32 bounds (PF/geom/_box.py:18-24)."""

    def __init__(self, lower, upper):
        self.lower = np.atleast_1d(np.asarray(lower, np.float32))
        self.upper = np.atleast_1d(np.asarray(upper, np.float32))
        if self.lower.shape != self.upper.shape:
            self.lower, self.upper = np.broadcast_arrays(self.lower, self.upper)
            self.lower, self.upper = self.lower.copy(), self.upper.copy()

    @staticmethod
    def to_box(value, resolution_hint=None):
        if value is None:
            return AABox(np.zeros(len(resolution_hint)), resolution_hint)
        if isinstance(value, AABox):
            return value
        return AABox(np.zeros(len(np.atleast_1d(value))), value)

    @property
    def size(self):
        return self.upper - self.lower

    @property
    def rank(self):
        return self.size.shape[-1]

    def global_to_local(self, x):
        return (x - self.lower) / self.size

    def local_to_global(self, x):
        return x * self.size + self.lower

    def __eq__(self, other):
        return isinstance(other, AABox) and np.array_equal(self.lower, other.lower) and np.array_equal(self.upper, other.upper)

    def __ne__(self, other):
        return not self == other

    def __hash__(self):
        return hash((tuple(self.lower.tolist()), tuple(self.upper.tolist())))

    def __repr__(self):
        return "AABox(%s to %s)" % (self.lower.tolist(), self.upper.tolist())


class _BoxFactory(object):
    """`box[0:1]`, `box[0:4, 0:5]` (PF/geom/_box.py:120-135)."""

    def __getitem__(self, item):
        item = item if isinstance(item, tuple) else (item,)
        lower, upper = [], []
        for s in item:
            assert isinstance(s, slice) and s.step is None, "box[...] takes slices lower:upper"
            lower.append(0 if s.start is None else s.start)
            upper.append(s.stop)
        return AABox(lower, upper)


box = _BoxFactory()


class Material(object):
    """Boundary material and its three extrapolation-mode maps (PF/physics/material.py:8-98)."""

    def __init__(self, name, solid=True, friction=0.0, periodic=False):
        self.name, self.solid, self.friction, self.periodic = name, bool(solid), friction, bool(periodic)

    @property
    def extrapolation_mode(self):
        """Regular (non-vector) fields: periodic | boundary (solid) | constant (open)."""
        return "periodic" if self.periodic else ("boundary" if self.solid else "constant")

    @property
    def accessible_extrapolation_mode(self):
        return "periodic" if self.periodic else ("constant" if self.solid else "boundary")

    @property
    def vector_extrapolation_mode(self):
        if self.periodic:
            return "periodic"
        if self.solid:
            assert self.friction in (0, 1)
            return "boundary" if self.friction == 0 else "constant"
        return "constant"

    def __repr__(self):
        return self.name


OPEN = Material("open", solid=False)
CLOSED = NO_STICK = SLIPPERY = Material("slippery", solid=True, friction=0)
NO_SLIP = STICKY = Material("sticky", solid=True, friction=1)
PERIODIC = Material("periodic", solid=False, periodic=True)


class Domain(object):
    """Resolution + boundaries + box (PF/physics/domain.py:13-35).  Only a single Material for all faces is supported
    (per-side mixed boundaries are outside the hot path: SURVEY 8f N3)."""

    def __init__(self, resolution, boundaries=OPEN, box=None):
        self.resolution = np.atleast_1d(np.asarray(resolution, np.int64))
        if not isinstance(boundaries, Material):
            raise NotImplementedError("per-axis / per-side boundaries are not on the B200 path: %s" % (boundaries,))
        self.boundaries = boundaries
        self.box = AABox.to_box(box, resolution_hint=self.resolution)
        assert self.box.rank == len(self.resolution), "box rank %d does not match resolution %s" % (self.box.rank, self.resolution)

    @staticmethod
    def as_domain(domain_like):
        if isinstance(domain_like, Domain):
            return domain_like
        if isinstance(domain_like, int):
            return Domain([domain_like])
        return Domain(domain_like)

    @property
    def rank(self):
        return len(self.resolution)

    @property
    def dx(self):
        return self.box.size.astype(np.float64) / self.resolution          # grid.py:77-79 (float64 under NumPy 2)

    def box_tuple(self):
        """(lower_y, lower_x, size_y, size_x) / (lower, size): the box arguments of the C-ABI."""
        return tuple(float(v) for v in self.box.lower) + tuple(float(v) for v in self.box.size)

    def centered_shape(self, batch_size=1):
        return (batch_size,) + tuple(int(r) for r in self.resolution)

    def __eq__(self, other):
        return isinstance(other, Domain) and np.array_equal(self.resolution, other.resolution) and self.box == other.box \
            and self.boundaries is other.boundaries

    def __hash__(self):
        return hash(tuple(self.resolution.tolist()))

    def __repr__(self):
        return "Domain(%s, %s, %s)" % (self.resolution.tolist(), self.boundaries, self.box)
