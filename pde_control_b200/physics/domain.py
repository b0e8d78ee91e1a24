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


def _collapse(values):
    """struct collapse (PF/struct/tensorop.py): (a, a) -> a per axis, and one value if every axis agrees."""
    per_axis = tuple(v[0] if v[0] == v[1] else (v[0], v[1]) for v in values)
    if all(isinstance(v, str) or isinstance(v, (bool, np.bool_)) for v in per_axis) and len(set(per_axis)) == 1:
        return per_axis[0]
    return per_axis


class MixedBoundaries(tuple):
    """Per-axis / per-side boundary materials, Domain(..., boundaries=[(lower_y, upper_y), x_material]) (PF/physics/domain.py:13-35).
    Offers Material's derived properties mapped over the sides (the reference's @mappable struct properties, material.py:60-98): a
    property is one value where all sides agree, else a tuple per axis whose entries are a value or a (lower, upper) pair."""

    def __new__(cls, boundaries, rank):
        axes = []
        assert len(boundaries) == rank, "one boundary entry per spatial dimension (highest dimension first)"
        for b in boundaries:
            if isinstance(b, Material):
                axes.append((b, b))
            else:
                lo, hi = b
                assert isinstance(lo, Material) and isinstance(hi, Material)
                axes.append((lo, hi))
        return tuple.__new__(cls, axes)

    def _map(self, attr):
        return _collapse([(getattr(lo, attr), getattr(hi, attr)) for lo, hi in self])

    extrapolation_mode = property(lambda self: self._map("extrapolation_mode"))
    accessible_extrapolation_mode = property(lambda self: self._map("accessible_extrapolation_mode"))
    vector_extrapolation_mode = property(lambda self: self._map("vector_extrapolation_mode"))

    @property
    def solid(self):
        """True only if EVERY side is solid (what np.all(Material.solid(boundaries)) asks, fluid.py:100)."""
        return all(lo.solid and hi.solid for lo, hi in self)

    @property
    def periodic(self):
        return any(lo.periodic or hi.periodic for lo, hi in self)

    @property
    def solid_sides(self):
        """Flat per-side flags, axis-major: (y-lower, y-upper, x-lower, x-upper) in 2-D -- the C-ABI's PDEC_CLOSED_SIDES order."""
        return tuple(bool(m.solid) for lo, hi in self for m in (lo, hi))

    def __repr__(self):
        return "[%s]" % ", ".join(repr(lo) if lo is hi else "(%r, %r)" % (lo, hi) for lo, hi in self)


def extrapolation_sides(extrapolation, rank):
    """An extrapolation (one mode, or per axis a mode or a (lower, upper) pair) -> [(lower_mode, upper_mode)] per axis."""
    if isinstance(extrapolation, str):
        return [(extrapolation, extrapolation)] * rank
    assert len(extrapolation) == rank
    return [(e, e) if isinstance(e, str) else (e[0], e[1]) for e in extrapolation]


class Domain(object):
    """Resolution + boundaries + box (PF/physics/domain.py:13-35).  `boundaries`: one Material for all faces, or per axis (highest
    dimension first) a Material or a (lower, upper) pair -- e.g. [(CLOSED, OPEN), CLOSED] = a box open at the top."""

    def __init__(self, resolution, boundaries=OPEN, box=None):
        self.resolution = np.atleast_1d(np.asarray(resolution, np.int64))
        if not isinstance(boundaries, Material):
            mixed = MixedBoundaries(boundaries, len(self.resolution))
            uniform = all(lo is mixed[0][0] and hi is mixed[0][0] for lo, hi in mixed)
            boundaries = mixed[0][0] if uniform else mixed
        self.boundaries = boundaries
        self.box = AABox.to_box(box, resolution_hint=self.resolution)
        assert self.box.rank == len(self.resolution), "box rank %d does not match resolution %s" % (self.box.rank, self.resolution)

    @property
    def solid_sides(self):
        if isinstance(self.boundaries, MixedBoundaries):
            return self.boundaries.solid_sides
        return (bool(self.boundaries.solid),) * (2 * self.rank)

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
            and (self.boundaries is other.boundaries or self.boundaries == other.boundaries)

    def __hash__(self):
        return hash(tuple(self.resolution.tolist()))

    def __repr__(self):
        return "Domain(%s, %s, %s)" % (self.resolution.tolist(), self.boundaries, self.box)
