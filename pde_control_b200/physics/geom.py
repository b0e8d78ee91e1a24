"""Geometries, analytic fields and the effect shorthands that feed the fluid step (host side; masks are static per scene).

Mirrors  Sphere, Union / union()      PF/geom/_sphere.py:8-29, PF/geom/_union.py:7-44
         GeometryMask (= mask)        PF/physics/field/mask.py:12-55 (antialias=False: 1 inside, 0 outside)
         ConstantField                PF/physics/field/constant.py:11-61
         mask * rate                  PF/physics/field/analytic.py:43-44, 77-103 (symbolic product, evaluated at the target field)
         Inflow, Accelerator, Fan, ConstantVelocity, HeatSource, ColdSource   PF/physics/field/effect.py:62-72

A mask is evaluated on the host in NumPy the way the reference does it (float32 sample points; the sphere test in float64
because the centre and radius are Python numbers there) and cached on the device per (geometry, grid).
"""
import warnings

import numpy as np
import torch

from .domain import AABox
from .field import ADD, FIX, GROW, CenteredGrid, Field, FieldEffect, StaggeredGrid


class Geometry(object):
    def lies_inside(self, location):
        raise NotImplementedError(self)

    def key(self):
        raise NotImplementedError(self)

    def __add__(self, other):
        return union(self, other)

    @property
    def rank(self):
        raise NotImplementedError(self)


def _box_lies_inside(self, location):
    """_box.py:58-61: bounds included, all axes."""
    location = np.asarray(location)
    return np.all((location >= self.lower) & (location <= self.upper), axis=-1, keepdims=True)


# AABox lives in domain.py (it is also the domain's box); give it the Geometry protocol here
AABox.lies_inside = _box_lies_inside
AABox.key = lambda self: ("box", tuple(self.lower.tolist()), tuple(self.upper.tolist()))
AABox.__add__ = lambda self, other: union(self, other)


class Sphere(Geometry):
    def __init__(self, center, radius):
        self.center = np.atleast_1d(np.asarray(center))
        self.radius = radius

    @property
    def rank(self):
        return len(self.center)

    def lies_inside(self, location):
        """_sphere.py:24-28: |x - c|^2 <= r^2."""
        location = np.asarray(location)
        return np.sum((location - self.center) ** 2, axis=-1, keepdims=True) <= self.radius ** 2

    def key(self):
        return ("sphere", tuple(float(c) for c in self.center), float(self.radius))

    def __repr__(self):
        return "Sphere(%s, %s)" % (self.center.tolist(), self.radius)


class Union(Geometry):
    def __init__(self, geometries):
        assert len(geometries) > 0
        self.geometries = tuple(geometries)

    @property
    def rank(self):
        return self.geometries[0].rank

    def lies_inside(self, location):
        inside = self.geometries[0].lies_inside(location)
        for g in self.geometries[1:]:
            inside = inside | g.lies_inside(location)
        return inside

    def key(self):
        return ("union",) + tuple(g.key() for g in self.geometries)


def union(*geometries):
    """_union.py:33-44: flattens nested unions."""
    if len(geometries) == 1 and isinstance(geometries[0], (tuple, list)):
        geometries = geometries[0]
    assert len(geometries) > 0, "the empty geometry is not on the B200 path"
    flat = ()
    for g in geometries:
        flat += g.geometries if isinstance(g, Union) else (g,)
    return Union(flat)


_MASKS = {}


def _grid_key(grid):
    return (tuple(int(r) for r in grid.resolution), tuple(grid.box.lower.tolist()), tuple(grid.box.upper.tolist()), str(grid.data.device))


def geometry_mask_tensor(geometry, grid):
    """float32 [1, spatial..., 1]: 1 where the centred grid's sample points lie inside the geometry."""
    key = (geometry.key(), _grid_key(grid))
    if key not in _MASKS:
        inside = geometry.lies_inside(grid.points.cpu().numpy())
        _MASKS[key] = torch.as_tensor(inside.astype(np.float32), device=grid.data.device)
    return _MASKS[key]


class AnalyticField(Field):
    """A field without sample points of its own: it is evaluated at the points of the field it meets (analytic.py:9-44)."""

    def __init__(self, name=None):
        Field.__init__(self, name=name)

    def _at_centered(self, grid, component):
        """float32 tensor broadcastable to grid.data: this field (its component `component`, or all of them when None) at grid.points."""
        raise NotImplementedError(self)

    def at(self, other):
        """Field.at (field.py:60-72) / broadcast_at (field.py:219-226): a staggered target is served component by component."""
        if isinstance(other, StaggeredGrid):
            return other.with_data([torch.zeros_like(c.data) + self._at_centered(c, d) for d, c in enumerate(other.data)])
        assert isinstance(other, CenteredGrid), other
        return other.copied_with(data=torch.zeros_like(other.data[..., :1]) + self._at_centered(other, None))

    def _dataop(self, other, op):
        assert not isinstance(other, Field), "products of two analytic fields are not on the B200 path"
        return _ScaledField(self, other, op)


class GeometryMask(AnalyticField):
    def __init__(self, geometries, name="mask"):
        AnalyticField.__init__(self, name=name)
        self.geometry = union(geometries) if isinstance(geometries, (tuple, list)) else geometries

    geometries = property(lambda self: self.geometry)
    component_count = 1

    def _at_centered(self, grid, component):
        return geometry_mask_tensor(self.geometry, grid)


mask = GeometryMask


class ConstantField(AnalyticField):
    """One value per component everywhere (constant.py:11-61); a scalar serves every component of a staggered target."""

    def __init__(self, value=1.0, name=None):
        AnalyticField.__init__(self, name=name)
        self.value = np.atleast_1d(np.asarray(value, np.float32))

    @property
    def component_count(self):
        return len(self.value)

    def _at_centered(self, grid, component):
        if component is None:
            return torch.as_tensor(self.value, device=grid.data.device)
        return float(self.value[component if len(self.value) > 1 else 0])


class _ScaledField(AnalyticField):
    """mask * rate (analytic.py:77-103): `factor` is a number or one number per component."""

    def __init__(self, field, factor, op):
        AnalyticField.__init__(self, name=field.name)
        self.field, self.op = field, op
        self.factor = factor if np.ndim(factor) == 0 else np.asarray(factor, np.float32)

    @property
    def component_count(self):
        return 1 if np.ndim(self.factor) == 0 else len(self.factor)

    def _at_centered(self, grid, component):
        base = self.field._at_centered(grid, component)
        if not isinstance(base, torch.Tensor):
            base = torch.as_tensor(base, dtype=torch.float32, device=grid.data.device)
        if np.ndim(self.factor) == 0:
            return self.op(base, float(self.factor))
        if component is None:
            return self.op(base, torch.as_tensor(self.factor, device=grid.data.device))
        return self.op(base, float(self.factor[component]))


def Inflow(geometry, rate=1.0, target="density"):
    return FieldEffect(GeometryMask(geometry, name="inflow") * rate, target, GROW, tags=("inflow", "effect"))


def Accelerator(geometry, acceleration):
    return FieldEffect(GeometryMask(geometry, name="fan") * acceleration, ("velocity",), GROW, tags=("fan", "effect"))


def Fan(*args, **kwargs):
    warnings.warn("'Fan' was renamed to 'Accelerator' in version 1.0.2.", DeprecationWarning)
    return Accelerator(*args, **kwargs)


def ConstantVelocity(geometry, velocity):
    return FieldEffect(ConstantField(velocity), ("velocity",), FIX, bounds=geometry, tags=("effect",))


def HeatSource(geometry, rate, name=None):
    return FieldEffect(GeometryMask(geometry, name="heat-source") * rate, ("temperature",), GROW, name=name)


def ColdSource(geometry, rate, name=None):
    return FieldEffect(GeometryMask(geometry, name="heat-source") * -rate, ("temperature",), GROW, name=name)
