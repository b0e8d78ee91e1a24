"""Field types on the hot path: CenteredGrid, StaggeredGrid, FieldEffect (data holders over CUDA tensors).

Mirrors  CenteredGrid   PF/physics/field/grid.py:24-216
         StaggeredGrid  PF/physics/field/staggered_grid.py:56-249
         Field operator overloads  PF/physics/field/field.py:129-167
         FieldEffect, GROW/ADD, effect_applied  PF/physics/field/effect.py:16-59
`.data` keeps the reference layout [batch, spatial..., channels]; the kernels see the channel-dropped views
(`.tensor`).  Field arithmetic (+ - *) is plain torch on the stored tensors; the physics (advection, diffusion,
projection) lives in ops.py / libpdec.so, not here.
"""
import numpy as np
import torch

from .domain import AABox, Domain, extrapolation_sides
from .world import State

GROW, ADD, FIX = "grow", "add", "fix"
_MODES = ("periodic", "constant", "boundary")


def _check_extrapolation(extrapolation):
    """One mode, or per axis a mode or a (lower, upper) pair of modes (mixed boundaries, PF/physics/domain.py:13-35)."""
    if isinstance(extrapolation, str):
        assert extrapolation in _MODES, extrapolation
        return extrapolation
    out = []
    for e in extrapolation:
        if isinstance(e, str):
            assert e in _MODES, e
            out.append(e)
        else:
            lo, hi = e
            assert lo in _MODES and hi in _MODES, e
            out.append((lo, hi))
    return tuple(out)


def _as_tensor(data, like=None):
    if isinstance(data, torch.Tensor):
        return data
    t = torch.as_tensor(np.asarray(data, np.float32))
    return t.to(like.device) if like is not None else t


class Field(State):
    def __init__(self, name=None, tags=(), age=0.0, batch_size=None):
        State.__init__(self, name=name, tags=tags, age=age, batch_size=batch_size)

    # ---- operator overloads: elementwise on the data (field.py:129-167) ----
    def _operand(self, other):
        if isinstance(other, Field):
            return other.data
        return other

    def __add__(self, other): return self._dataop(other, torch.add)
    def __radd__(self, other): return self._dataop(other, lambda a, b: b + a)
    def __sub__(self, other): return self._dataop(other, torch.sub)
    def __rsub__(self, other): return self._dataop(other, lambda a, b: b - a)
    def __mul__(self, other): return self._dataop(other, torch.mul)
    def __rmul__(self, other): return self._dataop(other, lambda a, b: b * a)
    def __truediv__(self, other): return self._dataop(other, torch.div)
    def __neg__(self): return self._dataop(-1.0, torch.mul)


class CenteredGrid(Field):
    """Regular grid sampled at cell centres; data [B, spatial..., C]."""

    def __init__(self, data, box=None, extrapolation="boundary", name=None, tags=(), age=0.0, batch_size=None):
        Field.__init__(self, name=name, tags=tags, age=age, batch_size=batch_size)
        data = _as_tensor(data)
        while data.dim() < 2:
            data = data.unsqueeze(0)
        self.data = data
        self.box = AABox.to_box(box, resolution_hint=self.resolution)
        self.extrapolation = _check_extrapolation(extrapolation)

    @staticmethod
    def sample(value, domain, batch_size=None, name=None, device="cuda"):
        """Constant value or tensor on `domain` with Material.extrapolation_mode (grid.py:43-58)."""
        assert isinstance(domain, Domain)
        shape = (batch_size or 1,) + tuple(int(r) for r in domain.resolution) + (1,)
        if isinstance(value, (int, float)):
            data = torch.full(shape, float(value), device=device)
        else:
            data = _as_tensor(value).to(device).reshape(shape)
        return CenteredGrid(data, box=domain.box, extrapolation=domain.boundaries.extrapolation_mode, name=name, batch_size=batch_size)

    @property
    def resolution(self):
        return np.asarray(self.data.shape[1:-1], np.int64)

    @property
    def rank(self):
        return self.data.dim() - 2

    @property
    def dx(self):
        return self.box.size.astype(np.float64) / self.resolution

    @property
    def component_count(self):
        return self.data.shape[-1]

    @property
    def tensor(self):
        """Channel-dropped contiguous view [B, spatial...] handed to the kernels (scalar fields only)."""
        assert self.data.shape[-1] == 1, "kernels take scalar fields"
        return self.data[..., 0]

    @property
    def points(self):
        """Cell centres [1, spatial..., rank] (CenteredGrid.getpoints, grid.py:176-181: float64 linspace cast to float32)."""
        axes = [np.linspace(0.5 / n, 1 - 0.5 / n, int(n)).astype(np.float32) for n in self.resolution]
        mesh = np.stack(np.meshgrid(*axes, indexing="ij"), -1)[None].astype(np.float32)
        return torch.as_tensor(mesh * self.box.size + self.box.lower, device=self.data.device)

    def with_data(self, data):
        if isinstance(data, CenteredGrid):
            data = data.data
        return self.copied_with(data=_as_tensor(data, self.data))

    def with_tensor(self, tensor):
        return self.copied_with(data=tensor.unsqueeze(-1))

    def compatible(self, other):
        return isinstance(other, (CenteredGrid, Domain)) and self.box == other.box and np.array_equal(self.resolution, other.resolution)

    def at(self, other):
        """Resampling to another field's sample points (grid.py:109-124).  Supported: a compatible centred grid (identity)
        and a staggered grid of the same domain: component d (or the single component) is sampled at the centres of the
        faces normal to axis d, which on the reference's fast path (padded by one cell with the field's extrapolation, then
        interpolate_linear at offset 0.5, grid.py:112-120 via field.py:219-226 broadcast_at) is the mean of the two adjacent cells."""
        if self.compatible(other):
            return self
        if isinstance(other, StaggeredGrid) and other.box == self.box and np.array_equal(other.resolution, self.resolution):
            if self.component_count not in (1, other.component_count):
                raise ValueError("Can only resample to staggered fields with same number of components")
            comps = []
            for d in range(other.component_count):
                c = self.data[..., d:d + 1] if self.component_count > 1 else self.data
                widths = [[0, 0]] * self.rank
                widths[d] = [1, 1]
                # a multi-component grid is unstacked first and CenteredGrid.unstack (grid.py:126-129) does not pass the
                # extrapolation on: its components are padded with the default 'boundary' whatever the field's own mode
                source = self.copied_with(data=c, extrapolation="boundary") if self.component_count > 1 else self.copied_with(data=c)
                padded = source.padded(widths).data
                lo = padded.narrow(1 + d, 0, padded.shape[1 + d] - 1)
                hi = padded.narrow(1 + d, 1, padded.shape[1 + d] - 1)
                comps.append(lo * 0.5 + hi * 0.5)                      # interpolate_linear: x0 * (1 - w) + x1 * w, w = 0.5 (nd.py:378-392)
            return other.with_data(comps)
        if isinstance(other, CenteredGrid) and self.component_count == 1:
            # generic path (Field.at, field.py:60-72): sample at the other grid's points
            return other.copied_with(data=self.sample_at(other.points))
        raise NotImplementedError("resampling of multi-component grids between different grids is not on the B200 path")

    def sample_at(self, points):
        """CenteredGrid.sample_at (grid.py:101-107): points [B | 1, ..., rank] in physical units -> [B, ..., 1]; local coordinates
        times the resolution minus one half, then the backend's resample with this field's extrapolation (ops.resample)."""
        from .. import ops
        from .._lib import MODE_OF_EXTRAPOLATION
        assert self.component_count == 1 and self.rank in (1, 2)
        if not isinstance(self.extrapolation, str):
            return self._padded_resample(points)
        points = _as_tensor(points, self.data).to(self.data.device)
        lower = torch.as_tensor(self.box.lower, device=points.device)
        size = torch.as_tensor(self.box.size, device=points.device)
        res = torch.as_tensor(np.asarray(self.resolution, np.float32), device=points.device)
        coords = (points - lower) / size * res - 0.5
        return ops.resample(self.tensor, coords, MODE_OF_EXTRAPOLATION[self.extrapolation]).unsqueeze(-1)

    def _padded_resample(self, points):
        """sample_at for a per-side extrapolation (grid.py:211-216): pad ONE cell on every side with the side's own mode, then resample
        the padded array with 'replicate' at coordinates shifted by that cell.  (Not the same as 'constant' sampling: between the last
        cell and a zero-padded ring the value is interpolated towards 0 instead of dropping to exactly 0.)"""
        from .. import ops
        from .._lib import MODE_REPLICATE
        padded = self.padded([[1, 1]] * self.rank).tensor.contiguous()
        points = _as_tensor(points, self.data).to(self.data.device)
        lower = torch.as_tensor(self.box.lower, device=points.device)
        size = torch.as_tensor(self.box.size, device=points.device)
        res = torch.as_tensor(np.asarray(self.resolution, np.float32), device=points.device)
        coords = (points - lower) / size * res + 0.5
        return ops.resample(padded, coords, MODE_REPLICATE).unsqueeze(-1)

    def padded(self, widths):
        """math.pad with the field's extrapolation (grid.py:164-170); plain torch, off the hot path.  Per-side modes are padded in the
        reference's pass order (backend_helper.py:22-36): all 'circular' sides first, then 'replicate', then 'constant'."""
        if isinstance(widths, int):
            widths = [[widths, widths]] * self.rank
        widths = [[int(lo), int(hi)] for lo, hi in widths]
        names = {"constant": "constant", "boundary": "replicate", "periodic": "circular"}
        sides = extrapolation_sides(self.extrapolation, self.rank)
        t = self.data.movedim(-1, 1)                           # [B, C, spatial...]
        for single in ("periodic", "boundary", "constant"):
            pad, any_width = [], False
            for ax in reversed(range(self.rank)):
                lo = widths[ax][0] if sides[ax][0] == single else 0
                hi = widths[ax][1] if sides[ax][1] == single else 0
                pad += [lo, hi]
                any_width = any_width or lo > 0 or hi > 0
            if not any_width:
                continue
            t = torch.nn.functional.pad(t, pad) if single == "constant" else torch.nn.functional.pad(t, pad, mode=names[single])
        w = np.asarray(widths, np.float64)
        new_box = AABox(self.box.lower - w[:, 0] * self.dx, self.box.upper + w[:, 1] * self.dx)
        return self.copied_with(data=t.movedim(1, -1), box=new_box)

    def unstack(self):
        return [self.copied_with(data=self.data[..., c:c + 1]) for c in range(self.component_count)]

    def _dataop(self, other, op):
        return self.copied_with(data=op(self.data, self._operand(other)))

    def __repr__(self):
        return "Grid[%s(%d), size=%s]" % ("x".join(str(r) for r in self.resolution), self.component_count, self.box.size)


def staggered_component_box(resolution, axis, box):
    """Box of staggered component `axis`: half a cell larger on both sides along its own axis (staggered_grid.py:49-53)."""
    unit = np.zeros(len(resolution))
    unit[axis] = float(box.size[axis]) / resolution[axis]
    return AABox(box.lower - unit / 2, box.upper + unit / 2)


class StaggeredGrid(Field):
    """Face-sampled vector field: `.data` is a tuple of per-axis CenteredGrids, component d has one more sample along
    axis d (staggered_grid.py:33-53, 78-92)."""

    def __init__(self, data, box=None, name=None, extrapolation="boundary", tags=(), age=0.0, batch_size=None):
        Field.__init__(self, name=name, tags=tags, age=age, batch_size=batch_size)
        self.extrapolation = _check_extrapolation(extrapolation)
        if isinstance(data, torch.Tensor) or isinstance(data, np.ndarray):
            components = unstack_staggered_tensor(_as_tensor(data))
        else:
            components = [c.data if isinstance(c, CenteredGrid) else _as_tensor(c) for c in data]
            components = [c if c.dim() == len(components) + 2 else c.unsqueeze(-1) for c in components]
        res = list(components[0].shape[1:-1])
        res[0] -= 1
        self._resolution = np.asarray(res, np.int64)
        self.box = AABox.to_box(box, resolution_hint=self._resolution)
        self.data = tuple(CenteredGrid(c, box=staggered_component_box(self._resolution, d, self.box), extrapolation=extrapolation,
                                       name="%s.%s" % (self.name, "yx"[d] if len(components) == 2 else d), batch_size=batch_size)
                          for d, c in enumerate(components))

    @staticmethod
    def sample(value, domain, batch_size=None, name=None, device="cuda"):
        B = batch_size or 1
        res = [int(r) for r in domain.resolution]
        comps = []
        for d in range(len(res)):
            shape = [B] + res + [1]
            shape[1 + d] += 1
            comps.append(torch.full(shape, float(value), device=device))
        return StaggeredGrid(comps, box=domain.box, name=name, extrapolation=domain.boundaries.vector_extrapolation_mode, batch_size=batch_size)

    @property
    def resolution(self):
        return self._resolution

    @property
    def rank(self):
        return len(self._resolution)

    @property
    def dx(self):
        return self.box.size.astype(np.float64) / self._resolution

    @property
    def component_count(self):
        return len(self.data)

    @property
    def tensors(self):
        """Channel-dropped component tensors (vy [B,ny+1,nx], vx [B,ny,nx+1]) handed to the kernels."""
        return tuple(c.tensor for c in self.data)

    def unstack(self):
        return self.data

    def with_data(self, data):
        return StaggeredGrid(data, box=self.box, name=self.name, extrapolation=self.extrapolation, tags=self.tags, age=self.age,
                             batch_size=self._batch_size)

    def with_tensors(self, tensors):
        return self.with_data([t.unsqueeze(-1) for t in tensors])

    def copied_with(self, **kwargs):
        if "extrapolation" in kwargs:
            new = Field.copied_with(self, **kwargs)
            new.data = tuple(c.copied_with(extrapolation=kwargs["extrapolation"]) for c in new.data)
            return new
        return Field.copied_with(self, **kwargs)

    def staggered_tensor(self):
        """[B, ny+1, nx+1, 2] (stack_staggered_components, staggered_grid.py:41-46)."""
        comps = []
        for d, c in enumerate(self.data):
            t = c.data.movedim(-1, 1)
            pad = []
            for ax in reversed(range(self.rank)):
                pad += [0, 0 if ax == d else 1]
            comps.append(torch.nn.functional.pad(t, pad).movedim(1, -1))
        return torch.cat(comps, -1)

    def compatible(self, other):
        return isinstance(other, StaggeredGrid) and self.box == other.box and np.array_equal(self.resolution, other.resolution)

    def _dataop(self, other, op):
        if isinstance(other, StaggeredGrid):
            assert self.compatible(other), "Fields are not compatible: %s and %s" % (self, other)
            return self.with_data([op(a.data, b.data) for a, b in zip(self.data, other.data)])
        return self.with_data([op(a.data, other) for a in self.data])

    def __repr__(self):
        return "StaggeredGrid[%s, size=%s]" % ("x".join(str(r) for r in self.resolution), self.box.size)


def unstack_staggered_tensor(tensor):
    """[B, ny+1, nx+1, 2] -> components with the surplus row/column removed (staggered_grid.py:33-39)."""
    rank = tensor.dim() - 2
    out = []
    for d in range(rank):
        sl = [slice(None)] + [slice(None) if ax == d else slice(None, -1) for ax in range(rank)] + [slice(d, d + 1)]
        out.append(tensor[tuple(sl)].contiguous())
    return out


class FieldEffect(State):
    """A field that acts on the states carrying one of `targets` as a tag (effect.py:16-45)."""

    def __init__(self, field, targets, mode=GROW, bounds=None, tags=("effect",), name=None, age=0.0, batch_size=None):
        targets = (targets,) if isinstance(targets, str) else tuple(targets)
        tags = tuple(tags) + tuple("%s_effect" % t for t in targets)
        State.__init__(self, name=name, tags=tags, age=age, batch_size=batch_size)
        assert mode in (GROW, ADD, FIX)
        self.field, self.targets, self.mode, self.bounds = field, targets, mode, bounds

    def __repr__(self):
        return "%s(%s to %s)" % (self.mode, self.field, self.targets)


def effect_applied(effect, field, dt):
    """effect_applied (effect.py:48-59).  GROW: field + effect * dt, ADD: field + effect, FIX: effect inside `bounds`, field outside."""
    resampled = effect.field.at(field)
    if effect.mode == GROW:
        return field + resampled * dt
    if effect.mode == ADD:
        return field + resampled
    if effect.mode == FIX:
        assert effect.bounds is not None
        from .geom import GeometryMask
        inside = GeometryMask([effect.bounds]).at(field)
        if isinstance(field, StaggeredGrid):
            return field.with_data([torch.where(m.data > 0, e.data, f.data) for m, e, f in zip(inside.data, resampled.data, field.data)])
        return field.copied_with(data=torch.where(inside.data > 0, resampled.data, field.data))
    raise ValueError("Invalid mode: %s" % effect.mode)
