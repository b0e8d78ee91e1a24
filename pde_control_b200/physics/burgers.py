"""BurgersVelocity state and Burgers physics over the fused CUDA step.

Mirrors  BurgersVelocity  PF/physics/burgers.py:11-23
         Burgers          PF/physics/burgers.py:26-48  (step_velocity = advect -> diffuse -> effects, one kernel here)
`tf_substeps`: the reference's `diffuse` collapses k substeps into one explicit step on TensorFlow and performs true
substepping on NumPy.  The default follows the named reference path (TensorFlow); pass
tf_substeps=False for the NumPy behaviour used by the notebook's data generation.
"""
import torch

from .. import ops
from .._lib import MODE_OF_EXTRAPOLATION, EFFECT_GROW, EFFECT_ADD
from .domain import Domain
from .field import CenteredGrid, FieldEffect, GROW, ADD, effect_applied
from .world import State, Physics, StateDependency


class DomainState(State):
    """State living on a Domain (PF/physics/domain.py:226-247)."""

    def __init__(self, domain, tags=(), name=None, age=0.0, batch_size=None):
        State.__init__(self, name=name, tags=tags, age=age, batch_size=batch_size)
        self.domain = Domain.as_domain(domain)

    @property
    def resolution(self):
        return self.domain.resolution

    @property
    def rank(self):
        return self.domain.rank

    def centered_grid(self, name, value, device="cuda"):
        if isinstance(value, CenteredGrid):
            grid = value
        else:
            grid = CenteredGrid.sample(value, self.domain, batch_size=self._batch_size, name=name, device=device)
        mode = self.domain.boundaries.extrapolation_mode
        return grid if grid.extrapolation == mode else grid.copied_with(extrapolation=mode)


class BurgersVelocity(DomainState):

    def __init__(self, domain, velocity=0, viscosity=0.1, tags=("burgers", "velocityfield"), name="burgers", age=0.0,
                 batch_size=None, device="cuda"):
        DomainState.__init__(self, domain, tags=tags, name=name, age=age, batch_size=batch_size)
        self.viscosity = viscosity
        self.velocity = self.centered_grid("velocity", velocity, device=device)

    def copied_with(self, **kwargs):
        if "velocity" in kwargs and not isinstance(kwargs["velocity"], CenteredGrid):
            v = kwargs["velocity"]
            v = v if isinstance(v, torch.Tensor) else torch.as_tensor(v, dtype=torch.float32, device=self.velocity.data.device)
            kwargs["velocity"] = self.velocity.copied_with(data=v.reshape(self.velocity.data.shape))
        return State.copied_with(self, **kwargs)

    def default_physics(self):
        return Burgers()


def fourier_diffuse(u, amount, dx):
    """diffuse, periodic branch (PF/physics/field/util.py:30-37) for u [B, n]: differentiable (torch.fft)."""
    import math
    import torch
    k = torch.fft.fftfreq(u.shape[1], device=u.device, dtype=torch.float32) / dx
    kernel = torch.exp(-(2.0 * math.pi) ** 2 * k * k * amount)
    return torch.fft.ifft(torch.fft.fft(u.to(torch.complex64), dim=1) * kernel, dim=1).real.contiguous()


class Burgers(Physics):

    def __init__(self, default_viscosity=0.1, viscosity=None, diffusion_substeps=1, tf_substeps=True):
        Physics.__init__(self, [StateDependency("effects", "velocity_effect", blocking=True)])
        if viscosity is not None:
            default_viscosity = viscosity
        self.default_viscosity = default_viscosity
        self.diffusion_substeps = diffusion_substeps
        self.tf_substeps = tf_substeps

    def step(self, v, dt=1.0, effects=()):
        if isinstance(v, BurgersVelocity):
            return v.copied_with(velocity=self.step_velocity(v.velocity, v.viscosity, dt, effects, self.diffusion_substeps,
                                                             self.tf_substeps), age=v.age + dt)
        return self.step_velocity(v, self.default_viscosity, dt, effects, self.diffusion_substeps, self.tf_substeps)

    @staticmethod
    def step_velocity(v, viscosity, dt, effects, diffusion_substeps, tf_substeps=True):
        """One fused kernel launch: semi-Lagrangian self-advection, explicit diffusion and the first GROW/ADD effect;
        further effects are added afterwards (effect.py:48-53)."""
        assert isinstance(v, CenteredGrid) and v.rank == 1, "the B200 Burgers path is 1-D"
        effects = tuple(effects)
        if v.extrapolation == "periodic":
            # the reference's diffuse switches to its FFT branch on periodic fields (PF/physics/field/util.py:30-37): circular
            # advection by the kernel (viscosity 0 adds exactly 0), then the spectrum times exp(-(2 pi k / dx)^2 nu dt) with cuFFT
            # through torch.fft (the reference calls the backend's fft the same way), then the effects
            out = ops.burgers_step(v.tensor, dt, 0.0, lower=float(v.box.lower[0]), size=float(v.box.size[0]),
                                   mode=MODE_OF_EXTRAPOLATION[v.extrapolation])
            v = v.with_tensor(fourier_diffuse(out, float(dt) * float(viscosity), float(v.dx[0])))
            for effect in effects:
                v = effect_applied(effect, v, dt)
            return v.copied_with(age=v.age + dt)
        force, effect_mode, rest = None, None, effects
        if effects and isinstance(effects[0], FieldEffect) and effects[0].mode in (GROW, ADD):
            force = effects[0].field.at(v).tensor
            if force.shape[0] != v.tensor.shape[0]:
                force = force.expand_as(v.tensor)
            effect_mode = EFFECT_GROW if effects[0].mode == GROW else EFFECT_ADD
            rest = effects[1:]
        out = ops.burgers_step(v.tensor, dt, viscosity, lower=float(v.box.lower[0]), size=float(v.box.size[0]),
                               mode=MODE_OF_EXTRAPOLATION[v.extrapolation], force=force, effect_mode=effect_mode,
                               substeps=diffusion_substeps, resubstep=not tf_substeps)
        v = v.with_tensor(out)
        for effect in rest:
            v = effect_applied(effect, v, dt)
        return v.copied_with(age=v.age + dt)
