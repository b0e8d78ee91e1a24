"""BurgersPDE (src/control/pde/burgers.py:5-42), ReplacePhysics (:83-89), GaussianClash / GaussianForce (:92-134).

The force loss -- for every consecutive frame pair `l2(s2 - Burgers().step(s1))` -- is ONE fused kernel launch over all
pairs and batch rows (pdec_burgers_pair_loss) that also produces dLoss/dframe; its L1 twin ("Total Force") comes out of
the same launch.
"""
import numpy as np
import torch

from .. import ops, datagen
from .._lib import MODE_OF_EXTRAPOLATION
from ..physics import BurgersVelocity, Physics, StateDependency, CenteredGrid
from .networks import OPResNet
from .pde_base import PDE


class BurgersPDE(PDE):

    def __init__(self, domain, viscosity, dt, periodic_diffusion="fft"):
        """periodic_diffusion: how the force loss diffuses on a PERIODIC domain.  'fft' (default) = what Burgers.step does there, as
        in the reference (diffuse switches to its spectral branch on periodic fields, PF/physics/field/util.py:30-37), so the loss
        equals sum_t l2(s[t+1] - Burgers().step(s[t])).  'stencil' = opt-in extension: the wrapped explicit 3-point stencil, all
        pairs in the one fused kernel launch (the north star's "explicit viscosity diffusion on a periodic 1D grid"); it is NOT
        the reference's periodic step.  Non-periodic domains always use the explicit stencil, like the reference."""
        PDE.__init__(self)
        assert periodic_diffusion in ("fft", "stencil")
        self.domain = domain
        self.viscosity = viscosity
        self.dt = dt
        self.periodic_diffusion = periodic_diffusion

    def create_pde(self, world, control_trainable, constant_prediction_offset):
        world.reset(world.batch_size, add_default_objects=False)
        u0 = BurgersVelocity(self.domain, viscosity=self.viscosity, batch_size=world.batch_size, name="burgers")
        world.add(u0, ReplacePhysics())

    def target_matching_loss(self, target_state, actual_state):
        diff = target_state.burgers.velocity.data - actual_state.burgers.velocity.data
        return (diff ** 2).sum() / 2 / diff.shape[0]                    # math.l2_loss (PF/math/nd.py:111-124)

    def total_force_loss(self, states):
        v0 = states[0].burgers.velocity
        frames = torch.stack([s.burgers.velocity.tensor for s in states])        # [T+1, B, N]
        if v0.extrapolation == "periodic" and self.periodic_diffusion == "fft":
            from ..physics.burgers import fourier_diffuse
            T, B, N = frames.shape[0] - 1, frames.shape[1], frames.shape[2]
            # all pairs at once: circular advection by the kernel (viscosity 0 adds exactly 0), spectral diffusion with cuFFT
            advected = ops.burgers_step(frames[:-1].reshape(T * B, N), self.dt, 0.0, lower=float(v0.box.lower[0]),
                                        size=float(v0.box.size[0]), mode=MODE_OF_EXTRAPOLATION[v0.extrapolation])
            stepped = fourier_diffuse(advected, float(self.dt) * float(self.viscosity), float(v0.dx[0]))
            d = frames[1:].reshape(T * B, N) - stepped
            self.scalars["Total Force"] = d.detach().abs().sum() / B
            return (d ** 2).sum() / 2 / B
        l2, l1 = ops.burgers_pair_loss(frames, self.dt, self.viscosity, lower=float(v0.box.lower[0]), size=float(v0.box.size[0]),
                                       mode=MODE_OF_EXTRAPOLATION[v0.extrapolation], substeps=1)
        self.scalars["Total Force"] = l1
        return l2

    def predict(self, n, initial_worldstate, target_worldstate, trainable):
        b1, b2 = initial_worldstate.burgers, target_worldstate.burgers
        net = self.network("OP%d" % n, OPResNet).to(b1.velocity.data.device)
        with torch.set_grad_enabled(bool(trainable) and torch.is_grad_enabled()):
            predicted = net(b1.velocity.data, b2.velocity.data)
        new_field = b1.copied_with(velocity=predicted, age=(b1.age + b2.age) / 2.)
        return initial_worldstate.state_replaced(new_field)


class ReplacePhysics(Physics):
    """The Burgers state is REPLACED by the OP prediction; the solver only runs inside the loss."""

    def __init__(self):
        Physics.__init__(self, dependencies=[StateDependency("next_state_prediction", "next_state_prediction", single_state=True, blocking=True)])

    def step(self, state, dt=1.0, next_state_prediction=None):
        return next_state_prediction.prediction.burgers


class _Analytic1D(object):
    """Stand-in for the reference's AnalyticField initialisers: sampled once at the cell centres of a domain."""

    def at(self, grid):
        x = grid.points[..., 0].cpu().numpy()
        data = self.sample_at(x).astype(np.float32)
        return grid.copied_with(data=torch.as_tensor(data, device=grid.data.device).unsqueeze(-1))


class GaussianClash(_Analytic1D):
    def __init__(self, batch_size):
        self.batch_size = batch_size

    def sample_at(self, x):
        n = x.shape[-1]
        return datagen.gaussian_clash(self.batch_size, n) if np.allclose(x[0], datagen.cell_centers(n)) else self._general(x)

    def _general(self, x):
        raise NotImplementedError("GaussianClash is defined on the unit box of the notebook")


class GaussianForce(_Analytic1D):
    def __init__(self, batch_size):
        self.loc = np.random.uniform(0.4, 0.6, batch_size)[:, None]
        self.amp = (np.random.uniform(-0.05, 0.05, batch_size) * 32)[:, None]
        self.sig = np.random.uniform(0.1, 0.4, batch_size)[:, None]

    def sample_at(self, x):
        return self.amp * np.exp(-0.5 * (x - self.loc) ** 2 / self.sig ** 2)
