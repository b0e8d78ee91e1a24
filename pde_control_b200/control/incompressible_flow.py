"""IncompressibleFluidPDE and ControlledIncompressibleFlow (src/control/pde/incompressible_flow.py:6-59).

CFE network -> scalar potential -> staggered_curl_2d -> density advection, all physics through the CUDA kernels;
the force loss self-advects the staggered velocity; the target loss (FFT-weighted L1) stays in torch.fft.
"""
import numpy as np
import torch

from ..physics import CenteredGrid, Physics, StateDependency, staggered_curl_2d, semi_lagrangian, staggered_force_loss, curl_advect
from .networks import UNet
from .pde_base import PDE


class IncompressibleFluidPDE(PDE):

    def __init__(self, domain, dt):
        PDE.__init__(self)
        self.domain = domain
        self.dt = dt

    def create_pde(self, world, control_trainable, constant_prediction_offset):
        world.reset(world.batch_size, add_default_objects=True)
        world.add(CenteredGrid.sample(0, self.domain, batch_size=world.batch_size, name="density"),
                  physics=ControlledIncompressibleFlow(self.domain, control_trainable, pde=self))

    def predict(self, n, initial_worldstate, target_worldstate, trainable):
        d1, d2 = initial_worldstate.density, target_worldstate.density
        net = self.network("OP%d" % n, lambda: UNet(2, 1, levels=3)).to(d1.data.device)
        with torch.set_grad_enabled(bool(trainable) and torch.is_grad_enabled()):
            pred = net(torch.cat([d1.data, d2.data], -1))
        density_pred = d1.copied_with(data=pred, age=(d1.age + d2.age) / 2.)
        return initial_worldstate.state_replaced(density_pred)

    def target_matching_loss(self, target_state, actual_state):
        """l1 of |fft(diff)|^2 weighted in frequency space (incompressible_flow.py:23-29).  `math.fftfreq(shape)` is in 'vector'
        mode (nd.py:317-323): k has one channel per axis, the Gaussian is taken per component and l1_loss sums both channels
        against the one-channel power spectrum -- the weight of mode (ky, kx) is G(ky) + G(kx), a cross, not G(ky) * G(kx)."""
        diff = target_state.density.data - actual_state.density.data
        f = torch.fft.fft2(diff[..., 0].to(torch.complex64))
        power = f.imag ** 2 + f.real ** 2
        ky = torch.fft.fftfreq(diff.shape[1], device=diff.device)
        kx = torch.fft.fftfreq(diff.shape[2], device=diff.device)
        gauss = lambda k: torch.exp(-0.5 * k ** 2 / 0.01 ** 2)
        weights = gauss(ky)[:, None] + gauss(kx)[None, :]
        return (power * weights).abs().sum() / diff.shape[0]

    def total_force_loss(self, states):
        loss, count = 0, 0
        for s1, s2 in zip(states[:-1], states[1:]):
            if hasattr(s1.density, "velocity") and hasattr(s2.density, "velocity"):
                v1, v2 = s1.density.velocity, s2.density.velocity
                loss = loss + staggered_force_loss(v1, v2, self.dt)     # l2_loss(semi_lagrangian(v1, v1, dt) - v2), one fused term
                count += 1
        return None if count == 0 else loss / count


class ControlledIncompressibleFlow(Physics):

    def __init__(self, domain, trainable, pde=None):
        Physics.__init__(self, dependencies=[StateDependency("pred", "next_state_prediction", single_state=True, blocking=True)])
        self.domain = domain
        self.trainable = trainable
        self.pde = pde if pde is not None else PDE()

    def step(self, density, dt=1.0, pred=None):
        net = self.pde.network("CFE", lambda: UNet(2, 1, levels=2)).to(density.data.device)
        with torch.set_grad_enabled(bool(self.trainable) and torch.is_grad_enabled()):
            potential = net(torch.cat([density.data, pred.prediction.density.data], -1))
        # the U-Net output is sampled like `density` (standard_networks.py:68): same box and extrapolation
        velocity_potential = CenteredGrid(potential, box=density.box, extrapolation=density.extrapolation, batch_size=density._batch_size)
        next_density, velocity = curl_advect(density, velocity_potential, dt)         # staggered_curl_2d + semi_lagrangian, one launch
        next_density = next_density.copied_with(age=density.age + dt)
        next_density.velocity = velocity
        return next_density
