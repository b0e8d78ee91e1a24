"""IncrementPDE: the toy problem of the reference's sanity script (src/control/pde/value.py:16-32, scripts/sanity_check.py).
A scalar per batch element moves towards its predicted next value: the effect equals (prediction - value) and the
physics adds it.  Runs on any device (plain torch, no kernels): used to exercise sequences / executor / ControlTraining.
"""
import torch

from ..physics import CenteredGrid, Domain, Physics, StateDependency
from .pde_base import PDE


class ScalarControl(Physics):
    """value += (predicted next value - value): ScalarEffectControl + FieldPhysics of the reference in one step."""

    def __init__(self):
        Physics.__init__(self, [StateDependency("pred", "next_state_prediction", single_state=True, blocking=True)])

    def step(self, scalar, dt=1.0, pred=None):
        force = pred.prediction.scalar.data - scalar.data
        return scalar.copied_with(data=scalar.data + force, age=scalar.age + dt)


class IncrementPDE(PDE):

    def __init__(self, device="cpu", learnable=False):
        PDE.__init__(self)
        self.device = device
        self.learnable = learnable

    def create_pde(self, world, control_trainable, constant_prediction_offset):
        world.reset(world.batch_size, add_default_objects=False)
        world.add(CenteredGrid.sample(0.0, Domain([1]), batch_size=world.batch_size, name="scalar", device=self.device), physics=ScalarControl())

    def target_matching_loss(self, target_state, actual_state):
        if not self.learnable:
            return None
        d = target_state.scalar.data - actual_state.scalar.data
        return (d ** 2).sum() / 2 / d.shape[0]

    def total_force_loss(self, states):
        return None

    def predict(self, n, initial, target, trainable):
        centre_age = (initial.scalar.age + target.scalar.age) / 2
        if self.learnable:          # a one-parameter "OP network": learns the blend weight (optimum 0.5)
            net = self.network("OP%d" % n, lambda: torch.nn.Linear(1, 1, bias=False)).to(initial.scalar.data.device)
            w = net.weight.reshape(())
            data = initial.scalar.data * (1 - w) + target.scalar.data * w
        else:
            data = (initial.scalar.data + target.scalar.data) * 0.5
        return initial.state_replaced(initial.scalar.copied_with(data=data, age=centre_age))
