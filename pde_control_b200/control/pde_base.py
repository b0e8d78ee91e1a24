"""PDE: the problem-definition interface of PDE-Control (src/control/pde/pde_base.py:8-99), eager-PyTorch edition.

`placeholder_state` keeps its name and role (a world state of correctly shaped tensors for time `age`); there is no
graph/placeholder distinction in eager mode -- ControlTraining writes the batch data into these states.
Networks are ordinary `torch.nn.Module`s kept per variable scope name ('OP2', 'OP4', ..., 'CFE') in `self.networks`.
"""
import torch


class PDE(object):

    def __init__(self):
        self.fields = {}
        self.scalars = {}
        self.networks = torch.nn.ModuleDict()

    def create_pde(self, world, control_trainable, constant_prediction_offset):
        raise NotImplementedError(self)

    def placeholder_state(self, world, age):
        """World state at time `age` with zero-filled tensors of the right shapes (pde_base.py:43-58)."""
        states = []
        for s in world.state.values():
            states.append(_zeros_like_state(s).copied_with(age=age))
        return type(world.state)(states)

    def predict(self, n, initial_worldstate, target_worldstate, trainable):
        raise NotImplementedError(self)

    def target_matching_loss(self, target_state, actual_state):
        raise NotImplementedError(self)

    def total_force_loss(self, states):
        raise NotImplementedError(self)

    def network(self, scope, factory):
        """Module of variable scope `scope`, created on first use (tf.variable_scope + AUTO_REUSE)."""
        if scope not in self.networks:
            self.networks[scope] = factory()
        return self.networks[scope]


def _zeros_like_state(state):
    from ..physics.field import CenteredGrid, StaggeredGrid
    new = state.copied_with()
    for key, value in list(vars(state).items()):
        if isinstance(value, CenteredGrid):
            setattr(new, key, value.copied_with(data=torch.zeros_like(value.data)))
        elif isinstance(value, StaggeredGrid):
            setattr(new, key, value.with_data([torch.zeros_like(c.data) for c in value.data]))
        elif key == "data" and isinstance(value, torch.Tensor):
            setattr(new, key, torch.zeros_like(value))
    return new
