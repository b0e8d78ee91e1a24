"""PDEExecutor: binds a frame sequence to a World (src/control/hierarchy.py:8-117).

execute_step: publish the predicted next state as a NextStatePrediction, call world.step(dt), read world.state.
partition:    ask the PDE's OP network for the centre frame.
load:         per-scope checkpoints with the OP_{n/2} -> OP_n fallback rule (hierarchy.py:66-88), on torch state_dicts.
"""
import os

import torch

from ..physics import State, StateProxy, StateCollection
from .sequences import SeqFrame, PartitioningExecutor, TYPE_KEYFRAME, TYPE_UNKNOWN, LinearSequence


class NextStatePrediction(State):
    """Carries the predicted world state of the next frame (hierarchy.py:105-117)."""

    def __init__(self, prediction, tags=("next_state_prediction",), name="next", age=0.0, batch_size=None):
        State.__init__(self, name=name, tags=tags, age=age, batch_size=batch_size)
        assert prediction is None or isinstance(prediction, StateCollection)
        self.prediction = prediction

    def __repr__(self):
        return self.__class__.__name__


class StateFrame(SeqFrame):
    def __init__(self, index, type):
        SeqFrame.__init__(self, index, type)
        self.worldstate = None

    def __getitem__(self, item):
        if isinstance(item, StateProxy):
            item = item.state
        return self.worldstate[item]


class PDEExecutor(PartitioningExecutor):

    def __init__(self, world, pde, target_state, trainable_networks, dt):
        self.world = world
        self.pde = pde
        self.worldsteps = 0
        self.next_state_prediction = NextStatePrediction(None)
        self.world.add(self.next_state_prediction)
        self.target_state = target_state
        self.trainable_networks = trainable_networks
        self.dt = dt

    def create_frame(self, index, step_count):
        frame = StateFrame(index, type=TYPE_KEYFRAME if index in (0, step_count) else TYPE_UNKNOWN)
        if index == 0:
            frame.worldstate = self.world.state
        if index == step_count:
            frame.worldstate = self.target_state
        return frame

    def execute_step(self, initial_frame, target_frame, sequence):
        PartitioningExecutor.execute_step(self, initial_frame, target_frame, sequence)
        assert initial_frame.index == self.worldsteps == target_frame.index - 1
        ws = initial_frame.worldstate
        if isinstance(sequence, LinearSequence):
            predicted_ws = self.target_state
        else:
            assert target_frame.worldstate is not None
            predicted_ws = target_frame.worldstate
        carrier = ws[self.next_state_prediction].copied_with(prediction=predicted_ws)
        initial_frame.worldstate = ws.state_replaced(carrier)
        self.world.state = initial_frame.worldstate
        self.world.step(dt=self.dt)
        self.worldsteps += 1
        if target_frame is sequence[-1]:
            self.world.remove(NextStatePrediction)
        target_frame.worldstate = self.world.state

    def partition(self, n, initial_frame, target_frame, center_frame):
        PartitioningExecutor.partition(self, n, initial_frame, target_frame, center_frame)
        center_frame.worldstate = self.pde.predict(n, initial_frame.worldstate, target_frame.worldstate,
                                                   trainable="OP%d" % n in self.trainable_networks)
        if center_frame.index == self.worldsteps + 1:
            assert center_frame.worldstate is not None
            self.next_state_prediction = self.next_state_prediction.copied_with(prediction=center_frame.worldstate)
            initial_frame.worldstate = self.world.state.state_replaced(self.next_state_prediction)

    def load(self, max_n, checkpoint_dict, preload_n, session=None, logf=print):
        """checkpoint_dict: scope name ('CFE', 'OP2', ...) -> path of a torch state_dict file."""
        def restore(path, scope):
            sd = torch.load(os.path.expanduser(path), map_location="cpu")
            if scope not in self.pde.networks:
                raise KeyError("network scope %s has not been built yet (run one step first)" % scope)
            self.pde.networks[scope].load_state_dict(sd)

        if "CFE" in checkpoint_dict:
            logf("Loading CFE from %s..." % checkpoint_dict["CFE"])
            restore(checkpoint_dict["CFE"], "CFE")
        n = 2
        while n <= max_n:
            if n == max_n and not preload_n:
                return
            i = n
            while i >= 2 and "OP%d" % i not in checkpoint_dict:
                i //= 2
            if i < 2:
                raise KeyError("no checkpoint for OP%d or any coarser OP network" % n)
            logf("Loading OP%d from %s%s..." % (n, checkpoint_dict["OP%d" % i], "" if i == n else " (OP%d weights)" % i))
            restore(checkpoint_dict["OP%d" % i], "OP%d" % n)
            n *= 2
