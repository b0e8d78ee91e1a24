"""PDEExecutor: the consumer of a frame schedule (sequences.py) that drives a World.

Behavioural contract = src/control/hierarchy.py:18-117: a simulated step publishes the state the controller should steer
towards as a `NextStatePrediction` in the world, calls `world.step(dt)` and stores `world.state` in the reached frame; a
partition event asks the PDE's OP network for the centre frame.  Own design: the executor keeps ONE cursor (`worldsteps`, the
index of the frame the world is at) and two small operations around it -- `_aim(worldstate)`, which rewrites the carrier state
in the world, and `_advance()`, which steps -- and both event handlers are written in terms of those.  Checkpoints are torch
`state_dict` files per network scope (the reference restores TF variable scopes).
"""
import os

import torch

from ..physics import State, StateProxy, StateCollection
from .sequences import SeqFrame, PartitioningExecutor, LinearSequence


class NextStatePrediction(State):
    """World state that carries the predicted world state of the next frame (what CFE steers towards)."""

    def __init__(self, prediction, tags=("next_state_prediction",), name="next", age=0.0, batch_size=None):
        State.__init__(self, name=name, tags=tags, age=age, batch_size=batch_size)
        assert prediction is None or isinstance(prediction, StateCollection)
        self.prediction = prediction

    def __repr__(self):
        return self.__class__.__name__


class StateFrame(SeqFrame):
    """A frame that holds the world state (real, predicted or target) of its time index."""

    def __init__(self, index, type):
        SeqFrame.__init__(self, index, type)
        self.worldstate = None

    def __getitem__(self, item):
        key = item.state if isinstance(item, StateProxy) else item
        return self.worldstate[key]


class PDEExecutor(PartitioningExecutor):

    def __init__(self, world, pde, target_state, trainable_networks, dt):
        self.world, self.pde, self.dt = world, pde, dt
        self.target_state = target_state
        self.trainable_networks = trainable_networks
        self.worldsteps = 0                                  # cursor: the world is at this frame
        self.next_state_prediction = NextStatePrediction(None)
        world.add(self.next_state_prediction)

    # ---- the two primitive operations -------------------------------------------------------------------------
    def _aim(self, base, prediction):
        """`base` with the carrier state pointing at `prediction`; remembers the new carrier."""
        self.next_state_prediction = base[self.next_state_prediction].copied_with(prediction=prediction)
        return base.state_replaced(self.next_state_prediction)

    def _advance(self, last):
        self.world.step(dt=self.dt)
        self.worldsteps += 1
        if last:
            self.world.remove(NextStatePrediction)           # the final frame carries no prediction
        return self.world.state

    # ---- schedule events --------------------------------------------------------------------------------------
    def create_frame(self, index, step_count):
        frame = StateFrame(index, PartitioningExecutor.create_frame(self, index, step_count).type)
        if index == step_count:
            frame.worldstate = self.target_state              # the end of the sequence is given
        elif index == 0:
            frame.worldstate = self.world.state               # the start is where the world is now
        return frame

    def execute_step(self, initial_frame, target_frame, sequence):
        PartitioningExecutor.execute_step(self, initial_frame, target_frame, sequence)
        assert (initial_frame.index, target_frame.index) == (self.worldsteps, self.worldsteps + 1), "steps must be executed in order"
        if isinstance(sequence, LinearSequence):
            goal = self.target_state                          # no OP networks: always steer towards the final target
        else:
            goal = target_frame.worldstate
            assert goal is not None, "frame %d has not been predicted" % target_frame.index
        initial_frame.worldstate = self.world.state = self._aim(initial_frame.worldstate, goal)
        target_frame.worldstate = self._advance(last=target_frame is sequence[-1])

    def partition(self, n, initial_frame, target_frame, center_frame):
        PartitioningExecutor.partition(self, n, initial_frame, target_frame, center_frame)
        trainable = ("OP%d" % n) in self.trainable_networks
        predicted = self.pde.predict(n, initial_frame.worldstate, target_frame.worldstate, trainable=trainable)
        center_frame.worldstate = predicted
        if center_frame.index == self.worldsteps + 1:
            # this prediction is the very next frame: it becomes the goal of the upcoming step right away
            assert predicted is not None
            initial_frame.worldstate = self._aim(self.world.state, predicted)

    # ---- checkpoints ------------------------------------------------------------------------------------------
    def load(self, max_n, checkpoint_dict, preload_n, session=None, logf=print):
        """checkpoint_dict: scope ('CFE', 'OP2', 'OP4', ...) -> path of a torch state_dict file.  An OP_n without a file of its
        own starts from the nearest coarser OP_{n/2^k} that has one (hierarchy.py:66-88)."""
        networks = self.pde.networks

        def restore(source_scope, scope):
            path = os.path.expanduser(checkpoint_dict[source_scope])
            note = "" if source_scope == scope else " (%s weights)" % source_scope
            logf("Loading %s from %s%s..." % (scope, path, note))
            if scope not in networks:
                raise KeyError("network scope %s has not been built yet (run one step first)" % scope)
            networks[scope].load_state_dict(torch.load(path, map_location="cpu"))

        if "CFE" in checkpoint_dict:
            restore("CFE", "CFE")
        lengths = []
        n = 2
        while n <= max_n:
            lengths.append(n)
            n *= 2
        if lengths and not preload_n and lengths[-1] == max_n:
            lengths.pop()                                     # OP_max_n is the one being trained from scratch
        for n in lengths:
            candidates = [i for i in lengths_down(n) if "OP%d" % i in checkpoint_dict]
            if not candidates:
                raise KeyError("no checkpoint for OP%d or any coarser OP network" % n)
            restore("OP%d" % candidates[0], "OP%d" % n)


def lengths_down(n):
    """n, n/2, n/4, ..., 2"""
    out = []
    while n >= 2:
        out.append(n)
        n //= 2
    return out
