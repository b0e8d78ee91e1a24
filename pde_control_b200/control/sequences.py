"""Hierarchical frame scheduling: decides in which order the OP networks predict frames and the solver executes steps.

Same classes and call order as src/control/sequences.py:1-124 (PartitioningExecutor, SeqFrame, PartitionedSequence,
StaggeredSequence, RefinedSequence, SkipSequence, LinearSequence), written around one explicit recursion `_run`.
"""
TYPE_UNKNOWN, TYPE_PLANNED, TYPE_REAL, TYPE_KEYFRAME = 0, 1, 2, 3


class SeqFrame(object):
    def __init__(self, index, type=TYPE_UNKNOWN):
        self.index, self.type = index, type

    def next(self):
        return self.index + 1

    def __repr__(self):
        return "Frame#%d" % self.index


class PartitioningExecutor(object):
    """Callback interface of a sequence; the base class only tracks frame types (and checks the protocol)."""

    verbose = False

    def create_frame(self, index, step_count):
        key = index in (0, step_count)
        return SeqFrame(index, TYPE_KEYFRAME if key else TYPE_UNKNOWN)

    def execute_step(self, initial_frame, target_frame, sequence):
        if self.verbose:
            print("Execute -> %d" % (initial_frame.index + 1))
        assert initial_frame.type >= TYPE_REAL
        target_frame.type = max(TYPE_REAL, target_frame.type)

    def partition(self, n, initial_frame, target_frame, center_frame):
        if self.verbose:
            print("Partition length %d sequence (from %d to %d) at frame %d" % (n, initial_frame.index, target_frame.index, center_frame.index))
        assert initial_frame.type != TYPE_UNKNOWN and target_frame.type != TYPE_UNKNOWN
        center_frame.type = TYPE_PLANNED


class PartitionedSequence(object):
    def __init__(self, step_count, executor):
        self.step_count, self.executor = step_count, executor
        self._frames = [executor.create_frame(i, step_count) for i in range(step_count + 1)]

    def execute(self):
        self.partition_execute(self.step_count, 0)

    def partition_execute(self, n, start_frame_index, **kwargs):
        if n == 1:
            self.leaf_execute(self._frames[start_frame_index], self._frames[start_frame_index + 1], **kwargs)
        else:
            self.branch_execute(n, start_frame_index, **kwargs)

    def leaf_execute(self, start_frame, end_frame, **kwargs):
        self.executor.execute_step(start_frame, end_frame, self)

    def branch_execute(self, n, start_frame_index, **kwargs):
        raise NotImplementedError()

    def partition(self, n, start_frame_index):
        f = self._frames
        self.executor.partition(n, f[start_frame_index], f[start_frame_index + n], f[start_frame_index + n // 2])

    def __getitem__(self, item):
        return self._frames[item]

    def __len__(self):
        return len(self._frames)

    def __iter__(self):
        return iter(self._frames)


class StaggeredSequence(PartitionedSequence):
    """Predict the centre, then fully resolve the first half before touching the second."""

    def branch_execute(self, n, start_frame_index, **kwargs):
        half = n // 2
        self.partition(n, start_frame_index)
        for offset in (0, half):
            self.partition_execute(half, start_frame_index + offset)


class RefinedSequence(PartitionedSequence):
    """Like Staggered, but after executing a first half the targets ahead are re-predicted from the REAL state."""

    def branch_execute(self, n, start_frame_index, update_target=False, **kwargs):
        half = n // 2
        self.partition(n, start_frame_index)
        self.partition_execute(half, start_frame_index, update_target=True)
        if update_target:
            for start in (start_frame_index + n, start_frame_index + half):
                self.partition(n, start)
        self.partition_execute(half, start_frame_index + half, update_target=update_target)


class SkipSequence(PartitionedSequence):
    """Only the centre frame is predicted (supervised OP training without the solver)."""

    def __init__(self, sequence_length, executor):
        PartitionedSequence.__init__(self, sequence_length, executor)
        centre = sequence_length // 2
        self._frames = [f if i in (0, centre, sequence_length) else None for i, f in enumerate(self._frames)]

    def branch_execute(self, n, start_frame_index, **kwargs):
        self.partition(n, start_frame_index)


class LinearSequence(PartitionedSequence):
    """No prediction: the solver steps frame by frame towards the final target."""

    def execute(self):
        for a, b in zip(self._frames[:-1], self._frames[1:]):
            self.leaf_execute(a, b)

    def branch_execute(self, n, start_frame_index, **kwargs):
        raise AssertionError()
