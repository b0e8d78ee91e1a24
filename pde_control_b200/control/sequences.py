"""Frame schedules: in which order the OP networks predict frames and the solver executes steps.

Behavioural contract (the executor call order) = src/control/sequences.py:40-124; the design is this repo's own.
A schedule is DATA: `schedule(kind, n)` produces the flat list of events

    ("partition", length, first_frame)      -> executor.partition(length, frames[first], frames[first + length], frames[centre])
    ("execute", frame)                      -> executor.execute_step(frames[frame], frames[frame + 1], sequence)

once, iteratively, from the step count alone (an explicit work stack, no recursion, no executor involved), and a sequence
object is a frame table plus a replay loop over that list.  Because the plan exists before anything runs, callers can
inspect it (`sequence.events`), count network evaluations per scope (`op_counts`), or capture the replay in a CUDA graph.
The order is pinned event for event against the reference module in tests/test_host_api.py.
"""
TYPE_UNKNOWN, TYPE_PLANNED, TYPE_REAL, TYPE_KEYFRAME = 0, 1, 2, 3

PARTITION, EXECUTE = "partition", "execute"

_STAGGERED, _REFINED, _SKIP, _LINEAR = "staggered", "refined", "skip", "linear"


def schedule(kind, n):
    """The event list of a length-n sequence of the given kind ('staggered', 'refined', 'skip', 'linear')."""
    if kind == _LINEAR:
        return [(EXECUTE, i) for i in range(n)]
    if kind == _SKIP:
        return [(PARTITION, n, 0)] if n > 1 else [(EXECUTE, 0)]
    if kind not in (_STAGGERED, _REFINED):
        raise ValueError("unknown schedule kind %r" % (kind,))
    events = []
    # work items, popped from the end: ("span", length, first, refresh) = resolve frames first .. first+length;
    # ("emit", event) = an event deferred until the items pushed after it are done
    todo = [("span", n, 0, False)]
    while todo:
        item = todo.pop()
        if item[0] == "emit":
            events.append(item[1])
            continue
        _, length, first, refresh = item
        if length == 1:
            events.append((EXECUTE, first))
            continue
        half = length // 2
        events.append((PARTITION, length, first))
        # pushed in reverse order of execution
        todo.append(("span", half, first + half, refresh if kind == _REFINED else False))
        if kind == _REFINED and refresh:
            # the first half has been executed: its end frame is REAL now, so the targets ahead are predicted again from it
            todo.append(("emit", (PARTITION, length, first + half)))
            todo.append(("emit", (PARTITION, length, first + length)))
        todo.append(("span", half, first, kind == _REFINED))
    return events


def op_counts(events):
    """{length: number of OP_length evaluations} of an event list."""
    counts = {}
    for e in events:
        if e[0] == PARTITION:
            counts[e[1]] = counts.get(e[1], 0) + 1
    return counts


class SeqFrame(object):
    def __init__(self, index, type=TYPE_UNKNOWN):
        self.index, self.type = index, type

    def next(self):
        return self.index + 1

    def __repr__(self):
        return "Frame#%d" % self.index


class PartitioningExecutor(object):
    """Consumer of a schedule.  The base class only tracks the frame types and checks the protocol: a step starts from a
    REAL (or key) frame, a prediction needs both ends at least PLANNED."""

    verbose = False

    def create_frame(self, index, step_count):
        return SeqFrame(index, TYPE_KEYFRAME if index in (0, step_count) else TYPE_UNKNOWN)

    def execute_step(self, initial_frame, target_frame, sequence):
        if self.verbose:
            print("Execute -> %d" % target_frame.index)
        assert initial_frame.type >= TYPE_REAL, "step from a frame that has not been simulated"
        if target_frame.type < TYPE_REAL:
            target_frame.type = TYPE_REAL

    def partition(self, n, initial_frame, target_frame, center_frame):
        if self.verbose:
            print("Partition length %d sequence (from %d to %d) at frame %d"
                  % (n, initial_frame.index, target_frame.index, center_frame.index))
        assert TYPE_UNKNOWN not in (initial_frame.type, target_frame.type), "prediction between unknown frames"
        center_frame.type = TYPE_PLANNED


class PartitionedSequence(object):
    """Frame table + replay of a precomputed schedule."""

    kind = None         # set by the subclasses

    def __init__(self, step_count, executor):
        self.step_count = step_count
        self.executor = executor
        self.events = schedule(self.kind, step_count)
        self._frames = [executor.create_frame(i, step_count) for i in range(step_count + 1)]
        self._keep(self._frames)

    def _keep(self, frames):
        """Hook: drop the frames a schedule never touches."""

    def execute(self):
        frames, executor = self._frames, self.executor
        for event in self.events:
            if event[0] == EXECUTE:
                i = event[1]
                executor.execute_step(frames[i], frames[i + 1], self)
            else:
                _, length, first = event
                executor.partition(length, frames[first], frames[first + length], frames[first + length // 2])

    def __getitem__(self, item):
        return self._frames[item]

    def __len__(self):
        return len(self._frames)

    def __iter__(self):
        return iter(self._frames)


class StaggeredSequence(PartitionedSequence):
    """Predict the centre, resolve the first half completely, then the second half."""
    kind = _STAGGERED


class RefinedSequence(PartitionedSequence):
    """Staggered, plus: once a first half has been simulated, the predictions ahead are redone from the real state."""
    kind = _REFINED


class SkipSequence(PartitionedSequence):
    """Only the centre frame is predicted, nothing is simulated (supervised OP training)."""
    kind = _SKIP

    def _keep(self, frames):
        kept = (0, self.step_count // 2, self.step_count)
        for i in range(len(frames)):
            if i not in kept:
                frames[i] = None


class LinearSequence(PartitionedSequence):
    """No prediction: the solver steps frame by frame towards the final target."""
    kind = _LINEAR
