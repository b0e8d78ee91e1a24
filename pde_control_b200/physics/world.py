"""Simulation-object layer: State / Physics / StateDependency / StateCollection / CollectivePhysics / World.

Host-side mirror of the reference interface the hot path is called through (no arithmetic here):
  State, StateDependency, Physics, Static     PF/physics/physics.py:8-151
  StateCollection, CollectivePhysics          PF/physics/collective.py:7-159
  StateProxy, World                           PF/physics/world.py:19-261
Same names, argument order, defaults and error behaviour (AssertionError on cyclic blocking dependencies,
AttributeError on item assignment).  States are immutable value objects: `copied_with` returns a new object.
`.data` tensors are torch.Tensor on CUDA.
"""
import copy
import itertools

_uid = itertools.count()


class State(object):
    """One configuration of a physical system (PF/physics/physics.py:8-68).  Attributes: name, tags, age."""

    def __init__(self, name=None, tags=(), age=0.0, batch_size=None):
        self._batch_size = batch_size
        self.name = str(name) if name is not None else "%s_%d" % (self.__class__.__name__.lower(), next(_uid))
        self.tags = tuple(tags)
        self.age = age

    def copied_with(self, **kwargs):
        new = copy.copy(self)
        for key, value in kwargs.items():
            if not hasattr(new, key):
                raise TypeError("%s has no attribute '%s'" % (type(self).__name__, key))
            setattr(new, key, value)
        return new

    def default_physics(self):
        return STATIC

    @property
    def state(self):
        return self

    def __repr__(self):
        return '%s[name="%s"]' % (self.__class__.__name__, self.name)


class Gravity(State):
    """Gravity state (PF/physics/field/effect.py:77-104): a scalar acts along -up (axis 0, physics_config.py:25-29), a vector
    as given.  A World adds one by default (world.py:128-130)."""

    def __init__(self, gravity=-9.81, name="gravity", **kwargs):
        assert gravity is not None
        State.__init__(self, name=name, tags=("gravity",), **kwargs)
        self.gravity = gravity


class StateDependency(object):
    """Request for other states, resolved by CollectivePhysics (PF/physics/physics.py:71-100)."""

    def __init__(self, parameter_name, tag, single_state=False, blocking=False, state_name=None):
        self.parameter_name = parameter_name
        self.tag = tag
        self.single_state = single_state
        self.blocking = blocking
        self.state_name = state_name
        if state_name is not None:
            assert single_state

    def __repr__(self):
        return "'%s' <- %s %s %s" % (self.parameter_name, "single" if self.single_state else "all",
                                     "updated" if self.blocking else "previous", self.state_name or self.tag)


class Physics(object):
    """Stateless law moving a State forward in time (PF/physics/physics.py:103-137)."""

    def __init__(self, dependencies=()):
        self.dependencies = tuple(dependencies)

    def step(self, state, dt=1.0, **dependent_states):
        raise NotImplementedError(self)

    @property
    def blocking_dependencies(self):
        return [d for d in self.dependencies if d.blocking]

    @property
    def non_blocking_dependencies(self):
        return [d for d in self.dependencies if not d.blocking]

    @property
    def dependency_names(self):
        return [d.parameter_name for d in self.dependencies]


class Static(Physics):
    """No natural evolution: only the age advances (PF/physics/physics.py:139-148)."""

    def step(self, state, dt=1.0, **dependent_states):
        return state.copied_with(age=state.age + dt)


STATIC = Static()


class StateCollection(dict):
    """Immutable name -> State mapping with attribute access (PF/physics/collective.py:7-108)."""

    def __init__(self, states=None):
        if states is None:
            states = {}
        elif not isinstance(states, dict):
            states = {s.name: s for s in states}
        dict.__init__(self, states)

    def __setitem__(self, key, value):
        raise AttributeError("StateCollections are immutable")

    def all_with_tag(self, tag):
        return [s for s in self.values() if tag in s.tags]

    def all_instances(self, cls):
        return [s for s in self.values() if isinstance(s, cls)]

    def state_added(self, state):
        assert state.name not in self, 'A state with name "%s" is already present. Use state_replaced() to replace it.' % state.name
        return StateCollection(list(self.values()) + [state])

    def state_replaced(self, new_state):
        assert new_state.name in self, 'No state found with name "%s"' % new_state.name
        return StateCollection([new_state if s.name == new_state.name else s for s in self.values()])

    def state_removed(self, state):
        name = state if isinstance(state, str) else state.name
        return StateCollection([s for s in self.values() if s.name != name])

    def __getitem__(self, item):
        if isinstance(item, str):
            return dict.__getitem__(self, item)
        if isinstance(item, (list, tuple)):
            return type(item)(self[i] for i in item)
        try:
            return dict.__getitem__(self, item.name)
        except AttributeError:
            raise ValueError("Illegal argument: %s" % (item,))

    def __getattr__(self, item):
        if item.startswith("__"):
            raise AttributeError(item)
        try:
            return dict.__getitem__(self, item)
        except KeyError:
            raise AttributeError(item)

    def __contains__(self, item):
        if isinstance(item, str):
            return dict.__contains__(self, item)
        if hasattr(item, "name"):
            return dict.__contains__(self, item.name)
        raise ValueError("Illegal type: %s" % type(item))

    def __hash__(self):
        return 0

    @property
    def states(self):
        return self

    def __repr__(self):
        return "[" + ", ".join(str(s) for s in self.values()) + "]"


CollectiveState = StateCollection


class CollectivePhysics(Physics):
    """Steps every state of a collection, honouring blocking dependencies (PF/physics/collective.py:111-190)."""

    def __init__(self):
        Physics.__init__(self, ())
        self.physics = {}            # state name -> Physics

    def for_(self, state):
        return self.physics.get(state.name) or state.default_physics()

    def add(self, name, physics):
        self.physics[name] = physics

    def remove(self, name):
        self.physics.pop(name, None)

    def step(self, state_collection, dt=1.0, **dependent_states):
        assert len(dependent_states) == 0
        if len(state_collection) == 0:
            return state_collection
        pending = list(state_collection.values())
        done = {}
        for _sweep in range(len(state_collection)):
            progressed = StateCollection(done)
            for state in tuple(pending):
                physics = self.for_(state)
                if self._fulfilled(physics.blocking_dependencies, state_collection, progressed):
                    nxt = self.substep(state, state_collection, dt, partial_next_state_collection=progressed)
                    assert isinstance(nxt, State), "step() called on %s did not return a State but '%s' for state '%s'" % (
                        type(physics).__name__, nxt, state)
                    assert nxt.name == state.name, "The state name must remain constant during step(). Caused by '%s' on state '%s'." % (
                        type(physics).__name__, state)
                    done[nxt.name] = nxt
                    pending.remove(state)
                    progressed = StateCollection(done)
            if not pending:
                return StateCollection([done[name] for name in state_collection.keys()])
        message = "Cyclic blocking_dependencies in simulation: %s" % pending
        for state in pending:
            physics = self.for_(state)
            message += '\nState "%s" with physics "%s" depends on %s' % (state, physics, physics.blocking_dependencies)
        raise AssertionError(message)

    def substep(self, state, state_collection, dt, override_physics=None, partial_next_state_collection=None):
        physics = self.for_(state) if override_physics is None else override_physics
        deps = {}
        self._gather(physics.dependencies, state_collection, deps)
        if partial_next_state_collection is not None:
            self._gather(physics.blocking_dependencies, partial_next_state_collection, deps)
        return physics.step(state, dt, **deps)

    @staticmethod
    def _matches(dep, collection):
        if dep.state_name is not None:
            return [collection[dep.state_name]] if dep.state_name in collection else []
        tags = dep.tag if isinstance(dep.tag, (list, tuple)) else [dep.tag]
        return [s for s in collection.values() if any(t in s.tags for t in tags)]

    def _gather(self, dependencies, collection, out):
        for dep in dependencies:
            found = self._matches(dep, collection)
            if dep.single_state:
                assert len(found) == 1, "Dependency %s requires 1 state but found %d" % (dep, len(found))
                out[dep.parameter_name] = found[0]
            else:
                out[dep.parameter_name] = tuple(found)

    def _fulfilled(self, dependencies, all_states, computed):
        for dep in dependencies:
            needed = self._matches(dep, all_states)
            if any(s.name not in computed for s in needed):
                return False
        return True


class StateProxy(object):
    """Live handle to a named state of a World (PF/physics/world.py:19-97): attribute reads see the current state,
    attribute writes replace the state in the world."""

    def __init__(self, world, state_name):
        object.__setattr__(self, "world", world)
        object.__setattr__(self, "state_name", state_name)

    @property
    def state(self):
        return self.world.state[self.state_name]

    @state.setter
    def state(self, state):
        assert state.name == self.state_name
        self.world.state = self.world.state.state_replaced(state)

    @property
    def physics(self):
        return self.world.physics.for_(self.state)

    @physics.setter
    def physics(self, phys):
        self.world.physics.add(self.state_name, phys)

    def step(self, dt=1.0, physics=None):
        self.world.step(self, dt=dt, physics=physics)

    def __getattr__(self, item):
        return getattr(self.state, item)

    def __setattr__(self, key, value):
        if key in ("state", "physics"):
            object.__setattr__(self, key, value)
        else:
            self.state = self.state.copied_with(**{key: value})


class World(object):
    """Container of states and their physics (PF/physics/world.py:104-261)."""

    def __init__(self, batch_size=None, add_default_objects=True):
        self._state = StateCollection()
        self.physics = CollectivePhysics()
        self.observers = set()
        self.batch_size = batch_size
        if add_default_objects:
            self.add(Gravity())                                # world.py:128-130

    @property
    def state(self):
        return self._state

    @state.setter
    def state(self, state):
        assert isinstance(state, StateCollection)
        self._state = state
        for observer in self.observers:
            observer(self)

    @property
    def age(self):
        return max([s.age for s in self._state.values()] or [0.0])

    def step(self, state=None, dt=1.0, physics=None):
        if state is None:
            if physics is None:
                physics = self.physics
            self.state = physics.step(self._state, dt)
            return self.state
        if isinstance(state, StateProxy):
            state = state.state
        nxt = self.physics.substep(state, self._state, dt, override_physics=physics)
        self.state = self._state.state_replaced(nxt)
        return nxt

    def stepped(self, state=None, dt=1.0, physics=None):
        if state is None:
            return (physics or self.physics).step(self._state, dt)
        if isinstance(state, StateProxy):
            state = state.state
        return self.physics.substep(state, self._state, dt, override_physics=physics)

    def add(self, state, physics=None):
        if isinstance(state, (tuple, list)):
            assert isinstance(physics, (tuple, list)) or physics is None
            physics = physics if physics is not None else [None] * len(state)
            return [self.add(s, p) for s, p in zip(state, physics)]
        assert isinstance(state, State)
        if state._batch_size is None:
            state._batch_size = self.batch_size
        self.state = self._state.state_added(state)
        if physics is not None:
            self.physics.add(state.name, physics)
        elif state.default_physics() is not None:
            self.physics.add(state.name, state.default_physics())
        return StateProxy(self, state.name)

    def remove(self, obj):
        if isinstance(obj, type):
            for s in self._state.all_instances(obj):
                self.remove(s)
            return
        if isinstance(obj, StateProxy):
            obj = obj.state
        self.state = self._state.state_removed(obj)
        self.physics.remove(obj if isinstance(obj, str) else obj.name)

    def reset(self, batch_size=None, add_default_objects=True):
        self._state = StateCollection()
        self.physics = CollectivePhysics()
        self.batch_size = batch_size
        if add_default_objects:
            self.add(Gravity())

    def clear(self):
        self.reset(self.batch_size)

    def get_physics(self, state):
        if isinstance(state, StateProxy):
            state = state.state
        return self.physics.for_(state)

    def __getattr__(self, item):
        if item.startswith("_") or item in ("physics", "observers", "batch_size"):
            raise AttributeError(item)
        if item in self._state:
            return StateProxy(self, item)
        raise AttributeError(item)
