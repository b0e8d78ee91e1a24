"""PhiFlow-1.4-compatible simulation objects over the B200 kernels (the names `from phi.flow import *` provides on the
hot path)."""
from .world import (State, StateDependency, Physics, Static, STATIC, StateCollection, CollectiveState, CollectivePhysics,
                    StateProxy, World, Gravity)
from .domain import AABox, box, Material, OPEN, CLOSED, NO_STICK, SLIPPERY, NO_SLIP, STICKY, PERIODIC, Domain
from .field import (Field, CenteredGrid, StaggeredGrid, FieldEffect, GROW, ADD, FIX, effect_applied, unstack_staggered_tensor,
                    staggered_component_box)
from .burgers import DomainState, BurgersVelocity, Burgers
from .fluid import (Fluid, IncompressibleFlow, INCOMPRESSIBLE_FLOW, SparseCG, poisson_solve, divergence_free, semi_lagrangian,
                    staggered_curl_2d, gravity_tensor, Obstacle, FourierSolver, SparseSciPy, staggered_force_loss, curl_advect)
from .geom import (Geometry, Sphere, Union, union, GeometryMask, mask, ConstantField, AnalyticField, Inflow, Accelerator, Fan,
                   ConstantVelocity, HeatSource, ColdSource)


class _Advect(object):
    semi_lagrangian = staticmethod(semi_lagrangian)
    advect = staticmethod(semi_lagrangian)


advect = _Advect()
