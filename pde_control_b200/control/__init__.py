"""PDE-Control layer (src/control/*) over the B200 physics: PDE definitions, frame sequences, executor, training app."""
from .pde_base import PDE
from .sequences import (SeqFrame, PartitioningExecutor, PartitionedSequence, StaggeredSequence, RefinedSequence, SkipSequence,
                        LinearSequence, TYPE_UNKNOWN, TYPE_PLANNED, TYPE_REAL, TYPE_KEYFRAME)
from .hierarchy import PDEExecutor, NextStatePrediction, StateFrame
from .burgers import BurgersPDE, ReplacePhysics, GaussianClash, GaussianForce
from .incompressible_flow import IncompressibleFluidPDE, ControlledIncompressibleFlow
from .value import IncrementPDE
from .networks import OPResNet, UNet
from .data import Dataset, write_sim
from .control_training import ControlTraining
