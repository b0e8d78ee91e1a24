"""pde_control_b200 -- B200-native (sm_100a) implementation of PDE-Control's differentiable-physics hot path.

Layout: csrc/ (CUDA kernels + C-ABI, built into libpdec.so), _lib.py (ctypes binding), ops.py (torch.autograd
Functions over the C-ABI), physics/ (PhiFlow-compatible Domain / grids / World / Burgers / IncompressibleFlow),
control/ (PDE, BurgersPDE, IncompressibleFluidPDE, sequences, ControlTraining).
"""
__version__ = "0.1.0"
