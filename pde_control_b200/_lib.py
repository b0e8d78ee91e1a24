"""ctypes binding of libpdec.so (the C-ABI CUDA library declared in include/pdec.h).

There is NO CPU fallback: if the library is missing or a call fails, a PdecError is raised.
"""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PDEC_LIB_PATH") or os.path.join(HERE, "libpdec.so")     # override: A/B builds (scripts/diag)

MODE_CONSTANT, MODE_REPLICATE, MODE_CIRCULAR = 0, 1, 2
EFFECT_NONE, EFFECT_GROW, EFFECT_ADD = 0, 1, 2
STATIC_MAX_PAIRS = 131072          # PDEC_STATIC_MAX_PAIRS of include/pdec.h (pdec_burgers_pair_loss_static)

# field extrapolation name -> resample/pad mode (reference: PF/physics/field/grid.py:240-249)
MODE_OF_EXTRAPOLATION = {"constant": MODE_CONSTANT, "boundary": MODE_REPLICATE, "periodic": MODE_CIRCULAR}


class PdecError(RuntimeError):
    def __init__(self, code, where):
        self.code = code
        msg = _lib().pdec_error_string(code).decode() if _LIB is not None else "library not loaded"
        RuntimeError.__init__(self, "%s failed: %s (code %d)" % (where, msg, code))


_LIB = None

_c = ctypes
_P = _c.c_void_p
_SIGNATURES = {
    "pdec_version": (_c.c_int, []),
    "pdec_error_string": (_c.c_char_p, [_c.c_int]),
    "pdec_device_count": (_c.c_int, []),
    "pdec_burgers_trajectory_fwd": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_double,
                                                _c.c_double, _c.c_float, _c.c_float, _c.c_int, _c.c_int, _c.c_int, _P]),
    "pdec_burgers_trajectory_bwd": (_c.c_int, [_P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int,
                                                _c.c_double, _c.c_double, _c.c_float, _c.c_float, _c.c_int, _c.c_int,
                                                _c.c_int, _P]),
    "pdec_burgers_pair_loss_workspace_bytes": (_c.c_size_t, [_c.c_int, _c.c_int, _c.c_int]),
    "pdec_burgers_pair_loss": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_double, _c.c_double,
                                           _c.c_float, _c.c_float, _c.c_int, _c.c_int, _c.c_float, _c.c_float, _P,
                                           _c.c_size_t, _P]),
    "pdec_burgers_pair_loss_static": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_double, _c.c_double,
                                                  _c.c_float, _c.c_float, _c.c_int, _c.c_int, _c.c_float, _c.c_float, _P]),
    "pdec_burgers_pair_loss_window_supported": (_c.c_int, [_c.c_int, _c.c_int, _c.c_float, _c.c_float, _c.c_int]),
    "pdec_burgers_pair_loss_window": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_double,
                                                  _c.c_double, _c.c_float, _c.c_float, _c.c_int, _c.c_int, _c.c_float, _c.c_float, _P,
                                                  _c.c_size_t, _P]),
    "pdec_burgers_pair_loss_window_strided": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_double,
                                                          _c.c_double, _c.c_float, _c.c_float, _c.c_int, _c.c_int, _c.c_float, _c.c_float,
                                                          _c.c_longlong, _P, _c.c_size_t, _P]),
    "pdec_host_pipeline_create": (_c.c_int, [_c.POINTER(_P), _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int]),
    "pdec_host_pipeline_destroy": (_c.c_int, [_P]),
    "pdec_burgers_pair_loss_host": (_c.c_int, [_P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_double, _c.c_double,
                                                _c.c_float, _c.c_float, _c.c_int, _c.c_int, _c.c_float, _c.c_float, _P]),
    "pdec_resample_fwd": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_longlong, _c.c_int, _P]),
    "pdec_resample_bwd_workspace_bytes": (_c.c_size_t, [_c.c_int, _c.c_int, _c.c_int, _c.c_int]),
    "pdec_resample_bwd": (_c.c_int, [_P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_longlong, _c.c_int, _P,
                                      _c.c_size_t, _P]),
    # ---- 2-D incompressible flow ----
    "pdec_curl_2d_fwd": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _P]),
    "pdec_curl_2d_bwd": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _P]),
    "pdec_advect_centered_2d_fwd": (_c.c_int, [_P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_double,
                                                _c.c_float, _c.c_float, _c.c_float, _c.c_float, _P]),
    "pdec_advect_staggered_2d_fwd": (_c.c_int, [_P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_double,
                                                 _c.c_float, _c.c_float, _c.c_float, _c.c_float, _P]),
    "pdec_advect_2d_workspace_bytes": (_c.c_size_t, [_c.c_int, _c.c_int, _c.c_int]),
    "pdec_advect_centered_2d_bwd": (_c.c_int, [_P, _P, _P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int,
                                                _c.c_double, _c.c_float, _c.c_float, _c.c_float, _c.c_float, _P, _c.c_size_t, _P]),
    "pdec_advect_staggered_2d_bwd": (_c.c_int, [_P, _P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int,
                                                 _c.c_double, _c.c_float, _c.c_float, _c.c_float, _c.c_float, _P, _c.c_size_t, _P]),
    "pdec_curl_advect_2d_fwd": (_c.c_int, [_P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_double,
                                            _c.c_float, _c.c_float, _c.c_float, _c.c_float, _P]),
    "pdec_curl_advect_2d_bwd": (_c.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_double,
                                            _c.c_float, _c.c_float, _c.c_float, _c.c_float, _P, _c.c_size_t, _P]),
    "pdec_staggered_force_loss_workspace_bytes": (_c.c_size_t, [_c.c_int, _c.c_int, _c.c_int]),
    "pdec_staggered_force_loss_fwd": (_c.c_int, [_P, _P, _P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_double,
                                                  _c.c_float, _c.c_float, _c.c_float, _c.c_float, _P, _c.c_size_t, _P]),
    "pdec_staggered_force_loss_bwd": (_c.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int,
                                                  _c.c_double, _c.c_float, _c.c_float, _c.c_float, _c.c_float, _P, _c.c_size_t, _P]),
    "pdec_mask_divergence_2d_fwd": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _P]),
    "pdec_mask_divergence_2d_bwd": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _P]),
    "pdec_subtract_pressure_gradient_2d_fwd": (_c.c_int, [_P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int,
                                                           _c.c_float, _c.c_float, _P]),
    "pdec_subtract_pressure_gradient_2d_bwd": (_c.c_int, [_P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int,
                                                           _c.c_float, _c.c_float, _P]),
    "pdec_normalize_2d_workspace_bytes": (_c.c_size_t, [_c.c_int, _c.c_int, _c.c_int]),
    "pdec_normalize_2d_fwd": (_c.c_int, [_P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_float, _P, _c.c_size_t, _P]),
    "pdec_normalize_2d_bwd": (_c.c_int, [_P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_float, _P, _c.c_size_t, _P]),
    "pdec_cg_poisson_2d_workspace_bytes": (_c.c_size_t, [_c.c_int, _c.c_int, _c.c_int]),
    "pdec_cg_poisson_2d": (_c.c_int, [_P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_float, _c.c_int, _c.c_int,
                                       _P, _c.c_size_t, _P]),
    "pdec_project_2d_supported": (_c.c_int, [_c.c_int, _c.c_int]),
    "pdec_project_2d_fwd": (_c.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_float, _c.c_float,
                                        _c.c_float, _c.c_int, _P]),
    "pdec_project_2d_bwd": (_c.c_int, [_P, _P, _P, _P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_float, _c.c_float,
                                        _c.c_float, _c.c_int, _P]),
    "pdec_poisson_direct_2d_spectral_bytes": (_c.c_size_t, []),
    "pdec_poisson_direct_2d_supported": (_c.c_int, [_c.c_int, _c.c_int]),
    "pdec_poisson_direct_2d_setup": (_c.c_int, [_P, _c.c_size_t, _c.c_int, _c.c_int, _c.c_int, _P]),
    "pdec_poisson_direct_2d": (_c.c_int, [_P, _P, _c.c_int, _c.c_int, _c.c_int, _P, _P]),
    "pdec_poisson_direct_tc_2d_spectral_bytes": (_c.c_size_t, []),
    "pdec_poisson_direct_tc_2d_setup": (_c.c_int, [_P, _c.c_size_t, _c.c_int, _c.c_int, _c.c_int, _P]),
    "pdec_poisson_direct_tc_2d": (_c.c_int, [_P, _P, _c.c_int, _c.c_int, _c.c_int, _P, _P]),
    "pdec_laplace_matrix_2d": (_c.c_int, [_P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _P]),
    "pdec_cg_poisson_2d_masked": (_c.c_int, [_P, _P, _P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_float, _c.c_int, _c.c_int,
                                              _P, _c.c_size_t, _P]),
    # ---- tiled / slab-decomposed CG ----
    "pdec_cg_slab_comm_bytes": (_c.c_size_t, [_c.c_int, _c.c_int]),
    "pdec_cg_slab_workspace_bytes": (_c.c_size_t, [_c.c_int, _c.c_int, _c.c_int]),
    "pdec_cg_slab_comm_init": (_c.c_int, [_P, _c.c_int, _c.c_int, _P]),
    "pdec_cg_poisson_2d_slab": (_c.c_int, [_P, _P, _P, _P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_float,
                                            _c.c_int, _P, _P, _c.c_size_t, _P]),
    "pdec_halo_inbox_bytes": (_c.c_size_t, [_c.c_int, _c.c_int, _c.c_int]),
    "pdec_halo_inbox_init": (_c.c_int, [_P, _c.c_int, _c.c_int, _c.c_int, _P]),
    "pdec_halo_push_rows": (_c.c_int, [_P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _P, _c.c_int, _c.c_uint, _c.c_int, _P]),
    "pdec_halo_pull_rows": (_c.c_int, [_P, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _P, _c.c_int, _c.c_uint, _c.c_int, _c.c_int,
                                        _P, _P]),
    "pdec_comm_alloc": (_c.c_int, [_c.c_size_t, _c.POINTER(_P), _c.POINTER(_c.c_ubyte)]),
    "pdec_comm_open": (_c.c_int, [_c.POINTER(_c.c_ubyte), _c.POINTER(_P)]),
    "pdec_comm_close": (_c.c_int, [_P]),
    "pdec_comm_free": (_c.c_int, [_P]),
}


class SlabCommStruct(_c.Structure):
    """pdec_slab_comm of include/pdec.h"""
    _fields_ = [("local", _P), ("peers", _P * 8), ("rank", _c.c_int), ("nranks", _c.c_int)]


def exported_symbols():
    """Names of every entry point include/pdec.h declares (used by the CPU symbol test)."""
    return sorted(_SIGNATURES)


def _lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libpdec.so is not built (%s). Run `python -m pde_control_b200.build`. "
                               "There is no CPU fallback." % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the symbol is missing: fail loudly
            fn.restype = res
            fn.argtypes = args
        _LIB = lib
    return _LIB


def lib():
    return _lib()


def check(code, where):
    if code != 0:
        raise PdecError(code, where)


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


_WORKSPACES = {}


def workspace(device, nbytes):
    """Caller-owned scratch for the C-ABI calls: one grow-only byte tensor per (device, stream).  Work on one
    stream is ordered, so consecutive ops may share it."""
    import torch
    key = (device.index if device.index is not None else torch.cuda.current_device(), torch.cuda.current_stream(device).cuda_stream)
    ws = _WORKSPACES.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(int(nbytes), 1 << 20), dtype=torch.uint8, device=device)
        _WORKSPACES[key] = ws
    return ws
