"""TEST INFRASTRUCTURE ONLY -- runtime shim that imports the UNMODIFIED reference
(PhiFlow 1.4.0 NumPy/SciPy backend) read-only from /root/reference.

In the build container the reference is /root/reference; on the GPU box it is the unmodified copy under baseline/_ref/
(git-ignored, staged by oracle/stage_reference.py at build time).
Used by oracle/make_golden.py to generate tests/golden/*.npz and by the CPU tests that
validate the restatement in oracle/restate.py against the real reference.

Shim items, zero edits under /root/reference:
  1. collections.{Iterable,Mapping,Sequence,Callable,...} aliases (scipy_backend.py:35)
  2. np.object/np.int/np.float/np.bool/np.complex aliases (scipy_backend.py:54, staggered_grid.py:195)
  3. scipy.signal.correlate forced to method='direct' 
  4. sys.path += /root/reference/PhiFlow
"""
import os
import sys

def _root():
    """/root/reference in the build container; on the GPU box the unmodified copy staged by oracle/stage_reference.py under the
    git-ignored baseline/_ref/ (travels with the gpurun snapshot)."""
    env = os.environ.get("PDEC_REFERENCE_ROOT")
    if env:
        return env
    if os.path.isdir("/root/reference/PhiFlow/phi"):
        return "/root/reference"
    return os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref")


REFERENCE_ROOT = _root()
_installed = False


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "PhiFlow", "phi"))


def install():
    """Apply the shim and make `import phi` resolve to the reference. Idempotent."""
    global _installed
    if _installed:
        return
    if not available():
        raise RuntimeError("reference not present at %s" % REFERENCE_ROOT)
    sys.dont_write_bytecode = True  # the mount is read-only
    import collections
    import collections.abc
    for name in ("Iterable", "Mapping", "MutableMapping", "Sequence", "Callable", "Hashable", "Sized", "Container"):
        if not hasattr(collections, name):
            setattr(collections, name, getattr(collections.abc, name))
    import numpy as np
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for name, typ in (("object", object), ("int", int), ("float", float), ("bool", bool), ("complex", complex)):
            if name not in np.__dict__:
                setattr(np, name, typ)
    import scipy.signal
    _orig = scipy.signal.correlate
    if not getattr(_orig, "_pdec_direct", False):
        def correlate(in1, in2, mode="full", method="auto"):
            return _orig(in1, in2, mode=mode, method="direct")
        correlate._pdec_direct = True
        scipy.signal.correlate = correlate
    path = os.path.join(REFERENCE_ROOT, "PhiFlow")
    if path not in sys.path:
        sys.path.insert(0, path)
    _installed = True


def flow():
    """Returns the reference's `phi.flow` namespace as a module."""
    install()
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        import phi.flow as pf
    return pf
