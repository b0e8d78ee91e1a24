"""CPU: libpdec.so loads and exports every symbol include/pdec.h declares (no compute calls without a GPU)."""
import os
import re

from conftest import ROOT


def _declared():
    text = open(os.path.join(ROOT, "include", "pdec.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pdec_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from pde_control_b200 import _lib
    lib = _lib.lib()
    declared = _declared()
    assert declared, "no declarations parsed"
    for name in declared:
        assert hasattr(lib, name), "libpdec.so does not export %s" % name
    # and the ctypes table covers the header exactly
    assert sorted(_lib.exported_symbols()) == declared


def test_version_and_error_strings():
    from pde_control_b200 import _lib
    lib = _lib.lib()
    assert lib.pdec_version() == 100
    assert lib.pdec_error_string(0) == b"ok"
    assert b"invalid" in lib.pdec_error_string(-1)
    assert lib.pdec_device_count() >= 0


def test_ops_fail_loudly_on_cpu_tensors():
    import pytest
    import torch
    from pde_control_b200 import ops
    with pytest.raises(RuntimeError):
        ops.burgers_step(torch.zeros(2, 16), 0.1, 0.1)
