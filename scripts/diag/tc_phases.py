"""Phase times of the tensor-core direct solver (CTA 0), library built with -DTC_TIMING:
scripts/diag/build_variants.sh poisson_direct.cu tctime:-DTC_TIMING ; PDEC_LIB_PATH=pde_control_b200/variants/libpdec_tctime.so"""
import os, sys, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from pde_control_b200 import ops, _lib
t = torch.randn(64, 128, 128, device="cuda"); out = torch.empty_like(t)
for _ in range(5): ops.poisson_direct_raw(t, out, False, tensor_cores=True)
torch.cuda.synchronize()
buf = (ctypes.c_ulonglong * 16)()
L = ctypes.CDLL(os.environ["PDEC_LIB_PATH"])
L.pdec_debug_tc_times(buf)
ts = np.array(list(buf), dtype=np.int64)
names = ["start", "R loaded"] + sum([["product %d complete" % (g + 1), "epilogue %d done" % (g + 1)] for g in range(4)], [])
for i in range(1, 10): print("%-22s +%6.2f us  (at %6.2f)" % (names[i], (ts[i] - ts[i - 1]) / 1e3, (ts[i] - ts[0]) / 1e3))
