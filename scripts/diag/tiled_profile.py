"""Diagnostic: where one iteration of cg_tiled_kernel spends its time (library built with -DPDEC_TILED_PROFILE)."""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from pde_control_b200 import ops, _lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
rhs = torch.randn(1, n, n, device="cuda"); out = torch.empty_like(rhs)
it = torch.zeros(1, dtype=torch.int32, device="cuda")
for _ in range(2):
    ops.cg_poisson_raw(rhs, out, False, 0.0, 1000, None, it, None)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 8)()
lib = ctypes.CDLL(_lib.LIB_PATH)
print("rc", lib.pdec_debug_tiled_profile(buf), "iterations", int(it[0]))
names = ["dots+reduce", "publish tile sums", "halo polls", "sum totals", "update+halo+matvec", "publish edges", "first halo look", "leader + rank sums poll"]
tot = sum(buf[i] for i in range(8))
for i, nm in enumerate(names):
    print("%-20s %8.0f cycles/iter  %5.1f %%" % (nm, buf[i] / 1001.0, 100.0 * buf[i] / tot))
print("total %.0f cycles/iter = %.2f us at 1.965 GHz" % (tot / 1001.0, tot / 1001.0 / 1965.0))
