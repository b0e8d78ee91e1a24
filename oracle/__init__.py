"""TEST INFRASTRUCTURE ONLY (never imported by the product path `pde_control_b200`).

oracle/restate.py   CPU restatement (PyTorch-CPU, fp32 or fp64) of the reference's TF-1.14 semantics for the
                    hot path; autograd of the restated op graph is the gradient oracle.
oracle/ref_shim.py  runtime shim importing the unmodified reference NumPy backend from /root/reference
                    (build container only).
oracle/make_golden.py  generates tests/golden/*.npz from the real reference through ref_shim.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this package.
"""
