import sys, os, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import torch.multiprocessing as mp
import test_slab_multi_gpu as T
if __name__ == "__main__":
    os.makedirs("/tmp/slabout", exist_ok=True)
    mp.spawn(T._step_worker, args=(2, 29539, "/tmp/slabout"), nprocs=2, join=True)
    out = np.load("/tmp/slabout/step.npy", allow_pickle=True)[0]
    for c in (False, True):
        r = out[c]; print(c, {k: r[k] for k in ("grad", "chained_state", "chained_grad", "chained_iterations", "status2")})
