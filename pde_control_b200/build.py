"""Builds libpdec.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

Usage: python -m pde_control_b200.build [--force]
The .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libpdec.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

COMMON = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
          "-Xcompiler", "-fPIC"]
# per-file extra flags: the physics kernels reproduce the reference's fp32 operation order -> no FMA contraction
SOURCES = {
    "abi.cu": [],
    "burgers.cu": ["-fmad=false"],
    "burgers_fast.cu": [],          # forward op order spelled out with __f*_rn intrinsics; adjoint may contract
    "burgers_fast2.cu": ["-fmad=false"],    # nvcc contracts __fmul2_rn + __fadd2_rn into FFMA2 (unlike the scalar _rn intrinsics): off; the adjoint calls fmaf / __ffma2_rn explicitly
    "fluid.cu": ["-fmad=false"],
    "cg.cu": [],
    "cg_tiled.cu": [],
    "resample.cu": ["-fmad=false"],
    "laplace_matrix.cu": [],
    "host_pipeline.cu": [],
    "poisson_direct.cu": [],
    "halo.cu": [],
    "noalloc.cu": [],               # static-scratch entry point: host wrapper only, no kernel of its own
}


def _sources():
    return [(os.path.join(CSRC, f), flags) for f, flags in SOURCES.items() if os.path.exists(os.path.join(CSRC, f))]


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [s for s, _ in _sources()] + [os.path.join(CSRC, "pdec_common.cuh"), os.path.join(CSRC, "cg_reduce.cuh"),
                                         os.path.join(HERE, "..", "include", "pdec.h"), __file__]
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force=False, verbose=False):
    if not force and not _stale():
        return LIB
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    objs = []
    headers = [os.path.join(CSRC, "pdec_common.cuh"), os.path.join(CSRC, "cg_reduce.cuh"),
               os.path.join(HERE, "..", "include", "pdec.h"), __file__]
    jobs = []
    for src, flags in _sources():
        obj = os.path.join(objdir, os.path.basename(src) + ".o")
        objs.append(obj)
        if not force and not verbose and os.path.exists(obj) and all(
                os.path.getmtime(d) <= os.path.getmtime(obj) for d in [src] + headers if os.path.exists(d)):
            continue        # object newer than its source and the shared headers
        cmd = [NVCC] + COMMON + flags + ["-c", src, "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        jobs.append((src, cmd))

    def compile_one(job):
        src, cmd = job
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
        return src, r.stderr

    # one nvcc per translation unit, all at once: the files are independent and the longest (cg.cu) bounds the build
    with ThreadPoolExecutor(max_workers=max(1, min(len(jobs), os.cpu_count() or 1))) as pool:
        for src, err in pool.map(compile_one, jobs):
            if verbose:
                print(src)
                print(err)
    cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
