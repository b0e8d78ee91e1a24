"""TEST / BASELINE INFRASTRUCTURE ONLY -- stages the UNMODIFIED reference for the GPU box.

/root/reference exists only in the build container.  `stage()` copies, byte for byte and without edits,
  /root/reference/PhiFlow/phi/**/*.py (+ VERSION)      -> baseline/_ref/PhiFlow/phi/        (PhiFlow 1.4.0, NumPy/SciPy backend)
  /root/reference/src/control/sequences.py             -> baseline/_ref/src/control/sequences.py   (TensorFlow-free scheduler module)
into the git-ignored directory baseline/_ref/ (listed in .gitignore, NOT in .gpurunignore: it travels with the gpurun snapshot like
the built libpdec.so, and never enters the history -- no reference source is committed to this repository).
oracle/ref_shim.py falls back to baseline/_ref when /root/reference is absent, so that on the GPU box
  * `bench.py --impl reference` times the reference's OWN forward path (Burgers.step / IncompressibleFlow.step, NumPy backend), and
  * the `reference`-marked tests run against the live reference there too.
Called from __graft_entry__.build(); a no-op when /root/reference is absent (the GPU box uses what was staged here).
"""
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
STAGED = os.path.join(os.path.dirname(HERE), "baseline", "_ref")        # the contract's place for the UNMODIFIED reference
SOURCE = "/root/reference"


def stage(source=SOURCE, verbose=False):
    phi = os.path.join(source, "PhiFlow", "phi")
    if not os.path.isdir(phi):
        return False
    dst_phi = os.path.join(STAGED, "PhiFlow", "phi")
    if os.path.isdir(dst_phi):
        shutil.rmtree(dst_phi)
    n = 0
    for root, dirs, files in os.walk(phi):
        dirs[:] = [d for d in dirs if d != "__pycache__"]
        rel = os.path.relpath(root, phi)
        for f in files:
            if f.endswith(".py") or f == "VERSION":
                os.makedirs(os.path.join(dst_phi, rel), exist_ok=True)
                shutil.copyfile(os.path.join(root, f), os.path.join(dst_phi, rel, f))
                n += 1
    seq = os.path.join(source, "src", "control", "sequences.py")
    if os.path.exists(seq):
        os.makedirs(os.path.join(STAGED, "src", "control"), exist_ok=True)
        shutil.copyfile(seq, os.path.join(STAGED, "src", "control", "sequences.py"))
        n += 1
    if verbose:
        print("staged %d reference files under %s" % (n, STAGED))
    return True


if __name__ == "__main__":
    print(stage(verbose=True))
