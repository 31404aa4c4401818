#!/usr/bin/env python
"""Put a copy of the reference's hot-path sources into oracle/_ref/ so that the UNMODIFIED reference can be
imported and timed on the GPU box, which has no /root/reference.

    python oracle/make_ref.py            (also called from __graft_entry__.build() when /root/reference exists)

TEST / BENCH INFRASTRUCTURE ONLY.  oracle/_ref/ is git-ignored (reference sources never enter the history of this
repository) but not gpurun-ignored, so it travels with the snapshot like the built libraries.  Only the files on the
path of SURVEY.md section 8a are taken: the engine (py_mc/mc.py, py_mc/mc_mpi.py) and the example models the
benchmarks name (examples/ising2D*.py, latgas*.py, analyze_ising2D_mpi.py, 2D_hardcore/).  Nothing under
py_mc_b200/ imports from there.
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
SOURCE = os.environ.get("PYMC_REFERENCE", "/root/reference")
FILES = [
    "py_mc/mc.py", "py_mc/mc_mpi.py",
    "examples/ising2D.py", "examples/ising2D_mpi.py", "examples/analyze_ising2D_mpi.py",
    "examples/latgas.py", "examples/latgas_move.py",
    "examples/2D_hardcore/model_setup.py", "examples/2D_hardcore/2D_hardcore.py",
    "LICENSE",
]


def make(verbose=False):
    """Copy the files; returns DEST, or None if the reference tree is not on this machine."""
    if not os.path.isdir(SOURCE):
        return DEST if os.path.isdir(DEST) else None
    for rel in FILES:
        src, dst = os.path.join(SOURCE, rel), os.path.join(DEST, rel)
        if not os.path.exists(src):
            continue
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not os.path.exists(dst) or os.path.getmtime(dst) < os.path.getmtime(src):
            shutil.copy2(src, dst)
            if verbose:
                print("oracle/_ref <-", rel)
    return DEST


def reference_root():
    """Where the unmodified reference can be imported from on this machine: oracle/_ref (travels to the GPU box),
    else /root/reference (build container), else None."""
    if os.path.exists(os.path.join(DEST, "py_mc", "mc.py")):
        return DEST
    if os.path.exists(os.path.join(SOURCE, "py_mc", "mc.py")):
        return SOURCE
    return None


if __name__ == "__main__":
    out = make(verbose=True)
    print(out if out else "no reference tree at %s" % SOURCE)
    sys.exit(0 if out else 1)
