"""Build recipe for the CPU oracle (test infrastructure, not product code).

    python oracle/build_oracle.py

compiles oracle/knn_oracle.c into oracle/_build/libknn_oracle.so with plain gcc.  The reference
itself is pure Python (Kdtree_Octree/lesson2/*.py, Registration/main.py): there is nothing to
compile from /root/reference, so no oracle/_ref is produced -- the Python reference is imported in
the build container only, by tests/golden/make_golden_*.py, to pin this restatement.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(HERE, "_build")
LIB = os.path.join(OUT_DIR, "libknn_oracle.so")
SRC = os.path.join(HERE, "knn_oracle.c")


def build(force: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    if (not force and os.path.exists(LIB)
            and os.path.getmtime(LIB) >= os.path.getmtime(SRC)):
        return LIB
    cmd = ["gcc", "-O2", "-ffp-contract=off", "-fopenmp", "-shared", "-fPIC", "-Wall",
           "-o", LIB, SRC, "-lm"]
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
