"""Build recipe for the CPU oracle (test infrastructure, not product code).

    python oracle/build_oracle.py

compiles oracle/knn_oracle.c into oracle/_build/libknn_oracle.so with plain gcc.  The kNN / radius / ICP
reference is pure Python (Kdtree_Octree/lesson2/*.py, Registration/main.py): nothing to compile there -- the
Python reference is imported in the build container only, by tests/golden/make_golden_*.py, to pin the
restatement.

oracle/_ref (git-ignored, travels to the GPU box): when /root/reference is present, the reference's OWN CUDA
sources of the two PointNet++ search ops are compiled where they lie --
    _ext-src/src/ball_query_gpu.cu, _ext-src/src/interpolate_gpu.cu
with oracle/ref_pointnet2/ first on the include path (a torch-free stand-in for the one helper header they include)
plus oracle/ref_pointnet2/export.cu (C entry points over their wrappers) -> oracle/_ref/libpointnet2_ref.so.
No reference source is copied into the repository.
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


REF_SRC = "/root/reference/Classification_ModelNet40/pointnet2_ops_lib/pointnet2_ops/_ext-src/src"
REF_DIR = os.path.join(HERE, "_ref")
REF_LIB = os.path.join(REF_DIR, "libpointnet2_ref.so")


def build_ref(force: bool = False):
    """-> path of oracle/_ref/libpointnet2_ref.so, or None when neither the reference tree nor a prebuilt copy exists."""
    srcs = [os.path.join(REF_SRC, "ball_query_gpu.cu"), os.path.join(REF_SRC, "interpolate_gpu.cu")]
    if not all(os.path.exists(s) for s in srcs):
        return REF_LIB if os.path.exists(REF_LIB) else None          # the GPU box: the prebuilt file travels
    stub = os.path.join(HERE, "ref_pointnet2")
    deps = srcs + [os.path.join(stub, "export.cu"), os.path.join(stub, "cuda_utils.h")]
    if not force and os.path.exists(REF_LIB) and all(os.path.getmtime(REF_LIB) >= os.path.getmtime(d) for d in deps):
        return REF_LIB
    os.makedirs(REF_DIR, exist_ok=True)
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O2", "-shared", "-Xcompiler", "-fPIC",
           "-I", stub, "-o", REF_LIB] + srcs + [os.path.join(stub, "export.cu")]
    subprocess.check_call(cmd)
    return REF_LIB


if __name__ == "__main__":
    print(build_ref(force="--force" in sys.argv))
    print(build(force="--force" in sys.argv))
