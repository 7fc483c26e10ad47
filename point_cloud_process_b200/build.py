"""In-tree build of libpcp_b200.so (CUDA, sm_100a) and the pybind11 module _pcp.

    python -m point_cloud_process_b200.build [--force]

Plain nvcc / g++ command lines (no torch, no JIT cache): the built .so files sit next to this
file, are git-ignored, and travel with the working tree.  A CMake + setup.py packaging in the
PCLKeypoints style (CMakeLists.txt / setup.py at the repo root) drives the same sources.
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_obj")
LIB = os.path.join(HERE, "libpcp_b200.so")
MOD = os.path.join(HERE, "_pcp" + sysconfig.get_config_var("EXT_SUFFIX"))

CU_SOURCES = ["pcp_capi.cu", "pcp_build.cu", "pcp_knn.cu", "pcp_knn_tile.cu", "pcp_brute.cu", "pcp_multi.cu", "pcp_pointnet.cu", "pcp_global.cu", "pcp_global_ransac.cu", "pcp_radius.cu", "pcp_radius_tile.cu", "pcp_icp.cu", "pcp_front.cu", "pcp_dbscan.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC,-fvisibility=hidden", "--expt-relaxed-constexpr",
              "-Xptxas", "-v", "-I", os.path.join(ROOT, "include"), "-I", CSRC]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd, log=None):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log is not None:
        with open(log, "w") as f:
            f.write(" ".join(cmd) + "\n" + r.stdout)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("build failed: " + " ".join(cmd))
    return r.stdout


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(ROOT, "include", "pcp.h"))
    objs = []
    procs = []
    for src in CU_SOURCES:
        s = os.path.join(CSRC, src)
        if not os.path.exists(s):
            continue
        o = os.path.join(OBJ, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _newer(o, [s] + headers):
            cmd = ["nvcc"] + NVCC_FLAGS + ["-c", s, "-o", o]
            log = open(o + ".log", "w")
            procs.append((cmd, subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT), o + ".log"))
    for cmd, p, log in procs:
        if p.wait() != 0:
            sys.stderr.write(open(log).read())
            raise RuntimeError("build failed: " + " ".join(cmd))
        if verbose:
            sys.stdout.write(open(log).read())
    if force or _newer(LIB, objs):
        _run(["nvcc", "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs +
             ["-Xlinker", "-soname,libpcp_b200.so", "-ldl"])
    bsrc = os.path.join(CSRC, "bindings.cpp")
    if force or _newer(MOD, [bsrc, LIB] + headers):
        import pybind11
        _run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-fvisibility=hidden",
              "-I", pybind11.get_include(), "-I", sysconfig.get_paths()["include"],
              "-I", os.path.join(ROOT, "include"), bsrc, "-o", MOD,
              "-L", HERE, "-lpcp_b200", "-Wl,-rpath,$ORIGIN"])
    return LIB, MOD


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
