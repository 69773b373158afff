"""Builds libfdgs.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

Usage: python build.py [--force]
The .so is git-ignored but travels to the GPU box with the repo snapshot.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libfdgs.so")
SOURCES = ["api.cu", "raster_fwd.cu", "raster_bwd.cu", "deform.cu", "deform_tc.cu", "deform_tc_bwd.cu", "glue.cu", "train_glue.cu", "init_kernels.cu"]
HEADERS = ["fdgs_common.cuh", "raster_math.cuh", "prof.cuh", "deform_common.cuh", "umma.cuh", os.path.join("..", "..", "include", "fdgs.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-DFDGS_ARCH=100",
] + os.environ.get("FDGS_EXTRA_NVCC", "").split()


def _newer(a, b):
    return os.path.getmtime(a) > os.path.getmtime(b)


def needs_build():
    if not os.path.exists(OUT):
        return True
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(_newer(d, OUT) for d in deps if os.path.exists(d))


def build(force=False, verbose=True):
    if not force and not needs_build():
        return OUT
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objs = []
    procs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    hdrs = [os.path.join(CSRC, h) for h in HEADERS] + [os.path.abspath(__file__)]
    stamp = os.path.join(HERE, "build", "flags.txt")          # objects are reused only under the same flags
    same_flags = os.path.exists(stamp) and open(stamp).read() == " ".join(NVCC_FLAGS)
    for s in SOURCES:
        o = os.path.join(HERE, "build", s.replace(".cu", ".o"))
        objs.append(o)
        deps = [os.path.join(CSRC, s)] + [h for h in hdrs if os.path.exists(h)]
        if not force and same_flags and os.path.exists(o) and not any(_newer(d, o) for d in deps):
            continue
        cmd = [nvcc] + NVCC_FLAGS + ["-c", os.path.join(CSRC, s), "-o", o]
        if verbose:
            print(" ".join(cmd), flush=True)
        procs.append((subprocess.Popen(cmd), cmd))
    for p, cmd in procs:
        if p.wait() != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
    with open(stamp, "w") as f:
        f.write(" ".join(NVCC_FLAGS))
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", OUT] + objs + ["-lcudart"]
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv)
    print(OUT)
