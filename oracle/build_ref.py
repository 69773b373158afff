"""TEST INFRASTRUCTURE — builds the UNMODIFIED reference rasterizer into oracle/_ref/.

Compiles the reference's own sources where they lie under /root/reference
(submodules/depth-diff-gaussian-rasterization: ext.cpp, rasterize_points.cu,
cuda_rasterizer/{forward,backward,rasterizer_impl}.cu) with nvcc/g++ directly (its setup.py /
CMake are not run) into the pybind module `oracle/_ref/ref_rasterizer/_C.so` for sm_100.
No reference source is copied; oracle/_ref/ is git-ignored but travels to the GPU box.

Used only by tests (primary GPU oracle) and by `bench.py --impl reference` (the reference arm).
"""
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("FDGS_REFERENCE", "/root/reference")
RAST = os.path.join(REF, "submodules", "depth-diff-gaussian-rasterization")
OUT_DIR = os.path.join(HERE, "_ref", "ref_rasterizer")
OUT = os.path.join(OUT_DIR, "_C.so")


def available():
    return os.path.exists(OUT)


def build(force=False, verbose=True):
    if available() and not force:
        return OUT
    if not os.path.isdir(RAST):
        raise FileNotFoundError(f"{RAST} not present (the reference only exists in the authoring container)")
    import torch
    from torch.utils import cpp_extension as ce
    os.makedirs(OUT_DIR, exist_ok=True)
    objdir = os.path.join(HERE, "_ref", "obj")
    os.makedirs(objdir, exist_ok=True)
    inc = []
    for p in ce.include_paths(device_type="cuda") + [sysconfig.get_paths()["include"]]:
        inc += ["-I", p]
    inc += ["-I", os.path.join(RAST, "third_party", "glm"), "-I", RAST]
    defs = ["-DTORCH_EXTENSION_NAME=_C", "-DTORCH_API_INCLUDE_EXTENSION_H",
            f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"]
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cu = ["cuda_rasterizer/forward.cu", "cuda_rasterizer/backward.cu", "cuda_rasterizer/rasterizer_impl.cu",
          "rasterize_points.cu"]
    procs, objs = [], []
    for s in cu:
        o = os.path.join(objdir, os.path.basename(s) + ".o")
        # same flags as the reference's setup.py (only -I glm; nvcc defaults otherwise), arch = sm_100
        cmd = [nvcc, "-std=c++17", "-O3", "-gencode", "arch=compute_100,code=sm_100", "-Xcompiler", "-fPIC",
               "-w", "--expt-relaxed-constexpr"] + defs + inc + ["-c", os.path.join(RAST, s), "-o", o]
        if verbose:
            print(" ".join(cmd), flush=True)
        procs.append((subprocess.Popen(cmd), cmd))
        objs.append(o)
    o = os.path.join(objdir, "ext.cpp.o")
    cmd = ["g++", "-std=c++17", "-O3", "-fPIC", "-w"] + defs + inc + ["-c", os.path.join(RAST, "ext.cpp"), "-o", o]
    if verbose:
        print(" ".join(cmd), flush=True)
    procs.append((subprocess.Popen(cmd), cmd))
    objs.append(o)
    for p, c in procs:
        if p.wait() != 0:
            raise RuntimeError("reference build failed: " + " ".join(c))
    libs = []
    for p in ce.library_paths(device_type="cuda"):
        libs += ["-L", p, "-Wl,-rpath," + p]
    cmd = ["g++", "-shared", "-o", OUT] + objs + libs + ["-lc10", "-ltorch_cpu", "-ltorch", "-ltorch_python",
                                                          "-lc10_cuda", "-ltorch_cuda", "-lcudart"]
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    with open(os.path.join(OUT_DIR, "__init__.py"), "w") as f:
        f.write('"""Built from the reference sources by oracle/build_ref.py (test infrastructure)."""\n'
                "import torch  # noqa: F401  (libtorch must be loaded first)\nfrom . import _C  # noqa: F401\n")
    return OUT


def load():
    """Imports the built reference module (needs a CUDA-capable torch to actually run it)."""
    if not available():
        raise ImportError("oracle/_ref not built: run `python oracle/build_ref.py` in the authoring container")
    p = os.path.join(HERE, "_ref")
    if p not in sys.path:
        sys.path.insert(0, p)
    import ref_rasterizer
    return ref_rasterizer._C


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
