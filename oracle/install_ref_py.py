"""TEST INFRASTRUCTURE / REFERENCE ARM — installs the UNMODIFIED reference Python files of the hot path into the
git-ignored `baseline/_ref/` so that they travel to the GPU box with the repo snapshot (like oracle/_ref, the reference
rasterizer built from its own sources).  Nothing under baseline/_ref is tracked, imported by the product, or edited:

    scene/hexplane.py  scene/grid.py  scene/deformation.py     the HexPlane deformation field  
    utils/graphics_utils.py                                    imported by scene/deformation.py:12
    scene/regulation.py                                        compute_plane_smoothness        

`oracle/ref_deform.py` loads them by file path (SURVEY App. C recipe) from /root/reference when that exists (authoring
container) and from baseline/_ref otherwise (GPU box).  Used by `bench.py --impl reference` and by the `-m gpu` tests
that compare with the real module on the same device; oracle/ref_port.py (the op-for-op port) stays as the fallback.
Run by `__graft_entry__.build()` in the authoring container; `python oracle/install_ref_py.py` re-runs it.
"""
import filecmp
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("FDGS_REFERENCE", "/root/reference")
DST = os.path.join(ROOT, "baseline", "_ref")
FILES = ("scene/hexplane.py", "scene/grid.py", "scene/deformation.py", "utils/graphics_utils.py", "scene/regulation.py")


def available():
    return all(os.path.exists(os.path.join(DST, f)) for f in FILES)


def install(verbose=True):
    if not os.path.isdir(os.path.join(REF, "scene")):
        raise FileNotFoundError(f"{REF} not present (the reference only exists in the authoring container)")
    for f in FILES:
        src, dst = os.path.join(REF, f), os.path.join(DST, f)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not (os.path.exists(dst) and filecmp.cmp(src, dst, shallow=False)):
            shutil.copyfile(src, dst)
            if verbose:
                print("installed", dst)
    return DST


if __name__ == "__main__":
    install()
    print(DST, "ok" if available() else "incomplete", file=sys.stderr)
