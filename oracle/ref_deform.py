"""TEST INFRASTRUCTURE — loads the REAL reference deformation module from /root/reference (authoring
container only; the GPU box has no /root/reference).  Recipe: SURVEY.md Appendix C — load
scene/{hexplane,grid,deformation}.py by file path with stub `tkinter` and `scene` packages so that
scene/__init__.py (plyfile/open3d/simple_knn) is never executed."""
import importlib.util
import os
import sys
import types

REF = os.environ.get("FDGS_REFERENCE", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REF, "scene"))


def load():
    if "scene.deformation" in sys.modules and getattr(sys.modules["scene.deformation"], "_fdgs_ref", False):
        return sys.modules["scene.deformation"]
    if not available():
        raise ImportError(f"{REF} not present")
    if REF not in sys.path:
        sys.path.insert(0, REF)                                   # utils.graphics_utils
    tk = types.ModuleType("tkinter"); tk.W = "w"; sys.modules.setdefault("tkinter", tk)   # deformation.py:5
    pkg = types.ModuleType("scene"); pkg.__path__ = [os.path.join(REF, "scene")]; sys.modules["scene"] = pkg
    for name in ("hexplane", "grid", "deformation"):
        spec = importlib.util.spec_from_file_location(f"scene.{name}", os.path.join(REF, "scene", f"{name}.py"))
        m = importlib.util.module_from_spec(spec)
        sys.modules[f"scene.{name}"] = m
        spec.loader.exec_module(m)
    sys.modules["scene.deformation"]._fdgs_ref = True
    return sys.modules["scene.deformation"]
