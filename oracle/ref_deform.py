"""TEST INFRASTRUCTURE — loads the REAL reference deformation module: from /root/reference in the authoring
container, else from the unmodified copy `oracle/install_ref_py.py` put under the git-ignored baseline/_ref/
(which travels to the GPU box; /root/reference does not exist there).  Recipe: SURVEY.md Appendix C — load
scene/{hexplane,grid,deformation}.py by file path with stub `tkinter` and `scene` packages so that
scene/__init__.py (plyfile/open3d/simple_knn) is never executed."""
import contextlib
import importlib.util
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
_ENV = os.environ.get("FDGS_REFERENCE", "/root/reference")
_INSTALLED = os.path.join(os.path.dirname(HERE), "baseline", "_ref")


def root():
    """Directory holding scene/ and utils/ of the reference, or None."""
    for r in (_ENV, _INSTALLED):
        if os.path.exists(os.path.join(r, "scene", "deformation.py")):
            return r
    return None


REF = root() or _ENV


def available():
    return root() is not None


def load():
    if "scene.deformation" in sys.modules and getattr(sys.modules["scene.deformation"], "_fdgs_ref", False):
        return sys.modules["scene.deformation"]
    ref = root()
    if ref is None:
        raise ImportError(f"neither {_ENV} nor {_INSTALLED} holds the reference's scene/deformation.py")
    if ref not in sys.path:
        sys.path.insert(0, ref)                                   # utils.graphics_utils
    tk = types.ModuleType("tkinter"); tk.W = "w"; sys.modules.setdefault("tkinter", tk)   # deformation.py:5
    pkg = types.ModuleType("scene"); pkg.__path__ = [os.path.join(ref, "scene")]; sys.modules["scene"] = pkg
    for name in ("hexplane", "grid", "deformation"):
        spec = importlib.util.spec_from_file_location(f"scene.{name}", os.path.join(ref, "scene", f"{name}.py"))
        m = importlib.util.module_from_spec(spec)
        sys.modules[f"scene.{name}"] = m
        spec.loader.exec_module(m)
    sys.modules["scene.deformation"]._fdgs_ref = True
    sys.modules["scene.deformation"]._fdgs_ref_root = ref
    return sys.modules["scene.deformation"]


def load_regulation():
    """scene/regulation.py (compute_plane_smoothness, compute_plane_tv) with a stub matplotlib (imported, never used
    by those two functions; not installed in this image)."""
    if "fdgs_ref_regulation" in sys.modules:
        return sys.modules["fdgs_ref_regulation"]
    ref = root()
    if ref is None or not os.path.exists(os.path.join(ref, "scene", "regulation.py")):
        raise ImportError("reference scene/regulation.py not available")
    if "matplotlib" not in sys.modules:
        try:
            import matplotlib.pyplot  # noqa: F401
        except ImportError:
            mpl = types.ModuleType("matplotlib"); plt = types.ModuleType("matplotlib.pyplot"); mpl.pyplot = plt
            sys.modules["matplotlib"] = mpl; sys.modules["matplotlib.pyplot"] = plt
    spec = importlib.util.spec_from_file_location("fdgs_ref_regulation", os.path.join(ref, "scene", "regulation.py"))
    m = importlib.util.module_from_spec(spec)
    sys.modules["fdgs_ref_regulation"] = m
    spec.loader.exec_module(m)
    return m


def make_network(args):
    """The reference's deform_network(args); its constructor / set_aabb print to stdout, which bench.py keeps for the
    ONE JSON line, so those prints go to stderr."""
    mod = load()
    with contextlib.redirect_stdout(sys.stderr):
        return mod.deform_network(args)


class quiet_stdout(contextlib.redirect_stdout):
    def __init__(self):
        super().__init__(sys.stderr)
