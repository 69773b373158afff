"""Import shim: exposes the package in `4dgaussians-fast-train_b200/` (not a valid identifier)
under the importable name `fdgs`."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "4dgaussians-fast-train_b200")
_spec = importlib.util.spec_from_file_location(
    "fdgs", os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["fdgs"] = _mod
_spec.loader.exec_module(_mod)
