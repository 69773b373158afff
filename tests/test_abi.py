"""CPU-side checks of the drop-in boundary: libfdgs.so loads, exports every symbol include/fdgs.h declares,
the ctypes mirrors have the C layout, and the host-only entry points / argument checks behave.
No kernel is launched here."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

import fdgs  # noqa: F401
from fdgs import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "fdgs.h")


def header_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fdgs_[a-z0-9_]+)\s*\(", src)))


def test_library_is_built_and_loads():
    assert os.path.exists(L.LIB_PATH), "run `python __graft_entry__.py` (build()) first"
    lib = L.load()
    assert lib.fdgs_abi_version() == 1
    assert lib.fdgs_compiled_arch() == 100


def test_every_header_symbol_is_exported_and_bound():
    syms = header_symbols()
    assert len(syms) >= 19
    raw = C.CDLL(L.LIB_PATH)
    for s in syms:
        assert hasattr(raw, s), f"{s} declared in include/fdgs.h but not exported by libfdgs.so"
        assert s in L.SYMBOLS, f"{s} has no ctypes binding in _lib.py"
    assert sorted(L.SYMBOLS) == syms


def test_exports_are_plain_c_no_torch_dependency():
    out = subprocess.run(["ldd", L.LIB_PATH], capture_output=True, text=True).stdout
    names = [ln.split()[0].lower() for ln in out.splitlines() if ln.strip()]   # library basenames, not install paths
    assert not [n for n in names if "torch" in n or "c10" in n or "python" in n], names
    nm = subprocess.run(["nm", "-D", "--defined-only", L.LIB_PATH], capture_output=True, text=True).stdout
    for s in header_symbols():
        assert re.search(rf"\sT {s}$", nm, flags=re.M), f"{s} is not an unmangled text symbol"


def test_fatbin_targets_sm_100a():
    out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-lelf", L.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out, out


STRUCTS = ["FdgsRasterForward", "FdgsRasterBackward", "FdgsDeformNet", "FdgsDeformForward", "FdgsDeformBackward",
           "FdgsActivateForward", "FdgsActivateBackward", "FdgsPlaneReg", "FdgsAdamTensor", "FdgsAdamStep"]


def test_ctypes_struct_layout_matches_header(tmp_path):
    """Compiles a C program that prints sizeof/offsetof from include/fdgs.h and compares with ctypes."""
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', "int main(void){"]
    for s in STRUCTS:
        cls = getattr(L, s)
        lines.append(f'printf("{s} %zu\\n", sizeof({s}));')
        for name, _ in cls._fields_:
            lines.append(f'printf("{s}.{name} %zu\\n", offsetof({s}, {name}));')
    lines.append("return 0;}")
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-std=c11", "-o", str(exe), str(src)])
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout
    got = dict(l.split() for l in out.splitlines())
    for s in STRUCTS:
        cls = getattr(L, s)
        assert int(got[s]) == C.sizeof(cls), s
        for name, _ in cls._fields_:
            assert int(got[f"{s}.{name}"]) == getattr(cls, name).offset, f"{s}.{name}"


def test_header_compiles_as_c_and_cxx(tmp_path):
    for comp, std in (("gcc", "-std=c99"), ("g++", "-std=c++11")):
        src = tmp_path / ("t.c" if comp == "gcc" else "t.cpp")
        src.write_text(f'#include "{HEADER}"\nint main(void){{ return FDGS_ABI_VERSION - 1; }}\n')
        subprocess.check_call([comp, std, "-Wall", "-Werror", "-pedantic", "-fsyntax-only", str(src)])


def test_error_strings():
    lib = L.load()
    assert lib.fdgs_error_string(0) == b"ok"
    for code in (-1, -2, -3, -4):
        assert lib.fdgs_error_string(code).startswith(b"fdgs:")
    assert b"unknown" in lib.fdgs_error_string(-99)


def test_workspace_size_functions():
    lib = L.load()
    g1, g2 = lib.fdgs_raster_geom_bytes(1000), lib.fdgs_raster_geom_bytes(300_000)
    assert g1 % 256 == 0 and g2 % 256 == 0 and 69 * 1000 <= g1 and 69 * 300_000 <= g2 <= 75 * 300_000
    im = lib.fdgs_raster_img_bytes(1352, 1014)
    T = 85 * 64
    assert im % 256 == 0 and im >= 8 * 1352 * 1014 + 20 * T
    b = lib.fdgs_raster_bin_bytes(1_931_440)
    assert b % 256 == 0 and 56 * 1_931_440 <= b <= 56 * 1_931_440 + 1024
    assert lib.fdgs_raster_geom_bytes(0) > 0 and lib.fdgs_raster_bin_bytes(0) > 0


def _dynerf_net():
    net = L.FdgsDeformNet()
    net.levels, net.channels, net.width, net.depth, net.head_mask = 2, 16, 128, 1, 31
    for l, m in enumerate((1, 2)):
        for i, r in enumerate((64 * m, 64 * m, 64 * m, 150)):
            net.res[l][i] = r
    return net


def test_deform_workspace_sizes_match_survey_plane_count():
    lib = L.load()
    net = _dynerf_net()
    assert lib.fdgs_deform_packed_bytes(C.byref(net)) == 2_365_440 * 4          # DyNeRF plane floats
    trow = lib.fdgs_deform_trow_bytes(C.byref(net))
    assert trow == 3 * 16 * (64 + 128) * 4
    assert lib.fdgs_deform_backward_ws_bytes(C.byref(net)) == 2_365_440 * 4 + trow
    bad = _dynerf_net(); bad.width = 100
    assert lib.fdgs_deform_packed_bytes(C.byref(bad)) == 0                         # unsupported -> 0, callers raise


def test_argument_errors_return_codes_without_touching_the_gpu():
    lib = L.load()
    assert lib.fdgs_raster_forward_plan(None, None) == -1
    assert lib.fdgs_raster_forward_render(None, 0, None) == -1
    assert lib.fdgs_raster_backward(None, None) == -1
    assert lib.fdgs_deform_forward(None, None) == -1
    assert lib.fdgs_deform_backward(None, None) == -1
    assert lib.fdgs_mark_visible(-1, None, None, None, None, None) == -1
    assert lib.fdgs_raster_copy_ranges(None, 16, 16, None, None) == -1
    a = L.FdgsRasterForward()
    a.P, a.W, a.H = 5, 0, 16
    assert lib.fdgs_raster_forward_plan(C.byref(a), None) == -1                    # W <= 0
    a.W = 16                                                                        # null pointers with P > 0
    assert lib.fdgs_raster_forward_plan(C.byref(a), None) == -1
    d = L.FdgsDeformForward()
    d.net = _dynerf_net(); d.net.width = 100
    assert lib.fdgs_deform_forward(C.byref(d), None) == -3                         # FDGS_E_UNSUPPORTED
    with pytest.raises(RuntimeError, match="bad argument"):
        L.check(-1, "x")


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    """No CPU fallback: without libfdgs.so the binding raises ImportError."""
    code = ("import sys, importlib.util; sys.path.insert(0, %r); import fdgs; from fdgs import _lib as L; "
            "L.LIB_PATH = %r; L._lib = None\n"
            "try:\n    L.load()\nexcept ImportError as e:\n    print('RAISED', 'no CPU fallback' in str(e))\n") % (ROOT, str(tmp_path / "nope.so"))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert "RAISED True" in out.stdout, out.stdout + out.stderr
