"""TEST INFRASTRUCTURE — ctypes wrapper of oracle/raster_oracle.cpp (CPU restatement of the reference
rasterizer).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "raster_oracle.cpp")
LIB = os.path.join(HERE, "_build", "libraster_oracle.so")


def build(force=False):
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-fPIC", "-shared", "-o", LIB, SRC])
    return LIB


class _Out(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("radii", "tiles_touched", "depths", "means2D", "cov3D", "conic_opacity", "rgb",
                                          "clamped", "keys", "point_list", "ranges")] + [("cap", C.c_int64)] + \
               [(n, C.c_void_p) for n in ("color", "depth", "final_T", "n_contrib", "dL_dpix", "dmean2D", "dconic",
                                          "dopacity", "dcolor", "dmeans3D", "dcov3D", "dsh", "dscale", "drot")]


_lib = None


def _load():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.fdgs_oracle_raster.restype = C.c_int64
        _lib.fdgs_oracle_raster.argtypes = [C.c_int] * 6 + [C.c_float] * 3 + [C.c_void_p] * 11 + [C.POINTER(_Out)]
    return _lib


def _f32(a):
    return None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def rasterize(*, W, H, tanfovx, tanfovy, bg, means3D, opacities, viewmatrix, projmatrix, campos, shs=None,
              colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, sh_degree=0, scale_modifier=1.0,
              dL_dpix=None, f64=False):
    """Forward (and, with dL_dpix, backward) of the reference rasterizer on the CPU.  numpy in / numpy out."""
    lib = _load()
    means3D = _f32(means3D)
    P = means3D.shape[0]
    shs, colors_precomp, scales, rotations, cov3D_precomp = map(_f32, (shs, colors_precomp, scales, rotations, cov3D_precomp))
    bg, opacities, viewmatrix, projmatrix, campos = map(_f32, (bg, opacities, viewmatrix, projmatrix, campos))
    M = shs.shape[1] if shs is not None else 0
    T = ((W + 15) // 16) * ((H + 15) // 16)
    o = dict(radii=np.zeros(P, np.int32), tiles_touched=np.zeros(P, np.uint32), depths=np.zeros(P, np.float32),
             means2D=np.zeros((P, 2), np.float32), cov3D=np.zeros((P, 6), np.float32),
             conic_opacity=np.zeros((P, 4), np.float32), rgb=np.zeros((P, 3), np.float32),
             clamped=np.zeros((P, 3), np.uint8), ranges=np.zeros((T, 2), np.uint32),
             color=np.zeros((3, H, W), np.float64), depth=np.zeros((1, H, W), np.float64),
             final_T=np.zeros((H, W), np.float64), n_contrib=np.zeros((H, W), np.uint32))
    if dL_dpix is not None:
        dL_dpix = _f32(dL_dpix)
        o.update(dmean2D=np.zeros((P, 3)), dconic=np.zeros((P, 4)), dopacity=np.zeros((P, 1)), dcolor=np.zeros((P, 3)),
                 dmeans3D=np.zeros((P, 3)), dcov3D=np.zeros((P, 6)), dsh=np.zeros((P, max(M, 1), 3)),
                 dscale=np.zeros((P, 3)), drot=np.zeros((P, 4)))
    cap = max(1024, 16 * P)
    while True:
        o["keys"] = np.zeros(cap, np.uint64)
        o["point_list"] = np.zeros(cap, np.uint32)
        out = _Out()
        for k, v in o.items():
            setattr(out, k, _p(v))
        out.cap = cap
        out.dL_dpix = _p(dL_dpix)
        R = lib.fdgs_oracle_raster(int(f64), P, int(sh_degree), M, W, H, float(tanfovx), float(tanfovy),
                                   float(scale_modifier), _p(bg), _p(means3D), _p(shs), _p(colors_precomp), _p(opacities),
                                   _p(scales), _p(rotations), _p(cov3D_precomp), _p(viewmatrix), _p(projmatrix), _p(campos),
                                   C.byref(out))
        if R >= 0:
            break
        cap = -R
        for k in ("dmean2D", "dconic", "dopacity", "dcolor"):
            if k in o:
                o[k][...] = 0
    o["keys"] = o["keys"][:R]
    o["point_list"] = o["point_list"][:R]
    o["num_rendered"] = int(R)
    if dL_dpix is not None and M == 0:
        o["dsh"] = o["dsh"][:, :0]
    return o
