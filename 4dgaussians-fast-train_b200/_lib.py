"""ctypes binding of libfdgs.so (C ABI declared in include/fdgs.h).

There is no CPU fallback: if the CUDA library is missing this module raises at
import of the first symbol.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FDGS_LIB") or os.path.join(_HERE, "libfdgs.so")    # FDGS_LIB: an A/B build of the same library

FDGS_MAX_LEVELS = 4
FDGS_NUM_HEADS = 5
HEAD_OUT = (3, 3, 4, 1, 48)

c_float_p = C.c_void_p  # raw device pointers are passed as integers


class FdgsRasterForward(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("W", C.c_int32), ("H", C.c_int32),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("scale_modifier", C.c_float),
        ("prefiltered", C.c_int32), ("debug", C.c_int32),
        ("background", C.c_void_p), ("means3D", C.c_void_p), ("shs", C.c_void_p),
        ("colors_precomp", C.c_void_p), ("opacities", C.c_void_p), ("scales", C.c_void_p),
        ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p), ("viewmatrix", C.c_void_p),
        ("projmatrix", C.c_void_p), ("campos", C.c_void_p),
        ("out_color", C.c_void_p), ("out_depth", C.c_void_p), ("radii", C.c_void_p),
        ("geom_ws", C.c_void_p), ("geom_bytes", C.c_size_t),
        ("img_ws", C.c_void_p), ("img_bytes", C.c_size_t),
        ("bin_ws", C.c_void_p), ("bin_bytes", C.c_size_t),
        ("counts_host", C.c_void_p),
        ("keys_out", C.c_void_p), ("point_list_out", C.c_void_p),
        ("activate", C.c_int32),
    ]


class FdgsRasterBackward(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("W", C.c_int32), ("H", C.c_int32),
        ("R", C.c_int64),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("scale_modifier", C.c_float),
        ("debug", C.c_int32),
        ("background", C.c_void_p), ("means3D", C.c_void_p), ("shs", C.c_void_p),
        ("colors_precomp", C.c_void_p), ("scales", C.c_void_p), ("rotations", C.c_void_p),
        ("cov3D_precomp", C.c_void_p), ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p),
        ("campos", C.c_void_p), ("radii", C.c_void_p),
        ("geom_ws", C.c_void_p), ("img_ws", C.c_void_p), ("bin_ws", C.c_void_p),
        ("dL_dpix", C.c_void_p),
        ("dL_dmeans2D", C.c_void_p), ("dL_dconic", C.c_void_p), ("dL_dopacity", C.c_void_p),
        ("dL_dcolors", C.c_void_p), ("dL_dmeans3D", C.c_void_p), ("dL_dcov3D", C.c_void_p),
        ("dL_dsh", C.c_void_p), ("dL_dscales", C.c_void_p), ("dL_drots", C.c_void_p),
        ("activate", C.c_int32),
    ]


class FdgsDeformNet(C.Structure):
    _fields_ = [
        ("levels", C.c_int32), ("channels", C.c_int32), ("width", C.c_int32), ("depth", C.c_int32),
        ("head_mask", C.c_int32), ("apply_rotation", C.c_int32),
        ("res", (C.c_int32 * 4) * FDGS_MAX_LEVELS),
        ("aabb", C.c_float * 6),
        ("planes", (C.c_void_p * 6) * FDGS_MAX_LEVELS),
        ("trunk_w", C.c_void_p * 8), ("trunk_b", C.c_void_p * 8),
        ("head_w1", C.c_void_p * FDGS_NUM_HEADS), ("head_b1", C.c_void_p * FDGS_NUM_HEADS),
        ("head_w2", C.c_void_p * FDGS_NUM_HEADS), ("head_b2", C.c_void_p * FDGS_NUM_HEADS),
    ]


class FdgsDeformForward(C.Structure):
    _fields_ = [
        ("net", FdgsDeformNet),
        ("P", C.c_int32), ("sh_coeffs", C.c_int32), ("time", C.c_float), ("time_dev", C.c_void_p),
        ("precision", C.c_int32), ("planes_packed_valid", C.c_int32),
        ("xyz", C.c_void_p), ("scales", C.c_void_p), ("rotations", C.c_void_p),
        ("opacity", C.c_void_p), ("shs", C.c_void_p),
        ("out_xyz", C.c_void_p), ("out_scales", C.c_void_p), ("out_rotations", C.c_void_p),
        ("out_opacity", C.c_void_p), ("out_shs", C.c_void_p),
        ("packed_ws", C.c_void_p), ("packed_bytes", C.c_size_t),
        ("trow_ws", C.c_void_p), ("trow_bytes", C.c_size_t),
        ("wprep_ws", C.c_void_p), ("wprep_bytes", C.c_size_t),
        ("save_ws", C.c_void_p), ("save_bytes", C.c_size_t),
        ("debug", C.c_int32),
    ]


class FdgsDeformBackward(C.Structure):
    _fields_ = [
        ("net", FdgsDeformNet),
        ("P", C.c_int32), ("time", C.c_float), ("time_dev", C.c_void_p), ("precision", C.c_int32),
        ("xyz", C.c_void_p),
        ("g_xyz", C.c_void_p), ("g_scales", C.c_void_p), ("g_rotations", C.c_void_p),
        ("g_opacity", C.c_void_p), ("g_shs", C.c_void_p),
        ("d_xyz_grid", C.c_void_p),
        ("d_planes", (C.c_void_p * 6) * FDGS_MAX_LEVELS),
        ("d_trunk_w", C.c_void_p * 8), ("d_trunk_b", C.c_void_p * 8),
        ("d_head_w1", C.c_void_p * FDGS_NUM_HEADS), ("d_head_b1", C.c_void_p * FDGS_NUM_HEADS),
        ("d_head_w2", C.c_void_p * FDGS_NUM_HEADS), ("d_head_b2", C.c_void_p * FDGS_NUM_HEADS),
        ("packed_ws", C.c_void_p), ("trow_ws", C.c_void_p),
        ("ws", C.c_void_p), ("ws_bytes", C.c_size_t),
        ("scratch", C.c_void_p), ("scratch_bytes", C.c_size_t),
        ("saved_ws", C.c_void_p), ("saved_bytes", C.c_size_t),
        ("debug", C.c_int32),
    ]


class FdgsActivateForward(C.Structure):
    _fields_ = [("P", C.c_int32), ("scales_raw", C.c_void_p), ("rotations_raw", C.c_void_p), ("opacity_raw", C.c_void_p),
                ("scales", C.c_void_p), ("rotations", C.c_void_p), ("opacity", C.c_void_p)]


class FdgsActivateBackward(C.Structure):
    _fields_ = [("P", C.c_int32), ("rotations_raw", C.c_void_p), ("scales", C.c_void_p), ("rotations", C.c_void_p),
                ("opacity", C.c_void_p), ("g_scales", C.c_void_p), ("g_rotations", C.c_void_p), ("g_opacity", C.c_void_p),
                ("d_scales_raw", C.c_void_p), ("d_rotations_raw", C.c_void_p), ("d_opacity_raw", C.c_void_p)]


class FdgsPlaneReg(C.Structure):
    _fields_ = [("levels", C.c_int32), ("channels", C.c_int32), ("res", (C.c_int32 * 4) * FDGS_MAX_LEVELS),
                ("planes", (C.c_void_p * 6) * FDGS_MAX_LEVELS), ("d_planes", (C.c_void_p * 6) * FDGS_MAX_LEVELS),
                ("plane_tv_weight", C.c_float), ("time_smoothness_weight", C.c_float), ("l1_time_planes_weight", C.c_float),
                ("grad_scale", C.c_void_p), ("loss", C.c_void_p)]


class FdgsAdamTensor(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("n", C.c_int64), ("lr", C.c_double), ("step", C.c_int64)]


class FdgsAdamStep(C.Structure):
    _fields_ = [("ntensors", C.c_int32), ("zero_grad", C.c_int32), ("beta1", C.c_double), ("beta2", C.c_double),
                ("eps", C.c_double), ("t", C.POINTER(FdgsAdamTensor))]


# every symbol include/fdgs.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "fdgs_abi_version": (C.c_int, []),
    "fdgs_error_string": (C.c_char_p, [C.c_int]),
    "fdgs_compiled_arch": (C.c_int, []),
    "fdgs_set_sm_limit": (None, [C.c_int]),
    "fdgs_profile_enable": (None, [C.c_int]),
    "fdgs_profile_collect": (C.c_int, [C.c_char_p, C.c_size_t]),
    "fdgs_launch_count": (C.c_ulonglong, []),
    "fdgs_raster_geom_bytes": (C.c_size_t, [C.c_int]),
    "fdgs_raster_img_bytes": (C.c_size_t, [C.c_int, C.c_int]),
    "fdgs_raster_bin_bytes": (C.c_size_t, [C.c_int64]),
    "fdgs_raster_forward_plan": (C.c_int, [C.POINTER(FdgsRasterForward), C.c_void_p]),
    "fdgs_raster_forward_render": (C.c_int, [C.POINTER(FdgsRasterForward), C.c_int64, C.c_void_p]),
    "fdgs_raster_forward": (C.c_int, [C.POINTER(FdgsRasterForward), C.c_int64, C.c_void_p]),
    "fdgs_raster_copy_ranges": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "fdgs_raster_count_pairs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "fdgs_raster_backward": (C.c_int, [C.POINTER(FdgsRasterBackward), C.c_void_p]),
    "fdgs_mark_visible": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fdgs_deform_packed_bytes": (C.c_size_t, [C.POINTER(FdgsDeformNet)]),
    "fdgs_deform_trow_bytes": (C.c_size_t, [C.POINTER(FdgsDeformNet)]),
    "fdgs_deform_wprep_bytes": (C.c_size_t, [C.POINTER(FdgsDeformNet)]),
    "fdgs_deform_forward_save_bytes": (C.c_size_t, [C.POINTER(FdgsDeformNet), C.c_int]),
    "fdgs_deform_forward": (C.c_int, [C.POINTER(FdgsDeformForward), C.c_void_p]),
    "fdgs_deform_backward_ws_bytes": (C.c_size_t, [C.POINTER(FdgsDeformNet)]),
    "fdgs_deform_backward_scratch_bytes": (C.c_size_t, [C.POINTER(FdgsDeformNet), C.c_int]),
    "fdgs_deform_backward": (C.c_int, [C.POINTER(FdgsDeformBackward), C.c_void_p]),
    "fdgs_activate_forward": (C.c_int, [C.POINTER(FdgsActivateForward), C.c_void_p]),
    "fdgs_activate_backward": (C.c_int, [C.POINTER(FdgsActivateBackward), C.c_void_p]),
    "fdgs_plane_regulation": (C.c_int, [C.POINTER(FdgsPlaneReg), C.c_void_p]),
    "fdgs_plane_smoothness": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fdgs_adam_step": (C.c_int, [C.POINTER(FdgsAdamStep), C.c_void_p]),
    "fdgs_densify_stats": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fdgs_knn_ws_bytes": (C.c_size_t, [C.c_int]),
    "fdgs_knn_mean_dist2": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "fdgs_knn_static_max_points": (C.c_int, []),
    "fdgs_knn_mean_dist2_static": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "fdgs_voxel_downsample_ws_bytes": (C.c_size_t, [C.c_int]),
    "fdgs_voxel_downsample": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                        C.c_void_p, C.c_size_t, C.c_void_p]),
    "fdgs_l1_loss_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_longlong, C.c_float, C.c_void_p, C.c_void_p]),
    "fdgs_l1_loss_backward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_longlong, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]),
}

_lib = None


def load():
    """Loads libfdgs.so; raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: the fdgs CUDA library is not built. Run "
            f"`python {os.path.join(_HERE, 'build.py')}` (needs nvcc); there is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    if lib.fdgs_abi_version() != 1:
        raise ImportError("libfdgs.so ABI version mismatch")
    _lib = lib
    return lib


def check(code, what):
    if code != 0:
        msg = load().fdgs_error_string(code).decode()
        raise RuntimeError(f"{what} failed: {msg} (code {code})")


def stream_ptr(dev=None):
    """Raw handle (c_void_p) of torch's current stream on `dev` (a torch.device; None = the current device).
    `torch.cuda.current_stream().cuda_stream` builds a Stream object through three Python helper layers (~20 us per call,
    seven calls per view); this is one C call."""
    import torch
    try:
        idx = dev.index if (dev is not None and dev.index is not None) else torch._C._cuda_getDevice()
        return C.c_void_p(torch._C._cuda_getCurrentRawStream(idx))
    except AttributeError:                                   # a torch build without these private entry points
        return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


class _NullContext:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


_NULL_CONTEXT = _NullContext()


def on_device(dev):
    """`with on_device(dev):` = `with torch.cuda.device(dev):`, without the device switch (and its Python overhead) when
    `dev` is already the current device - the usual case."""
    import torch
    try:
        if dev.index is None or torch._C._cuda_getDevice() == dev.index:
            return _NULL_CONTEXT
    except AttributeError:
        pass
    return torch.cuda.device(dev)


def ptr(t):
    """Device pointer of a torch tensor, or None for an absent/empty tensor."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def profile_enable(on=True):
    load().fdgs_profile_enable(int(bool(on)))


def profile_collect():
    """{kernel_name: (launches, total_ms)} since the last collect (synchronises the device)."""
    buf = C.create_string_buffer(1 << 16)
    load().fdgs_profile_collect(buf, len(buf))
    out = {}
    for line in buf.value.decode().splitlines():
        name, n, ms = line.split()
        out[name] = (int(n), float(ms))
    return out


def launch_count():
    return int(load().fdgs_launch_count())
