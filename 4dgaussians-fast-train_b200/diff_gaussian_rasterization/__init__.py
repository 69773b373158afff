"""Drop-in replacement for the reference's `diff_gaussian_rasterization` package.

Same public surface, argument order, keyword names, return arity, dtypes and shapes as
submodules/depth-diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py
(:21-42 rasterize_gaussians, :44-155 _RasterizeGaussians, :157-169 settings tuple,
:171-220 GaussianRasterizer) — but the native side is libfdgs.so (hand-written sm_100a
kernels behind the C ABI in include/fdgs.h) instead of the pybind `_C` module.

Usable either as `<package>.diff_gaussian_rasterization` or, with the package directory on
sys.path, as the top-level `diff_gaussian_rasterization` the reference code imports.
"""
from typing import NamedTuple
import ctypes as C
import importlib.util
import os
import sys

import torch
import torch.nn as nn

try:  # imported as a sub-package
    from .. import _lib as L
except ImportError:  # imported top-level (drop-in mode)
    _p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_lib.py")
    _spec = importlib.util.spec_from_file_location("fdgs_lib_standalone", _p)
    L = sys.modules.get("fdgs_lib_standalone")
    if L is None:
        L = importlib.util.module_from_spec(_spec)
        sys.modules["fdgs_lib_standalone"] = L
        _spec.loader.exec_module(L)

TILE = 16
_ALIGN = 256


def _au(x):
    return (x + _ALIGN - 1) & ~(_ALIGN - 1)


def cpu_deep_copy_tuple(input_tuple):
    copied_tensors = [item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple]
    return tuple(copied_tensors)


_last_R = 0
_keep_last = False       # diagnostics only: keep a reference to the last forward's workspaces (see count_pairs)
_last_ws = None

# ---- one-shot ("speculative") forward: no blocking read of num_rendered --------------------------------------------------
# The reference stalls the host once per render to read num_rendered (rasterizer_impl.cu:282) because its binning buffer is
# sized from it; the default path here does the same (fdgs_raster_forward_plan / _render).  With speculative mode ON the
# instance workspace is sized from the largest count recently seen for this (device, P, W, H) x 1.3 and every kernel of the
# forward is queued at once (fdgs_raster_forward); the real count comes back asynchronously and is examined by
# check_pending().  If it did not fit, the device rendered nothing for that call (background image, zero gradients) and
# check_pending() returns False: the caller must discard that step and redo it with speculative mode off.
# fdgs.dp.train_iteration does exactly that.  Off by default: a bare GaussianRasterizer call behaves like the reference's.
_spec_on = False
_caps = {}               # (device index, P, W, H) -> largest recent num_rendered
_pending = []            # [(event, pinned counts, capacity, key)] of one-shot calls not yet examined
_slots = {}              # device index -> free pinned int32[4] tensors
_overflow_seen = False
SPEC_HEADROOM, SPEC_SLACK = 1.3, 65536


class speculative_mode:
    """Context manager: `with speculative_mode(True): ...` renders without the per-call host sync (see above)."""

    def __init__(self, on=True):
        self.on = bool(on)

    def __enter__(self):
        global _spec_on
        self.prev, _spec_on = _spec_on, self.on
        return self

    def __exit__(self, *exc):
        global _spec_on
        _spec_on = self.prev
        return False


def _note_count(key, R):
    _caps[key] = max(int(_caps.get(key, 0) * 0.999), int(R))


def check_pending(block=True):
    """Examines the counts of the one-shot calls issued so far (all of them with block=True - this waits for the device to
    pass their scan kernels; only the completed ones otherwise).  Returns False if any of them overflowed its workspace
    since the last call of this function, i.e. produced an empty render."""
    global _overflow_seen, _last_R
    keep = []
    for item in _pending:
        ev, counts, cap, key = item
        if block:
            ev.synchronize()
        elif not ev.query():
            keep.append(item)
            continue
        R = int(counts[0])
        _last_R = R
        _note_count(key, R)
        if int(counts[3]) or R > cap:
            _overflow_seen = True
        _slots.setdefault(key[0], []).append(counts)
    _pending[:] = keep
    ok = not _overflow_seen
    _overflow_seen = False
    return ok


def _take_slot(device):
    free = _slots.setdefault(device.index, [])
    if free:
        return free.pop()
    return torch.zeros(4, dtype=torch.int32).pin_memory()


def last_num_rendered():
    """num_rendered (R) of the most recent forward call in this process (bench / diagnostics)."""
    return _last_R


def _f32(t, name, device):
    """contiguous FP32 tensor on `device`, or None when the argument is an empty placeholder."""
    if t is None or t.numel() == 0:
        return None
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32, got {t.dtype}")
    if t.device != device:
        raise RuntimeError(f"{name} must be on {device}, got {t.device}")
    return t.contiguous()


def _stream():
    return L.stream_ptr()


class _Workspaces:
    """The three caller-owned workspaces of one forward call (kept alive for backward)."""
    __slots__ = ("geom", "img", "bin")


def _forward_native(bg, means3D, colors_precomp, opacities, scales, rotations, scale_modifier,
                    cov3Ds_precomp, viewmatrix, projmatrix, tanfovx, tanfovy, image_height,
                    image_width, sh, degree, campos, prefiltered, debug, want_debug_lists=False, activate=False):
    """Mirror of RasterizeGaussiansCUDA (rasterize_points.cu:35-117).
    Returns (num_rendered, color, depth, radii, geomBuffer, binningBuffer, imgBuffer[, extras])."""
    lib = L.load()
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise RuntimeError("fdgs rasterizer needs CUDA tensors (there is no CPU fallback)")
    dev = means3D.device
    P = means3D.size(0)
    H, W = int(image_height), int(image_width)
    with L.on_device(dev):
        color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
        depth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
        radii = torch.empty((P,), dtype=torch.int32, device=dev)
        geom = torch.empty((lib.fdgs_raster_geom_bytes(P),), dtype=torch.uint8, device=dev)
        img = torch.empty((lib.fdgs_raster_img_bytes(W, H),), dtype=torch.uint8, device=dev)
        if P == 0:
            color.zero_(); depth.zero_()
            binb = torch.empty((0,), dtype=torch.uint8, device=dev)
            out = (0, color, depth, radii, geom, binb, img)
            return out + ({},) if want_debug_lists else out
        sh_c = _f32(sh, "sh", dev)
        M = sh.size(1) if (sh is not None and sh.numel() != 0 and sh.dim() >= 2 and sh.size(0) != 0) else 0
        a = L.FdgsRasterForward()
        a.P, a.D, a.M, a.W, a.H = P, int(degree), int(M), W, H
        a.tanfovx, a.tanfovy, a.scale_modifier = float(tanfovx), float(tanfovy), float(scale_modifier)
        a.prefiltered, a.debug = int(bool(prefiltered)), int(bool(debug))
        a.activate = int(bool(activate))
        keep = [
            _f32(bg, "bg", dev), _f32(means3D, "means3D", dev), sh_c, _f32(colors_precomp, "colors_precomp", dev),
            _f32(opacities, "opacities", dev), _f32(scales, "scales", dev), _f32(rotations, "rotations", dev),
            _f32(cov3Ds_precomp, "cov3Ds_precomp", dev), _f32(viewmatrix, "viewmatrix", dev),
            _f32(projmatrix, "projmatrix", dev), _f32(campos, "campos", dev),
        ]
        (a.background, a.means3D, a.shs, a.colors_precomp, a.opacities, a.scales, a.rotations, a.cov3D_precomp,
         a.viewmatrix, a.projmatrix, a.campos) = [L.ptr(t) for t in keep]
        if keep[8] is None or keep[8].numel() != 16 or keep[9] is None or keep[9].numel() != 16:
            raise RuntimeError("viewmatrix / projmatrix must hold 16 floats")
        a.out_color, a.out_depth, a.radii = color.data_ptr(), depth.data_ptr(), radii.data_ptr()
        a.geom_ws, a.geom_bytes = geom.data_ptr(), geom.numel()
        a.img_ws, a.img_bytes = img.data_ptr(), img.numel()
        st = _stream()
        key = (dev.index, int(P), W, H)
        global _last_R, _last_ws
        if _spec_on and key in _caps and not want_debug_lists and not debug:
            # one-shot path: workspace sized from the recent maximum, counts read back asynchronously (check_pending)
            cap = int(_caps[key] * SPEC_HEADROOM) + SPEC_SLACK
            binb = torch.empty((lib.fdgs_raster_bin_bytes(cap),), dtype=torch.uint8, device=dev)
            a.bin_ws, a.bin_bytes = binb.data_ptr(), binb.numel()
            counts = _take_slot(dev)
            a.counts_host = counts.data_ptr()
            L.check(lib.fdgs_raster_forward(C.byref(a), cap, st), "fdgs_raster_forward")
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            _pending.append((ev, counts, cap, key))
            if len(_pending) > 256:
                check_pending(block=False)
            if _keep_last:
                _last_ws = (img, binb, cap, W, H)
            return (cap, color, depth, radii, geom, binb, img)
        # exact path: a pinned slot of this call's own (two threads rendering on one device must not share the counts)
        counts = _take_slot(dev)
        try:
            a.counts_host = counts.data_ptr()
            L.check(lib.fdgs_raster_forward_plan(C.byref(a), st), "fdgs_raster_forward_plan")
            R = int(counts[0])
            _last_R = R
            _note_count(key, R)
            binb = torch.empty((lib.fdgs_raster_bin_bytes(R),), dtype=torch.uint8, device=dev)
            a.bin_ws, a.bin_bytes = binb.data_ptr(), binb.numel()
            extras = {}
            if want_debug_lists:
                extras["keys"] = torch.empty((R,), dtype=torch.int64, device=dev)
                extras["point_list"] = torch.empty((R,), dtype=torch.int32, device=dev)
                a.keys_out = extras["keys"].data_ptr() if R else None
                a.point_list_out = extras["point_list"].data_ptr() if R else None
            L.check(lib.fdgs_raster_forward_render(C.byref(a), R, st), "fdgs_raster_forward_render")
        finally:
            _slots.setdefault(dev.index, []).append(counts)   # both calls have read it on the host by now
        if _keep_last:
            _last_ws = (img, binb, R, W, H)
        out = (R, color, depth, radii, geom, binb, img)
        if want_debug_lists:
            T = ((W + TILE - 1) // TILE) * ((H + TILE - 1) // TILE)
            ranges = torch.empty((T, 2), dtype=torch.int32, device=dev)
            L.check(lib.fdgs_raster_copy_ranges(img.data_ptr(), W, H, ranges.data_ptr(), st), "fdgs_raster_copy_ranges")
            extras["ranges"] = ranges
            extras.update(_geom_views(geom, P))
            extras.update(_img_views(img, W, H))
            return out + (extras,)
        return out


def count_pairs(render_fn, cam, pc, bg):
    """Diagnostic for the blend kernels' roofline (bench.py): renders one view forward-only and counts, on the device,
    the pixel x Gaussian pairs a per-pixel walk tests (sum of n_contrib) and those that contribute (alpha >= 1/255)."""
    global _keep_last, _last_ws
    lib = L.load()
    _keep_last = True
    try:
        with torch.no_grad():
            render_fn(cam, pc, bg_color=bg, stage="fine")
        img, binb, R, W, H = _last_ws
    finally:
        _keep_last = False
        _last_ws = None
    counts = torch.zeros(2, dtype=torch.int64, device=img.device)
    with L.on_device(img.device):
        L.check(lib.fdgs_raster_count_pairs(img.data_ptr(), binb.data_ptr() if binb.numel() else None, int(R), W, H,
                                            counts.data_ptr(), _stream()), "fdgs_raster_count_pairs")
    c = counts.cpu().tolist()
    return {"num_rendered": int(R), "tested": int(c[0]), "contributing": int(c[1])}


def _geom_views(geom, P):
    """Typed views into the geometry workspace (layout: csrc/fdgs_common.cuh FdgsGeom)."""
    p = max(P, 1)
    off = 0
    out = {}
    for name, dtype, per, shape in (("depths", torch.float32, 4, (P,)), ("means2D", torch.float32, 8, (P, 2)),
                                    ("cov3D", torch.float32, 24, (P, 6)), ("conic_opacity", torch.float32, 16, (P, 4)),
                                    ("rgb", torch.float32, 12, (P, 3)), ("clamped", torch.uint8, 1, (P,)),
                                    ("tiles_touched", torch.int32, 4, (P,))):
        out[name] = geom[off:off + P * per].view(dtype).view(shape)
        off += _au(p * per)
    return out


def _img_views(img, W, H):
    hw = max(W * H, 1)
    return {"final_T": img[0:W * H * 4].view(torch.float32).view(H, W),
            "n_contrib": img[_au(hw * 4):_au(hw * 4) + W * H * 4].view(torch.int32).view(H, W)}


def _backward_native(bg, means3D, radii, colors_precomp, scales, rotations, scale_modifier, cov3Ds_precomp,
                     viewmatrix, projmatrix, tanfovx, tanfovy, grad_out_color, sh, degree, campos,
                     geomBuffer, num_rendered, binningBuffer, imgBuffer, debug, activate=False):
    """Mirror of RasterizeGaussiansBackwardCUDA (rasterize_points.cu:119-198)."""
    lib = L.load()
    dev = means3D.device
    P = means3D.size(0)
    H, W = grad_out_color.size(1), grad_out_color.size(2)
    M = sh.size(1) if (sh is not None and sh.numel() != 0 and sh.dim() >= 2 and sh.size(0) != 0) else 0
    with L.on_device(dev):
        f = dict(dtype=torch.float32, device=dev)
        dL_dmeans3D = torch.empty((P, 3), **f)
        dL_dmeans2D = torch.empty((P, 3), **f)
        dL_dcolors = torch.empty((P, 3), **f)
        dL_dconic = torch.empty((P, 2, 2), **f)
        dL_dopacity = torch.empty((P, 1), **f)
        dL_dcov3D = torch.empty((P, 6), **f)
        dL_dsh = torch.empty((P, M, 3), **f)
        dL_dscales = torch.empty((P, 3), **f)
        dL_drotations = torch.empty((P, 4), **f)
        if P != 0:
            a = L.FdgsRasterBackward()
            a.P, a.D, a.M, a.W, a.H, a.R = P, int(degree), int(M), W, H, int(num_rendered)
            a.tanfovx, a.tanfovy, a.scale_modifier = float(tanfovx), float(tanfovy), float(scale_modifier)
            a.debug = int(bool(debug))
            a.activate = int(bool(activate))
            keep = [
                _f32(bg, "bg", dev), _f32(means3D, "means3D", dev), _f32(sh, "sh", dev),
                _f32(colors_precomp, "colors_precomp", dev), _f32(scales, "scales", dev),
                _f32(rotations, "rotations", dev), _f32(cov3Ds_precomp, "cov3Ds_precomp", dev),
                _f32(viewmatrix, "viewmatrix", dev), _f32(projmatrix, "projmatrix", dev), _f32(campos, "campos", dev),
                _f32(grad_out_color, "grad_out_color", dev),
            ]
            (a.background, a.means3D, a.shs, a.colors_precomp, a.scales, a.rotations, a.cov3D_precomp,
             a.viewmatrix, a.projmatrix, a.campos, a.dL_dpix) = [L.ptr(t) for t in keep]
            a.radii = radii.data_ptr()
            a.geom_ws, a.img_ws = geomBuffer.data_ptr(), imgBuffer.data_ptr()
            a.bin_ws = binningBuffer.data_ptr() if binningBuffer.numel() else None
            a.dL_dmeans2D, a.dL_dconic, a.dL_dopacity = dL_dmeans2D.data_ptr(), dL_dconic.data_ptr(), dL_dopacity.data_ptr()
            a.dL_dcolors, a.dL_dmeans3D, a.dL_dcov3D = dL_dcolors.data_ptr(), dL_dmeans3D.data_ptr(), dL_dcov3D.data_ptr()
            a.dL_dsh = dL_dsh.data_ptr() if M else None
            a.dL_dscales, a.dL_drots = dL_dscales.data_ptr(), dL_drotations.data_ptr()
            L.check(lib.fdgs_raster_backward(C.byref(a), _stream()), "fdgs_raster_backward")
    return dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations


def _mark_visible(positions, viewmatrix, projmatrix):
    """Mirror of markVisible (rasterize_points.cu:200-219)."""
    lib = L.load()
    P = positions.size(0)
    present = torch.zeros((P,), dtype=torch.bool, device=positions.device)
    if P != 0:
        with L.on_device(positions.device):
            pos = _f32(positions, "positions", positions.device)
            vm = _f32(viewmatrix, "viewmatrix", positions.device)
            pm = _f32(projmatrix, "projmatrix", positions.device)
            L.check(lib.fdgs_mark_visible(P, pos.data_ptr(), vm.data_ptr(), pm.data_ptr(), present.data_ptr(), _stream()),
                    "fdgs_mark_visible")
    return present


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        args = (
            raster_settings.bg, means3D, colors_precomp, opacities, scales, rotations,
            raster_settings.scale_modifier, cov3Ds_precomp, raster_settings.viewmatrix,
            raster_settings.projmatrix, raster_settings.tanfovx, raster_settings.tanfovy,
            raster_settings.image_height, raster_settings.image_width, sh, raster_settings.sh_degree,
            raster_settings.campos, raster_settings.prefiltered, raster_settings.debug,
        )
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)  # copy them before they can be corrupted
            try:
                num_rendered, color, depth, radii, geomBuffer, binningBuffer, imgBuffer = _forward_native(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            num_rendered, color, depth, radii, geomBuffer, binningBuffer, imgBuffer = _forward_native(*args)

        ctx.raster_settings = raster_settings
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              geomBuffer, binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        return color, radii, depth

    @staticmethod
    def backward(ctx, grad_out_color, grad_radii, grad_depth):
        # depth is forward-only in the reference too (grad_depth is dropped, __init__.py:101-129)
        num_rendered = ctx.num_rendered
        raster_settings = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
         geomBuffer, binningBuffer, imgBuffer) = ctx.saved_tensors
        args = (raster_settings.bg, means3D, radii, colors_precomp, scales, rotations,
                raster_settings.scale_modifier, cov3Ds_precomp, raster_settings.viewmatrix,
                raster_settings.projmatrix, raster_settings.tanfovx, raster_settings.tanfovy,
                grad_out_color, sh, raster_settings.sh_degree, raster_settings.campos,
                geomBuffer, num_rendered, binningBuffer, imgBuffer, raster_settings.debug)
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp,
                 grad_sh, grad_scales, grad_rotations) = _backward_native(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp,
             grad_sh, grad_scales, grad_rotations) = _backward_native(*args)

        def _like(g, x):
            # placeholders (empty CPU tensors) never require grad; shape the rest like their input
            if x is None or x.numel() == 0:
                return None
            return g.view(x.shape) if g.numel() == x.numel() else g

        grads = (
            grad_means3D,
            grad_means2D,
            _like(grad_sh, sh),
            _like(grad_colors_precomp, colors_precomp),
            grad_opacities,
            _like(grad_scales, scales),
            _like(grad_rotations, rotations),
            _like(grad_cov3Ds_precomp, cov3Ds_precomp),
            None,
        )
        return grads


class _RasterizeGaussiansRaw(torch.autograd.Function):
    """The rasterizer fed with the RAW deformation outputs (log-scales, un-normalised quaternions, opacity logits): the
    activations of gaussian_renderer/__init__.py:97-99 and their VJPs run inside the preprocess kernels
    (FdgsRasterForward.activate), so the activated tensors never exist.  Same outputs as
    rasterizer(..., opacities=sigmoid(o), scales=exp(s), rotations=normalize(r)); gradients are w.r.t. the raw inputs."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, opacity_logits, log_scales, raw_rotations, raster_settings):
        e = torch.Tensor([])
        s = raster_settings
        num_rendered, color, depth, radii, geomBuffer, binningBuffer, imgBuffer = _forward_native(
            s.bg, means3D, e, opacity_logits, log_scales, raw_rotations, s.scale_modifier, e, s.viewmatrix, s.projmatrix, s.tanfovx,
            s.tanfovy, s.image_height, s.image_width, sh, s.sh_degree, s.campos, s.prefiltered, s.debug, activate=True)
        ctx.raster_settings = raster_settings
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(means3D, log_scales, raw_rotations, radii, sh, geomBuffer, binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        return color, radii, depth

    @staticmethod
    def backward(ctx, grad_out_color, grad_radii, grad_depth):
        s = ctx.raster_settings
        means3D, log_scales, raw_rotations, radii, sh, geomBuffer, binningBuffer, imgBuffer = ctx.saved_tensors
        e = torch.Tensor([])
        (grad_means2D, _gc, grad_opacities, grad_means3D, _gcov, grad_sh, grad_scales, grad_rotations) = _backward_native(
            s.bg, means3D, radii, e, log_scales, raw_rotations, s.scale_modifier, e, s.viewmatrix, s.projmatrix, s.tanfovx,
            s.tanfovy, grad_out_color, sh, s.sh_degree, s.campos, geomBuffer, ctx.num_rendered, binningBuffer, imgBuffer,
            s.debug, activate=True)
        return grad_means3D, grad_means2D, grad_sh.view(sh.shape), grad_opacities, grad_scales, grad_rotations, None


def rasterize_gaussians_raw(means3D, means2D, shs, opacity_logits, log_scales, raw_rotations, raster_settings):
    """(rendered_image, radii, depth) from the raw deformation outputs; see _RasterizeGaussiansRaw."""
    return _RasterizeGaussiansRaw.apply(means3D, means2D, shs, opacity_logits, log_scales, raw_rotations, raster_settings)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        # Mark visible points (based on frustum culling for camera) with a boolean
        with torch.no_grad():
            raster_settings = self.raster_settings
            visible = _mark_visible(positions, raster_settings.viewmatrix, raster_settings.projmatrix)
        return visible

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        raster_settings = self.raster_settings

        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')

        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])

        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, raster_settings)


def rasterize_debug(raster_settings, means3D, opacities, shs=None, colors_precomp=None, scales=None,
                    rotations=None, cov3D_precomp=None):
    """Forward-only call that also returns the internal state the parity tests compare bit-for-bit:
    sorted keys / point_list / tile ranges / per-Gaussian geometry / final_T / n_contrib."""
    e = torch.Tensor([])
    s = raster_settings
    out = _forward_native(s.bg, means3D, e if colors_precomp is None else colors_precomp, opacities,
                          e if scales is None else scales, e if rotations is None else rotations, s.scale_modifier,
                          e if cov3D_precomp is None else cov3D_precomp, s.viewmatrix, s.projmatrix, s.tanfovx,
                          s.tanfovy, s.image_height, s.image_width, e if shs is None else shs, s.sh_degree, s.campos,
                          s.prefiltered, s.debug, want_debug_lists=True)
    num_rendered, color, depth, radii, geom, binb, img, extras = out
    extras.update(num_rendered=num_rendered, color=color, depth=depth, radii=radii, _keepalive=(geom, binb, img))
    return extras
