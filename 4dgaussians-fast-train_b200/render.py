"""Caller-side glue: mirror of gaussian_renderer.render() (gaussian_renderer/__init__.py:18-138),
the function that feeds the deformation field's outputs into the rasterizer.  Same return dict.

Differences, none observable in the results:
  * the per-call `.cuda()` copies of the camera matrices (:45-48) are skipped when the tensors are
    already on the device;
  * `time` is passed as one device scalar instead of a [P,1] repeat (:52) — the fused deformation
    kernel takes one timestamp per view;
  * when the model binds the reference's three activations (exp / normalize / sigmoid, :97-99) they run INSIDE the
    rasterizer's preprocess kernels, forward and backward (`rasterize_gaussians_raw`; the activated [P,3] / [P,4] / [P,1]
    tensors never exist in memory) - or, with `fold_activations = False`, as one fused kernel each way
    (fdgs.activations); any other callables are used as they are.
"""
import math

import torch

from .activations import fused_activations, is_reference_activations
from .diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians_raw

fold_activations = True      # False: keep the activations as their own fused kernel pair (fdgs.activations); A/B and parity tests


def render(viewpoint_camera, pc, pipe=None, bg_color=None, scaling_modifier=1.0, override_color=None,
           stage="fine", cam_type=None):
    xyz = pc.get_xyz
    dev = xyz.device
    screenspace_points = torch.zeros_like(xyz, dtype=xyz.dtype, requires_grad=True, device=dev) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    means3D = xyz
    debug = bool(getattr(pipe, "debug", False)) if pipe is not None else False
    for flag in ("compute_cov3D_python", "convert_SHs_python"):
        # the reference's render() crashes on the first and only reaches the second from the GUI
        if pipe is not None and getattr(pipe, flag, False):
            raise NotImplementedError(f"pipe.{flag}=True is not supported by fdgs.render (covariances and SH colours are "
                                      "computed inside the rasterizer kernels); pass override_color for precomputed colours")
    if cam_type != "PanopticSports":
        tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
        tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
        raster_settings = GaussianRasterizationSettings(
            image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
            tanfovx=tanfovx, tanfovy=tanfovy, bg=bg_color, scale_modifier=scaling_modifier,
            viewmatrix=viewpoint_camera.world_view_transform.to(dev),
            projmatrix=viewpoint_camera.full_proj_transform.to(dev),
            sh_degree=pc.active_sh_degree, campos=viewpoint_camera.camera_center.to(dev),
            prefiltered=False, debug=debug)
        time = viewpoint_camera.time
    else:
        raster_settings = viewpoint_camera['camera']
        time = viewpoint_camera['time']
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)

    means2D = screenspace_points
    opacity = pc._opacity
    shs = pc.get_features
    scales = pc._scaling
    rotations = pc._rotation
    if "coarse" in stage:
        means3D_final, scales_final, rotations_final, opacity_final, shs_final = means3D, scales, rotations, opacity, shs
    elif "fine" in stage:
        means3D_final, scales_final, rotations_final, opacity_final, shs_final = pc._deformation(
            means3D, scales, rotations, opacity, shs, time)
    else:
        raise NotImplementedError

    colors_precomp = override_color
    if (scales_final.is_cuda and is_reference_activations(pc) and colors_precomp is None and not debug
            and opacity_final.dim() == 2 and fold_activations):
        # exp / normalize / sigmoid (:97-99) and their VJPs run INSIDE the rasterizer's preprocess kernels
        rendered_image, radii, depth = rasterize_gaussians_raw(means3D_final, means2D, shs_final, opacity_final, scales_final,
                                                               rotations_final, raster_settings)
        return {"render": rendered_image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
                "radii": radii, "depth": depth}
    if scales_final.is_cuda and is_reference_activations(pc):
        scales_final, rotations_final, opacity = fused_activations(scales_final, rotations_final, opacity_final)
    else:
        scales_final = pc.scaling_activation(scales_final)
        rotations_final = pc.rotation_activation(rotations_final)
        opacity = pc.opacity_activation(opacity_final)

    rendered_image, radii, depth = rasterizer(
        means3D=means3D_final, means2D=means2D, shs=shs_final if colors_precomp is None else None,
        colors_precomp=colors_precomp, opacities=opacity, scales=scales_final, rotations=rotations_final,
        cov3D_precomp=None)
    return {"render": rendered_image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
            "radii": radii, "depth": depth}
