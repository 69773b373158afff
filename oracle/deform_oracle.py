"""TEST INFRASTRUCTURE — CPU restatement of the reference deformation path (never imported by the product).

Restates, with plain tensor arithmetic and explicit gathers (no F.grid_sample, no nn.Module), what
    deform_network.forward_dynamic            scene/deformation.py:198-212
    Deformation.query_time / forward_dynamic  scene/deformation.py:67-83, 97-148
    HexPlaneField.get_density                 scene/hexplane.py:160-175
    interpolate_ms_features                   scene/hexplane.py:73-106
    grid_sample_wrapper                       scene/hexplane.py:21-46
compute, with the bilinear rule of ATen's grid_sampler_2d (align_corners=True, padding_mode='border',
ATen/native/GridSampler.h:27-36,58-84,143-203 of torch 2.11).  Gradients come from autograd over this
restatement (float64 copy = arbiter).

Pinned: `oracle/make_golden.py` runs the REAL reference module (imported from /root/reference in the
authoring container) on seeded inputs and stores its outputs and gradients in tests/golden/; the
`-m "not gpu"` tests check this restatement against those fixtures.
"""
import itertools

import torch

HEADS = ("pos_deform", "scales_deform", "rotations_deform", "opacity_deform", "shs_deform")
PAIRS = list(itertools.combinations(range(4), 2))   # (0,1),(0,2),(0,3),(1,2),(1,3),(2,3)  hexplane.py:80-82


def _unnormalize_clip(q, size):
    # grid_sampler_unnormalize (align_corners) + clip_coordinates:  GridSampler.h:27-36, 58-60
    ix = ((q + 1) / 2) * (size - 1)
    return torch.clamp(ix, 0, size - 1)


def bilinear_border(plane, qx, qy):
    """plane [C,H,W]; qx indexes W, qy indexes H (hexplane.py:58-63).  Returns [N,C]."""
    Cc, H, W = plane.shape
    ix = _unnormalize_clip(qx, W)
    iy = _unnormalize_clip(qy, H)
    x0 = torch.floor(ix)
    y0 = torch.floor(iy)
    x1 = x0 + 1
    y1 = y0 + 1
    nw = (x1 - ix) * (y1 - iy)
    ne = (ix - x0) * (y1 - iy)
    sw = (x1 - ix) * (iy - y0)
    se = (ix - x0) * (iy - y0)
    x0i, y0i, x1i, y1i = x0.long(), y0.long(), x1.long(), y1.long()
    flat = plane.reshape(Cc, H * W)

    def tap(xi, yi, w):
        ok = (xi >= 0) & (xi < W) & (yi >= 0) & (yi < H)          # within_bounds_2d, GridSampler.h:143-172
        idx = (yi.clamp(0, H - 1) * W + xi.clamp(0, W - 1))
        v = flat[:, idx]                                            # [C,N]
        return (v * (w * ok.to(w.dtype)).unsqueeze(0))

    out = tap(x0i, y0i, nw) + tap(x1i, y0i, ne) + tap(x0i, y1i, sw) + tap(x1i, y1i, se)
    return out.t()


def hexplane_features(sd, cfg, xyz, t):
    """sd: state_dict-like mapping; returns feat [N, C*L] (hexplane.py:160-175, 73-106)."""
    aabb = sd["deformation_net.grid.aabb"].to(xyz.dtype)
    pts = (xyz - aabb[0]) * (2.0 / (aabb[1] - aabb[0])) - 1.0       # normalize_aabb, hexplane.py:19-20
    pts = torch.cat((pts, t), dim=-1)                               # time is NOT normalised (:164)
    feats = []
    for l in range(len(cfg["multires"])):
        interp = 1.0
        for k, (a, b) in enumerate(PAIRS):
            plane = sd[f"deformation_net.grid.grids.{l}.{k}"][0].to(xyz.dtype)
            interp = interp * bilinear_border(plane, pts[:, a], pts[:, b])
        feats.append(interp)
    return torch.cat(feats, dim=-1)


def _linear(sd, name, x):
    return x @ sd[name + ".weight"].to(x.dtype).t() + sd[name + ".bias"].to(x.dtype)


def deform_forward(sd, cfg, point, scales, rotations, opacity, shs, times_sel):
    """Returns the reference's 5-tuple (pre-activation)."""
    feat = hexplane_features(sd, cfg, point[:, :3], times_sel[:, :1])
    depth = max(int(cfg["defor_depth"]), 1)
    h = _linear(sd, "deformation_net.feature_out.0", feat)           # deformation.py:55-60, 80
    for d in range(1, depth):
        h = _linear(sd, f"deformation_net.feature_out.{2 * d}", torch.relu(h))

    def head(name):                                                  # Sequential(ReLU, Linear, ReLU, Linear) :61-65
        z = _linear(sd, f"deformation_net.{name}.1", torch.relu(h))
        return _linear(sd, f"deformation_net.{name}.3", torch.relu(z))

    pts = point[:, :3] if cfg.get("no_dx", False) else point[:, :3] + head("pos_deform")         # :106-111 (mask == 1)
    sc = scales[:, :3] if cfg.get("no_ds", False) else scales[:, :3] + head("scales_deform")      # :112-119
    if cfg.get("no_dr", False):
        rot = rotations[:, :4]
    else:
        dr = head("rotations_deform")
        if cfg.get("apply_rotation", False):                                                       # :127-128
            q1, q2 = rotations, dr
            w = q1[:, 0] * q2[:, 0] - q1[:, 1] * q2[:, 1] - q1[:, 2] * q2[:, 2] - q1[:, 3] * q2[:, 3]
            x = q1[:, 0] * q2[:, 1] + q1[:, 1] * q2[:, 0] + q1[:, 2] * q2[:, 3] - q1[:, 3] * q2[:, 2]
            y = q1[:, 0] * q2[:, 2] - q1[:, 1] * q2[:, 3] + q1[:, 2] * q2[:, 0] + q1[:, 3] * q2[:, 1]
            z = q1[:, 0] * q2[:, 3] + q1[:, 1] * q2[:, 2] - q1[:, 2] * q2[:, 1] + q1[:, 3] * q2[:, 0]
            q3 = torch.stack((w, x, y, z), dim=1)
            rot = q3 / torch.norm(q3, dim=1, keepdim=True)
        else:
            rot = rotations[:, :4] + dr
    op = opacity[:, :1] if cfg.get("no_do", True) else opacity[:, :1] + head("opacity_deform")    # :132-138
    if cfg.get("no_dshs", True):
        sh = shs
    else:
        sh = shs + head("shs_deform").reshape(shs.shape[0], 16, 3)                                 # :139-146 (broadcasts M=1)
    return pts, sc, rot, op, sh


def cfg_from_args(args):
    return dict(multires=list(args.multires), defor_depth=int(args.defor_depth), no_dx=bool(args.no_dx),
                no_ds=bool(args.no_ds), no_dr=bool(args.no_dr), no_do=bool(args.no_do), no_dshs=bool(args.no_dshs),
                apply_rotation=bool(args.apply_rotation))
