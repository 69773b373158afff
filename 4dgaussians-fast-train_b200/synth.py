"""Seeded synthetic scenes of the BASELINE.json shapes: random Gaussian clouds,
pinhole cameras with the reference's matrix conventions (scene/cameras.py:55-66,
utils/graphics_utils.py:38-71), and a small parameter container shaped like GaussianModel."""
import math
from types import SimpleNamespace

import numpy as np
import torch


def projection_matrix(znear, zfar, fovx, fovy):
    """utils/graphics_utils.py:51-71 (z sign +1)."""
    tx, ty = math.tan(fovx / 2), math.tan(fovy / 2)
    top, right = ty * znear, tx * znear
    Pm = np.zeros((4, 4), dtype=np.float64)
    Pm[0, 0] = 2.0 * znear / (2 * right)
    Pm[1, 1] = 2.0 * znear / (2 * top)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    return Pm


def make_camera(W, H, fovx, angle=0.0, dist=4.0, time=0.0, device="cpu", znear=0.01, zfar=100.0):
    """Camera on a circle of radius `dist` around the y axis, looking at the origin.
    Returns the fields render() reads from a reference `Camera` (scene/cameras.py)."""
    focal = W / (2.0 * math.tan(fovx / 2))
    fovy = 2.0 * math.atan(H / (2.0 * focal))
    Cw = np.array([dist * math.sin(angle), 0.0, dist * math.cos(angle)])
    f = -Cw / np.linalg.norm(Cw)                  # camera +z looks at the origin
    down = np.array([0.0, -1.0, 0.0])
    r = np.cross(down, f); r /= np.linalg.norm(r)
    d = np.cross(f, r)
    R = np.stack([r, d, f], axis=0)               # world -> camera rotation (rows = camera axes)
    W2V = np.eye(4)
    W2V[:3, :3] = R
    W2V[:3, 3] = -R @ Cw
    Pm = projection_matrix(znear, zfar, fovx, fovy)
    wvt = torch.tensor(W2V.T, dtype=torch.float32)                      # world_view_transform  (cameras.py:61)
    proj = torch.tensor(Pm.T, dtype=torch.float32)                      # projection_matrix     (:63)
    full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0)         # full_proj_transform   (:64)
    center = wvt.inverse()[3, :3]                                       # camera_center         (:65)
    return SimpleNamespace(image_width=W, image_height=H, FoVx=fovx, FoVy=fovy, time=float(time),
                           world_view_transform=wvt.to(device), full_proj_transform=full.to(device),
                           camera_center=center.to(device))


def make_cameras(W, H, fovx, n, device="cpu", time_den=None):
    den = time_den if time_den else n
    return [make_camera(W, H, fovx, angle=2 * math.pi * i / max(n, 1), time=i / den, device=device) for i in range(n)]


def make_gaussians(P, extent=1.3, logscale=-4.6, sh_degree=3, seed=0, device="cpu"):
    """Random cloud with the SURVEY §8d distributions.  Returns a dict of FP32 tensors."""
    rng = np.random.default_rng(seed)
    M = (sh_degree + 1) ** 2
    xyz = rng.uniform(-extent, extent, size=(P, 3))
    scaling = logscale + rng.normal(0, 0.6, size=(P, 1)) + rng.normal(0, 0.5, size=(P, 3))
    rotation = rng.normal(0, 1, size=(P, 4))
    opacity = rng.normal(0, 2, size=(P, 1))
    f_dc = rng.normal(0, 1, size=(P, 1, 3))
    f_rest = rng.normal(0, 0.1, size=(P, M - 1, 3))
    t = lambda a: torch.tensor(a, dtype=torch.float32, device=device)
    return dict(xyz=t(xyz), scaling=t(scaling), rotation=t(rotation), opacity=t(opacity), f_dc=t(f_dc), f_rest=t(f_rest))


class SynthGaussians(torch.nn.Module):
    """The per-Gaussian parameters render() reads from the reference's GaussianModel
    (scene/gaussian_model.py:52-64, 108-131): _xyz, _scaling (log), _rotation (raw), _opacity (logit),
    _features_dc / _features_rest, plus `_deformation`."""

    def __init__(self, tensors, deformation, sh_degree=3):
        super().__init__()
        P = torch.nn.Parameter
        self._xyz = P(tensors["xyz"].clone())
        self._scaling = P(tensors["scaling"].clone())
        self._rotation = P(tensors["rotation"].clone())
        self._opacity = P(tensors["opacity"].clone())
        self._features_dc = P(tensors["f_dc"].clone())
        self._features_rest = P(tensors["f_rest"].clone())
        self._deformation = deformation
        self.active_sh_degree = sh_degree
        self.max_sh_degree = sh_degree

    @property
    def get_xyz(self):
        return self._xyz

    @property
    def get_features(self):
        return torch.cat((self._features_dc, self._features_rest), dim=1)

    scaling_activation = staticmethod(torch.exp)
    opacity_activation = staticmethod(torch.sigmoid)
    rotation_activation = staticmethod(torch.nn.functional.normalize)


def perturb_deformation(net, seed=1, time_sigma=0.05):
    """reference init plus N(0, 0.05^2) on the time planes so the field is non-trivial;
    biases of the Linear layers keep torch's default init."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for planes in net.deformation_net.grid.grids:
            for k in (2, 4, 5):
                p = planes[k]
                p.add_(torch.randn(p.shape, generator=g).to(p.device) * time_sigma)
    return net


def seeded_deformation(make_net, args, seed=1234, device="cpu", time_sigma=0.05):
    """Builds a deform_network (ours or the reference's) with identical seeded parameters."""
    torch.manual_seed(seed)
    net = make_net(args)
    perturb_deformation(net, seed=seed + 1, time_sigma=time_sigma)
    return net.to(device)
