"""fdgs — B200-native (sm_100a) 4D-Gaussian training hot path.

Sub-modules
  diff_gaussian_rasterization : drop-in GaussianRasterizer / GaussianRasterizationSettings
  deformation                 : drop-in deform_network (fused HexPlane sampler + MLP heads)
  render                      : the caller-side glue (mirror of gaussian_renderer.render)
  dp                          : view-parallel data parallelism (one view per GPU, NCCL all-reduce)
  synth                       : seeded synthetic scenes of the BASELINE.json shapes

The directory name is not a Python identifier; import it through the repo-root `fdgs.py` shim
(`import fdgs`) or put this directory on sys.path for the drop-in `diff_gaussian_rasterization`.
"""
from . import _lib  # noqa: F401

__all__ = ["_lib"]
