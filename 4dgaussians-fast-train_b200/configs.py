"""The deformation-field / image shapes of the BASELINE.json configs, as the
`args` namespace the reference's deform_network takes (arguments/__init__.py:79-107 defaults,
overlaid with arguments/{dnerf,hypernerf,dynerf}/*.py)."""
from types import SimpleNamespace


def hidden_params(**over):
    d = dict(net_width=64, timebase_pe=4, defor_depth=1, posebase_pe=10, scale_rotation_pe=2, opacity_pe=2,
             timenet_width=64, timenet_output=32, bounds=1.6, plane_tv_weight=0.0001, time_smoothness_weight=0.01,
             l1_time_planes=0.0001,
             kplanes_config={'grid_dimensions': 2, 'input_coordinate_dim': 4, 'output_coordinate_dim': 32,
                             'resolution': [64, 64, 64, 25]},
             multires=[1, 2, 4, 8], no_dx=False, no_grid=False, no_ds=False, no_dr=False, no_do=True, no_dshs=True,
             empty_voxel=False, grid_pe=0, static_mlp=False, apply_rotation=False)
    d.update(over)
    d["kplanes_config"] = dict(d["kplanes_config"])
    return SimpleNamespace(**d)


def _kp(c, res):
    return {'grid_dimensions': 2, 'input_coordinate_dim': 4, 'output_coordinate_dim': c, 'resolution': list(res)}


CONFIGS = {
    # name: (args factory, W, H, default P, views, fovx, position half-extent, log-scale mean)
    "tiny": dict(args=lambda: hidden_params(kplanes_config=_kp(8, [16, 16, 16, 20]), multires=[1], defor_depth=0,
                                            net_width=32, no_do=False, no_dshs=False),
                 W=200, H=120, P=2000, views=1, fovx=0.6911, extent=1.3, logscale=-3.5),
    # the DyNeRF net (C=16, L=2, Wn=128, five heads: the tcgen05 kernels) on small planes and a small image
    "dynerf_small": dict(args=lambda: hidden_params(kplanes_config=_kp(16, [16, 16, 16, 20]), multires=[1, 2], defor_depth=0,
                                                    net_width=128, no_do=False, no_dshs=False),
                         W=200, H=120, P=3000, views=2, fovx=0.6911, extent=1.3, logscale=-3.5),
    "dnerf": dict(args=lambda: hidden_params(kplanes_config=_kp(32, [64, 64, 64, 75]), multires=[1, 2], defor_depth=0,
                                             net_width=64, bounds=1.6),
                  W=800, H=800, P=200_000, views=1, fovx=0.6911, extent=1.3, logscale=-4.828),
    "dnerf50k": dict(args=lambda: hidden_params(kplanes_config=_kp(32, [64, 64, 64, 75]), multires=[1, 2],
                                                defor_depth=0, net_width=64, bounds=1.6),
                     W=800, H=800, P=50_000, views=1, fovx=0.6911, extent=1.3, logscale=-4.605),
    "hypernerf": dict(args=lambda: hidden_params(kplanes_config=_kp(16, [64, 64, 64, 150]), multires=[1, 2, 4],
                                                 defor_depth=1, net_width=128),
                      W=536, H=960, P=300_000, views=2, fovx=0.9, extent=2.0, logscale=-4.828),
    "dynerf": dict(args=lambda: hidden_params(kplanes_config=_kp(16, [64, 64, 64, 150]), multires=[1, 2],
                                              defor_depth=0, net_width=128, no_do=False, no_dshs=False),
                   W=1352, H=1014, P=300_000, views=8, fovx=1.2, extent=2.0, logscale=-4.828),
    "infer1080p": dict(args=lambda: hidden_params(kplanes_config=_kp(16, [64, 64, 64, 150]), multires=[1, 2],
                                                  defor_depth=0, net_width=128, no_do=False, no_dshs=False),
                       W=1920, H=1080, P=1_000_000, views=300, fovx=1.2, extent=2.0, logscale=-5.298),
}
