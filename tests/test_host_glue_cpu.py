"""Host-side glue of the deformation operator that needs no GPU: the cached parameter look-up and the cached FdgsDeformNet
block (fdgs/deformation.py: _flat, _fill_net) must follow the module through deepcopy / pickle / parameter replacement."""
import copy
import ctypes as C
import pickle

import torch

import fdgs  # noqa: F401
from fdgs import _lib as L
from fdgs.configs import CONFIGS
from fdgs.deformation import deform_network


def _block(dn):
    a = L.FdgsDeformForward()
    keep = dn._fill_net(a.net)
    return a, C.string_at(C.addressof(a.net), C.sizeof(a.net)), keep


def test_fill_net_cache_is_exact_and_follows_the_parameters():
    net = deform_network(CONFIGS["dynerf_small"]["args"]())
    dn = net.deformation_net
    a1, b1, _ = _block(dn)                      # fills field by field and caches
    a2, b2, keep = _block(dn)                   # served from the cache
    assert b1 == b2 and keep == [] and "_net_cache" in dn.__dict__
    assert a2.net.head_w1[0] == dn.pos_deform[1].weight.data_ptr()
    assert a2.net.planes[0][0] == dn.grid.grids[0][0].data_ptr()
    # new storage for one parameter: the cached block must not be served any more
    with torch.no_grad():
        dn.pos_deform[1].weight.data = dn.pos_deform[1].weight.data.clone()
    a3, b3, _ = _block(dn)
    assert a3.net.head_w1[0] == dn.pos_deform[1].weight.data_ptr() and b3 != b1
    # a replaced Parameter object invalidates the look-up cache as well
    dn.pos_deform[1].weight = torch.nn.Parameter(dn.pos_deform[1].weight.detach().clone())
    a4, _, _ = _block(dn)
    assert a4.net.head_w1[0] == dn.pos_deform[1].weight.data_ptr()
    # aabb is part of the block
    dn.set_aabb([2.0, 2.0, 2.0], [-2.0, -2.0, -2.0])
    a5, _, _ = _block(dn)
    assert [a5.net.aabb[i] for i in range(6)] == [2.0, 2.0, 2.0, -2.0, -2.0, -2.0]


def test_caches_survive_deepcopy_and_pickle():
    net = deform_network(CONFIGS["tiny"]["args"]())
    _block(net.deformation_net)
    twin = copy.deepcopy(net)
    a, _, _ = _block(twin.deformation_net)
    assert a.net.head_w1[0] == twin.deformation_net.pos_deform[1].weight.data_ptr()
    assert a.net.head_w1[0] != net.deformation_net.pos_deform[1].weight.data_ptr()
    back = pickle.loads(pickle.dumps(net))
    b, _, _ = _block(back.deformation_net)
    assert b.net.head_w1[0] == back.deformation_net.pos_deform[1].weight.data_ptr()
    # state_dict round trip keeps the reference's keys
    assert set(back.state_dict().keys()) == set(net.state_dict().keys())


def test_non_contiguous_parameter_is_not_cached():
    net = deform_network(CONFIGS["tiny"]["args"]())
    dn = net.deformation_net
    w = dn.pos_deform[1].weight
    with torch.no_grad():
        w.data = w.data.t().contiguous().t()     # same values, column-major storage
    assert not w.is_contiguous()
    a, _, keep = _block(dn)
    assert len(keep) == 1 and a.net.head_w1[0] == keep[0].data_ptr() and "_net_cache" not in dn.__dict__
