"""Drop-in `deform_network` backed by the fused sm_100a HexPlane-sampler + MLP kernels.

Mirrors the reference's module tree so that state_dict keys, parameter names (Adam groups split on
"grid" in the name), `.get_mlp_parameters()`, `.get_grid_parameters()`, `.deformation_net.set_aabb`,
`.get_aabb` and `.deformation_net.grid.grids` keep working unchanged:

  scene/deformation.py:161-216  deform_network     -> deform_network
  scene/deformation.py:16-160   Deformation        -> Deformation
  scene/hexplane.py:109-183     HexPlaneField      -> HexPlaneField
  scene/hexplane.py:48-70       init_grid_param    -> init_grid_param
  scene/deformation.py:218-224  initialize_weights -> initialize_weights

The arithmetic (grid_sample x 6L, products, concat, 3-11 Linear layers; ~200 ATen launches forward
and ~400 backward per view in the reference) runs in ONE CUDA kernel per direction through the
C ABI in include/fdgs.h (fdgs_deform_forward / fdgs_deform_backward).  There is no PyTorch/CPU
fallback: CPU tensors raise.
"""
import ctypes as C
import itertools

import torch
import torch.nn as nn
import torch.nn.init as init

from . import _lib as L

HEADS = ("pos_deform", "scales_deform", "rotations_deform", "opacity_deform", "shs_deform")
HEAD_OUT = (3, 3, 4, 1, 48)


def normalize_aabb(pts, aabb):
    return (pts - aabb[0]) * (2.0 / (aabb[1] - aabb[0])) - 1.0


def init_grid_param(grid_nd, in_dim, out_dim, reso, a=0.1, b=0.5):
    assert in_dim == len(reso), "Resolution must have same number of elements as input-dimension"
    has_time_planes = in_dim == 4
    assert grid_nd <= in_dim
    coo_combs = list(itertools.combinations(range(in_dim), grid_nd))
    grid_coefs = nn.ParameterList()
    for coo_comb in coo_combs:
        new_grid_coef = nn.Parameter(torch.empty([1, out_dim] + [reso[cc] for cc in coo_comb[::-1]]))
        if has_time_planes and 3 in coo_comb:  # time planes start at 1
            nn.init.ones_(new_grid_coef)
        else:
            nn.init.uniform_(new_grid_coef, a=a, b=b)
        grid_coefs.append(new_grid_coef)
    return grid_coefs


class HexPlaneField(nn.Module):
    """Parameter container of the 6 x L k-planes (scene/hexplane.py:109-159).  Sampling is fused
    into the deformation kernels, so `forward` here exists only for API completeness."""

    def __init__(self, bounds, planeconfig, multires):
        super().__init__()
        aabb = torch.tensor([[bounds, bounds, bounds], [-bounds, -bounds, -bounds]], dtype=torch.float32)
        self.aabb = nn.Parameter(aabb, requires_grad=False)
        self.grid_config = [planeconfig]
        self.multiscale_res_multipliers = multires
        self.concat_features = True
        self.grids = nn.ModuleList()
        self.feat_dim = 0
        self.level_res = []
        for res in self.multiscale_res_multipliers:
            config = self.grid_config[0].copy()
            config["resolution"] = [r * res for r in config["resolution"][:3]] + config["resolution"][3:]
            gp = init_grid_param(grid_nd=config["grid_dimensions"], in_dim=config["input_coordinate_dim"],
                                 out_dim=config["output_coordinate_dim"], reso=config["resolution"])
            self.feat_dim += gp[-1].shape[1]
            self.grids.append(gp)
            self.level_res.append(list(config["resolution"]))

    @property
    def get_aabb(self):
        return self.aabb[0], self.aabb[1]

    def set_aabb(self, xyz_max, xyz_min):
        aabb = torch.tensor([xyz_max, xyz_min], dtype=torch.float32)
        self.aabb = nn.Parameter(aabb.to(self.aabb.device), requires_grad=False)
        self._aabb_host = None

    def aabb_host(self):
        """(max xyz, min xyz) as 6 host floats, cached (one D2H per set_aabb / load_state_dict)."""
        key = (self.aabb.data_ptr(), self.aabb._version)
        cached = getattr(self, "_aabb_host", None)
        if cached is None or cached[0] != key:
            vals = [float(v) for v in self.aabb.detach().reshape(-1).cpu().tolist()]
            self._aabb_host = (key, vals)
        return self._aabb_host[1]


class _DeformFunction(torch.autograd.Function):
    """out_h = in_h + head_h(trunk(hexplane(xyz, t)))  for the active heads."""

    @staticmethod
    def forward(ctx, mod, time_dev, time_host, point, scales, rotations, opacity, shs, *params):
        lib = L.load()
        net_mod = mod
        dev = point.device
        P = point.shape[0]
        f = dict(dtype=torch.float32, device=dev)
        hm = net_mod.head_mask
        point_c = point.contiguous()
        ins = [point_c,
               scales.contiguous() if (hm >> 1) & 1 else None,
               rotations.contiguous() if ((hm >> 2) & 1 and not net_mod.args.apply_rotation) else None,
               opacity.contiguous() if (hm >> 3) & 1 else None,
               shs.contiguous() if (hm >> 4) & 1 else None]
        outs = [torch.empty((P, 3), **f) if (hm >> 0) & 1 else None,
                torch.empty((P, 3), **f) if (hm >> 1) & 1 else None,
                torch.empty((P, 4), **f) if (hm >> 2) & 1 else None,
                torch.empty((P, 1), **f) if (hm >> 3) & 1 else None,
                torch.empty((P, 16, 3), **f) if (hm >> 4) & 1 else None]
        with L.on_device(dev):
            a = L.FdgsDeformForward()
            keep = net_mod._fill_net(a.net)
            a.P = P
            a.sh_coeffs = int(shs.shape[1]) if ((hm >> 4) & 1) else 16
            a.time = float(time_host) if time_host is not None else 0.0
            a.time_dev = time_dev.data_ptr() if time_dev is not None else None
            a.precision = int(net_mod.precision)
            packed, valid = net_mod._packed_planes(lib, a.net)
            a.planes_packed_valid = int(valid)
            trow = torch.empty((lib.fdgs_deform_trow_bytes(C.byref(a.net)),), dtype=torch.uint8, device=dev)
            a.xyz, a.scales, a.rotations, a.opacity, a.shs = [L.ptr(t) for t in ins]
            a.out_xyz, a.out_scales, a.out_rotations, a.out_opacity, a.out_shs = [L.ptr(t) for t in outs]
            a.packed_ws, a.packed_bytes = packed.data_ptr(), packed.numel()
            a.trow_ws, a.trow_bytes = trow.data_ptr(), trow.numel()
            wprep = net_mod._wprep(lib, a.net, dev)
            a.wprep_ws, a.wprep_bytes = wprep.data_ptr(), wprep.numel()
            # training: the kernel keeps feat / relu(h) / relu(z_j) for the backward (no recompute there)
            saved = None
            if net_mod.save_activations and any(ctx.needs_input_grad) and int(net_mod.precision) != 2:
                nsave = lib.fdgs_deform_forward_save_bytes(C.byref(a.net), int(P))
                if nsave:
                    saved = torch.empty((nsave,), dtype=torch.uint8, device=dev)
                    a.save_ws, a.save_bytes = saved.data_ptr(), saved.numel()
            a.debug = 0
            L.check(lib.fdgs_deform_forward(C.byref(a), L.stream_ptr()),
                    "fdgs_deform_forward")
            net_mod._mark_packed_valid()
            del keep
        ctx.mod = net_mod
        ctx.time_host = time_host
        # the backward re-reads the live parameter storage and the module-level packed-plane cache through raw pointers,
        # which autograd's version counters cannot see: remember the versions and refuse a stale backward ourselves
        ctx.param_versions = tuple(p._version for p in params)
        ctx.params = params
        ctx.plane_key = net_mod._packed_key
        ctx.save_for_backward(point_c, packed, trow, *([time_dev] if time_dev is not None else []))
        ctx.saved_act = saved
        ctx.has_time_dev = time_dev is not None
        ctx.n_params = len(params)
        ctx.sh_coeffs = a.sh_coeffs
        return tuple(o for o in outs if o is not None)

    @staticmethod
    def backward(ctx, *gouts):
        lib = L.load()
        mod = ctx.mod
        if tuple(p._version for p in ctx.params) != ctx.param_versions or mod._packed_key != ctx.plane_key:
            raise RuntimeError("one of the variables needed for gradient computation has been modified by an inplace "
                               "operation: a deformation parameter (or the packed-plane cache) changed between this "
                               "forward and its backward (fdgs re-reads parameters through raw pointers; run backward "
                               "before optimizer.step(), see INTEGRATION.md)")
        saved = ctx.saved_tensors
        point_c, packed, trow = saved[0], saved[1], saved[2]
        time_dev = saved[3] if ctx.has_time_dev else None
        dev = point_c.device
        P = point_c.shape[0]
        hm = mod.head_mask
        g = [None] * 5
        it = iter(gouts)
        for h in range(5):
            if (hm >> h) & 1:
                gh = next(it)
                g[h] = gh.contiguous() if gh is not None else None
        # identity paths (mask == 1, scene/deformation.py:104-146): the upstream gradients of the scale / rotation /
        # opacity / SH outputs ARE the gradients of the matching inputs.  A caller that owns the destination buffers
        # (fdgs.dp: the all-reduce bucket) may take them here, BEFORE the deformation kernels are queued, and start
        # reducing them while those kernels run; it returns the set of head indices it consumed.
        consumed = ()
        sink = getattr(mod, "identity_sink", None)
        if sink is not None and not mod.args.apply_rotation and ctx.sh_coeffs == 16:
            consumed = sink(g) or ()
        with L.on_device(dev):
            a = L.FdgsDeformBackward()
            keep = mod._fill_net(a.net)
            a.P = P
            a.time = float(ctx.time_host) if ctx.time_host is not None else 0.0
            a.time_dev = time_dev.data_ptr() if time_dev is not None else None
            a.precision = int(mod.precision)
            a.xyz = point_c.data_ptr()
            a.g_xyz, a.g_scales, a.g_rotations, a.g_opacity, a.g_shs = [L.ptr(t) for t in g]
            d_xyz = torch.empty((P, 3), dtype=torch.float32, device=dev)
            a.d_xyz_grid = d_xyz.data_ptr()
            pgrads = mod._alloc_param_grads(a, g)
            ws = torch.empty((lib.fdgs_deform_backward_ws_bytes(C.byref(a.net)),), dtype=torch.uint8, device=dev)
            a.packed_ws, a.trow_ws = packed.data_ptr(), trow.data_ptr()
            a.ws, a.ws_bytes = ws.data_ptr(), ws.numel()
            scratch = mod._bwd_scratch(lib, a.net, P, dev) if int(mod.precision) != 2 else None
            if scratch is not None:
                a.scratch, a.scratch_bytes = scratch.data_ptr(), scratch.numel()
            if ctx.saved_act is not None and scratch is not None:
                a.saved_ws, a.saved_bytes = ctx.saved_act.data_ptr(), ctx.saved_act.numel()
            a.debug = 0
            L.check(lib.fdgs_deform_backward(C.byref(a), L.stream_ptr()),
                    "fdgs_deform_backward")
            ctx.saved_act = None      # stream-ordered: the caching allocator may hand the block out again
            del keep
        # identity paths (mask == 1, scene/deformation.py:104-146)
        d_point = d_xyz + g[0] if g[0] is not None else d_xyz
        d_scales = g[1] if 1 not in consumed else None
        d_rot = g[2] if (not mod.args.apply_rotation and 2 not in consumed) else None
        d_opacity = g[3] if 3 not in consumed else None
        d_shs = None
        if g[4] is not None and 4 not in consumed:
            d_shs = g[4] if ctx.sh_coeffs == 16 else g[4].sum(dim=1, keepdim=True)
        return (None, None, None, d_point, d_scales, d_rot, d_opacity, d_shs) + tuple(pgrads)


class Deformation(nn.Module):
    def __init__(self, D=8, W=256, input_ch=27, input_ch_time=9, grid_pe=0, skips=[], args=None):
        super().__init__()
        self.D = D
        self.W = W
        self.input_ch = input_ch
        self.input_ch_time = input_ch_time
        self.skips = skips
        self.grid_pe = grid_pe
        self.no_grid = args.no_grid
        for flag in ("no_grid", "empty_voxel", "static_mlp"):
            if getattr(args, flag, False):
                raise NotImplementedError(f"{flag}=True is a 'useless' flag of the reference (arguments/__init__.py:99-107) "
                                          "and is not supported by the fused kernels")
        if grid_pe not in (0, 1):
            raise NotImplementedError("grid_pe > 1 is not supported by the fused kernels")
        if D > 1:
            raise NotImplementedError("defor_depth > 1 is not supported by the fused kernels (every shipped config of the "
                                      "reference uses 0 or 1, arguments/*/*.py): the backward has no multi-layer trunk")
        self.grid = HexPlaneField(args.bounds, args.kplanes_config, args.multires)
        self.args = args
        self.ratio = 0
        self.precision = 0      # 0: FP32-equivalent (3xTF32 on tcgen05), 1: single-pass TF32, 2: FP32 CUDA-core kernels
        self.grads_in_place = False    # see _alloc_param_grads
        self.save_activations = True   # training forward keeps the hidden activations (~3.2 KB/point) instead of recomputing
        self.identity_sink = None      # see _DeformFunction.backward (set by fdgs.dp.train_iteration for its last view)
        self.validate_time = False     # True: check (one device sync) that times_sel is one constant per call, see forward_dynamic
        self._packed = None
        self._packed_key = None
        self.create_net()

    @property
    def get_aabb(self):
        return self.grid.get_aabb

    def set_aabb(self, xyz_max, xyz_min):
        self.grid.set_aabb(xyz_max, xyz_min)

    def create_net(self):
        grid_out_dim = self.grid.feat_dim
        layers = [nn.Linear(grid_out_dim, self.W)]
        for _ in range(self.D - 1):
            layers.append(nn.ReLU())
            layers.append(nn.Linear(self.W, self.W))
        self.feature_out = nn.Sequential(*layers)
        self.pos_deform = nn.Sequential(nn.ReLU(), nn.Linear(self.W, self.W), nn.ReLU(), nn.Linear(self.W, 3))
        self.scales_deform = nn.Sequential(nn.ReLU(), nn.Linear(self.W, self.W), nn.ReLU(), nn.Linear(self.W, 3))
        self.rotations_deform = nn.Sequential(nn.ReLU(), nn.Linear(self.W, self.W), nn.ReLU(), nn.Linear(self.W, 4))
        self.opacity_deform = nn.Sequential(nn.ReLU(), nn.Linear(self.W, self.W), nn.ReLU(), nn.Linear(self.W, 1))
        self.shs_deform = nn.Sequential(nn.ReLU(), nn.Linear(self.W, self.W), nn.ReLU(), nn.Linear(self.W, 16 * 3))

    # ---- glue to the C ABI --------------------------------------------------------------------
    @property
    def head_mask(self):
        a = self.args
        return ((not a.no_dx) << 0) | ((not a.no_ds) << 1) | ((not a.no_dr) << 2) | ((not a.no_do) << 3) | \
               ((not a.no_dshs) << 4)

    def _trunk_linears(self):
        return [m for m in self.feature_out if isinstance(m, nn.Linear)]

    def _flat(self):
        """(planes[l][k], [(trunk w, b)], [(head w1, b1, w2, b2)]) - the Parameter objects the kernels read, looked up once:
        the module structure is static, and walking ModuleList / Sequential / nn.Module.__getattr__ on every call was a
        measurable part of the host time per view (hundreds of attribute look-ups).  Re-derived if a parameter object was
        replaced (two identity checks)."""
        c = self.__dict__.get("_flat_cache")
        if c is not None and c[0][0][0] is self.grid.grids[0][0] and c[2][0][0] is getattr(self, HEADS[0])[1].weight:
            return c
        planes = [[p for p in pl] for pl in self.grid.grids]
        trunk = [(lin.weight, lin.bias) for lin in self._trunk_linears()]
        heads = []
        for name in HEADS:
            seq = getattr(self, name)
            heads.append((seq[1].weight, seq[1].bias, seq[3].weight, seq[3].bias))
        c = (planes, trunk, heads)
        self.__dict__["_flat_cache"] = c
        return c

    def _fill_net(self, net):
        """Fills an FdgsDeformNet with shapes and raw device pointers; returns tensors to keep alive.
        The filled block is cached and copied (one memmove) while the parameters' storage, the aabb and the head flags are
        unchanged: writing ~60 fields of nested ctypes arrays one by one took ~80 us per call, twice per view."""
        planes_f, trunk_f, heads_f = self._flat()
        flat = [p for pl in planes_f for p in pl] + [p for wb in trunk_f for p in wb] + [p for hp in heads_f for p in hp]
        g = self.grid
        aabb = g.aabb_host()
        key = (tuple(p.data_ptr() for p in flat), tuple(aabb), self.head_mask, bool(self.args.apply_rotation), self.W, self.D)
        cached = self.__dict__.get("_net_cache")
        if cached is not None and cached[0] == key:
            C.memmove(C.addressof(net), cached[1], C.sizeof(net))
            return []
        keep = self._fill_net_fields(net)
        if not keep and all(p.is_contiguous() for p in flat):
            # kept as bytes: a ctypes block with pointer fields would make the module un-picklable / un-deepcopy-able
            self.__dict__["_net_cache"] = (key, C.string_at(C.addressof(net), C.sizeof(net)))
        return keep

    def _fill_net_fields(self, net):
        keep = []
        g = self.grid
        net.levels = len(g.grids)
        net.channels = g.grids[0][0].shape[1]
        net.width = self.W
        net.depth = max(self.D, 1)
        net.head_mask = self.head_mask
        net.apply_rotation = int(bool(self.args.apply_rotation))
        for l, res in enumerate(g.level_res):
            for i in range(4):
                net.res[l][i] = int(res[i])
        for i, v in enumerate(g.aabb_host()):
            net.aabb[i] = v
        planes_f, trunk_f, heads_f = self._flat()

        def ptr(p):                      # a Parameter is normally contiguous: its own storage, nothing to keep alive
            if p.is_contiguous():
                return p.data_ptr()
            t = p.detach().contiguous()
            keep.append(t)
            return t.data_ptr()

        for l, planes in enumerate(planes_f):
            for k, p in enumerate(planes):
                net.planes[l][k] = ptr(p)
        for d, (w, b) in enumerate(trunk_f):
            net.trunk_w[d], net.trunk_b[d] = ptr(w), ptr(b)
        for h, (w1, b1, w2, b2) in enumerate(heads_f):
            net.head_w1[h], net.head_b1[h] = ptr(w1), ptr(b1)
            net.head_w2[h], net.head_b2[h] = ptr(w2), ptr(b2)
        return keep

    def _plane_key(self):
        return tuple((p.data_ptr(), p._version) for planes in self._flat()[0] for p in planes)

    def _packed_planes(self, lib, net):
        """Channel-last copy of the planes, rebuilt only when a plane parameter changed."""
        key = self._plane_key()
        dev = self.grid.grids[0][0].device
        nbytes = lib.fdgs_deform_packed_bytes(C.byref(net))
        if nbytes == 0:
            raise RuntimeError("fdgs: deformation configuration not supported by the fused kernels "
                               "(channels in {4,8,16,..,64}, net_width in {64,128}, C*L <= 128)")
        if self._packed is None or self._packed.numel() != nbytes or self._packed.device != dev:
            self._packed = torch.empty((nbytes,), dtype=torch.uint8, device=dev)
            self._packed_key = None
        self._pending_key = key
        return self._packed, (self._packed_key == key)

    def _wprep(self, lib, net, dev):
        """Scratch for the tensor-core formatted MLP weights (rewritten by every call)."""
        n = lib.fdgs_deform_wprep_bytes(C.byref(net))
        w = getattr(self, "_wprep_ws", None)
        if w is None or w.numel() < n or w.device != dev:
            w = torch.empty((max(n, 256),), dtype=torch.uint8, device=dev)
            self._wprep_ws = w
        return w

    def _bwd_scratch(self, lib, net, P, dev):
        """Scratch of the tensor-core backward (activations exchanged between its two kernels; dead after the call,
        so one buffer serves every view).  None when this net runs the CUDA-core backward."""
        n = lib.fdgs_deform_backward_scratch_bytes(C.byref(net), int(P))
        if n == 0:
            return None
        w = getattr(self, "_bwd_scratch_ws", None)
        if w is None or w.numel() < n or w.device != dev:
            self._bwd_scratch_ws = None
            w = torch.empty((n,), dtype=torch.uint8, device=dev)
            self._bwd_scratch_ws = w
        return w

    def _mark_packed_valid(self):
        self._packed_key = self._pending_key

    def _param_list(self):
        """Parameters that receive gradients, in the order `_alloc_param_grads` returns them."""
        planes_f, trunk_f, heads_f = self._flat()
        ps = [p for planes in planes_f for p in planes]
        for w, b in trunk_f:
            ps += [w, b]
        hm = self.head_mask
        for h, hp in enumerate(heads_f):
            if (hm >> h) & 1:
                ps += list(hp)
        return ps

    def _alloc_param_grads(self, a, g):
        """Gradient buffers of every parameter, in `_param_list` order.  The kernels ACCUMULATE (+=) into them.  With
        `grads_in_place` (set by fdgs.dp.GradBucket users) a parameter whose .grad already exists - e.g. a view into
        the flat all-reduce bucket - receives its gradient directly and autograd gets None for it: no zero-fill and
        no `grad += new` pass per parameter and view."""
        grads = []

        def buf(p):
            if self.grads_in_place and p.grad is not None and p.grad.is_contiguous() and p.grad.dtype == torch.float32 \
                    and p.grad.shape == p.shape:
                grads.append(None)
                return p.grad.data_ptr()
            t = torch.zeros_like(p, memory_format=torch.contiguous_format)
            grads.append(t)
            return t.data_ptr()

        planes_f, trunk_f, heads_f = self._flat()
        for l, planes in enumerate(planes_f):
            for k, p in enumerate(planes):
                a.d_planes[l][k] = buf(p)
        for d, (w, b) in enumerate(trunk_f):
            a.d_trunk_w[d], a.d_trunk_b[d] = buf(w), buf(b)
        hm = self.head_mask
        for h, (w1, b1, w2, b2) in enumerate(heads_f):
            if (hm >> h) & 1:
                a.d_head_w1[h], a.d_head_b1[h] = buf(w1), buf(b1)
                a.d_head_w2[h], a.d_head_b2[h] = buf(w2), buf(b2)
        return grads

    # ---- reference API ------------------------------------------------------------------------
    def forward(self, rays_pts_emb, scales_emb=None, rotations_emb=None, opacity=None, shs_emb=None,
                time_feature=None, time_emb=None):
        if time_emb is None:
            raise NotImplementedError("forward_static needs static_mlp, a 'useless' flag of the reference")
        return self.forward_dynamic(rays_pts_emb, scales_emb, rotations_emb, opacity, shs_emb, time_feature, time_emb)

    def forward_dynamic(self, rays_pts_emb, scales_emb, rotations_emb, opacity_emb, shs_emb, time_feature, time_emb):
        if not rays_pts_emb.is_cuda:
            raise RuntimeError("fdgs deformation needs CUDA tensors (there is no CPU fallback)")
        if rays_pts_emb.dtype != torch.float32:
            raise RuntimeError("fdgs deformation needs float32 inputs")
        point = rays_pts_emb[:, :3]
        hm = self.head_mask
        # the fused kernel uses ONE timestamp per call: render() builds time as scalar.repeat(P,1)
        # (gaussian_renderer/__init__.py:52).  A python float / 0-d tensor is accepted too.
        # Per-point timestamps (which the reference's API shape would allow) are NOT supported: a CPU tensor is checked
        # here, a CUDA tensor only with `validate_time` (the comparison costs a device sync).
        time_dev, time_host = None, None
        if isinstance(time_emb, torch.Tensor):
            if time_emb.numel() > 1 and (self.validate_time or not time_emb.is_cuda):
                flat = time_emb.detach().reshape(-1)
                if not bool((flat == flat[0]).all()):
                    raise ValueError("fdgs deformation takes ONE timestamp per call (times_sel must be a constant column, "
                                     "as gaussian_renderer/__init__.py:52 builds it); got per-point times")
            if time_emb.is_cuda:
                time_dev = time_emb.detach().reshape(-1)[:1].to(torch.float32)
            else:
                time_host = float(time_emb.reshape(-1)[0])
        else:
            time_host = float(time_emb)
        scales = scales_emb[:, :3] if scales_emb is not None else None
        rotations = rotations_emb[:, :4] if rotations_emb is not None else None
        opacity_in = opacity_emb[:, :1] if opacity_emb is not None else None
        outs = _DeformFunction.apply(self, time_dev, time_host, point, scales, rotations, opacity_in, shs_emb,
                                     *self._param_list())
        it = iter(outs)
        pts = next(it) if (hm >> 0) & 1 else point
        scales_o = next(it) if (hm >> 1) & 1 else scales
        if (hm >> 2) & 1:
            dr = next(it)
            if self.args.apply_rotation:
                from .quat import batch_quaternion_multiply
                rot_o = batch_quaternion_multiply(rotations_emb, dr)
            else:
                rot_o = dr
        else:
            rot_o = rotations
        opacity_o = next(it) if (hm >> 3) & 1 else opacity_in
        shs_o = next(it) if (hm >> 4) & 1 else shs_emb
        return pts, scales_o, rot_o, opacity_o, shs_o

    def get_mlp_parameters(self):
        return [param for name, param in self.named_parameters() if "grid" not in name]

    def get_grid_parameters(self):
        return [param for name, param in self.named_parameters() if "grid" in name]


class deform_network(nn.Module):
    def __init__(self, args):
        super().__init__()
        net_width = args.net_width
        timebase_pe = args.timebase_pe
        defor_depth = args.defor_depth
        posbase_pe = args.posebase_pe
        scale_rotation_pe = args.scale_rotation_pe
        opacity_pe = args.opacity_pe
        timenet_width = args.timenet_width
        timenet_output = args.timenet_output
        grid_pe = args.grid_pe
        times_ch = 2 * timebase_pe + 1
        # kept for state_dict parity; never used by the reference's forward either (deformation.py:203-204)
        self.timenet = nn.Sequential(nn.Linear(times_ch, timenet_width), nn.ReLU(),
                                     nn.Linear(timenet_width, timenet_output))
        self.deformation_net = Deformation(W=net_width, D=defor_depth, input_ch=(3) + (3 * (posbase_pe)) * 2,
                                           grid_pe=grid_pe, input_ch_time=timenet_output, args=args)
        self.register_buffer('time_poc', torch.FloatTensor([(2 ** i) for i in range(timebase_pe)]))
        self.register_buffer('pos_poc', torch.FloatTensor([(2 ** i) for i in range(posbase_pe)]))
        self.register_buffer('rotation_scaling_poc', torch.FloatTensor([(2 ** i) for i in range(scale_rotation_pe)]))
        self.register_buffer('opacity_poc', torch.FloatTensor([(2 ** i) for i in range(opacity_pe)]))
        self.apply(initialize_weights)

    def forward(self, point, scales=None, rotations=None, opacity=None, shs=None, times_sel=None):
        return self.forward_dynamic(point, scales, rotations, opacity, shs, times_sel)

    @property
    def get_aabb(self):
        return self.deformation_net.get_aabb

    @property
    def get_empty_ratio(self):
        return self.deformation_net.ratio

    def forward_dynamic(self, point, scales=None, rotations=None, opacity=None, shs=None, times_sel=None):
        # The reference builds sin/cos embeddings of point/scales/rotations here and then only reads
        # their first 3/3/4 columns (deformation.py:200-202, :73, :107-130): dead compute, skipped.
        return self.deformation_net(point, scales, rotations, opacity, shs, None, times_sel)

    def get_mlp_parameters(self):
        return self.deformation_net.get_mlp_parameters() + list(self.timenet.parameters())

    def get_grid_parameters(self):
        return self.deformation_net.get_grid_parameters()


def initialize_weights(m):
    if isinstance(m, nn.Linear):
        init.xavier_uniform_(m.weight, gain=1)
        if m.bias is not None:
            init.xavier_uniform_(m.weight, gain=1)   # sic: the reference re-inits the weight, never the bias


def poc_fre(input_data, poc_buf):
    input_data_emb = (input_data.unsqueeze(-1) * poc_buf).flatten(-2)
    input_data_sin = input_data_emb.sin()
    input_data_cos = input_data_emb.cos()
    return torch.cat([input_data, input_data_sin, input_data_cos], -1)
