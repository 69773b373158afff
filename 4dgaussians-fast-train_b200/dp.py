"""View-parallel data parallelism for the training hot path.

The reference renders the B views of a batch in a Python loop and calls one backward on the mean L1 of
the concatenated images (train.py:197-262); views share only parameters.  Here view i of the batch goes
to rank i mod G, every rank holds a full replica of the Gaussians + HexPlane + MLP, and the ONE exchange
step of an iteration is a sum all-reduce of a flat FP32 gradient bucket.

The bucket is allocated once; every parameter's `.grad` (and the view-space gradient accumulator the
densifier reads, train.py:266-268) is a VIEW into it, so autograd accumulates straight into the bucket
and there is no pack / unpack pass around the collective.  Parameters that never receive a gradient
(timenet.*, grid.aabb, disabled heads) are not in the bucket.
"""
import torch
import torch.distributed as dist


def shard_views(n_views, rank, world):
    """Indices of the batch's views owned by `rank` (view i -> rank i mod world)."""
    return [i for i in range(n_views) if i % world == rank]


class GradBucket:
    """One flat FP32 buffer holding the gradients of `params` plus named extra slots."""

    def __init__(self, params, extra=None, device=None):
        self.params = [p for p in params if p.requires_grad]
        if not self.params and not extra:
            raise ValueError("GradBucket needs at least one tensor")
        dev = device if device is not None else self.params[0].device
        sizes = [p.numel() for p in self.params]
        self.extra_shapes = dict(extra or {})
        for shape in self.extra_shapes.values():
            n = 1
            for s in shape:
                n *= int(s)
            sizes.append(n)
        # 64-float (256 B) aligned slots keep every view vector-load friendly
        self.offsets = []
        off = 0
        for n in sizes:
            self.offsets.append(off)
            off += (n + 63) // 64 * 64
        self.flat = torch.zeros(off, dtype=torch.float32, device=dev)
        self.views = []
        for p, o in zip(self.params, self.offsets):
            v = self.flat[o:o + p.numel()].view(p.shape)
            self.views.append(v)
        self.extra = {}
        for (name, shape), o in zip(self.extra_shapes.items(), self.offsets[len(self.params):]):
            n = 1
            for s in shape:
                n *= int(s)
            self.extra[name] = self.flat[o:o + n].view(tuple(shape))
        self.attach()

    def attach(self):
        """(Re)points every parameter's .grad at its bucket slot."""
        for p, v in zip(self.params, self.views):
            p.grad = v

    def zero(self):
        self.flat.zero_()
        # an optimizer's zero_grad(set_to_none=True) would have dropped the views
        for p, v in zip(self.params, self.views):
            if p.grad is None or p.grad.data_ptr() != v.data_ptr():
                p.grad = v

    def nbytes(self):
        return self.flat.numel() * 4

    def span(self, params):
        """(lo, hi) float offsets of the contiguous run of bucket slots holding `params`' gradients, or None if they are
        not adjacent slots in that order."""
        idx = []
        for q in params:
            k = [j for j, p in enumerate(self.params) if p is q]
            if not k:
                return None
            idx.append(k[0])
        if idx != list(range(idx[0], idx[0] + len(idx))):
            return None
        return self.offsets[idx[0]], self.offsets[idx[-1]] + self.params[idx[-1]].numel()

    def all_reduce_range(self, lo, hi, group=None, async_op=False):
        """Sum over ranks of flat[lo:hi] only (see train_iteration(overlap_sh=True))."""
        if hi <= lo:
            return None
        return dist.all_reduce(self.flat[lo:hi], op=dist.ReduceOp.SUM, group=group, async_op=async_op)

    def all_reduce(self, group=None, async_op=False):
        """Sum over ranks; a no-op outside an initialised process group or at world size 1."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)


def grads_in_place(module, on=True):
    """Lets the fused deformation backward accumulate parameter gradients straight into the existing .grad tensors
    (the bucket slots) instead of returning fresh tensors that autograd then adds - the kernels' epilogues write the
    buffer NCCL reduces.  Only valid while every .grad is owned by a GradBucket (never replaced, zeroed by bucket.zero())."""
    for m in module.modules():
        if hasattr(m, "grads_in_place"):
            m.grads_in_place = bool(on)
    return module


def _sm_limit(on, spare=None):
    """Caps the persistent fdgs deformation kernels at (SM count - spare) CTAs while an all-reduce is in flight."""
    try:
        from . import _lib
        lib = _lib.load()
    except Exception:
        return
    if on:
        import os
        spare = int(os.environ.get("FDGS_OVERLAP_SPARE_SMS", "8")) if spare is None else spare
        n = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count
        lib.fdgs_set_sm_limit(max(1, n - spare))
    else:
        lib.fdgs_set_sm_limit(0)


def l1_view_loss(image, gt, n_views_total):
    """This view's share of the reference's loss: l1_loss(cat(images), cat(gts)) over B equal-size views
    (utils/loss_utils.py l1_loss = mean|a-b|, train.py:231-244) == sum_i mean|img_i - gt_i| / B."""
    return (image - gt).abs().mean() / float(n_views_total)


class ViewStager:
    """Host -> device staging of a batch's ground-truth images for train_iteration (the reference's loop does
    `viewpoint_cam.original_image.cuda()` per view on the compute stream, train.py:222-226).

    The images are copied from pinned host memory into persistent device buffers on a separate copy stream, one event per
    view; `stager[i]` (what train_iteration's `gts[i]` evaluates) makes the current stream wait for view i's copy only, so
    the upload of view i+1 runs under the kernels of view i instead of in front of them."""

    def __init__(self, like, device):
        self.stream = torch.cuda.Stream(device=device)
        self.buf = [None if t is None else torch.empty(t.shape, dtype=t.dtype, device=device) for t in like]
        self.events = [None if t is None else torch.cuda.Event() for t in like]
        self.nbytes = 0

    def upload(self, host, views):
        """Starts this step's copies (views: the indices this rank renders, in render order).  Only the first view's copy is
        queued here; view k+1's is queued when view k's image is fetched, so that the step's first kernels are not held up
        behind the host work of issuing every copy.  Copies start once everything already queued on the current stream -
        i.e. the previous step's reads of the buffers - has finished."""
        self.stream.wait_stream(torch.cuda.current_stream())
        self.nbytes = 0
        self._host = host
        self._pending = list(views)
        self._issued = set()
        self._issue_next()
        return self

    def _issue_next(self):
        if not self._pending:
            return
        i = self._pending.pop(0)
        with torch.cuda.stream(self.stream):
            self.buf[i].copy_(self._host[i], non_blocking=True)
            self.events[i].record(self.stream)
        self._issued.add(i)
        self.nbytes += self._host[i].numel() * self._host[i].element_size()

    def __len__(self):
        return len(self.buf)

    def __getitem__(self, i):
        while i not in self._issued and self._pending:       # out-of-order access: issue up to and including view i
            self._issue_next()
        torch.cuda.current_stream().wait_event(self.events[i])
        self._issue_next()
        return self.buf[i]


class _SharedFeatures:
    """Stand-in for `pc` during one iteration: everything is `pc`'s, except that `get_features` hands out ONE
    [P,16,3] leaf for all of the rank's views instead of a fresh cat(_features_dc, _features_rest) per render()
    (scene/gaussian_model.py:123-127 cats on every access; the parameters cannot change inside an iteration)."""

    def __init__(self, pc, leaf):
        object.__setattr__(self, "_pc", pc)
        object.__setattr__(self, "_leaf", leaf)

    @property
    def get_features(self):
        return self._leaf

    def __getattr__(self, name):
        return getattr(self._pc, name)

    def __setattr__(self, name, value):
        setattr(self._pc, name, value)


def train_iteration(render_fn, pc, cams, gts, bg, bucket, rank=0, world=1, read_loss=False, loss_fn=None,
                    share_features=True, overlap_sh=False, reduce_stats=True, speculative=False):
    """One fwd+bwd iteration over the batch `cams` (all B views; this rank renders its shard).
    loss_fn(image, gt, B): per-view loss; default = the ATen l1_view_loss above (fdgs.losses.l1_view_loss is the fused one).
    share_features: concatenate the SH coefficients once per iteration, let the views' SH gradients sum in one [P,16,3]
    buffer and split it into _features_dc / _features_rest gradients once (same additions, in the same view order, as the
    reference's per-view cat backward; saves a cat and a strided two-way gradient accumulation per view).
    overlap_sh: the gradients of scale / rotation / opacity / SH (56 of 59 floats per Gaussian, 80 % of the bucket) are final
    as soon as the LAST view's rasterizer backward has run, before its deformation backward; their part of the
    all-reduce is started there (asynchronously, on the process group's stream) and the rest of the bucket follows at
    the end - the sums are the same, in three collectives instead of one.  The collective SCHEDULE depends only on
    rank-independent inputs (overlap_sh, the process group, the bucket layout): a rank that owns no view (world > B) issues
    the same three range all-reduces.  At world size 1 the same code path runs without the collectives.
    reduce_stats (world > 1): the densifier's inputs are made identical on every rank - `radii` is all-reduced with MAX
    (train.py:229 takes the max over ALL B views) and visibility is derived from it (train.py:230: any over views ==
    max radius > 0), so that replicas take the same split / prune decisions and keep the same P.
    speculative: render through the fdgs rasterizer's one-shot forward (no per-view host sync on num_rendered, see
    fdgs.diff_gaussian_rasterization.speculative_mode): the host runs ahead of the device over views and iterations.  The
    counts are examined where the step synchronises anyway - right after the loss read-back with read_loss=True; a step
    whose instance workspace turned out too small is then REDONE with the exact path before anything consumed its
    gradients.  Without read_loss the examination is non-blocking and a late finding raises (single process) - call
    fdgs.diff_gaussian_rasterization.check_pending() at your own synchronisation point to be sure earlier.
    Returns (local loss sum, radii max over all B views [P] int32, visibility [P] bool) - what train.py:229-232,346-349
    feed to max_radii2D / add_densification_stats."""
    if speculative:
        from . import diff_gaussian_rasterization as _rast
        with _rast.speculative_mode(True):
            out = train_iteration(render_fn, pc, cams, gts, bg, bucket, rank, world, read_loss, loss_fn, share_features,
                                  overlap_sh, reduce_stats, speculative=False)
        if not _rast.check_pending(block=bool(read_loss)):
            if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
                raise RuntimeError("fdgs: a one-shot render overflowed its instance workspace on this rank; the replicas "
                                   "cannot redo a step independently - rerun the iteration with speculative=False")
            if not read_loss:
                raise RuntimeError("fdgs: an earlier one-shot render overflowed its instance workspace (found late: "
                                   "read_loss=False); its gradients were incomplete")
            return train_iteration(render_fn, pc, cams, gts, bg, bucket, rank, world, read_loss, loss_fn, share_features,
                                   overlap_sh, reduce_stats, speculative=False)
        return out
    B = len(cams)
    bucket.zero()
    loss_sum = None
    radii_max = None
    mine = shard_views(B, rank, world)
    feats = leaf = None
    pc_it = pc
    has_feats = isinstance(getattr(type(pc), "get_features", None), property)
    if share_features and len(mine) > 1 and has_feats:
        feats = pc.get_features
        if feats.requires_grad:
            leaf = feats.detach().requires_grad_(True)
            pc_it = _SharedFeatures(pc, leaf)
    distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
    # ---- early part of the all-reduce (overlap_sh) ---------------------------------------------------------------------
    # The gradients of _scaling, _rotation, _opacity, _features_dc, _features_rest (56 of the 59 floats per Gaussian; adjacent
    # bucket slots in that order) pass through the deformation's identity paths unchanged, so they are final as soon as the
    # LAST view's rasterizer backward has run.  With an fdgs deformation (`identity_sink`) they are added into their bucket
    # slots right there, BEFORE that view's deformation-backward kernels are queued, and their all-reduce starts on the
    # process group's stream while those kernels run (which leave it a few SMs: fdgs_set_sm_limit).  Without the sink
    # (another deformation module) the same three collectives are issued after the loop: same sums, no overlap.
    span = None
    early_params = None
    if overlap_sh and has_feats and hasattr(pc, "_features_dc") and hasattr(pc, "_features_rest"):
        if all(hasattr(pc, n) for n in ("_scaling", "_rotation", "_opacity")):
            early_params = [pc._scaling, pc._rotation, pc._opacity, pc._features_dc, pc._features_rest]
            span = bucket.span(early_params)                        # rank-independent: same bucket layout everywhere
        if span is None:                                            # other slot order: the SH slots alone (2/3 of the bucket)
            early_params = [pc._features_dc, pc._features_rest]
            span = bucket.span(early_params)
    early = []          # [async work handle or None] once the early part of the all-reduce is in flight
    sink_mod = getattr(getattr(pc, "_deformation", None), "deformation_net", None) if span is not None else None
    sink_state = {"leaf_done": False}

    def _start_early():
        early.append(bucket.all_reduce_range(*span, async_op=True) if distributed else None)

    def _sink(g):
        """Called by the deformation backward of the rank's LAST view with the upstream gradients of its five outputs."""
        took = []
        with torch.no_grad():
            if len(early_params) == 5:
                for h, p in ((1, pc._scaling), (2, pc._rotation), (3, pc._opacity)):
                    if g[h] is not None and p.grad is not None:
                        p.grad.add_(g[h].view(p.grad.shape))
                        took.append(h)
            if g[4] is not None and pc._features_dc.grad is not None and pc._features_rest.grad is not None:
                n_dc = pc._features_dc.shape[1]
                tot = g[4] if (leaf is None or leaf.grad is None) else leaf.grad.add_(g[4])   # earlier views' sum + this view
                pc._features_dc.grad.add_(tot[:, :n_dc])
                pc._features_rest.grad.add_(tot[:, n_dc:])
                sink_state["leaf_done"] = True
                took.append(4)
        if 4 in took and (len(early_params) == 2 or len(took) == 4):
            _start_early()
            if distributed:
                _sm_limit(True)
        return took

    for i in mine:
        last = span is not None and i == mine[-1]
        if last and sink_mod is not None and hasattr(sink_mod, "identity_sink"):
            sink_mod.identity_sink = _sink
        try:
            out = render_fn(cams[i], pc_it, bg_color=bg, stage="fine")
            loss = (loss_fn or l1_view_loss)(out["render"], gts[i], B)
            loss.backward()
        finally:
            if last and sink_mod is not None and hasattr(sink_mod, "identity_sink"):
                sink_mod.identity_sink = None
                _sm_limit(False)
        vs = out["viewspace_points"].grad
        if vs is not None and "viewspace" in bucket.extra:
            bucket.extra["viewspace"].add_(vs)
        loss_sum = loss.detach() if loss_sum is None else loss_sum + loss.detach()
        radii_max = out["radii"] if radii_max is None else torch.maximum(radii_max, out["radii"])
    if leaf is not None and leaf.grad is not None and not sink_state["leaf_done"]:
        feats.backward(leaf.grad)               # the once-per-iteration cat's backward: split into _features_dc / _features_rest
    if span is not None:
        # every rank issues the same three collectives in the same order: [early span], [0, lo), [hi, end)
        if not early:
            _start_early()                      # no sink on this rank (it owns no view / foreign deformation module)
        bucket.all_reduce_range(0, span[0]) if distributed else None
        bucket.all_reduce_range(span[1], bucket.flat.numel()) if distributed else None
        if early[0] is not None:
            early[0].wait()
    else:
        bucket.all_reduce()
    if radii_max is None:           # this rank owns no view of the batch
        P = bucket.extra["viewspace"].shape[0] if "viewspace" in bucket.extra else pc.get_xyz.shape[0]
        radii_max = torch.zeros((P,), dtype=torch.int32, device=bucket.flat.device)
    if distributed and reduce_stats:
        radii_max = radii_max.contiguous()
        dist.all_reduce(radii_max, op=dist.ReduceOp.MAX)
    visibility = radii_max > 0
    if loss_sum is None:
        loss_sum = torch.zeros((), dtype=torch.float32, device=bucket.flat.device)
    if read_loss:
        return float(loss_sum), radii_max, visibility
    return loss_sum, radii_max, visibility
