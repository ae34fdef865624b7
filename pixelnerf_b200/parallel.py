"""Ray-sharded multi-GPU rendering: one process per GPU over ``torch.distributed``.

The reference's multi-GPU path is ``torch.nn.DataParallel(_RenderWrapper, gpus, dim=1)``
(src/render/nerf.py:367-371): per call it re-broadcasts every parameter and the cached latent,
scatters the rays along dim 1 and gathers the outputs on gpus[0].  Rays are independent given the
encoded scene, so here each rank keeps its weights resident, the encoder rank broadcasts the encoded
scene once per ``encode`` (``broadcast_encoded``) and every call only moves results: a contiguous
ray range per rank in, one all-gather of the rendered values out (16 B/ray for rgb+depth).
There is no collective inside the data path.  Training uses the same sharding: the gather is differentiable
(each rank back-propagates its own ray range) and ``allreduce_gradients`` sums the weight gradients.

The wrapper is backend agnostic (NCCL on the GPUs; the CPU tests drive the same code over gloo with
a stand-in render function).

Two forms, chosen by ``NeRFRenderer.bind_parallel(net, gpus)`` when ``len(gpus) > 1``:

* ``RaySharded``    -- one process per GPU under ``torchrun`` (the fast path: no GIL, no per-call replica check);
* ``DeviceSharded`` -- ONE process driving every GPU of ``gpus``, which is what the reference's scripts do
  (``renderer.bind_parallel(net, args.gpu_id)`` from a single process, eval/gen_video.py:110, train/train.py:93):
  a resident replica of the model per extra device (re-synchronised only when a parameter or the encoded scene
  changed -- DataParallel re-broadcasts all of it on every call), one host thread per device like DataParallel's
  ``parallel_apply``, and peer copies of the rays in / the rendered values out.
"""
import copy
import threading

import torch
import torch.distributed as dist


def ray_range(n_rays, world, rank):
    """Contiguous, balanced split of ``n_rays``: the first ``n_rays % world`` ranks get one extra ray."""
    base, extra = divmod(n_rays, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_draws(draws, n_rays_per_obj, sb, world, rank):
    """Slice injected random draws (dict of (SB*B, k) tensors) with the same ray ranges, so the
    sharded render consumes exactly the draws the single-GPU render would."""
    if draws is None:
        return None
    s, e = ray_range(n_rays_per_obj, world, rank)
    return {k: v.reshape(sb, n_rays_per_obj, *v.shape[1:])[:, s:e].reshape(sb * (e - s), *v.shape[1:]).contiguous()
            for k, v in draws.items()}


def _gather_many_dim1(ts, counts, group):
    """All-gather a LIST of tensors that differ between ranks only in dim 1 (ragged ray ranges) and concatenate
    each along dim 1 -- with ONE collective: every rank packs its tensors side by side into one
    (SB, width, features) buffer (width = the largest range: only the tail of the shorter ranges is padding) and a
    single ``all_gather_into_tensor`` fills the preallocated (world, SB, width, features) result."""
    world, width = len(counts), max(counts)
    if not ts:
        return []
    sb, mine = ts[0].shape[0], ts[0].shape[1]
    dtype = ts[0].dtype
    feats = [int(torch.Size(t.shape[2:]).numel()) for t in ts]      # features per ray (1 for (SB, rays) tensors)
    if any(t.dtype != dtype or t.shape[0] != sb or t.shape[1] != mine for t in ts):
        raise RuntimeError("ray-sharded gather: outputs of one call must share dtype and the (SB, rays) leading shape")
    total = sum(feats)
    send = ts[0].new_zeros(sb, width, total) if mine < width else ts[0].new_empty(sb, width, total)
    off = 0
    for t, f in zip(ts, feats):
        send[:, :mine, off:off + f] = t.reshape(sb, mine, f)
        off += f
    recv = send.new_empty(world * send.numel())     # flat: the concatenating form every backend accepts
    dist.all_gather_into_tensor(recv, send.reshape(-1), group=group)
    recv = recv.view(world, sb, width, total)
    if min(counts) == width:       # equal ranges: one strided copy
        full = recv.permute(1, 0, 2, 3).reshape(sb, world * width, total)
    else:
        full = torch.cat([recv[r, :, :c] for r, c in enumerate(counts)], dim=1)
    outs, off = [], 0
    for t, f in zip(ts, feats):
        outs.append(full[:, :, off:off + f].reshape(sb, full.shape[1], *t.shape[2:]).contiguous())
        off += f
    return outs


def _gather_dim1(t, counts, group):
    """All-gather one tensor that differs between ranks only in dim 1 and concatenate along it."""
    return _gather_many_dim1([t], counts, group)[0]


class _GatherDim1Grad(torch.autograd.Function):
    """_gather_dim1 with a backward: every rank evaluates the same loss on the same gathered tensor, so the
    gradient of this rank's slice is the matching slice of the upstream gradient.  The parameter gradients
    each rank then holds are its rays' share of the single-process gradient; ``allreduce_gradients`` sums them."""

    @staticmethod
    def forward(ctx, t, counts, rank, group):
        ctx.start, ctx.count = sum(counts[:rank]), counts[rank]
        return _gather_dim1(t, counts, group)

    @staticmethod
    def backward(ctx, g):
        return g[:, ctx.start:ctx.start + ctx.count].contiguous(), None, None, None


def allreduce_gradients(modules, group=None, latent=None, latent_src=0):
    """Training, SURVEY 8e: after ``loss.backward()`` on every rank (each over its own ray range of the same
    batch), sum the parameter gradients over the ranks -- one flat bucket per dtype (the two ResnetFC weight sets
    are 27.5 MB of fp32; the replicated encoder's gradients ride along when it is trained) and one all-reduce each.
    Parameters without a gradient on some rank (an empty ray range) contribute zeros.  If ``latent`` is a
    non-parameter tensor with a gradient (an encoder that lives on rank ``latent_src`` only), its gradient is
    reduced to that rank.  Returns the number of bytes all-reduced."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return 0
    if isinstance(modules, torch.nn.Module):
        modules = [modules]
    params = [q for m in modules for q in m.parameters() if q.requires_grad]
    nbytes = 0
    by_type = {}
    for q in params:
        by_type.setdefault((q.dtype, q.device), []).append(q)
    for (dtype, dev), qs in by_type.items():
        flat = torch.zeros(sum(q.numel() for q in qs), dtype=dtype, device=dev)
        off = 0
        for q in qs:
            if q.grad is not None:
                flat[off:off + q.numel()].copy_(q.grad.reshape(-1))
            off += q.numel()
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        nbytes += flat.numel() * flat.element_size()
        off = 0
        for q in qs:
            g = flat[off:off + q.numel()].view_as(q)
            if q.grad is None:
                q.grad = g.clone()
            else:
                q.grad.copy_(g)
            off += q.numel()
    if latent is not None and latent.grad is not None:
        dist.reduce(latent.grad, dst=latent_src, op=dist.ReduceOp.SUM, group=group)
        nbytes += latent.grad.numel() * latent.grad.element_size()
    return nbytes


def _map_tensors(obj, fn):
    if torch.is_tensor(obj):
        return fn(obj)
    if isinstance(obj, dict):
        return {k: _map_tensors(v, fn) for k, v in obj.items()}
    if isinstance(obj, (tuple, list)):
        return type(obj)(_map_tensors(v, fn) for v in obj)
    return obj


class RaySharded(torch.nn.Module):
    """Wraps a ``_RenderWrapper``-like module ``f(rays (SB,B,8), want_weights) -> tensors with a
    (SB, B, ...) leading shape`` (a tuple or a nested dict of them)."""

    def __init__(self, wrapped, group=None):
        super().__init__()
        self.wrapped = wrapped
        self.group = group

    def forward(self, rays, want_weights=False):
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError(
                "multi-GPU rendering runs one process per GPU: launch with torchrun (or init "
                "torch.distributed) and pass every rank the same rays; bind_parallel(net, [gpu]) "
                "with a single id renders on one device")
        world, rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        if world == 1:
            return self.wrapped(rays, want_weights)
        n = rays.shape[1]
        counts = [ray_range(n, world, r)[1] - ray_range(n, world, r)[0] for r in range(world)]
        s, e = ray_range(n, world, rank)
        renderer = getattr(self.wrapped, "renderer", None)
        if renderer is not None and getattr(renderer, "injected_draws", None) is not None:
            renderer.injected_draws = shard_draws(renderer.injected_draws, n, rays.shape[0], world, rank)
        if renderer is not None and hasattr(renderer, "seed_offset"):
            # every rank draws the same call seed from its identically seeded CPU generator; offsetting it by the
            # range's first ray gives every shard its own jitter (the chunk loop offsets by ray index the same way)
            renderer.seed_offset = s
        mine = rays[:, s:e]
        if not mine.is_cuda and torch.cuda.is_available() and dist.get_backend(self.group) == "nccl":
            # rays may stay in (pinned) host memory: only this rank's range crosses PCIe (1/world of the frame)
            dev = next((q.device for q in self.wrapped.parameters() if q.is_cuda), None)
            if dev is not None:
                mine = mine.to(dev, non_blocking=True)
        try:
            local = self.wrapped(mine.contiguous(), want_weights)
        finally:
            if renderer is not None and hasattr(renderer, "seed_offset"):
                renderer.seed_offset = 0
        flat = []
        _map_tensors(local, lambda t: flat.append(t))
        if any(t.requires_grad for t in flat):   # training: keep the autograd graph through the gather
            return _map_tensors(local, lambda t: _GatherDim1Grad.apply(t, counts, rank, self.group)
                                if t.requires_grad else _gather_dim1(t, counts, self.group))
        gathered = iter(_gather_many_dim1(flat, counts, self.group))
        return _map_tensors(local, lambda t: next(gathered))


def broadcast_encoded(net, src=0, group=None):
    """After ``net.encode(...)`` on rank ``src``: ship the encoded scene (latent, world-to-camera
    poses, focal, principal point, image shape, view count) to every rank and refresh the kernel-side
    camera block.  One broadcast per ``encode`` replaces DataParallel's per-call module replication."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    rank = dist.get_rank(group)
    dev = net.poses.device
    meta = torch.zeros(8, dtype=torch.int64, device=dev)
    if rank == src:
        lat = net.encoder.latent
        meta[:4] = torch.tensor(lat.shape, device=dev)
        meta[4], meta[5] = net.num_views_per_obj, net.num_objs
        meta[6], meta[7] = net.focal.shape[0], net.c.shape[0]
    dist.broadcast(meta, src, group=group)
    v, ch, hl, wl, ns, nobj, nf, nc = (int(x) for x in meta.tolist())
    if rank != src:
        net.encoder.latent = torch.empty(v, ch, hl, wl, device=dev)
        net.poses = torch.empty(v, 3, 4, device=dev)
        net.focal = torch.empty(nf, 2, device=dev)
        net.c = torch.empty(nc, 2, device=dev)
        net.encoder.latent_scaling = torch.empty(2, device=dev)
        net.num_views_per_obj, net.num_objs = ns, nobj
    for t in (net.encoder.latent, net.poses, net.focal, net.c, net.image_shape, net.encoder.latent_scaling):
        buf = t.detach().contiguous().float()
        dist.broadcast(buf, src, group=group)
        if rank != src:
            t.copy_(buf)
    net._refresh_camera()


# --------------------------------------------------------------------------------------------------------------
# One process, several devices (what ``bind_parallel(net, args.gpu_id)`` means in the reference's scripts)
# --------------------------------------------------------------------------------------------------------------
_RENDERER_ATTRS = ("n_coarse", "n_fine", "n_fine_depth", "using_fine", "noise_std", "depth_std", "white_bkgd",
                   "lindisp", "eval_batch_size", "sched", "max_rows_per_launch", "max_rows_train", "max_rows_train_tc",
                   "train_stash_bytes", "depth_sample_grad", "training")


def _without_caches(net):
    """Context: the model with its kernel-side caches (ctypes structs, packed weights, tables, a non-leaf latent)
    set aside, so that ``copy.deepcopy`` sees plain parameters and buffers."""
    class _Ctx:
        def __enter__(self):
            self.saved = {}
            for name in ("_cam", "_latent_nhwc", "_tables", "_train_stash", "_param_proxies"):
                if hasattr(net, name):
                    self.saved[name] = getattr(net, name)
            net._cam, net._latent_nhwc, net._tables = None, {}, {}
            self.packed = []
            for m in (getattr(net, "mlp_coarse", None), getattr(net, "mlp_fine", None)):
                if m is not None and hasattr(m, "_packed"):
                    self.packed.append((m, m._packed))
                    m._packed = {}
            enc = getattr(net, "encoder", None)
            self.latent = getattr(enc, "latent", None)
            self.latents = getattr(enc, "latents", None)
            if self.latent is not None:
                enc.latent = self.latent.detach()
            if self.latents is not None:
                enc.latents = None

        def __exit__(self, *exc):
            for name, v in self.saved.items():
                setattr(net, name, v)
            for m, pk in self.packed:
                m._packed = pk
            enc = getattr(net, "encoder", None)
            if self.latent is not None:
                enc.latent = self.latent
            if self.latents is not None:
                enc.latents = self.latents
    return _Ctx()


class DeviceSharded(torch.nn.Module):
    """``wrapped`` is the ``_RenderWrapper`` of the primary device ``gpus[0]`` (it owns ``net`` and ``renderer``);
    calls split the rays of dim 1 into ``len(gpus)`` contiguous ranges (src/render/nerf.py:367-371 scatters the
    same axis), render range ``i`` on ``gpus[i]`` and return everything on the primary device."""

    def __init__(self, wrapped, gpus):
        super().__init__()
        self.wrapped = wrapped
        self.gpus = [int(g) for g in gpus]
        self._replicas = None      # [(device, wrapper)] for gpus[1:], built on first use
        self._weights_key = None
        self._scene_key = None

    # ------------------------------------------------------------------ replicas
    def _build(self):
        from .render.nerf import _RenderWrapper
        net, renderer = self.wrapped.net, self.wrapped.renderer
        self._replicas = []
        for g in self.gpus[1:]:
            dev = torch.device("cuda", g)
            with _without_caches(net):
                rnet = copy.deepcopy(net)
            rnet = rnet.to(dev)
            rren = copy.deepcopy(renderer).to(dev)
            self._replicas.append((dev, _RenderWrapper(rnet, rren, self.wrapped.simple_output)))
        self._weights_key = self._scene_key = None

    @staticmethod
    def _mlps(net):
        return [m for m in (net.mlp_coarse, net.mlp_fine) if m is not None]

    def _sync(self, differentiable):
        """Bring the replicas up to date with the primary: parameters when any of them changed (an optimiser step,
        load_weights), the encoded scene when ``encode`` ran again, the renderer's mutable settings always."""
        net, renderer = self.wrapped.net, self.wrapped.renderer
        if self._replicas is None:
            self._build()
        params = [q for m in self._mlps(net) for q in m.parameters()]
        wkey = tuple((q.data_ptr(), q._version, q.requires_grad) for q in params) + (net.mlp_fine is None,)
        lat = net.encoder.latent
        skey = (lat.data_ptr(), lat._version, tuple(lat.shape), net.poses.data_ptr(), net.poses._version,
                net.focal.data_ptr(), net.c.data_ptr(), net.num_views_per_obj, net.num_objs)
        for dev, wrap in self._replicas:
            rnet, rren = wrap.net, wrap.renderer
            if wkey != self._weights_key:
                if net.mlp_fine is None:
                    rnet.mlp_fine = None
                with torch.no_grad():
                    for q, rq in zip(params, [x for m in self._mlps(rnet) for x in m.parameters()]):
                        rq.copy_(q, non_blocking=True)       # peer copy; bumps the version -> kernels repack
                        rq.requires_grad_(q.requires_grad)
            if skey != self._scene_key or differentiable or rnet.encoder.latent.requires_grad:
                src = lat if (differentiable and lat.requires_grad and not net.stop_encoder_grad) else lat.detach()
                rnet.encoder.latent = src.to(dev, non_blocking=True)
                rnet.encoder.latent_scaling = net.encoder.latent_scaling.to(dev)
                rnet.poses, rnet.focal, rnet.c = net.poses.to(dev), net.focal.to(dev), net.c.to(dev)
                rnet.image_shape = net.image_shape.to(dev)
                rnet.num_views_per_obj, rnet.num_objs = net.num_views_per_obj, net.num_objs
                rnet._refresh_camera()
            rnet.stop_encoder_grad = net.stop_encoder_grad
            rnet.mlp_precision = net.mlp_precision
            rnet.train(net.training)
            for a in _RENDERER_ATTRS:
                if hasattr(renderer, a):
                    setattr(rren, a, getattr(renderer, a))
            if renderer.sched is not None:
                rren.last_sched.copy_(renderer.last_sched)
                rren.iter_idx.copy_(renderer.iter_idx)
            wrap.simple_output = self.wrapped.simple_output
            # training: the replica computes with its own (value-synchronised) parameters, but autograd has to
            # carry the gradients home: differentiable peer copies of the primary's parameters stand in as the
            # render node's inputs (render/nerf.py reads ``_param_proxies``)
            rnet._param_proxies = None
            if differentiable:
                from . import _lib
                names = _lib.weight_names()
                prox = []
                for m in (net.mlp_coarse, net.mlp_fine):
                    if m is None:
                        prox.append([])
                        continue
                    wd = m.weight_dict()
                    prox.append([wd[k].to(dev) for k in names])
                rnet._param_proxies = tuple(prox)
        self._weights_key, self._scene_key = wkey, skey

    # ------------------------------------------------------------------- forward
    def forward(self, rays, want_weights=False):
        n_dev = len(self.gpus)
        if rays.dim() != 3 or rays.shape[1] < n_dev or not rays.is_cuda:
            return self.wrapped(rays, want_weights)     # tiny / empty / CPU input: the primary's own behaviour
        net, renderer = self.wrapped.net, self.wrapped.renderer
        primary = torch.device("cuda", self.gpus[0])
        if rays.device != primary:
            rays = rays.to(primary)
        grad = torch.is_grad_enabled()
        differentiable = grad and any(q.requires_grad for m in self._mlps(net) for q in m.parameters())
        self._sync(differentiable)
        sb, n = rays.shape[0], rays.shape[1]
        ranges = [ray_range(n, n_dev, i) for i in range(n_dev)]
        draws, forced = renderer.injected_draws, renderer.injected_z
        renderer.injected_draws = renderer.injected_z = None
        if forced:
            raise NotImplementedError("teacher-forced sample positions (a test hook) on the multi-device wrapper")
        seed = renderer._seed()
        wrappers = [(primary, self.wrapped)] + self._replicas
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(primary))
        results, errors = [None] * n_dev, [None] * n_dev

        def work(i):
            dev, wrap = wrappers[i]
            s, e = ranges[i]
            try:
                with torch.cuda.device(dev), torch.set_grad_enabled(grad):
                    torch.cuda.current_stream(dev).wait_event(ready)
                    part = rays[:, s:e].to(dev, non_blocking=True).contiguous()
                    ren = wrap.renderer
                    if draws is not None:
                        ren.injected_draws = {k: v.to(dev, non_blocking=True)
                                              for k, v in shard_draws(draws, n, sb, n_dev, i).items()}
                    ren.seed_override = seed + s     # one call seed, offset by the range's first ray
                    out = wrap(part, want_weights)
                    done = torch.cuda.Event()
                    done.record(torch.cuda.current_stream(dev))
                    results[i] = (out, done)
            except BaseException as exc:   # re-raised on the calling thread
                errors[i] = exc

        threads = [threading.Thread(target=work, args=(i,), daemon=True) for i in range(1, n_dev)]
        for t in threads:
            t.start()
        work(0)
        for t in threads:
            t.join()
        for exc in errors:
            if exc is not None:
                raise exc
        main = torch.cuda.current_stream(primary)
        homes = []
        for out, done in results:
            main.wait_event(done)
            homes.append(_map_tensors(out, lambda t: t if t.device == primary else t.to(primary, non_blocking=True)))
        return _cat_dim1(homes)


def _cat_dim1(parts):
    first = parts[0]
    if torch.is_tensor(first):
        return torch.cat(parts, dim=1)
    if isinstance(first, dict):
        return {k: _cat_dim1([p[k] for p in parts]) for k in first}
    if isinstance(first, (tuple, list)):
        return type(first)(_cat_dim1([p[i] for p in parts]) for i in range(len(first)))
    return first
