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
"""
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


def _gather_dim1(t, counts, group):
    """All-gather tensors that differ only in dim 1 (ragged ray ranges) and concatenate along it."""
    world = len(counts)
    width = max(counts)
    pad = t.new_zeros(t.shape[0], width, *t.shape[2:])
    pad[:, : t.shape[1]] = t
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad.contiguous(), group=group)
    return torch.cat([o[:, :c] for o, c in zip(out, counts)], dim=1)


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
        local = self.wrapped(rays[:, s:e].contiguous(), want_weights)

        def gather(t):
            if t.requires_grad:   # training: keep the autograd graph through the gather
                return _GatherDim1Grad.apply(t, counts, rank, self.group)
            return _gather_dim1(t, counts, self.group)
        return _map_tensors(local, gather)


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
