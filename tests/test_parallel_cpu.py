"""Multi-process logic of the ray-sharded renderer on CPU (gloo, world_size 2 and 3): ray ranges,
draw slicing, ragged all-gather of tuple and nested-dict outputs, and the encoded-scene broadcast.
The render function is a deterministic stand-in (the kernels need a GPU); what is tested is that the
N-rank result equals the 1-rank result bit for bit."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pixelnerf_b200 import parallel


class _FakeRenderer:
    injected_draws = None
    using_fine = True


class _FakeWrapper(torch.nn.Module):
    """rays (SB,B,8) -> (rgb (SB,B,3), depth (SB,B)) or a nested dict; mixes in the injected draws."""

    def __init__(self, as_dict):
        super().__init__()
        self.renderer = _FakeRenderer()
        self.as_dict = as_dict

    def forward(self, rays, want_weights=False):
        sb, b, _ = rays.shape
        d = self.renderer.injected_draws
        self.renderer.injected_draws = None
        u = d["u"].reshape(sb, b, 4) if d is not None else torch.zeros(sb, b, 4)
        rgb = rays[..., :3] * 2.0 + u[..., :3]
        depth = rays[..., 6] + rays[..., 7] + u[..., 3]
        if not self.as_dict:
            return rgb, depth
        out = {"coarse": {"rgb": rgb, "depth": depth}, "fine": {"rgb": rgb + 1, "depth": depth + 1}}
        if want_weights:
            out["coarse"]["weights"] = u.repeat(1, 1, 2)
        return out


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_rays, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)
        sb = 2
        rays = torch.randn(sb, n_rays, 8, generator=g)
        draws = {"u": torch.rand(sb * n_rays, 4, generator=g)}
        res = {}
        for as_dict in (False, True):
            single = _FakeWrapper(as_dict)
            single.renderer.injected_draws = dict(draws)
            ref = single(rays, want_weights=True)
            w = _FakeWrapper(as_dict)
            w.renderer.injected_draws = dict(draws)
            out = parallel.RaySharded(w)(rays, want_weights=True)
            flat = lambda o: [o[0], o[1]] if isinstance(o, tuple) else [o["coarse"]["rgb"], o["coarse"]["depth"],
                                                                        o["coarse"]["weights"], o["fine"]["rgb"]]
            res[as_dict] = all(torch.equal(a, b) for a, b in zip(flat(out), flat(ref)))
        q.put((rank, res[False], res[True]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_rays", [(2, 10), (2, 7), (3, 8), (3, 2)])
def test_ray_sharded_equals_single(world, n_rays):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_rays, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    got = sorted(q.get(timeout=10) for _ in range(world))
    assert got == [(r, True, True) for r in range(world)]


def test_ray_range_partition():
    for n in (0, 1, 5, 16384, 120000):
        for world in (1, 2, 3, 4, 8):
            edges = [parallel.ray_range(n, world, r) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            sizes = [e - s for s, e in edges]
            assert max(sizes) - min(sizes) <= 1


def test_unsharded_call_without_process_group_raises():
    with pytest.raises(RuntimeError, match="one process per GPU"):
        parallel.RaySharded(_FakeWrapper(False))(torch.zeros(1, 4, 8))


class _DiffWrapper(torch.nn.Module):
    """Differentiable stand-in renderer with parameters: rgb = tanh(rays W + b), depth = |rays| . v."""

    def __init__(self):
        super().__init__()
        g = torch.Generator().manual_seed(3)
        self.lin = torch.nn.Linear(8, 3)
        self.v = torch.nn.Parameter(torch.randn(8, generator=g))
        self.unused = torch.nn.Parameter(torch.zeros(2))   # never receives a gradient
        with torch.no_grad():
            self.lin.weight.copy_(torch.randn(3, 8, generator=g))
            self.lin.bias.copy_(torch.randn(3, generator=g))
        self.renderer = _FakeRenderer()

    def forward(self, rays, want_weights=False):
        return {"coarse": {"rgb": torch.tanh(self.lin(rays)), "depth": (rays.abs() * self.v).sum(-1)}}


def _grad_worker(rank, world, port, n_rays, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(1)
        rays = torch.randn(2, n_rays, 8, generator=g).double()
        target = torch.rand(2, n_rays, 3, generator=g).double()

        def loss_of(out):
            return ((out["coarse"]["rgb"] - target) ** 2).mean() + out["coarse"]["depth"].mean()

        single = _DiffWrapper().double()
        loss_of(single(rays)).backward()
        ref = {n: p.grad.clone() for n, p in single.named_parameters() if p.grad is not None}
        w = _DiffWrapper().double()
        sharded = parallel.RaySharded(w)
        loss = loss_of(sharded(rays))
        loss.backward()
        local_only = {n: p.grad.clone() for n, p in w.named_parameters() if p.grad is not None}
        nbytes = parallel.allreduce_gradients(w)
        ok = all(torch.allclose(dict(w.named_parameters())[n].grad, gr, rtol=1e-12, atol=1e-14) for n, gr in ref.items())
        # before the all-reduce a rank holds only its rays' share
        partial = any(not torch.allclose(local_only[n], ref[n]) for n in ref)
        zero_unused = float(w.unused.grad.abs().max()) == 0.0
        q.put((rank, bool(ok), bool(partial or world == 1), zero_unused, nbytes == 8 * (24 + 3 + 8 + 2)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_rays", [(2, 9), (3, 2)])
def test_sharded_training_gradients_equal_single(world, n_rays):
    """SURVEY 8e training: differentiable gather + allreduce_gradients == single-process gradient (an empty ray
    range on one rank included: world 3, 2 rays)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_grad_worker, args=(r, world, port, n_rays, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    got = sorted(q.get(timeout=10) for _ in range(world))
    assert got == [(r, True, True, True, True) for r in range(world)]
