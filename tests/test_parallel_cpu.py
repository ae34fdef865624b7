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
