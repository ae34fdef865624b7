"""BASELINE.json's configurations at their full sizes, through properties that do not need the CPU oracle
(which would take minutes to hours there): the render of a ray range does not depend on how the frame is
chunked or sharded (bit exact), the two MLP paths agree on the coarse pass to the reduced-precision
tolerance, and the outputs satisfy the compositing invariants.  Small-size parity against the oracle and
the golden vectors is in test_gpu_render.py / test_gpu_mlp_tc.py."""
import pytest
import torch

from pixelnerf_b200 import synthetic as syn
from pixelnerf_b200.testing import build_net

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _render(ren, net, rays, draws, max_rows=None):
    ren.injected_draws = {k: v for k, v in draws.items()}
    old = ren.max_rows_per_launch
    if max_rows is not None:
        ren.max_rows_per_launch = max_rows
    try:
        with torch.no_grad():
            return ren(net, rays[None], want_weights=True)
    finally:
        ren.max_rows_per_launch = old


def _scene(name, per_ray_bounds=False):
    sc = syn.scene(name)
    rays = sc["rays"].clone()
    if per_ray_bounds:   # near/far columns differ per ray
        g = torch.Generator().manual_seed(4)
        rays[:, 6] = 0.1 + torch.rand(rays.shape[0], generator=g) * 0.4
        rays[:, 7] = 5.0 - torch.rand(rays.shape[0], generator=g) * 0.5
    return sc, rays.to(DEV)


@pytest.mark.parametrize("name,per_ray", [("srn_car", False), ("srn_chair", False), ("dtu", True)])
def test_full_frame_chunk_and_shard_invariance(name, per_ray):
    """Config 2 (16 384 rays, NS=1), config 3 (one 128x128 frame of the video, NS=2), config 4 (120 000 rays,
    NS=3, per-ray bounds) on the tensor-core path: whole frame vs ragged chunks vs a ray sub-range."""
    sc, rays = _scene(name, per_ray)
    net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), torch.device(DEV), mlp_precision="fp16")
    n = rays.shape[0]
    d = {k: v.to(DEV) for k, v in syn.draws(3, n, 64, 16, 16).items()}
    whole = _render(ren, net, rays, d)
    rows_per_ray = 96 * sc["ns"]
    chunked = _render(ren, net, rays, d, max_rows=rows_per_ray * 4099)      # ragged: 4099 does not divide n
    for a, b in ((whole.fine.rgb, chunked.fine.rgb), (whole.fine.depth, chunked.fine.depth),
                 (whole.coarse.weights, chunked.coarse.weights), (whole.fine.weights, chunked.fine.weights)):
        assert torch.equal(a, b)
    lo, hi = n // 3 + 7, n // 3 + 7 + 5000                                   # what another rank would render
    part = _render(ren, net, rays[lo:hi], {k: v[lo:hi] for k, v in d.items()})
    assert torch.equal(part.fine.rgb[0], whole.fine.rgb[0, lo:hi])
    assert torch.equal(part.fine.depth[0], whole.fine.depth[0, lo:hi])
    # compositing invariants (src/render/nerf.py:228-249)
    for p in (whole.coarse, whole.fine):
        w = p.weights[0]
        assert torch.isfinite(p.rgb).all() and torch.isfinite(p.depth).all() and torch.isfinite(w).all()
        assert (w >= 0).all() and float(w.sum(-1).max()) <= 1.0 + 1e-4
        assert float(p.rgb.min()) >= -1e-5 and float(p.rgb.max()) <= 1.0 + 1e-4
        assert (p.depth[0] <= rays[:, 7] * (1 + 1e-5)).all()


def test_full_frame_fp16_vs_fp32_coarse_pass():
    """Config 2 at full size: the coarse pass does not depend on any sampling decision, so the tensor-core
    path must match the fp32 parity path (itself 1e-4 from the reference) to north_star's 2e-3 on every ray."""
    sc, rays = _scene("srn_car")
    wc, wf = syn.mlp_weights(1), syn.mlp_weights(2)
    d = {k: v.to(DEV) for k, v in syn.draws(5, rays.shape[0], 64, 16, 16).items()}
    outs = {}
    for prec in ("fp16", "fp32"):
        net, ren = build_net(sc, wc, wf, torch.device(DEV), mlp_precision=prec)
        outs[prec] = _render(ren, net, rays, d).coarse
    span = float((rays[:, 7] - rays[:, 6]).max())
    assert float((outs["fp16"].rgb - outs["fp32"].rgb).abs().max()) < 2e-3
    assert float((outs["fp16"].weights - outs["fp32"].weights).abs().max()) < 2e-3
    assert float((outs["fp16"].depth - outs["fp32"].depth).abs().max()) < 2e-3 * max(1.0, span)


def test_dense_point_query_chunks():
    """the reference's sigma-volume extraction calls ``net.forward`` on 65 536-point chunks of
    a regular grid (src/util/recon.py:12-78, train/train.py:412-425) -- same gather + MLP kernels, no sampling,
    no compositing.  A 64^3 grid: chunked = unchunked bit for bit, tensor-core path within north_star's
    reduced-precision tolerance of the fp32 path, densities non-negative and colours in [0, 1]."""
    sc = syn.scene("srn_car")
    wc, wf = syn.mlp_weights(1), syn.mlp_weights(2)
    t = torch.linspace(-0.6, 0.6, 64)
    xyz = torch.stack(torch.meshgrid(t, t, t, indexing="ij"), -1).reshape(1, -1, 3).to(DEV)
    dirs = torch.nn.functional.normalize(torch.randn(xyz.shape, generator=torch.Generator().manual_seed(2)), dim=-1).to(DEV)
    outs = {}
    for prec in ("fp16", "fp32"):
        net, _ = build_net(sc, wc, wf, torch.device(DEV), mlp_precision=prec)
        with torch.no_grad():
            whole = net(xyz, coarse=True, viewdirs=dirs)
            parts = torch.cat([net(xyz[:, s:s + 65536], coarse=True, viewdirs=dirs[:, s:s + 65536])
                               for s in range(0, xyz.shape[1], 65536)], dim=1)
        assert whole.shape == (1, 64 ** 3, 4) and torch.equal(whole, parts)
        assert float(whole[..., 3].min()) >= 0.0 and 0.0 <= float(whole[..., :3].min()) and float(whole[..., :3].max()) <= 1.0
        outs[prec] = whole
    assert float((outs["fp16"][..., :3] - outs["fp32"][..., :3]).abs().max()) < 2e-3
    sig = outs["fp32"][..., 3]
    assert float((outs["fp16"][..., 3] - sig).abs().max()) < 2e-3 * max(1.0, float(sig.max()))
