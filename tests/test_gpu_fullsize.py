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
    assert float((outs["fp16"].rgb - outs["fp32"].rgb).abs().max()) < 2e-3
    assert float((outs["fp16"].weights - outs["fp32"].weights).abs().max()) < 2e-3
    assert float((outs["fp16"].depth - outs["fp32"].depth).abs().max()) < 2e-3


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


@pytest.mark.parametrize("prec,tol", [("fp32", 1e-4), ("fp16x3", 1e-4), ("fp16x2", 2e-3), ("fp16", 2e-3)])
def test_real_size_dtu_slice_vs_live_reference(prec, tol):
    """BASELINE config 4 at its REAL size against the live reference (tests/golden/full_dtu_full_ns3.npz: 300 x 400
    images, 150 x 200 latent, three views, per-ray near / far, 2 048 rays = a contiguous run + a frame-wide
    lattice), stage-wise with teacher forcing, in two chunks whose boundary falls inside the ray range
    (``max_rows_per_launch`` splits at NS = 3): sample positions bit exact, rgb / weights / depth (world units) within
    north_star's tolerance of the path -- except fp16 depth on this scene class, asserted at its measured, stated
    value (helpers.TC_DEPTH_TOL_DTU; BASELINE.md "Parity"); the two-pass chain (fp16x2) meets 2e-3 here too."""
    import numpy as np
    from helpers import full_case_inputs, golden_files, load, max_abs
    from test_gpu_model import _net_for
    fx = load(golden_files("full_")[0])
    inp = full_case_inputs(fx)
    net, ren = _net_for(fx, inp, precision=prec)
    rays = inp["rays"].to(DEV)
    n = rays.shape[1]
    ren.injected_draws = {k: v.to(DEV) for k, v in inp["draws"].items()}
    ren.injected_z = dict(w_coarse=torch.from_numpy(fx["w_coarse"]).reshape(n, 64).to(DEV),
                          depth_coarse=torch.from_numpy(fx["depth_coarse"]).reshape(n).to(DEV))
    ren.max_rows_per_launch = 3 * 96 * 1300          # 1 300 rays per chunk: 2 048 rays cross one boundary
    with torch.no_grad():
        out = ren(net, rays, want_weights=True)
    z_c, z_f = ren.last_z
    assert np.array_equal(z_c[0, ::16].cpu().numpy(), fx["z_coarse_sub"])        # bit exact
    assert np.array_equal(z_f[0, ::4].cpu().numpy(), fx["z_fine_sub"])           # bit exact
    errs = dict(rgb_c=max_abs(out.coarse.rgb, fx["rgb_coarse"]), w_c=max_abs(out.coarse.weights, fx["w_coarse"]),
                depth_c=max_abs(out.coarse.depth, fx["depth_coarse"]), rgb_f=max_abs(out.fine.rgb, fx["rgb_fine"]),
                depth_f=max_abs(out.fine.depth, fx["depth_fine"]),
                w_f=max_abs(out.fine.weights[0, ::8], fx["w_fine_sub"]))
    if prec == "fp16":      # the one stated deviation: depth on DTU-class scenes (helpers.TC_DEPTH_TOL_DTU)
        from helpers import TC_DEPTH_TOL_DTU
        assert errs.pop("depth_c") < TC_DEPTH_TOL_DTU and errs.pop("depth_f") < TC_DEPTH_TOL_DTU, errs
    assert all(v < tol for v in errs.values()), errs
