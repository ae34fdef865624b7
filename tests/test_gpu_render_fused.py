"""pn_render_rays (the whole no-grad render in one C call) against the staged path that calls the same kernels
one by one from Python: identical inputs through identical kernels, so every output must be bit-identical --
with injected draws, with in-kernel draws from a seed, NS = 1 and 2, several objects, coarse-only, mlp_fine = None
(reference eval/eval.py:58-59), an empty batch, and a bad-argument error through the C ABI."""
import ctypes as C

import pytest
import torch

from pixelnerf_b200 import _lib, synthetic as syn
from pixelnerf_b200.testing import build_net, scene_camera, set_encoded

pytestmark = pytest.mark.gpu


def _both(ren, net, rays, draws, monkeypatch, seed=None):
    outs = []
    for staged in ("0", "1"):
        monkeypatch.setenv("PIXELNERF_B200_STAGED", staged)
        ren.injected_draws = dict(draws) if draws else None
        if seed is not None:
            torch.manual_seed(seed)
        with torch.no_grad():
            o = ren(net, rays, want_weights=True)
        outs.append((o, ren.last_z))
    return outs


def _same(a, b, fine=True):
    (oa, za), (ob, zb) = a, b
    for part in ("coarse",) + (("fine",) if fine else ()):
        for k in ("rgb", "depth", "weights"):
            assert torch.equal(oa[part][k], ob[part][k]), (part, k)
    assert torch.equal(za[0], zb[0])
    if fine:
        assert torch.equal(za[1], zb[1])


@pytest.mark.parametrize("scene,n", [("sn64", 777), ("srn_chair", 500), ("dtu", 333)])
def test_fused_equals_staged_injected_draws(scene, n, monkeypatch):
    sc = syn.scene(scene)
    net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), torch.device("cuda"), mlp_precision="fp16")
    rays = sc["rays"][torch.linspace(0, sc["rays"].shape[0] - 1, n).long()].cuda()[None]
    d = {k: v.cuda() for k, v in syn.draws(5, n, 64, 16, 16).items()}
    a, b = _both(ren, net, rays, d, monkeypatch)
    _same(a, b)
    assert a[0].fine.rgb.shape == (1, n, 3) and a[0].fine.weights.shape == (1, n, 96)


def test_fused_equals_staged_seeded_and_chunked(monkeypatch):
    sc = syn.scene("sn64")
    net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), torch.device("cuda"), mlp_precision="fp16")
    rays = sc["rays"][:1000].cuda()[None]
    a, b = _both(ren, net, rays, None, monkeypatch, seed=11)
    _same(a, b)
    ren.max_rows_per_launch = 96 * 128   # 8 chunks, the last one ragged
    c, d = _both(ren, net, rays, None, monkeypatch, seed=11)
    _same(c, d)


def test_fused_multi_object_coarse_only_and_no_fine_mlp(monkeypatch):
    sc = syn.scene("sn64")
    dev = torch.device("cuda")
    net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), dev, mlp_precision="fp16")
    cam = scene_camera(sc)
    sb = 3
    lat = syn.latent_maps(4, sb, sc["hl"], sc["wl"])
    set_encoded(net, lat, cam["poses"].expand(sb, -1, -1).contiguous(), cam["focal"], cam["c"], cam["image_shape"], 1)
    rays = sc["rays"][torch.randint(0, 4096, (sb, 200), generator=torch.Generator().manual_seed(0))].cuda()
    d = {k: v.cuda() for k, v in syn.draws(6, sb * 200, 64, 16, 16).items()}
    a, b = _both(ren, net, rays, d, monkeypatch)
    _same(a, b)
    ren.max_rows_per_launch = 96 * sb * 64   # chunks cut along the per-object ray axis
    c, e = _both(ren, net, rays, d, monkeypatch)
    _same(c, e)
    _same(a, c)
    net.mlp_fine = None                      # eval/eval.py:58-59: the fine pass runs through mlp_coarse
    f, g = _both(ren, net, rays, d, monkeypatch)
    _same(f, g)
    assert not torch.equal(f[0].fine.rgb, a[0].fine.rgb)
    ren.n_fine, ren.using_fine = 0, False    # eval/eval.py:60: callers switch the fine pass off after construction
    h, i = _both(ren, net, rays, d, monkeypatch)
    _same(h, i, fine=False)
    assert len(h[0].fine) == 0


def test_fused_wrapper_empty_and_c_abi_errors():
    sc = syn.scene("sn64")
    net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), torch.device("cuda"), mlp_precision="fp16")
    par = ren.bind_parallel(net, [0], simple_output=True).eval()
    rgb, depth = par(torch.zeros(0, 8, device="cuda"))
    assert rgb.shape == (0, 3) and depth.shape == (0,)
    lib = _lib.lib()
    cfg = _lib.RenderConfig(64, 16, 16, 0.01, 0, 1, 0)
    assert lib.pn_render_rays_workspace_bytes(0, 1, C.byref(cfg)) % 256 == 0
    assert lib.pn_render_rays_workspace_bytes(100, 2, C.byref(cfg)) > lib.pn_render_rays_workspace_bytes(100, 1, C.byref(cfg))
    # undersized workspace -> PN_ERR_INVALID and a message, nothing launched
    with pytest.raises(RuntimeError, match="workspace too small"):
        real = lib.pn_render_rays_workspace_bytes

        class Small:
            def __call__(self, *a):
                return 256
        try:
            _lib._lib.pn_render_rays_workspace_bytes = Small()
            mc = (net.mlp_coarse.packed("fp16"), net.latent_tables(True))
            _lib.render_rays(net._cam, sc["hl"], sc["wl"], mc, None, sc["rays"][:64].cuda(), 64, 16, 16, 0.01, False, True)
        finally:
            _lib._lib.pn_render_rays_workspace_bytes = real


@pytest.mark.parametrize("prec,tol", [("fp32", 1e-4), ("fp16", 2e-3)])
def test_eval_override_64_coarse_128_fine(prec, tol, monkeypatch):
    """The sampling resolution the reference's eval scripts switch to after construction (eval/eval_approx.py:79-82,
    eval/gen_video.py:192-195): n_coarse = 64, n_fine = 128 (112 importance + 16 depth samples, K2 = 192: the 256-wide
    sort and the 32-lane composite).  Stage-wise against the oracle: fine sampling fed the oracle's coarse weights and
    depth must reproduce its 192 sorted positions bit for bit, both passes within the path's tolerance."""
    from pixelnerf_b200.testing import oracle_render
    sc = syn.scene("sn64")
    wc, wf = syn.mlp_weights(1), syn.mlp_weights(2)
    n = 300
    rays = sc["rays"][torch.linspace(0, 4095, n).long()]
    draws = syn.draws(8, n, 64, 112, 16)
    ref = oracle_render(sc, wc, wf, rays, draws, 64, 128, 16)
    net, ren = build_net(sc, wc, wf, torch.device("cuda"), mlp_precision=prec)
    ren.n_fine, ren.using_fine = 128, True
    d = {k: v.cuda() for k, v in draws.items()}
    ren.injected_draws = dict(d)
    ren.injected_z = dict(w_coarse=ref["w_coarse"].cuda(), depth_coarse=ref["depth_coarse"].cuda())
    with torch.no_grad():
        out = ren(net, rays[None].cuda(), want_weights=True)
    z_c, z_f = ren.last_z
    assert torch.equal(z_c[0].cpu(), ref["z_coarse"]) and torch.equal(z_f[0].cpu(), ref["z_fine"])
    assert out.fine.weights.shape == (1, n, 192)
    for part, key in (("coarse", "coarse"), ("fine", "fine")):
        assert (out[part].rgb[0].cpu() - ref[f"rgb_{key}"]).abs().max().item() < tol
        assert (out[part].weights[0].cpu() - ref[f"w_{key}"]).abs().max().item() < tol
        assert (out[part].depth[0].cpu() - ref[f"depth_{key}"]).abs().max().item() < tol    # world units, absolute
    if prec == "fp16":   # and the one-call path equals the staged one at this resolution too
        a, b = _both(ren, net, rays[None].cuda(), d, monkeypatch)
        _same(a, b)
