"""Out-of-bounds hardening of the kernels with ragged tails (compute-sanitizer is not available on this pool): every
output lives in the middle of a larger sentinel-filled allocation and the margins must come back untouched, at row /
sample counts that leave partial warps, partial float4s and partial tiles."""
import ctypes as C

import pytest
import torch

from pixelnerf_b200 import _lib, synthetic as syn
from pixelnerf_b200.testing import build_net

pytestmark = pytest.mark.gpu
PAD = 4096   # elements on each side
SENT = -12345.0


class Guarded:
    """A tensor of `shape` carved out of the middle of a sentinel-filled buffer."""

    def __init__(self, shape, dtype=torch.float32):
        n = 1
        for s in shape:
            n *= s
        self.buf = torch.full((n + 2 * PAD,), SENT, device="cuda", dtype=torch.float32).to(dtype)
        self.n = n
        self.t = self.buf[PAD:PAD + n].view(*shape)

    def ptr(self):
        return self.t.data_ptr()

    def intact(self):
        ref = torch.tensor(SENT, dtype=torch.float32).to(self.buf.dtype).item()
        return bool((self.buf[:PAD] == ref).all()) and bool((self.buf[PAD + self.n:] == ref).all())


def _rays(r, seed=0):
    g = torch.Generator().manual_seed(seed)
    rays = torch.zeros(r, 8)
    rays[:, 3:6] = torch.nn.functional.normalize(torch.randn(r, 3, generator=g), dim=-1)
    rays[:, 2] = 1.3
    rays[:, 6] = 0.8
    rays[:, 7] = 1.8
    return rays.cuda()


@pytest.mark.parametrize("r,k", [(1, 1), (5, 33), (13, 64), (7, 96), (3, 129), (9, 250), (2, 512)])
def test_composite_writes_stay_inside(r, k):
    lib = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    rays = _rays(r)
    z = torch.sort(0.8 + torch.rand(r, k, device="cuda"), -1)[0].contiguous()
    rs = torch.rand(r, k, 4, device="cuda")
    w, rgb, d = Guarded((r, k)), Guarded((r, 3)), Guarded((r,))
    _lib._check(lib.pn_composite_fwd(rs.data_ptr(), z.data_ptr(), rays.data_ptr(), w.ptr(), rgb.ptr(), d.ptr(), r, k, 1,
                                     st), "fwd")
    g_rgb = torch.rand(r, 3, device="cuda")
    g_rs, g_z = Guarded((r, k, 4)), Guarded((r, k))
    _lib._check(lib.pn_composite_bwd(rs.data_ptr(), z.data_ptr(), rays.data_ptr(), g_rgb.data_ptr(), None, None,
                                     g_rs.ptr(), g_z.ptr(), r, k, 1, st), "bwd")
    torch.cuda.synchronize()
    for gd in (w, rgb, d, g_rs, g_z):
        assert gd.intact()
        assert torch.isfinite(gd.t).all() and not (gd.t == SENT).any()    # and every element was written
    w2, rgb2, d2 = _lib.composite_fwd(rs, z, rays, True)
    assert torch.equal(w2, w.t) and torch.equal(rgb2, rgb.t) and torch.equal(d2, d.t)


@pytest.mark.parametrize("r,kc,kf,kd", [(1, 64, 16, 16), (5, 64, 16, 16), (3, 48, 7, 3), (4, 128, 16, 16),
                                        (2, 64, 112, 16), (3, 32, 0, 5), (6, 96, 20, 12)])
def test_sampling_writes_stay_inside(r, kc, kf, kd):
    lib = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    rays = _rays(r, 1)
    lin = torch.linspace(0, 1 - 1.0 / kc, kc).cuda()
    zc = Guarded((r, kc))
    _lib._check(lib.pn_sample_coarse(rays.data_ptr(), None, lin.data_ptr(), zc.ptr(), r, kc, 0, 5, st), "coarse")
    wts = torch.rand(r, kc, device="cuda")
    dep = torch.full((r,), 1.2, device="cuda")
    k2 = kc + kf + kd
    zs, zi, zd = Guarded((r, k2)), Guarded((r, max(kf, 1))), Guarded((r, max(kd, 1)))
    _lib._check(lib.pn_sample_fine(rays.data_ptr(), zc.ptr(), wts.data_ptr(), dep.data_ptr(), None, None, None, zs.ptr(),
                                   zi.ptr() if kf else None, zd.ptr() if kd else None, r, kc, kf, kd, 0.01, 0, 9, st),
                "fine")
    torch.cuda.synchronize()
    assert zc.intact() and zs.intact() and zi.intact() and zd.intact()
    out = zs.t
    assert not (out == SENT).any() and bool((out[:, 1:] >= out[:, :-1]).all())
    # the merged row is exactly the multiset {coarse, importance, depth}
    parts = [zc.t] + ([zi.t[:, :kf]] if kf else []) + ([zd.t[:, :kd]] if kd else [])
    assert torch.equal(out, torch.sort(torch.cat(parts, -1), -1)[0])


@pytest.mark.parametrize("n_rays", [1, 3, 11, 345])
def test_point_taps_and_features_writes_stay_inside(n_rays):
    sc = syn.scene("sn64")
    net, _ = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), torch.device("cuda"), mlp_precision="fp16")
    lib = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    k = 96
    rays = sc["rays"][:n_rays].cuda().contiguous()
    z = torch.sort(1.2 + 2.8 * torch.rand(n_rays, k, device="cuda"), -1)[0].contiguous()
    cb = net._cam
    lat = net.latent_nhwc(torch.float16)
    cam = _lib.Camera(cb["poses"].data_ptr(), cb["focal"].data_ptr(), cb["c"].data_ptr(), cb["scale_x"], cb["scale_y"],
                      lat.shape[1], lat.shape[2], cb["n_objects"], cb["n_views"])
    rows = n_rays * k
    fx = Guarded((rows, 64), torch.float16)
    _lib._check(lib.pn_point_taps(C.byref(cam), rays.data_ptr(), z.data_ptr(), k, None, None, rows, fx.ptr(), st), "taps")
    fz, fx2 = Guarded((rows, 512), torch.float16), Guarded((rows, 64), torch.float16)
    _lib._check(lib.pn_point_features(C.byref(cam), lat.data_ptr(), 1, rays.data_ptr(), z.data_ptr(), k, None, None, rows,
                                      fz.ptr(), fx2.ptr(), 64, 1, st), "features")
    torch.cuda.synchronize()
    assert fx.intact() and fz.intact() and fx2.intact()
    assert torch.isfinite(fz.t.float()).all() and not (fz.t.float() == SENT).any()
    assert torch.equal(fx.t[:, :48], fx2.t[:, :48])      # same positional code from both kernels


@pytest.mark.parametrize("rows", [1, 100, 257, 1000])
def test_tma_store_epilogues_stay_inside(rows):
    """The TMA-store / TMA-load epilogues of the rows GEMM (dgrad forms) at ragged row counts: the bf16 output box
    of the last tile is clipped by the tensor map, in place and out of place."""
    lib = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    g = torch.Generator().manual_seed(rows)
    a = torch.randn(rows, 512, generator=g).bfloat16().cuda()
    w = (torch.randn(512, 512, generator=g) * 0.05).bfloat16().cuda()
    mask = torch.relu(torch.randn(rows, 512, generator=g)).bfloat16().cuda()
    out = Guarded((rows, 512), torch.bfloat16)
    _lib._check(lib.pn_gemm_tc_rows(a.data_ptr(), 1, w.data_ptr(), 1, rows, 512, 512, mask.data_ptr(), None, 0, None,
                                    out.ptr(), 1.0, None, None, None, st), "fc_1 form")
    res = Guarded((rows, 512), torch.bfloat16)
    r0 = torch.randn(rows, 512, generator=g).bfloat16().cuda()
    res.t.copy_(r0)
    _lib._check(lib.pn_gemm_tc_rows(a.data_ptr(), 1, w.data_ptr(), 1, rows, 512, 512, mask.data_ptr(), None, 0, res.ptr(),
                                    res.ptr(), 1.0, None, None, None, st), "fc_0 form, in place")
    torch.cuda.synchronize()
    assert out.intact() and res.intact()
    ref = (a.double() @ w.double()) * (mask.double() > 0)
    scale = max(1.0, ref.abs().max().item())
    assert (out.t.double() - ref).abs().max().item() < 2 ** -7 * scale
    assert (res.t.double() - (ref + r0.double())).abs().max().item() < 2 ** -7 * max(scale, (ref + r0.double()).abs().max().item())


@pytest.mark.parametrize("ns,ppo", [(1, 77), (2, 130)])
def test_training_forward_saved_operands_stay_inside(ns, ppo):
    """pn_mlp_forward_tc_train: the 11 saved-operand slots are written by TMA stores with a tile-padded pitch; the
    buffer's margins and the other outputs' margins stay untouched at ragged row counts."""
    from helpers import make_taps
    from test_gpu_mlp_tc import _inputs, _tables
    lib = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    w = syn.mlp_weights(42)
    packed = _lib.pack_weights_tc({k: v.cuda() for k, v in w.items()})
    n = ns * ppo
    idx, wt, latent, fx = _inputs(n)
    tables = _tables(w, latent)
    fx_t = make_taps(idx, wt, fx).cuda()
    abytes = lib.pn_mlp_tc_acts_bytes(n)
    acts = Guarded((abytes // 4,))
    out = Guarded((ppo, 4))
    nbytes = lib.pn_mlp_tc_workspace_bytes(n, ns)
    ws = torch.empty(nbytes, device="cuda", dtype=torch.uint8)
    assert acts.ptr() % 128 == 0
    _lib._check(lib.pn_mlp_forward_tc_train(packed.data_ptr(), tables.data_ptr(), tables.shape[1], fx_t.data_ptr(), 1, ns,
                                            ppo, out.ptr(), None, ws.data_ptr(), nbytes, acts.ptr(), abytes, st), "train fwd")
    torch.cuda.synchronize()
    assert acts.intact() and out.intact()
    ref = _lib.mlp_forward_tc(packed, tables, fx_t, 1, ns, ppo)
    assert torch.equal(ref, out.t)
    a = _lib.tc_acts_view(acts.t.view(torch.uint8), n)
    assert torch.isfinite(a[:6].float()).all() and float(a[0].float().abs().max()) > 0


def test_dgrad_epilogue_pipeline_is_deterministic():
    """The TMA-input pipeline of the rows GEMM (two buffers per warp, requests two chunks ahead across items, stores
    through one staging buffer): 25 runs over 300 tiles (four items per CTA pair and more) give bit-identical results,
    in place and out of place -- a phase or buffer-reuse error would show as a run-to-run difference."""
    lib = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    rows = 76800 + 37
    g = torch.Generator().manual_seed(1)
    a = torch.randn(rows, 512, generator=g).bfloat16().cuda()
    w = (torch.randn(512, 512, generator=g) * 0.05).bfloat16().cuda()
    mask = torch.relu(torch.randn(rows, 512, generator=g)).bfloat16().cuda()
    r0 = torch.randn(rows, 512, generator=g).bfloat16().cuda()
    cs = torch.zeros(512, device="cuda")
    first = None
    for rep in range(25):
        out = torch.empty(rows, 512, device="cuda", dtype=torch.bfloat16)
        res = r0.clone()
        _lib._check(lib.pn_gemm_tc_rows(a.data_ptr(), 1, w.data_ptr(), 1, rows, 512, 512, mask.data_ptr(), None, 0, None,
                                        out.data_ptr(), 1.0, cs.data_ptr(), None, None, st), "fc_1 form")
        _lib._check(lib.pn_gemm_tc_rows(out.data_ptr(), 1, w.data_ptr(), 1, rows, 512, 512, mask.data_ptr(), None, 0,
                                        res.data_ptr(), res.data_ptr(), 1.0, cs.data_ptr(), cs.data_ptr(), None, st),
                    "fc_0 form")
        if first is None:
            first = (out.clone(), res.clone())
        else:
            assert torch.equal(out, first[0]) and torch.equal(res, first[1]), rep


@pytest.mark.parametrize("ns,ppo", [(1, 128 * 160 + 5), (2, 128 * 90 + 3)])
def test_training_forward_is_deterministic(ns, ppo):
    """The chain kernel that saves its operands (staging buffer shared between the table gather and the TMA stores)
    and the two-launch NS > 1 chain (fp16 hand-off, L2 prefetch): 12 runs over more tiles than CTA pairs give
    bit-identical outputs and saved operands."""
    from helpers import make_taps
    from test_gpu_mlp_tc import _inputs, _tables
    w = syn.mlp_weights(42)
    packed = _lib.pack_weights_tc({k: v.cuda() for k, v in w.items()})
    n = ns * ppo
    idx, wt, latent, fx = _inputs(n)
    tables = _tables(w, latent)
    fx_t = make_taps(idx, wt, fx).cuda()
    first = None
    for rep in range(12):
        out, acts = _lib.mlp_forward_tc_train(packed, tables, fx_t, 1, ns, ppo)
        a = _lib.tc_acts_view(acts, n)
        got = (out.clone(), a[:6].clone(), a[6:, :ppo].clone())
        if first is None:
            first = got
        else:
            for x, y in zip(got, first):
                assert torch.equal(x, y), rep
    assert torch.equal(first[0], _lib.mlp_forward_tc(packed, tables, fx_t, 1, ns, ppo))
