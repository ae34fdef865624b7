"""Tensor-core (tcgen05) ResnetFC chain on the GPU.

Two oracles:
  * the float64 emulation of the chain's own arithmetic (fp16-rounded GEMM operands, exact sums,
    fp32-exact residual and biases; helpers.emulate_tc_chain) -- checks the kernel does what it
    claims, layer by layer, tightly;
  * the reference's fp32 path (golden vectors of the live reference / the oracle) -- checks the
    fp16-operand path against north_star's reduced-precision tolerance: 2e-3 max-abs on
    rgb / weights / depth (depth in world units, absolute).
"""
import numpy as np
import pytest
import torch

import oracle
from oracle import model as omodel
from helpers import emulate_tc_chain, golden_files, grouped_row_index, load, make_taps, render_case_inputs, tc_depth_tol, to_grouped
from test_gpu_model import _net_for

pytestmark = pytest.mark.gpu

TC_TOL = 2e-3  # north_star: "2e-3 in the bf16/TF32 MLP path"


N_TEX = 8 * 24


def _inputs(n, seed=0):
    """A (1,8,24,512) channels-last latent, n rows of random taps (convex weights) and positional codes."""
    g = torch.Generator().manual_seed(seed)
    latent = torch.rand(N_TEX, 512, generator=g)
    idx = torch.randint(0, N_TEX, (n, 4), generator=g)
    wt = torch.rand(n, 4, generator=g)
    wt = wt / wt.sum(1, keepdim=True)
    fx = torch.zeros(n, 64)
    fx[:, :42] = torch.randn(n, 42, generator=g)
    return idx, wt, latent, fx.half()


def _xlo_rows(code, fx16):
    """The second operand row of the precise chains: x_lo = fp16(x - fp16(x)) in columns [0, 42) of a (n, 64) fp16 row."""
    lo = torch.zeros(code.shape[0], 64, dtype=torch.float16)
    lo[:, :42] = (code[:, :42] - fx16[:, :42].float()).half()
    return lo


def _tables(w, latent):
    from pixelnerf_b200 import _lib
    return _lib.prepare_latent_tables({k: v.cuda() for k, v in w.items()}, latent.reshape(1, 8, 24, 512).cuda())


@pytest.mark.parametrize("ns,ppo,sb", [(1, 1, 1), (1, 300, 1), (2, 130, 2), (3, 77, 1), (1, 128 * 150 + 5, 1)])
def test_chain_vs_emulation(ns, ppo, sb):
    """Ragged row counts, more tiles than CTA pairs, NS = 1..3: final pre-activations and outputs."""
    from pixelnerf_b200 import _lib, synthetic as syn
    w = syn.mlp_weights(42)
    packed = _lib.pack_weights_tc({k: v.cuda() for k, v in w.items()})
    n = sb * ns * ppo
    idx, wt, latent, fx = _inputs(n)
    out, pre = _lib.mlp_forward_tc(packed, _tables(w, latent), make_taps(idx, wt, fx).cuda(), sb, ns, ppo,
                                   want_pre=True)
    torch.cuda.synchronize()
    # emulate object by object (the view mean is per object)
    refs = []
    for o in range(sb):
        sl = slice(o * ns * ppo, (o + 1) * ns * ppo)
        refs.append(emulate_tc_chain(w, idx[sl], wt[sl], latent, fx[sl].float(), ns, ppo)[1])
    ref = torch.cat(refs)
    # the kernel accumulates in fp32 and rounds operands like the emulation; what is left are rare fp16
    # rounding-boundary flips of single activations / table entries (one ulp = 2^-7 at the hidden stream's
    # magnitude of ~8, whatever the size of the four outputs), hence the floor on the scale
    scale = max(8.0, ref.abs().max().item())
    assert (pre.cpu().double() - ref).abs().max().item() < 1.5e-3 * scale
    assert (out.cpu()[:, :3].double() - torch.sigmoid(ref[:, :3])).abs().max().item() < 1e-3


@pytest.mark.parametrize("stage", [0, 2, 5])
def test_residual_stream_stages(stage):
    """h as held in tensor memory (plus cumulative bias) at the start, middle and end of the chain."""
    from pixelnerf_b200 import _lib, synthetic as syn
    w = syn.mlp_weights(7)
    packed = _lib.pack_weights_tc({k: v.cuda() for k, v in w.items()})
    idx, wt, latent, fx = _inputs(200, seed=3)
    _, dbg = _lib.mlp_forward_tc(packed, _tables(w, latent), make_taps(idx, wt, fx).cuda(), 1, 1, 200,
                                 debug_stage=stage)
    ref = emulate_tc_chain(w, idx, wt, latent, fx.float(), 1, 200)[0][stage]
    # stage 0 has no rounded activations upstream: what is left are fp16 rounding-boundary flips of single
    # table entries (the tables come from an fp32 GEMM, the emulation from float64), one ulp = 2^-9 at |v| in [2,4)
    tol = 2.5e-3 if stage == 0 else 5e-3
    assert (dbg.cpu().double() - ref).abs().max().item() < tol


@pytest.mark.parametrize("path", golden_files("render_"))
def test_net_forward_fp16_vs_golden(path):
    fx = load(path)
    inp = render_case_inputs(fx)
    net, _ = _net_for(fx, inp, precision="fp16")
    xyz, dirs = torch.from_numpy(fx["probe_xyz"]).cuda(), torch.from_numpy(fx["probe_dirs"]).cuda()
    with torch.no_grad():
        out_c = net(xyz, coarse=True, viewdirs=dirs).cpu().numpy()
        out_f = net(xyz, coarse=False, viewdirs=dirs).cpu().numpy()
    for out, ref in ((out_c, fx["probe_coarse"]), (out_f, fx["probe_fine"])):
        assert np.abs(out[..., :3] - ref[..., :3]).max() < TC_TOL
        assert np.abs(out[..., 3] - ref[..., 3]).max() < TC_TOL * max(1.0, np.abs(ref[..., 3]).max())


@pytest.mark.parametrize("prec", ["fp16", "fp16x2"])
@pytest.mark.parametrize("path", golden_files("render_"))
def test_render_stagewise_fp16(path, prec):
    """Teacher-forced stage-wise render (see test_gpu_render.py) with the tensor-core MLP: the fast chain (fp16) and
    the two-pass chain (fp16x2, which meets north_star's 2e-3 on depth on every scene class, DTU included)."""
    fx = load(path)
    inp = render_case_inputs(fx)
    sb, ns = inp["sb"], inp["ns"]
    net, ren = _net_for(fx, inp, precision=prec)
    rays = inp["rays"]
    draws = {k: v.cuda() for k, v in inp["draws"].items()}

    def net_fn(xyz, dirs, coarse):
        return omodel.net_forward(inp["wc"] if coarse else inp["wf"], inp["latent"], inp["cam"], ns, xyz, dirs)
    with torch.no_grad():
        ref = oracle.render_rays(net_fn, rays.reshape(-1, 8), sb, 64, 32, 16, inp["draws"], white_bkgd=inp["white"],
                                 lindisp=inp["lindisp"])
    ren.injected_draws = draws
    ren.injected_z = dict(w_coarse=ref["w_coarse"].cuda(), depth_coarse=ref["depth_coarse"].cuda())
    with torch.no_grad():
        out = ren(net, rays.cuda(), want_weights=True)
    z_c, z_f = ren.last_z
    # sampling does not depend on the MLP precision when teacher-forced: still bit-exact
    assert np.array_equal(z_c.reshape(-1, 64).cpu().numpy(), ref["z_coarse"].numpy())
    assert np.array_equal(z_f.reshape(-1, 96).cpu().numpy(), ref["z_fine"].numpy())
    for part, tag in ((out.coarse, "coarse"), (out.fine, "fine")):
        assert np.abs(part.rgb.cpu().numpy() - fx["rgb_" + tag]).max() < TC_TOL, tag
        assert np.abs(part.weights.cpu().numpy() - fx["w_" + tag]).max() < TC_TOL, tag
        # world units, absolute; 2e-3 except the stated DTU-class deviation (helpers.TC_DEPTH_TOL_DTU)
        depth_tol = TC_TOL if prec == "fp16x2" else tc_depth_tol(fx["name"])
        assert np.abs(part.depth.cpu().numpy() - fx["depth_" + tag]).max() < depth_tol, tag


def test_unimplemented_operand_format_raises():
    from pixelnerf_b200 import _lib, synthetic as syn
    with pytest.raises(NotImplementedError):
        _lib.pack_weights_tc({k: v.cuda() for k, v in syn.mlp_weights(1).items()}, precision="bf16")


@pytest.mark.parametrize("ns,ppo", [(1, 1), (1, 77), (1, 128 * 150 + 5), (2, 128 * 40 + 3), (1, 128 * 74 * 3 + 17)])
def test_chain_is_deterministic(ns, ppo):
    """The chain has no atomics and a fixed schedule per row tile: repeated launches must agree bit for bit,
    whatever mix of 4-CTA and 2-CTA clusters the GPU forms (a protocol race would show up here)."""
    from pixelnerf_b200 import _lib, synthetic as syn
    w = syn.mlp_weights(5)
    packed = _lib.pack_weights_tc({k: v.cuda() for k, v in w.items()})
    n = ns * ppo
    idx, wt, latent, fx = _inputs(n, seed=11)
    tables = _tables(w, latent)
    fxt = make_taps(idx, wt, fx).cuda()
    first = None
    for _ in range(12):
        out, pre = _lib.mlp_forward_tc(packed, tables, fxt, 1, ns, ppo, want_pre=True)
        torch.cuda.synchronize()
        if first is None:
            first = pre.clone()
            assert torch.isfinite(first).all()
        else:
            assert torch.equal(pre, first)


@pytest.mark.parametrize("texels", [(1, 8, 24), (4, 64, 64), (3, 15, 20)])
def test_tensor_core_tables_match_fp32_tables(texels):
    """Training mode rebuilds the per-texel lin_z tables every step on the tensor cores (fp16 latent, fp16 hi + lo
    weights, fp32 sums): against the fp32 CUDA-core tables of eval mode and against a float64 product of the same
    fp16-rounded latent.  Texel counts that are and are not multiples of the 256-row GEMM tile."""
    from pixelnerf_b200 import _lib, synthetic as syn
    w = {k: v.cuda() for k, v in syn.mlp_weights(42).items()}
    v, hl, wl = texels
    lat = syn.latent_maps(3, v, hl, wl).permute(0, 2, 3, 1).contiguous().cuda()     # channels-last fp32
    t32 = _lib.prepare_latent_tables(w, lat).float()
    t16 = _lib.prepare_latent_tables(w, lat.half()).float()
    assert t16.shape == t32.shape == (3, v * hl * wl, 512)
    scale = t32.abs().max().item()
    # vs the fp32 tables: the latent's fp16 rounding (5e-4 relative per element, summed over 512 terms) + fp16 output
    assert (t16 - t32).abs().max().item() < 2e-3 * scale
    # vs float64 on the SAME fp16 latent: only the weight split's residue (2^-22) and the fp16 output rounding remain
    l16 = lat.half().double().reshape(-1, 512)
    for b in range(3):
        folded = w["lin_in.bias"] if b == 0 else w[f"blocks.{b - 1}.fc_1.bias"]
        ref = l16 @ w[f"lin_z.{b}.weight"].double().T + w[f"lin_z.{b}.bias"].double() + folded.double()
        assert (t16[b].double() - ref).abs().max().item() < 6e-4 * scale   # half an fp16 ulp at the largest entry + slack


def test_training_mode_uses_tensor_core_tables_and_stays_in_tolerance():
    from pixelnerf_b200 import synthetic as syn
    from pixelnerf_b200.testing import build_net, oracle_render
    sc = syn.scene("srn_chair")
    wc, wf = syn.mlp_weights(1), syn.mlp_weights(2)
    net, ren = build_net(sc, wc, wf, torch.device("cuda"), mlp_precision="fp16")
    rays = sc["rays"][torch.linspace(0, sc["rays"].shape[0] - 1, 200).long()]
    draws = syn.draws(3, 200, 64, 16, 16)
    ref = oracle_render(sc, wc, wf, rays, draws)
    outs = {}
    for mode in ("eval", "train"):
        getattr(net, mode)()
        ren.injected_draws = {k: v.cuda() for k, v in draws.items()}
        with torch.no_grad():
            outs[mode] = ren(net, rays[None].cuda(), want_weights=True)
        assert (outs[mode].coarse.rgb[0].cpu() - ref["rgb_coarse"]).abs().max().item() < 2e-3
        assert (outs[mode].coarse.weights[0].cpu() - ref["w_coarse"]).abs().max().item() < 2e-3
    assert not torch.equal(outs["eval"].coarse.weights, outs["train"].coarse.weights)   # different table builds
    net.eval()


@pytest.mark.parametrize("ns,ppo,sb", [(2, 130, 2), (2, 1, 1), (3, 77, 1), (3, 1301, 2), (4, 500, 1), (5, 97, 1),
                                       (2, 64 * 150 + 5, 1), (3, 40 * 74 * 2 + 11, 1)])
def test_grouped_chain_vs_emulation_and_two_launch_form(ns, ppo, sb):
    """n_views > 1 in ONE launch (view mean on chip, pn_mlp_forward_tc_grouped): against the float64 emulation of its
    arithmetic (mean of the fp32 h over the views, one fp16 rounding) and against the two-launch form (which rounds
    every view's h to fp16 before the mean).  Ragged point counts, view counts that do and do not divide 32, several
    objects, more super-tiles than CTA pairs; repeated launches agree bit for bit."""
    from pixelnerf_b200 import _lib, synthetic as syn
    w = syn.mlp_weights(42)
    packed = _lib.pack_weights_tc({k: v.cuda() for k, v in w.items()})
    n = sb * ns * ppo
    idx, wt, latent, fx = _inputs(n, seed=ns)
    tables = _tables(w, latent)
    rows = make_taps(idx, wt, fx)                       # (object, view, point) order
    grouped = to_grouped(rows, sb, ns, ppo).cuda()
    assert grouped.shape[0] == int(_lib.lib().pn_point_taps_grouped_rows(sb * ppo, ns))
    out, pre = _lib.mlp_forward_tc_grouped(packed, tables, grouped, sb, ns, ppo, want_pre=True)
    torch.cuda.synchronize()
    assert torch.isfinite(pre).all() and pre.shape == (sb * ppo, 4)
    if n <= 20000:
        refs = []
        for o in range(sb):
            sl = slice(o * ns * ppo, (o + 1) * ns * ppo)
            refs.append(emulate_tc_chain(w, idx[sl], wt[sl], latent, fx[sl].float(), ns, ppo, fused_mean=True)[1])
        ref = torch.cat(refs)
        scale = max(8.0, ref.abs().max().item())
        assert (pre.cpu().double() - ref).abs().max().item() < 1.5e-3 * scale
        assert (out.cpu()[:, :3].double() - torch.sigmoid(ref[:, :3])).abs().max().item() < 1e-3
    two, pre_two = _lib.mlp_forward_tc(packed, tables, rows.cuda(), sb, ns, ppo, want_pre=True)
    scale = max(8.0, pre_two.abs().max().item())
    assert (pre - pre_two).abs().max().item() < 3e-3 * scale       # same arithmetic up to where the fp16 rounding of h sits
    assert (out[:, :3] - two[:, :3]).abs().max().item() < 2e-3
    for _ in range(6):
        again = _lib.mlp_forward_tc_grouped(packed, tables, grouped, sb, ns, ppo, want_pre=True)[1]
        assert torch.equal(again, pre)


@pytest.mark.parametrize("name", ["srn_chair", "dtu"])
def test_grouped_taps_are_a_permutation_of_the_ordinary_taps(name):
    """pn_point_taps_grouped writes exactly the rows of pn_point_taps, in grouped order, zero rows elsewhere."""
    from pixelnerf_b200 import _lib, synthetic as syn
    from pixelnerf_b200.testing import build_net
    sc = syn.scene(name)
    net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), torch.device("cuda"), mlp_precision="fp16")
    n_rays, k = 333, 96
    rays = sc["rays"][torch.linspace(0, sc["rays"].shape[0] - 1, n_rays).long()].cuda()
    z = ren.sample_coarse(rays, u=torch.rand(n_rays, 64, device="cuda"))
    z = torch.sort(torch.cat([z, z[:, :32] + 1e-3], -1), -1)[0].contiguous()
    _, hl, wl, _ = net.latent_nhwc(torch.float32).shape
    ppo = n_rays * k
    plain = _lib.point_taps(net._cam, hl, wl, ppo, rays=rays, z=z)
    grouped = _lib.point_taps_grouped(net._cam, hl, wl, ppo, rays=rays, z=z)
    idx = grouped_row_index(1, sc["ns"], ppo).cuda()
    assert grouped.shape[0] == idx.shape[0]
    got = grouped.view(torch.int32)
    assert torch.equal(got[idx >= 0], plain.view(torch.int32)[idx[idx >= 0]])
    assert int(got[idx < 0].abs().sum()) == 0


@pytest.mark.parametrize("ns,ppo,sb", [(1, 1, 1), (1, 300, 1), (1, 128 * 150 + 5, 1), (2, 130, 2), (3, 1301, 2), (5, 97, 1)])
def test_two_pass_chain_vs_emulation(ns, ppo, sb):
    """mlp_precision = fp16x2 (pn_mlp_forward_tc_x2): the precise chain without its third operand pass -- weights as
    fp16 hi + lo pairs, activations rounded to fp16 ONCE, tables / lin_in / head in fp32.  Against the float64
    emulation of exactly that arithmetic (helpers.emulate_tc_chain: weight_pairs, fp32_front) at the fast chain's bar,
    and it must sit between the fast chain's arithmetic and the exact evaluation: closer to exact float64 than the
    emulation of the fast chain is.  Repeated launches agree bit for bit."""
    from oracle import model as omodel
    from pixelnerf_b200 import _lib, synthetic as syn
    w = syn.mlp_weights(42)
    wd = {k: v.cuda() for k, v in w.items()}
    packed = _lib.pack_weights_x3(wd)
    n = sb * ns * ppo
    idx, wt, latent, _ = _inputs(n, seed=ns + 20)
    g = torch.Generator().manual_seed(6)
    code = torch.zeros(n, 44)
    code[:, :42] = torch.randn(n, 42, generator=g)                    # NOT representable in fp16
    fx = torch.zeros(n, 64, dtype=torch.float16)
    fx[:, :42] = code[:, :42].half()
    rows = make_taps(idx, wt, fx)
    tables32 = _lib.prepare_latent_tables_f32(wd, latent.reshape(1, 8, 24, 512).cuda())
    xlo = _xlo_rows(code, fx)
    if ns > 1:
        rows_dev, code_dev = to_grouped(rows, sb, ns, ppo).cuda(), to_grouped(xlo, sb, ns, ppo).cuda().contiguous()
    else:
        rows_dev, code_dev = rows.cuda(), xlo.cuda()
    out, pre = _lib.mlp_forward_tc_x3(packed, tables32, rows_dev, code_dev, sb, ns, ppo, want_pre=True, passes=2)
    torch.cuda.synchronize()
    assert torch.isfinite(pre).all() and pre.shape == (sb * ppo, 4)
    pts = min(ppo, 3000)                                   # emulate a prefix of every object's points
    refs, fast, exact, got = [], [], [], []
    for o in range(sb):
        sl = torch.cat([torch.arange(pts) + (o * ns + v) * ppo for v in range(ns)])
        refs.append(emulate_tc_chain(w, idx[sl], wt[sl], latent, code[sl], ns, pts, fused_mean=True, weight_pairs=True,
                                     fp32_front=True)[1])
        fast.append(emulate_tc_chain(w, idx[sl], wt[sl], latent, fx[sl].float(), ns, pts, fused_mean=True)[1])
        z = (latent.double()[idx[sl].long()] * wt[sl].double().unsqueeze(-1)).sum(1)
        w64 = {k: v.double() for k, v in w.items()}
        w64["_points_per_obj"] = pts
        exact.append(omodel.resnetfc_forward(w64, torch.cat([z, code[sl, :42].double()], 1), ns))
        got.append(pre[o * ppo: o * ppo + pts].cpu().double())
    ref, fast, exact, got = torch.cat(refs), torch.cat(fast), torch.cat(exact), torch.cat(got)
    scale = max(8.0, ref.abs().max().item())
    assert (got - ref).abs().max().item() < 1.5e-3 * scale
    if ppo >= 300:
        assert (got - exact).abs().mean().item() < 0.8 * (fast - exact).abs().mean().item()
    again = _lib.mlp_forward_tc_x3(packed, tables32, rows_dev, code_dev, sb, ns, ppo, want_pre=True, passes=2)[1]
    assert torch.equal(again, pre)


@pytest.mark.parametrize("ns,ppo,sb", [(1, 1, 1), (1, 300, 1), (1, 128 * 150 + 5, 1), (2, 130, 2), (3, 1301, 1)])
def test_precise_chain_vs_float64(ns, ppo, sb):
    """mlp_precision = fp16x3 (pn_mlp_forward_tc_x3: three fp16 hi/lo operand passes per 512-wide GEMM, everything
    else fp32) against an EXACT float64 evaluation of ResnetFC on the same unrounded inputs -- no emulation of any
    rounding.  Bar: 4e-5 of the largest pre-activation (the sigma row carries a gain of 10, where fp32 arithmetic itself
    is ~1e-5 from float64) and 1e-5 on the colours -- ~40x tighter than the fast chain's bar against its own emulation."""
    from oracle import model as omodel
    from pixelnerf_b200 import _lib, synthetic as syn
    w = syn.mlp_weights(42)
    wd = {k: v.cuda() for k, v in w.items()}
    packed = _lib.pack_weights_x3(wd)
    n = sb * ns * ppo
    idx, wt, latent, fx16 = _inputs(n, seed=ns + 10)
    g = torch.Generator().manual_seed(5)
    code = torch.zeros(n, 44)
    code[:, :42] = torch.randn(n, 42, generator=g)                    # NOT representable in fp16
    fx = torch.zeros(n, 64, dtype=torch.float16)
    fx[:, :42] = code[:, :42].half()
    rows = make_taps(idx, wt, fx)
    tables32 = _lib.prepare_latent_tables_f32(wd, latent.reshape(1, 8, 24, 512).cuda())
    xlo = _xlo_rows(code, fx)
    if ns > 1:
        rows_dev, code_dev = to_grouped(rows, sb, ns, ppo).cuda(), to_grouped(xlo, sb, ns, ppo).cuda().contiguous()
    else:
        rows_dev, code_dev = rows.cuda(), xlo.cuda()
    out, pre = _lib.mlp_forward_tc_x3(packed, tables32, rows_dev, code_dev, sb, ns, ppo, want_pre=True)
    torch.cuda.synchronize()
    assert torch.isfinite(pre).all()
    m = min(n, 6000 * ns) // ns * ns if sb == 1 else n       # float64 reference on a prefix of the points (whole objects when sb > 1)
    z = (latent.double()[idx.long()] * wt.double().unsqueeze(-1)).sum(1)
    refs = []
    for o in range(sb):
        pts = ppo if sb > 1 else m // ns
        sl = torch.cat([torch.arange(pts) + (o * ns + v) * ppo for v in range(ns)])
        w64 = {k: v.double() for k, v in w.items()}
        w64["_points_per_obj"] = pts
        refs.append(omodel.resnetfc_forward(w64, torch.cat([z[sl], code[sl, :42].double()], 1), ns))
    ref = torch.cat(refs)
    got = pre.cpu().double() if sb > 1 else pre.cpu().double()[: m // ns]
    scale = max(1.0, ref.abs().max().item())
    err = (got - ref).abs().max().item()
    assert err < 4e-5 * scale, (err, scale)
    o_got = out.cpu().double() if sb > 1 else out.cpu().double()[: m // ns]
    assert (o_got[:, :3] - torch.sigmoid(ref[:, :3])).abs().max().item() < 1e-5
    again = _lib.mlp_forward_tc_x3(packed, tables32, rows_dev, code_dev, sb, ns, ppo, want_pre=True)[1]
    assert torch.equal(again, pre)
