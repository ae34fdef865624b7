"""Tensor-core training path (forward chain that keeps its bf16 GEMM operands + dgrad / wgrad GEMMs with bf16
gradients, csrc/pn_train_tc.cu) against float64 autograd over the CPU oracle, i.e. against what the reference's
``loss.backward()`` computes (train/train.py:271-281) on identical inputs.

Tolerance: gradients travel between layers as bf16 (2^-9 relative rounding per element, unbiased, fp32 sums) on
top of the forward's fp16 operands, so the bar is reduced-precision: every tensor within 2e-2 of its largest
entry, the typical tensor within 5e-3 (the fp32 path's bars are 1e-2 / 1e-3, test_gpu_backward.py)."""
import pytest
import torch

from oracle import model as omodel
from pixelnerf_b200 import _lib, synthetic as syn
from helpers import emulate_tc_backward, make_taps
from test_gpu_backward import run_render_backward
from test_gpu_mlp_tc import _inputs, _tables

pytestmark = pytest.mark.gpu


def _rel(got, ref):
    return (got.cpu().double() - ref.double()).abs().max().item() / max(ref.abs().max().item(), 1e-12)


def _case(ns, ppo, sb, positive):
    w = syn.mlp_weights(11)
    n = sb * ns * ppo
    idx, wt, latent, fx = _inputs(n, seed=5)
    g_out = torch.randn(sb * ppo, 4, generator=torch.Generator().manual_seed(1))
    if positive:
        g_out = g_out.abs()
    wd = {k: v.cuda() for k, v in w.items()}
    fx_t = make_taps(idx, wt, fx).cuda()
    lat = latent.reshape(1, 8, 24, 512).cuda().contiguous()
    packed = _lib.pack_weights_tc(wd)
    out, acts = _lib.mlp_forward_tc_train(packed, _tables(w, latent), fx_t, sb, ns, ppo)
    plain = _lib.mlp_forward_tc(packed, _tables(w, latent), fx_t, sb, ns, ppo)
    assert torch.equal(out, plain)          # keeping the activations does not change the forward
    grads = {k: torch.zeros_like(v) for k, v in wd.items()}
    g_fz = _lib.mlp_backward_tc(_lib.pack_weights_fp32(wd), _lib.pack_weights_bf16(wd), fx_t, lat, sb, ns, ppo, acts,
                                out, g_out.cuda(), grads, True)
    torch.cuda.synchronize()
    z64 = (wt.double()[:, :, None] * latent.double()[idx]).sum(1)
    return dict(w=w, fx=fx, z64=z64, g_out=g_out, out=out.cpu(), acts=acts, grads=grads, g_fz=g_fz, n=n)


def _report(worst):
    print({k: f"{v:.1e}" for k, v in sorted(worst.items(), key=lambda kv: -kv[1])[:6]})
    vals = sorted(worst.values())
    return vals[-1], vals[len(vals) // 2]


@pytest.mark.parametrize("ns,ppo,sb", [(1, 300, 1), (2, 130, 2), (3, 77, 1), (1, 5000, 1)])
def test_mlp_backward_tc_vs_emulation(ns, ppo, sb):
    """The kernels' arithmetic, tightly: float64 backward given the forward's own saved operands (so every relu
    decision is the kernels'), random-sign upstream gradient.  What is left is the bf16 rounding of the
    gradients between layers (the residual-stream gradient included): measured 2e-3..7e-3 of each tensor's
    largest entry."""
    c = _case(ns, ppo, sb, positive=False)
    n, m = c["n"], sb * ppo
    a = _lib.tc_acts_view(c["acts"], n)
    al = [a[i].cpu().double() if i < 6 or ns == 1 else a[i, :m].cpu().double() for i in range(11)]
    ref, gz = emulate_tc_backward(c["w"], al, c["z64"], c["fx"][:, :42].double(), c["out"], c["g_out"], sb, ns, ppo)
    worst = {k: _rel(c["grads"][k], ref[k]) for k in _lib.weight_names()}
    worst["g_feat_z"] = _rel(c["g_fz"], gz)
    hi, med = _report(worst)
    assert hi < 1.5e-2 and med < 8e-3, worst


@pytest.mark.parametrize("positive", [True, False])
@pytest.mark.parametrize("ns,ppo,sb", [(1, 300, 1), (2, 130, 2), (1, 5000, 1)])
def test_mlp_backward_tc_vs_autograd(ns, ppo, sb, positive):
    """End to end against float64 autograd over the oracle MLP + head.  The forward's fp16 operands flip the
    relu decision of activations within ~1e-3 of zero (about 0.2 % of the units here); with a random-sign
    upstream gradient every parameter gradient is a cancelling sum in which one flipped unit shows at full
    size (measured 2-9 % of the largest entry, an unbiased noise), with a one-signed upstream gradient it
    shows in proportion (typical tensor below 1 %, lin_in.weight -- sums over random-sign code entries -- 5 %)."""
    c = _case(ns, ppo, sb, positive)
    w, fx, g_out = c["w"], c["fx"], c["g_out"]
    torch.set_default_dtype(torch.float64)
    try:
        wr = {k: v.double().requires_grad_(True) for k, v in w.items()}
        zr = c["z64"].clone().requires_grad_(True)
        outs = []
        for o in range(sb):
            sl = slice(o * ns * ppo, (o + 1) * ns * ppo)
            wd = dict(wr)
            wd["_points_per_obj"] = ppo
            pre = omodel.resnetfc_forward(wd, torch.cat([zr[sl], fx[sl, :42].double()], -1), ns)
            outs.append(torch.cat([torch.sigmoid(pre[:, :3]), torch.relu(pre[:, 3:4])], -1))
        out_ref = torch.cat(outs)
        (out_ref * g_out.double()).sum().backward()
    finally:
        torch.set_default_dtype(torch.float32)
    assert (c["out"].double() - out_ref.detach()).abs().max() < 2e-3 * max(1.0, out_ref.abs().max().item())
    worst = {k: _rel(c["grads"][k], wr[k].grad) for k in _lib.weight_names()}
    # dL/dz is per row, not a sum over rows: a flipped relu decision changes that row's entries at full size, so
    # its max error is the worst single row of thousands (measured 9e-2); the bar is on the rms over all entries
    dz = c["g_fz"].cpu().double() - zr.grad
    worst["g_feat_z (rms)"] = (dz.pow(2).mean().sqrt() / zr.grad.pow(2).mean().sqrt()).item()
    hi, med = _report(worst)
    if positive:    # lin_in / fc_0 still sum over random-sign inputs: 5e-2 measured on lin_in.weight
        assert hi < 8e-2 and med < 2e-2, worst
    else:
        assert hi < 0.15 and med < 0.06, worst


@pytest.mark.parametrize("positive", [True, False])
@pytest.mark.parametrize("scene,n_rays,stash", [("sn64", 96, True), ("srn_chair", 64, False), ("srn_chair", 64, True)])
def test_render_backward_tc_vs_oracle(scene, n_rays, stash, positive):
    """Through the public NeRFRenderer API (both passes, composite backward, gather backward), same two
    regimes as above; with the forward's operands kept for the backward (stash) and with the backward
    re-running the forward chunk by chunk."""
    loss, loss_ref, worst = run_render_backward(scene, n_rays, mlp_precision="fp16", positive=positive, stash=stash,
                                                depth_grad=False)
    assert abs(loss - loss_ref) < 2e-3 * max(1.0, abs(loss_ref))
    hi, med = _report(worst)
    if positive:
        assert hi < 2e-2 and med < 1e-2, worst
    else:
        assert hi < 0.15 and med < 0.06, worst


@pytest.mark.parametrize("scene,n_rays,stash", [("sn64", 96, True), ("srn_chair", 64, False)])
def test_depth_sample_gradient_tc_vs_fp32_path(scene, n_rays, stash, monkeypatch):
    """The gradient the depth-guided sample positions carry to the coarse depth (pn_depth_sample_bwd, reference
    nerf.py:150-161 + 290-293) is a derivative of the fine colour w.r.t. a sample position: a cancelling sum of a
    compositing, a positional-code and a grid-gradient term (measured on the tensor-core path -- fp16 forward
    operands, bf16 gradients -- in relative l2 against the fp32 path: sn64 NS=1 each term 2 %, their sum 8 %;
    srn_chair NS=2 compositing term 8 % from the forward's fp16 operands alone, code term 3 %, sum 22 %).  It is therefore
    checked against the fp32 kernel path (itself 3e-3 from float64 autograd, test_gpu_backward.py) per ray in
    relative l2, and the fine MLP's gradients -- which do not pass through it -- keep the reduced-precision bars."""
    got = {}
    orig = _lib.depth_sample_bwd
    for prec in ("fp32", "fp16"):
        parts = []

        def hook(cam, latent, rays, z_sorted, z_depth, g_z, g_h0, g_fz, w_in, g_depth, _parts=parts):
            orig(cam, latent, rays, z_sorted, z_depth, g_z, g_h0, g_fz, w_in, g_depth)
            _parts.append(g_depth.clone())
        monkeypatch.setattr(_lib, "depth_sample_bwd", hook)
        loss, loss_ref, worst = run_render_backward(scene, n_rays, mlp_precision=prec, positive=True, stash=stash,
                                                    fine_only=True)
        got[prec] = torch.cat(parts)
        if prec == "fp16":
            fine = [v for k, v in worst.items() if k.startswith("mlp_fine.")]
            assert max(fine) < 2e-2, worst
    monkeypatch.setattr(_lib, "depth_sample_bwd", orig)
    a, b = got["fp32"], got["fp16"]
    assert a.shape == b.shape and float(a.abs().max()) > 0
    rel = float((a - b).norm() / a.norm())
    print(f"dL/d depth_coarse through the depth samples: tensor-core vs fp32 path, relative l2 {rel:.3f}")
    assert rel < 0.35
