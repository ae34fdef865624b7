"""Training with the tensor-core path tracks training with the fp32 path.

The per-tensor gradient bars of the tensor-core backward are wide by construction (bf16 gradients, relu decisions of
an fp16-operand forward: tests/test_gpu_backward_tc.py).  What makes them acceptable is shown here: the same student,
from the same initial weights, with the same rays and the same in-kernel draws at every step, trained for 150 Adam
steps against a teacher's rendering (the loss of train/train.py:264-281: MSE of the coarse and of the fine rgb) once
with ``mlp_precision = fp16`` and once with ``fp32``.  Both must learn (loss down by a fifth or more; measured: 0.178 ->
0.130, then a plateau) and the two loss curves must stay within a band of each other: every 10-step window mean within
1 % (measured: 0.02 %, the curves agree to four digits in all fifteen windows)."""
import numpy as np
import pytest
import torch

from pixelnerf_b200 import synthetic as syn
from pixelnerf_b200.testing import build_net

pytestmark = pytest.mark.gpu

STEPS, WINDOW = 150, 10


def _curve(prec, sc, rays, target):
    dev = torch.device("cuda")
    net, ren = build_net(sc, syn.mlp_weights(11, sigma_bias=0.0), syn.mlp_weights(12, sigma_bias=0.0), dev, mlp_precision=prec)
    net.train()
    ren.train()
    params = [q for n, q in net.named_parameters() if n.startswith("mlp_")]
    for q in params:
        q.requires_grad_(True)
    opt = torch.optim.Adam(params, lr=2e-4)
    losses = []
    for step in range(STEPS):
        torch.manual_seed(1000 + step)            # the call seed of the in-kernel draws: same for both precisions
        opt.zero_grad(set_to_none=True)
        out = ren(net, rays, want_weights=True)
        loss = ((out.coarse.rgb - target) ** 2).mean() + ((out.fine.rgb - target) ** 2).mean()
        loss.backward()
        opt.step()
        losses.append(float(loss.detach()))
    return np.asarray(losses)


def test_fp16_training_tracks_fp32_training():
    sc = syn.scene("sn64")
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(3)
    rays = sc["rays"][torch.randint(0, sc["rays"].shape[0], (512,), generator=g)].to(dev)[None]
    teacher, tren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), dev, mlp_precision="fp32")
    with torch.no_grad():
        torch.manual_seed(7)
        target = tren(teacher, rays).fine.rgb.clone()
    c32 = _curve("fp32", sc, rays, target)
    c16 = _curve("fp16", sc, rays, target)
    assert np.isfinite(c32).all() and np.isfinite(c16).all()
    w32 = c32.reshape(-1, WINDOW).mean(1)
    w16 = c16.reshape(-1, WINDOW).mean(1)
    print("fp32 window means", np.round(w32, 5).tolist())
    print("fp16 window means", np.round(w16, 5).tolist())
    # both learn ...
    assert w32[-1] < 0.8 * w32[0], (w32[0], w32[-1])
    assert w16[-1] < 0.8 * w16[0], (w16[0], w16[-1])
    # ... and along the same curve
    rel = np.abs(w16 - w32) / w32
    assert rel.max() < 0.01, (rel.round(5).tolist())
    assert abs(c16[0] - c32[0]) < 2e-3 * max(1.0, c32[0])      # step 0: same weights, forward-only difference
