"""Determinism stress of the tensor-core chain on a GPU (diagnostics, not a test): repeated launches of
NS = 1..3 problems whose tile counts exercise both the 4-CTA and the 2-CTA clusters must agree bit for bit.
    python tests/tc_stress.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from pixelnerf_b200 import _lib, synthetic as syn
from helpers import make_taps
import test_gpu_mlp_tc as T
w = syn.mlp_weights(5)
packed = _lib.pack_weights_tc({k: v.cuda() for k, v in w.items()})
bad = 0
for ns, ppo, sb in [(3, 64 * 117, 2), (2, 128 * 40 + 3, 1), (3, 9001, 2), (1, 128 * 300 + 1, 1), (3, 64 * 300, 1)]:
    n = ns * ppo * sb
    idx, wt, latent, fx = T._inputs(n, seed=ns + ppo)
    tables = T._tables(w, latent)
    fxt = make_taps(idx, wt, fx).cuda()
    first = None
    for it in range(40):
        out, pre = _lib.mlp_forward_tc(packed, tables, fxt, sb, ns, ppo, want_pre=True)
        torch.cuda.synchronize()
        if first is None:
            first = pre.clone()
        elif not torch.equal(pre, first):
            bad += 1
            d = (pre - first).abs()
            print("MISMATCH", ns, ppo, sb, it, float(d.max()), int((d > 0).sum()), (d.max(1).values > 0).nonzero().flatten()[:10].tolist())
    print("done", ns, ppo, sb, "finite", bool(torch.isfinite(first).all()), flush=True)
print("bad", bad)
