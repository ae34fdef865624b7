"""Stage-by-stage probe of the tensor-core ResnetFC chain on a GPU (diagnostics, not a test):
    python tests/tc_probe.py [rows] [ns]
prints, for every debug stage, the max error against the float64 emulation of the chain and where
the worst rows / columns are, so a wrong descriptor or layout shows up as a pattern."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import emulate_tc_chain, make_taps  # noqa: E402
from pixelnerf_b200 import _lib, synthetic as syn  # noqa: E402


def main():
    rows = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    ns = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    stages = [int(s) for s in sys.argv[3].split(",")] if len(sys.argv) > 3 else [0, 1, 2, 3, 4, 5]
    ppo = rows
    n = ns * ppo
    w = syn.mlp_weights(42)
    g = torch.Generator().manual_seed(0)
    n_tex = 8 * 24
    latent = torch.rand(n_tex, 512, generator=g)
    idx = torch.randint(0, n_tex, (n, 4), generator=g)
    wt = torch.rand(n, 4, generator=g)
    wt = wt / wt.sum(1, keepdim=True)
    fx = torch.zeros(n, 64)
    fx[:, :42] = torch.randn(n, 42, generator=g)
    fx = fx.half()
    ref_stages, ref_pre = emulate_tc_chain(w, idx, wt, latent, fx.float(), ns, ppo)
    wd = {k: v.cuda() for k, v in w.items()}
    packed = _lib.pack_weights_tc(wd)
    tables = _lib.prepare_latent_tables(wd, latent.reshape(1, 8, 24, 512).cuda())
    fxt = make_taps(idx, wt, fx).cuda()
    torch.cuda.synchronize()
    print("packed ok", flush=True)
    for st in stages:
        (out, pre), dbg = _lib.mlp_forward_tc(packed, tables, fxt, 1, ns, ppo, want_pre=True, debug_stage=st)
        torch.cuda.synchronize()
        ref = ref_stages[st]
        d = (dbg.cpu().double() - ref).abs()
        print(f"stage {st}: max|err| {d.max().item():.3e}  ref scale {ref.abs().max().item():.2f}  "
              f"mean|err| {d.mean().item():.3e}", flush=True)
        if d.max().item() > 1e-2:
            bad_rows = (d.max(1).values > 1e-2).nonzero().flatten()
            bad_cols = (d.max(0).values > 1e-2).nonzero().flatten()
            print("   bad rows:", bad_rows[:16].tolist(), "... n =", len(bad_rows))
            print("   bad cols:", bad_cols[:16].tolist(), "... n =", len(bad_cols))
            print("   got[0,:8] ", dbg[0, :8].tolist())
            print("   ref[0,:8] ", ref[0, :8].tolist())
            print("   got[0,128:136] ", dbg[0, 128:136].tolist())
            print("   ref[0,128:136] ", ref[0, 128:136].tolist())
    dp = (pre.cpu().double() - ref_pre).abs()
    print(f"lin_out pre-activation: max|err| {dp.max().item():.3e} (scale {ref_pre.abs().max().item():.2f})")
    print("pre[0]", pre[0].tolist(), "ref", ref_pre[0].tolist())


if __name__ == "__main__":
    main()
