"""Profiling target (not a test): render whole frames of one BASELINE scene through the public API so that ncu can
capture the launches of ONE frame.  `--frames F` frames are rendered; the first F-1 are warm-up.

    python tests/ncu_frame.py --scene srn_car|srn_chair|dtu|sn64 [--frames 3] [--precision fp16]

Chain launches per frame: 2 at NS = 1 (coarse + fine pass), 4 at NS > 1 (per-view segment + merged segment per pass).
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pixelnerf_b200 import _lib, synthetic as syn  # noqa: E402
from pixelnerf_b200.testing import build_net  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="srn_car")
    ap.add_argument("--frames", type=int, default=3)
    ap.add_argument("--precision", default="fp16")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    sc = syn.scene(args.scene)
    net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), dev, mlp_precision=args.precision)
    par = ren.bind_parallel(net, [0], simple_output=True).eval()
    rays = sc["rays"].clone()
    if args.scene == "dtu":
        g = torch.Generator().manual_seed(4)
        rays[:, 6] = 0.1 + torch.rand(rays.shape[0], generator=g) * 0.4
        rays[:, 7] = 5.0 - torch.rand(rays.shape[0], generator=g) * 0.5
    rays = rays[None].to(dev)
    torch.manual_seed(0)
    n0 = _lib.launch_count()
    with torch.no_grad():
        for _ in range(args.frames):
            rgb, depth = par(rays)
    torch.cuda.synchronize()
    assert torch.isfinite(rgb).all() and torch.isfinite(depth).all()
    print(f"{args.scene}: {rays.shape[1]} rays x {args.frames} frames, NS={sc['ns']}, "
          f"{(_lib.launch_count() - n0) // args.frames} kernel launches per frame, rgb mean {float(rgb.mean()):.4f}")


if __name__ == "__main__":
    main()
