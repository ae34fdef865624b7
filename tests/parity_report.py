"""Measured parity table (not a test): max-abs error of rgb / weights / depth (depth in WORLD UNITS, absolute)
of both MLP paths against the references, per configuration.

  * golden fixtures of the LIVE reference (tests/golden/render_*.npz, full_*.npz), stage-wise with teacher forcing;
  * BASELINE.json's configurations at their real sizes against the CPU oracle on a ray sample (stage-wise), plus the
    end-to-end statistics of the same sample (share of rays within the tolerance).

    python tests/parity_report.py [--rays 256] [--out gpurun_out/r02_parity.json]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))

import oracle  # noqa: E402
from oracle import model as omodel  # noqa: E402
from helpers import full_case_inputs, golden_files, load, max_abs, render_case_inputs  # noqa: E402
from pixelnerf_b200 import synthetic as syn  # noqa: E402
from pixelnerf_b200.testing import build_net, oracle_render  # noqa: E402
from test_gpu_model import _net_for  # noqa: E402

TOL = {"fp32": 1e-4, "fp16x3": 1e-4, "fp16x2": 2e-3, "fp16": 2e-3}


def stagewise(net, ren, rays, draws, ref):
    """ref: dict with rgb_/w_/depth_ coarse+fine and (optionally) z_coarse / z_fine of the reference side."""
    n = rays.shape[0] * rays.shape[1]
    ren.injected_draws = {k: v.cuda() for k, v in draws.items()}
    ren.injected_z = dict(w_coarse=torch.as_tensor(ref["w_coarse"]).reshape(n, -1).cuda(),
                          depth_coarse=torch.as_tensor(ref["depth_coarse"]).reshape(n).cuda())
    with torch.no_grad():
        out = ren(net, rays.cuda(), want_weights=True)
    z_c, z_f = ren.last_z
    res = {}
    for part, tag in ((out.coarse, "coarse"), (out.fine, "fine")):
        res[f"rgb_{tag}"] = max_abs(part.rgb, ref["rgb_" + tag])
        res[f"depth_{tag}"] = max_abs(part.depth, ref["depth_" + tag])
        if "w_" + tag in ref:
            res[f"weights_{tag}"] = max_abs(part.weights, ref["w_" + tag])
    if "z_fine" in ref:
        res["z_bit_exact"] = bool(np.array_equal(z_c.reshape(n, -1).cpu().numpy(), np.asarray(ref["z_coarse"])) and
                                  np.array_equal(z_f.reshape(n, -1).cpu().numpy(), np.asarray(ref["z_fine"])))
    return res


def end_to_end(net, ren, rays, draws, ref, tol):
    ren.injected_draws = {k: v.cuda() for k, v in draws.items()}
    with torch.no_grad():
        out = ren(net, rays.cuda(), want_weights=True)
    e_rgb = (out.fine.rgb.reshape(-1, 3).cpu() - torch.as_tensor(ref["rgb_fine"]).reshape(-1, 3)).abs().max(-1)[0].numpy()
    e_dep = (out.fine.depth.reshape(-1).cpu() - torch.as_tensor(ref["depth_fine"]).reshape(-1)).abs().numpy()
    return dict(rays=int(e_rgb.size), rgb_within_tol=float(np.mean(e_rgb < tol)), depth_within_tol=float(np.mean(e_dep < tol)),
                rgb_q999=float(np.quantile(e_rgb, 0.999)), depth_q999=float(np.quantile(e_dep, 0.999)),
                rgb_max=float(e_rgb.max()), depth_max=float(e_dep.max()))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=256)
    ap.add_argument("--out", default=os.path.join(os.path.dirname(HERE), "gpurun_out", "r02_parity.json"))
    args = ap.parse_args()
    torch.set_num_threads(os.cpu_count() or 8)
    report = {"tolerance": TOL, "unit": "max-abs; depth in world units (absolute)", "golden": {}, "configs": {}}

    for path in golden_files("render_"):
        fx = load(path)
        inp = render_case_inputs(fx)
        ref = {k: fx[k] for k in ("rgb_coarse", "w_coarse", "depth_coarse", "rgb_fine", "w_fine", "depth_fine")}
        span = float((inp["rays"][..., 7] - inp["rays"][..., 6]).max())
        for prec in ("fp32", "fp16x3", "fp16x2", "fp16"):
            net, ren = _net_for(fx, inp, precision=prec)
            r = stagewise(net, ren, inp["rays"], inp["draws"], ref)
            r["e2e"] = end_to_end(net, ren, inp["rays"], inp["draws"], ref, TOL[prec])
            r["ray_span"] = span
            report["golden"].setdefault(str(fx["name"]), {})[prec] = r
            print("golden", fx["name"], prec, json.dumps(r), flush=True)

    fx = load(golden_files("full_")[0])
    inp = full_case_inputs(fx)
    ref = {k: fx[k] for k in ("rgb_coarse", "w_coarse", "depth_coarse", "rgb_fine", "depth_fine")}
    for prec in ("fp32", "fp16x3", "fp16x2", "fp16"):
        net, ren = _net_for(fx, inp, precision=prec)
        r = stagewise(net, ren, inp["rays"], inp["draws"], ref)
        r["e2e"] = end_to_end(net, ren, inp["rays"], inp["draws"], ref, TOL[prec])
        r["ray_span"] = float((inp["rays"][..., 7] - inp["rays"][..., 6]).max())
        report["golden"].setdefault("dtu_full_ns3 (300x400, 2048 rays)", {})[prec] = r
        print("golden dtu_full", prec, json.dumps(r), flush=True)

    for name, per_ray in (("sn64", False), ("srn_car", False), ("srn_chair", False), ("dtu", True)):
        sc = syn.scene(name)
        rays = sc["rays"].clone()
        if per_ray:
            g = torch.Generator().manual_seed(4)
            rays[:, 6] = 0.1 + torch.rand(rays.shape[0], generator=g) * 0.4
            rays[:, 7] = 5.0 - torch.rand(rays.shape[0], generator=g) * 0.5
        rays = rays[torch.linspace(0, rays.shape[0] - 1, args.rays).long()]
        wc, wf = syn.mlp_weights(1), syn.mlp_weights(2)
        draws = syn.draws(3, args.rays, 64, 16, 16)
        ref = oracle_render(sc, wc, wf, rays, draws)
        for prec in ("fp32", "fp16x3", "fp16x2", "fp16"):
            net, ren = build_net(sc, wc, wf, torch.device("cuda"), mlp_precision=prec)
            r = stagewise(net, ren, rays[None], draws, ref)
            r["e2e"] = end_to_end(net, ren, rays[None], draws, ref, TOL[prec])
            r["ray_span"] = float((rays[:, 7] - rays[:, 6]).max())
            r["sum_w_fine_mean"] = float(ref["w_fine"].sum(-1).mean())
            report["configs"].setdefault(name, {})[prec] = r
            print("config", name, prec, json.dumps(r), flush=True)

    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(report, f, indent=1)
    print("wrote", args.out)


if __name__ == "__main__":
    main()
