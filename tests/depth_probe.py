"""Probe (not a test): dL/d depth_coarse through the depth-guided samples, fp32 path vs tensor-core path."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pixelnerf_b200 import _lib, synthetic as syn
from pixelnerf_b200.testing import build_net
from test_gpu_backward import run_render_backward

caps = {}
orig = _lib.depth_sample_bwd


def run(prec, scene=os.environ.get("SCENE", "sn64"), n=int(os.environ.get("NRAYS", "96"))):
    got = []

    def hook(cam, latent, rays, z_sorted, z_depth, g_z, g_h0, g_fz, w_in, g_depth):
        orig(cam, latent, rays, z_sorted, z_depth, g_z, g_h0, g_fz, w_in, g_depth)
        only_gz = torch.zeros_like(g_depth)
        orig(cam, latent, rays, z_sorted, z_depth, g_z, torch.zeros_like(g_h0), torch.zeros_like(g_fz), w_in, only_gz)
        only_h0 = torch.zeros_like(g_depth)
        orig(cam, latent, rays, z_sorted, z_depth, None, g_h0, torch.zeros_like(g_fz), w_in, only_h0)
        got.append((g_depth.clone(), only_gz, only_h0, g_h0.float().clone(), g_fz.clone()))
    _lib.depth_sample_bwd = hook
    try:
        loss, loss_ref, worst = run_render_backward(scene, n, mlp_precision=prec, fine_only=True, positive=True)
    finally:
        _lib.depth_sample_bwd = orig
    print(prec, "loss", loss, loss_ref)
    print({k: f"{v:.1e}" for k, v in sorted(worst.items(), key=lambda kv: -kv[1])[:8]})
    return got


a = run("fp32")
b = run("fp16")
for j, name in enumerate(("total", "composite", "code", "g_h0", "g_fz")):
    u = torch.cat([x[j] for x in a]); v = torch.cat([x[j] for x in b])
    print(name, tuple(u.shape), "max|fp32|", float(u.abs().max()), "max diff", float((u - v).abs().max()),
          "rel l2", float((u - v).norm() / u.norm()))
