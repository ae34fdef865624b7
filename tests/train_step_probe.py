"""Where a 512-ray training step (BASELINE config 5: 4 objects x 128 rays) spends its time: wall clock per step,
GPU-busy time and the top kernels / CPU ops from torch.profiler.  Not a test.
    python tests/train_step_probe.py [rays_per_object]"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pixelnerf_b200 import synthetic as syn  # noqa: E402
from pixelnerf_b200.testing import build_net, scene_camera, set_encoded  # noqa: E402

n_obj_rays = int(sys.argv[1]) if len(sys.argv) > 1 else 128
dev = torch.device("cuda", 0)
sb = 4
sc = syn.scene("srn_car")
net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), dev, mlp_precision="fp16")
cam = scene_camera(sc)
lat = syn.latent_maps(21, sb, sc["hl"], sc["wl"])
set_encoded(net, lat, cam["poses"].expand(sb, -1, -1).contiguous(), cam["focal"], cam["c"], cam["image_shape"], 1)
net.train(); ren.train()
for q in net.parameters():
    q.requires_grad_(True)
opt = torch.optim.Adam([q for n, q in net.named_parameters() if n.startswith("mlp_")], lr=1e-4)
g = torch.Generator().manual_seed(0)
idx = torch.randint(0, sc["rays"].shape[0], (sb, n_obj_rays), generator=g)
rays = sc["rays"][idx].to(dev)
target = torch.rand(sb, n_obj_rays, 3, generator=g).to(dev)


def step():
    opt.zero_grad(set_to_none=True)
    out = ren(net, rays, want_weights=True)
    loss = ((out.coarse.rgb - target) ** 2).mean() + ((out.fine.rgb - target) ** 2).mean()
    loss.backward()
    opt.step()


for _ in range(5):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20):
    step()
torch.cuda.synchronize()
print(f"{sb * n_obj_rays} rays/step: {(time.perf_counter() - t0) / 20 * 1e3:.3f} ms per step (wall)")
# phases (wall, with a sync after each)
ph = {"zero_grad": 0.0, "forward": 0.0, "backward": 0.0, "adam": 0.0}
for _ in range(10):
    torch.cuda.synchronize(); a = time.perf_counter()
    opt.zero_grad(set_to_none=True)
    torch.cuda.synchronize(); b = time.perf_counter()
    out = ren(net, rays, want_weights=True)
    loss = ((out.coarse.rgb - target) ** 2).mean() + ((out.fine.rgb - target) ** 2).mean()
    torch.cuda.synchronize(); c = time.perf_counter()
    loss.backward()
    torch.cuda.synchronize(); d = time.perf_counter()
    opt.step()
    torch.cuda.synchronize(); e = time.perf_counter()
    for k, v in zip(ph, (b - a, c - b, d - c, e - d)):
        ph[k] += v / 10 * 1e3
print("phases, synced (ms):", {k: round(v, 3) for k, v in ph.items()})
from torch.profiler import ProfilerActivity, profile  # noqa: E402
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(5):
        step()
    torch.cuda.synchronize()
ev = prof.key_averages()
gpu = sorted([(e.device_time_total / 5, e.count / 5, e.key) for e in ev if e.device_time_total > 0], reverse=True)
print(f"GPU busy per step: {sum(t for t, _, _ in gpu) / 1e3:.3f} ms in {sum(c for _, c, _ in gpu):.0f} kernels")
for t, c, k in gpu[:14]:
    print(f"  {t:9.1f} us  x{c:5.1f}  {k[:90]}")
cpu = sorted([(e.self_cpu_time_total / 5, e.count / 5, e.key) for e in ev], reverse=True)
print("top CPU self time per step:")
for t, c, k in cpu[:12]:
    print(f"  {t:9.1f} us  x{c:5.1f}  {k[:90]}")
# host side only: python-level profile of the step (where the launch gaps come from)
import cProfile  # noqa: E402
import pstats  # noqa: E402
pr = cProfile.Profile()
pr.enable()
for _ in range(20):
    step()
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
