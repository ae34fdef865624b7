"""Probe (not a test): where the host time of one end-to-end call goes (cProfile around bench.py's e2e step)."""
import cProfile, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pixelnerf_b200 import synthetic as syn
from pixelnerf_b200.testing import build_net

dev = torch.device("cuda", 0)
sc = syn.scene("srn_car")
net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), dev, mlp_precision="fp16")
par = ren.bind_parallel(net, [0], simple_output=True).eval()
rays_host = sc["rays"].reshape(1, -1, 8).contiguous().pin_memory()
out_host = torch.empty(rays_host.shape[1], 4).pin_memory()


def e2e():
    with torch.no_grad():
        rgb, depth = par(rays_host.to(dev, non_blocking=True))
    out_host[:, :3].copy_(rgb[0], non_blocking=True)
    out_host[:, 3].copy_(depth[0], non_blocking=True)
    torch.cuda.current_stream().synchronize()


def launch_only():
    with torch.no_grad():
        par(rays_dev)


rays_dev = rays_host.to(dev)
for _ in range(5):
    e2e()
t0 = time.perf_counter()
for _ in range(50):
    e2e()
print("e2e ms/call", (time.perf_counter() - t0) / 50 * 1e3)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(50):
    launch_only()
t1 = time.perf_counter()
torch.cuda.synchronize()
print("host ms/call to enqueue one frame", (t1 - t0) / 50 * 1e3, "; with GPU drain", (time.perf_counter() - t0) / 50 * 1e3)
pr = cProfile.Profile()
pr.enable()
for _ in range(50):
    launch_only()
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
