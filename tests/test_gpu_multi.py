"""N-GPU ray-sharded render == 1-GPU render, bit for bit (needs >= 2 visible GPUs; skipped otherwise)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_sharded_equals_single_gpu():
    n = min(torch.cuda.device_count(), 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", "29511", os.path.join(ROOT, "tests", "multi_gpu_check.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_sharded_training_gradients():
    n = min(torch.cuda.device_count(), 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", "29512", os.path.join(ROOT, "tests", "multi_gpu_train_check.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_render_on_second_device_while_first_is_current():
    """The reference's scripts pick a device with util.get_cuda(args.gpu_id[0]) and never call
    torch.cuda.set_device: with --gpu_id 1 every tensor lives on cuda:1 while cuda:0 stays current.  The kernels must
    run on the tensors' device (and stream) and give the same result as on cuda:0."""
    from pixelnerf_b200 import synthetic as syn
    from pixelnerf_b200.testing import build_net
    sc = syn.scene("sn64")
    wc, wf = syn.mlp_weights(1), syn.mlp_weights(2)
    rays = sc["rays"][:500]
    draws = syn.draws(3, 500, 64, 16, 16)
    outs = []
    torch.cuda.set_device(0)
    for dev in ("cuda:0", "cuda:1"):
        for prec in ("fp16", "fp32"):
            net, ren = build_net(sc, wc, wf, torch.device(dev), mlp_precision=prec)
            ren.injected_draws = {k: v.to(dev) for k, v in draws.items()}
            assert torch.cuda.current_device() == 0
            with torch.no_grad():
                o = ren(net, rays[None].to(dev), want_weights=True)
            assert o.fine.rgb.device == torch.device(dev)
            outs.append((o.fine.rgb.cpu(), o.fine.depth.cpu(), o.coarse.weights.cpu()))
    for a, b in zip(outs[:2], outs[2:]):
        assert all(torch.equal(x, y) for x, y in zip(a, b))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
@pytest.mark.parametrize("simple", [True, False])
def test_in_process_multi_gpu_bind_parallel_equals_single(simple):
    """bind_parallel(net, [0, 1, ...]) from ONE process (the reference's scripts pass args.gpu_id straight in,
    eval/gen_video.py:110): the rays of a call are split over resident replicas and the result equals the
    single-device render bit for bit on the same injected draws; a second call after new weights / a new encode
    re-synchronises the replicas."""
    from pixelnerf_b200 import synthetic as syn
    from pixelnerf_b200.parallel import DeviceSharded
    from pixelnerf_b200.testing import build_net, scene_camera, set_encoded
    gpus = list(range(min(torch.cuda.device_count(), 4)))
    sc = syn.scene("srn_chair")
    wc, wf = syn.mlp_weights(1), syn.mlp_weights(2)
    dev = torch.device("cuda:0")
    net, ren = build_net(sc, wc, wf, dev, mlp_precision="fp16")
    n = 3001
    rays = sc["rays"][:n].to(dev)[None]
    draws = {k: v.to(dev) for k, v in syn.draws(5, n, 64, 16, 16).items()}
    single = ren.bind_parallel(net, [0], simple_output=simple).eval()
    multi = ren.bind_parallel(net, gpus, simple_output=simple).eval()
    assert isinstance(multi, DeviceSharded)

    def flat(o):
        return list(o) if isinstance(o, tuple) else [o["coarse"]["rgb"], o["coarse"]["depth"], o["coarse"]["weights"],
                                                     o["fine"]["rgb"], o["fine"]["depth"], o["fine"]["weights"]]

    def compare():
        with torch.no_grad():
            ren.injected_draws = dict(draws)
            a = flat(single(rays, want_weights=True))
            ren.injected_draws = dict(draws)
            b = flat(multi(rays, want_weights=True))
        assert all(y.device == dev and torch.equal(x, y) for x, y in zip(a, b))

    compare()
    with torch.no_grad():          # an optimiser step / load_weights: replicas must follow
        net.mlp_fine.lin_out.bias.add_(0.05)
        net.mlp_coarse.blocks[2].fc_1.weight.mul_(1.01)
    compare()
    cam = scene_camera(sc)         # a new encode
    set_encoded(net, syn.latent_maps(77, 2, 64, 64), cam["poses"], cam["focal"], cam["c"], cam["image_shape"], 2)
    compare()
    ren.n_fine, ren.using_fine = 0, False      # callers mutate the renderer after binding (eval/eval.py:60)
    with torch.no_grad():
        out = multi(rays) if simple else multi(rays)["coarse"]["rgb"]
    assert (out[0] if simple else out).shape == (1, n, 3)
    # without injected draws every shard must get its own jitter (not one pattern repeated per shard)
    ren.n_fine, ren.using_fine = 32, True
    with torch.no_grad():
        multi(rays)
    zs = [w.renderer.last_z[0] for _, w in [(None, multi.wrapped)] + multi._replicas]
    m = min(z.shape[1] for z in zs)
    assert not torch.equal(zs[0][0, :m].cpu(), zs[1][0, :m].cpu())


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
@pytest.mark.parametrize("prec,tol", [("fp32", 1e-4), ("fp16", 3e-2)])
def test_in_process_multi_gpu_training_gradients(prec, tol):
    """Training through the in-process wrapper: the replicas' gradients reach the primary's parameters through
    differentiable peer copies, and their sum equals the single-device gradient (fp32 path: to rounding; tensor-core
    path: to the bf16 gradient rounding, as in tests/multi_gpu_train_check.py)."""
    from pixelnerf_b200 import synthetic as syn
    from pixelnerf_b200.testing import build_net
    gpus = list(range(min(torch.cuda.device_count(), 4)))
    sc = syn.scene("sn64")
    dev = torch.device("cuda:0")
    n = 402
    rays = sc["rays"][:n].to(dev)[None]
    draws = {k: v.to(dev) for k, v in syn.draws(7, n, 64, 16, 16).items()}
    target = torch.rand(1, n, 3, generator=torch.Generator().manual_seed(2)).to(dev)
    grads = {}
    for tag, ids in (("single", [0]), ("multi", gpus)):
        net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), dev, mlp_precision=prec)
        net.train(); ren.train()
        par = ren.bind_parallel(net, ids).train()
        ren.injected_draws = dict(draws)
        out = par(rays, want_weights=True)
        loss = ((out["coarse"]["rgb"] - target) ** 2).mean() + ((out["fine"]["rgb"] - target) ** 2).mean()
        loss.backward()
        grads[tag] = {k: q.grad.clone() for k, q in net.named_parameters() if q.grad is not None}
    assert set(grads["single"]) == set(grads["multi"]) and len(grads["single"]) >= 60
    for k, g in grads["single"].items():
        assert float((grads["multi"][k] - g).abs().max()) <= tol * max(float(g.abs().max()), 1e-12), k
