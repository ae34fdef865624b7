"""A whole training step captured in a CUDA graph (the reference's 512-ray step is launch bound).

train/train.py:224-281 per step: ``net.encode(images, poses, focal, c)`` -> ``render_par(rays, want_weights=True)`` ->
MSE of the coarse and fine rgb -> ``backward`` -> ``optim.step``.  To replay that from a graph every kernel argument
must stay valid: the draws' call seed therefore lives in device memory (``NeRFRenderer.use_device_seed``,
``pn_set_device_seed``) and is advanced by a captured add, and ``encode`` reads nothing back from the device.

* the device seed: a given word gives the same sample positions, the next call different ones;
* the captured step: N replays from the same start (weights, Adam state, seed word) as N eager steps give the same
  loss curve -- to the accumulation-order noise of the split-K weight gradients (fp32 atomics), 2e-3 relative over
  12 steps -- and the replayed draws differ from step to step.
"""
import numpy as np
import pytest
import torch

from pixelnerf_b200 import synthetic as syn
from pixelnerf_b200.testing import build_net, default_conf, load_mlp_weights

pytestmark = pytest.mark.gpu


def test_device_seed_is_the_call_seed_and_advances():
    sc = syn.scene("sn64")
    dev = torch.device("cuda")
    net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), dev, mlp_precision="fp16")
    rays = sc["rays"][:300].to(dev)[None]
    ren.use_device_seed(dev)
    assert ren.device_seed.dtype == torch.int64 and ren.device_seed.is_cuda
    zs = []
    with torch.no_grad():
        for start in (1234, 1234, 99):
            ren.device_seed.fill_(start)
            ren(net, rays)
            zs.append((ren.last_z[0].clone(), ren.last_z[1].clone()))
            assert int(ren.device_seed) == start + ren._SEED_STRIDE       # advanced once per call, on the device
        ren(net, rays)                                                    # no reset: the word moved on
        zs.append((ren.last_z[0].clone(), ren.last_z[1].clone()))
    assert torch.equal(zs[0][0], zs[1][0]) and torch.equal(zs[0][1], zs[1][1])          # same word, same draws
    assert not torch.equal(zs[0][0], zs[2][0]) and not torch.equal(zs[2][0], zs[3][0])  # other word, other draws
    # and the host seed is back in charge once the word is dropped
    ren.use_device_seed(enable=False)
    with torch.no_grad():
        torch.manual_seed(5)
        ren(net, rays)
        a = ren.last_z[0].clone()
        torch.manual_seed(5)
        ren(net, rays)
    assert torch.equal(a, ren.last_z[0])


def _trainer(seed_word):
    import pixelnerf_b200
    model, render, util = pixelnerf_b200.install()
    dev = torch.device("cuda")
    torch.manual_seed(0)
    conf = default_conf(mlp_precision="fp16")
    net = model.make_model(conf["model"]).to(device=dev)
    # conditioned MLP weights (the reference's init leaves sigma at relu(<= 0) = 0 for most points: nothing to learn from)
    load_mlp_weights(net, syn.mlp_weights(11, sigma_bias=0.0), syn.mlp_weights(12, sigma_bias=0.0))
    ren = render.NeRFRenderer.from_conf(conf["renderer"], lindisp=False).to(device=dev)
    par = ren.bind_parallel(net, [0]).train()
    net.train()
    net.encoder.eval()                         # batch-norm statistics frozen: the curve depends on the MLPs alone
    ren.use_device_seed(dev)
    ren.device_seed.fill_(seed_word)
    # the encoder takes part in the step (forward and backward are captured) but only the MLPs are updated: Adam on
    # a ResNet amplifies the accumulation-order noise of the atomics into a visibly different trajectory within a few
    # steps, which would say nothing about the graph
    opt = torch.optim.Adam([q for n, q in net.named_parameters() if n.startswith("mlp_")], lr=1e-4, capturable=True)
    h = w = 32
    sb, nv, n_rays = 2, 3, 128
    g = torch.Generator().manual_seed(4)
    images = torch.stack([syn.images(i, 1, nv, h, w)[0] for i in range(sb)]).to(dev)            # (SB, NV, 3, H, W)
    poses = torch.stack([torch.stack([util.pose_spherical(a + 15.0 * i, -10.0, 1.3) for a in (0.0, 40.0, 80.0)])
                         for i in range(sb)]).to(dev)
    focal = torch.tensor([36.0, 36.0]).to(dev)
    cam_rays = util.gen_rays(poses.reshape(-1, 4, 4), w, h, focal[:1], 0.8, 1.8).reshape(sb, nv * h * w, 8)
    pick = torch.randint(0, nv * h * w, (sb, n_rays), generator=g).to(dev)
    rays = torch.gather(cam_rays, 1, pick[..., None].expand(-1, -1, 8)).contiguous()
    target = torch.rand(sb, n_rays, 3, generator=g).to(dev)
    src = torch.tensor([0, 2], device=dev)

    def step():
        opt.zero_grad(set_to_none=True)
        net.encode(images[:, src], poses[:, src], focal)
        out = par(rays, want_weights=True)          # a plain dict, like the reference's wrapper (nerf.py:22-42)
        loss = ((out["coarse"]["rgb"] - target) ** 2).mean() + ((out["fine"]["rgb"] - target) ** 2).mean()
        loss.backward()
        opt.step()
        return loss

    return net, ren, step


N_STEPS = 12


def test_training_step_replayed_from_a_cuda_graph_matches_eager_steps():
    # eager: 3 warm-up steps (as the capture protocol below needs them) + N steps
    net, ren, step = _trainer(777)
    for _ in range(3):
        step()
    eager = [float(step()) for _ in range(N_STEPS)]
    z_eager = ren.last_z[0].clone()

    net, ren, step = _trainer(777)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):              # torch's capture protocol: warm up on a side stream
        for _ in range(3):
            step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        loss = step()
        z_keep = ren.last_z[0]                 # a graph-pool tensor: rewritten by every replay
    # capturing does not run anything: the first replay is step 1
    replayed, zs = [], []
    for _ in range(N_STEPS):
        graph.replay()
        replayed.append(float(loss))
        zs.append(z_keep.clone())
    assert np.isfinite(replayed).all()
    assert not torch.equal(zs[0], zs[1]) and not torch.equal(zs[1], zs[2])      # new jitter on every replay
    assert torch.equal(zs[-1], z_eager)                                          # ... and the same as the eager run's
    rel = np.abs(np.asarray(replayed) - np.asarray(eager)) / np.abs(np.asarray(eager))
    print("eager   ", np.round(eager, 6).tolist())
    print("replayed", np.round(replayed, 6).tolist())
    assert rel.max() < 2e-3, rel.tolist()
    assert replayed[-1] < replayed[0]          # it trains
