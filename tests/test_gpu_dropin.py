"""The drop-in surface on the GPU: the call sequence of the reference's eval/gen_video.py:63-237 through
``pixelnerf_b200.install()`` (no reference tree on the GPU box: the mirrors are registered as model / render / util;
the overlay on the live reference is exercised by tests/test_install_cpu.py), the two mirror methods nothing in the
render path calls (``SpatialEncoder.index``, ``PositionalEncoding.forward``) and the video helper."""
import numpy as np
import pytest
import torch

from pixelnerf_b200 import synthetic as syn
from pixelnerf_b200.testing import build_net, default_conf

pytestmark = pytest.mark.gpu


def test_gen_video_call_sequence_through_install():
    import pixelnerf_b200
    model, render, util = pixelnerf_b200.install()
    conf = default_conf(mlp_precision="fp16")
    device = util.get_cuda(0)
    net = model.make_model(conf["model"]).to(device=device)
    renderer = render.NeRFRenderer.from_conf(conf["renderer"], lindisp=False, eval_batch_size=5000).to(device=device)
    render_par = renderer.bind_parallel(net, [0], simple_output=True).eval()
    h = w = 32
    nv = 4
    render_poses = torch.stack([util.pose_spherical(a, -10.0, 1.3) for a in np.linspace(-180, 180, nv + 1)[:-1]], 0)
    focal = torch.tensor(36.0)[None]
    render_rays = util.gen_rays(render_poses, w, h, focal, 0.8, 1.8, c=None).to(device=device)     # host poses
    dev_rays = util.gen_rays(render_poses.to(device), w, h, focal, 0.8, 1.8, c=None)               # device kernel
    assert float((render_rays - dev_rays).abs().max()) < 1e-6
    images = syn.images(0, 1, 3, h, w)[0]
    poses = torch.stack([util.pose_spherical(a, -10.0, 1.3) for a in (0.0, 40.0, 80.0)])
    src = torch.tensor([1])
    with torch.no_grad():
        net.encode(images[src].unsqueeze(0), poses[src].unsqueeze(0).to(device=device), focal.to(device), c=None)
        all_rgb = []
        for rays in torch.split(render_rays.view(-1, 8), 1500, dim=0):
            rgb, _depth = render_par(rays[None])
            all_rgb.append(rgb[0])
        frames = torch.cat(all_rgb).view(-1, h, w, 3)
    assert frames.shape == (nv, h, w, 3) and torch.isfinite(frames).all()
    assert 0.0 <= float(frames.min()) and float(frames.max()) <= 1.0 + 1e-4
    # the video helper: same loop with device ray generation and the 8-bit conversion before the copy
    from pixelnerf_b200.video import render_video
    torch.manual_seed(0)
    with torch.no_grad():
        u8 = render_video(render_par, render_poses, w, h, focal, 0.8, 1.8, ray_batch_size=1500)
    assert u8.dtype == torch.uint8 and u8.shape == (nv, h, w, 3) and not u8.is_cuda


def test_encoder_index_matches_grid_sample():
    """SpatialEncoder.index (src/model/encoder.py:80-109) against torch's grid_sample on the same latent."""
    from pixelnerf_b200.model.encoder import SpatialEncoder
    enc = SpatialEncoder(pretrained=False).cuda()
    lat = syn.latent_maps(3, 2, 24, 40).cuda()
    enc.latent = lat
    ls = torch.tensor([40.0, 24.0], device="cuda")
    enc.latent_scaling = ls / (ls - 1) * 2.0
    g = torch.Generator().manual_seed(0)
    uv = (torch.rand(2, 333, 2, generator=g) * torch.tensor([90.0, 60.0]) - 5.0).cuda()   # some outside the image
    image_size = torch.tensor([80.0, 48.0], device="cuda")
    got = enc.index(uv, None, image_size)
    grid = (uv * (enc.latent_scaling / image_size) - 1.0).unsqueeze(2)
    ref = torch.nn.functional.grid_sample(lat, grid, align_corners=True, mode="bilinear", padding_mode="border")[..., 0]
    assert got.shape == ref.shape == (2, 512, 333)
    assert float((got - ref).abs().max()) < 2e-5
    one = enc.index(uv[:1], None, image_size)          # a single uv set is broadcast over the maps (encoder.py:91-92)
    assert one.shape == (2, 512, 333) and torch.equal(one[0], got[0])


def test_positional_encoding_matches_oracle():
    from oracle import model as omodel
    from pixelnerf_b200.model.code import PositionalEncoding
    code = PositionalEncoding(num_freqs=6, d_in=3, freq_factor=1.5, include_input=True).cuda()
    x = (torch.rand(1000, 3, generator=torch.Generator().manual_seed(1)) * 6 - 3)
    got = code(x.cuda())
    ref = omodel.positional_encoding(x, 6, 1.5)
    assert got.shape == ref.shape == (1000, 39)
    assert float((got.cpu() - ref).abs().max()) < 2e-5


def test_parameter_change_between_forward_and_backward_raises():
    """The backward computes against what the forward captured; an in-place parameter update in between is an
    error (torch's autograd reports the same situation the same way)."""
    sc = syn.scene("sn64")
    net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), torch.device("cuda"), mlp_precision="fp16")
    net.train(); ren.train()
    rays = sc["rays"][:64].cuda()[None]
    out = ren(net, rays)
    loss = out.fine.rgb.mean()
    with torch.no_grad():
        net.mlp_coarse.lin_out.bias.add_(1.0)
    with pytest.raises(RuntimeError, match="modified by an inplace operation"):
        loss.backward()


def test_encode_between_forward_and_backward_uses_the_forwards_scene():
    """Two encodes before one backward (losses of two scenes summed): each render's backward must use its own
    camera / latent, so the summed gradient equals the sum of the two separate gradients."""
    sc = syn.scene("sn64")
    wc, wf = syn.mlp_weights(1), syn.mlp_weights(2)
    dev = torch.device("cuda")
    from pixelnerf_b200.testing import scene_camera, set_encoded
    rays = sc["rays"][100:164].cuda()[None]
    lat2 = syn.latent_maps(99, 1, 32, 32)
    cam = scene_camera(sc)

    def grads(which):
        net, ren = build_net(sc, wc, wf, dev, mlp_precision="fp32")
        net.train(); ren.train()
        total = 0.0
        for i, lat in enumerate((sc["latent"], lat2)):
            if which not in ("both", i):
                continue
            set_encoded(net, lat, cam["poses"], cam["focal"], cam["c"], cam["image_shape"], 1)
            ren.injected_draws = {k: v.cuda() for k, v in syn.draws(5, 64, 64, 16, 16).items()}
            total = total + ren(net, rays).fine.rgb.square().mean()
        total.backward()
        return net.mlp_fine.lin_out.weight.grad.clone(), net.mlp_coarse.lin_in.weight.grad.clone()

    a, b, both = grads(0), grads(1), grads("both")
    for x, y, z in zip(a, b, both):
        assert float((x + y - z).abs().max()) < 1e-5 * max(1.0, float(z.abs().max()))
        assert float((x - z).abs().max()) > 1e-4 * float(z.abs().max())     # the second scene does contribute
