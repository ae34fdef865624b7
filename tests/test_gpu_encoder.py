"""Encoder hand-off: pn_encoder_latent = the tail of SpatialEncoder.forward (reference
src/model/encoder.py:150-163: F.interpolate(bilinear, align_corners=True) of every ResNet level to level 0's size,
torch.cat over channels) written channels-last, against torch's own ops; and encode() through it against encode()
through the torch tail (same ResNet activations), down to the rendered frame."""
import pytest
import torch
import torch.nn.functional as F

from pixelnerf_b200 import _lib, synthetic as syn
from pixelnerf_b200.testing import default_conf
from pixelnerf_b200.model import make_model
from pixelnerf_b200.render import NeRFRenderer

pytestmark = pytest.mark.gpu


def _levels(v, sizes, seed):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(v, c, h, w, generator=g).cuda() for c, (h, w) in zip((64, 64, 128, 256), sizes)]


@pytest.mark.parametrize("v,sizes", [
    (1, [(64, 64), (32, 32), (16, 16), (8, 8)]),          # SRN 128x128
    (2, [(32, 32), (32, 32), (16, 16), (8, 8)]),          # sn64, use_first_pool=False
    (3, [(150, 200), (75, 100), (38, 50), (19, 25)]),     # DTU 300x400: non-square, odd level sizes
    (1, [(5, 33), (3, 17), (2, 9), (1, 5)]),              # ragged tiles, a 1-row level
])
def test_encoder_latent_matches_torch(v, sizes):
    lv = _levels(v, sizes, 7)
    ref = torch.cat([F.interpolate(l, sizes[0], mode="bilinear", align_corners=True) for l in lv], dim=1)
    out = _lib.encoder_latent(lv, torch.float32)
    assert out.shape == (v, sizes[0][0], sizes[0][1], 512) and out.is_contiguous()
    err = (out.permute(0, 3, 1, 2) - ref).abs().max().item()
    assert err < 2e-6 * max(1.0, ref.abs().max().item()), err   # fp32 rounding of the 4-tap blend only
    assert torch.equal(out[..., :64], lv[0].permute(0, 2, 3, 1))   # level 0 is copied exactly
    half = _lib.encoder_latent(lv, torch.float16)
    assert torch.equal(half, out.half())
    # fewer levels (num_layers < 4)
    out2 = _lib.encoder_latent(lv[:2], torch.float32)
    assert torch.equal(out2, out[..., :128].contiguous())


def test_encode_fused_tail_equals_torch_tail_through_the_render():
    dev = torch.device("cuda")
    torch.manual_seed(0)
    conf = default_conf(mlp_precision="fp16")
    net = make_model(conf["model"]).to(dev).eval()
    for blk in list(net.mlp_coarse.blocks) + list(net.mlp_fine.blocks):   # reference init zeroes fc_1
        torch.nn.init.kaiming_normal_(blk.fc_1.weight, a=0, mode="fan_in")
    ren = NeRFRenderer.from_conf(conf["renderer"]).to(dev).eval()
    sc = syn.scene("srn_chair")
    images = syn.images(3, 1, 2, 128, 128).to(dev)
    poses = sc["src_poses"][None].to(dev)
    focal = torch.tensor(131.25, device=dev)
    with torch.no_grad():
        net.encode(images, poses, focal)
    lat_fused = net.encoder.latent
    assert lat_fused.shape == (2, 512, 64, 64) and not lat_fused.is_contiguous()
    assert net.latent_nhwc(torch.float32).data_ptr() == lat_fused.data_ptr()   # no transposed copy was made
    rays = sc["rays"][::16].to(dev)[None]
    zs = torch.linspace(1.25, 2.75, 64, device=dev).expand(rays.shape[1], 64).contiguous()
    fz_a, fx_a = _lib.point_features(net._cam, net.latent_nhwc(torch.float32), rays.shape[1] * 64, rays=rays[0], z=zs)
    d = {k: v.to(dev) for k, v in syn.draws(2, rays.shape[1], 64, 16, 16).items()}
    ren.injected_draws = dict(d)
    with torch.no_grad():
        a = ren(net, rays, want_weights=True)
    # the torch tail: with grad mode on and a trainable encoder the fused tail is not taken
    net.encode(images, poses, focal)
    lat_torch = net.encoder.latent.detach()
    assert lat_torch.is_contiguous()
    scale = lat_torch.abs().max().item()
    assert (lat_fused - lat_torch).abs().max().item() < 2e-6 * scale
    net.encoder.latent = lat_torch
    net._refresh_camera()
    assert net.latent_nhwc(torch.float32).data_ptr() != lat_torch.data_ptr()   # this one is a transposed copy
    # the gathered latent rows (the MLP's input) agree to fp32 rounding, the positional code exactly
    fz_b, fx_b = _lib.point_features(net._cam, net.latent_nhwc(torch.float32), rays.shape[1] * 64, rays=rays[0], z=zs)
    assert (fz_a - fz_b).abs().max().item() < 4e-6 * scale and torch.equal(fx_a, fx_b)
    ren.injected_draws = dict(d)
    with torch.no_grad():
        b = ren(net, rays, want_weights=True)
    # a random-init network is razor sharp (sigma in the thousands), so single rays amplify the 1e-6 latent
    # differences: the frame agrees in the median, not in the max
    assert (a.coarse.rgb - b.coarse.rgb).abs().median().item() < 1e-5
    assert (a.fine.rgb - b.fine.rgb).abs().median().item() < 1e-5
