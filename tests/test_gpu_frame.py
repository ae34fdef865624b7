"""Either side of the render path: device ray generation against the host restatement of
util.gen_rays (src/util/util.py:238-276) and the 8-bit frame conversion against numpy's
``(frames * 255).astype(np.uint8)`` (eval/gen_video.py:234-236)."""
import numpy as np
import pytest
import torch

from helpers import golden_files, load
from pixelnerf_b200 import _lib
from pixelnerf_b200.util import util

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("path", golden_files("rays_"))
def test_gen_rays_kernel_vs_live_reference(path):
    """pn_gen_rays against rays_*.npz = the reference's own util.gen_rays on its own util.pose_spherical poses
    (src/util/util.py:238-276, 309-323; tests/golden/make_golden.py: rays_cases): origins and bounds are copies (bit
    exact), unit directions within 1e-6 (the normalisation and the 3x3 product round in a different order)."""
    fx = load(path)
    w, h, rows = int(fx["w"]), int(fx["h"]), int(fx["pose_rows"])
    focal = fx["focal"].tolist() if fx["focal"].ndim else float(fx["focal"])
    c = fx["c"].tolist() if fx["c"].size else None
    poses = torch.from_numpy(fx["poses"])[:, :rows].contiguous().cuda()
    got = util.gen_rays(poses, w, h, focal, float(fx["near"]), float(fx["far"]), c=c)
    assert got.is_cuda and got.shape == (poses.shape[0], h, w, 8)
    sub = got[:, torch.from_numpy(fx["ys"]).cuda()][:, :, torch.from_numpy(fx["xs"]).cuda()].cpu().numpy()
    assert np.array_equal(sub[..., :3], fx["rays"][..., :3]) and np.array_equal(sub[..., 6:], fx["rays"][..., 6:])
    assert np.abs(sub[..., 3:6] - fx["rays"][..., 3:6]).max() < 1e-6


@pytest.mark.parametrize("w,h,focal,c,rows", [
    (128, 128, 131.25, None, 4),                       # SRN: scalar focal, centre = image centre
    (400, 300, (360.0, 355.5), (201.3, 148.9), 4),     # DTU: per-axis focal, off-centre principal point
    (7, 5, 3.0, None, 3),                              # tiny, 3x4 poses
])
def test_gen_rays_vs_host(w, h, focal, c, rows):
    poses = torch.stack([util.pose_spherical(a, -30.0, 1.3) for a in np.linspace(-180, 180, 7)[:-1]])
    poses = poses[:, :rows].contiguous()
    ref = util.gen_rays(poses, w, h, focal, 0.8, 1.8, c=c)
    got = util.gen_rays(poses.cuda(), w, h, focal, 0.8, 1.8, c=c)
    assert got.is_cuda and got.shape == ref.shape == (6, h, w, 8)
    # origins and bounds are copies; directions differ only by fp32 contraction order in the 3x3 product
    assert torch.equal(got[..., :3].cpu(), ref[..., :3]) and torch.equal(got[..., 6:].cpu(), ref[..., 6:])
    assert float((got[..., 3:6].cpu() - ref[..., 3:6]).abs().max()) < 1e-6
    assert float((got[..., 3:6].norm(dim=-1) - 1).abs().max()) < 1e-6


def test_gen_rays_rejects_bad_input():
    with pytest.raises(ValueError):
        _lib.gen_rays(torch.zeros(2, 4, 3, device="cuda"), 8, 8, 1.0, 0.1, 1.0)
    with pytest.raises(RuntimeError):
        _lib.gen_rays(torch.eye(4, device="cuda")[None], 8, 8, 0.0, 0.1, 1.0)


@pytest.mark.parametrize("n", [0, 1, 3, 128 * 128 * 3, 1000003])
def test_frame_to_u8_vs_numpy(n):
    g = torch.Generator().manual_seed(n)
    x = torch.rand(n, generator=g) * 1.0001          # white background can reach 1 + 1e-4
    if n > 2:
        x[0], x[1], x[2] = 0.0, 1.0, 254.999 / 255
    ref = (np.clip(x.numpy() * np.float32(255), 0, 255)).astype(np.uint8)
    got = _lib.frame_to_u8(x.cuda())
    assert got.dtype == torch.uint8 and np.array_equal(got.cpu().numpy(), ref)
    if n > 8:      # unaligned start: the vector path must not be taken
        got = _lib.frame_to_u8(x.cuda()[1:])
        assert np.array_equal(got.cpu().numpy(), ref[1:])


@pytest.mark.parametrize("scene,rows,c,chunk_rays", [
    ("sn64", 4, None, None),                 # NS = 1, whole call in one chunk
    ("sn64", 3, (30.5, 33.25), 3000),        # 3x4 poses, off-centre principal point, chunk boundaries inside a frame
    ("srn_chair", 4, None, 5000),            # NS = 2: grouped taps + one-launch chain under the fused ray source
])
def test_render_frames_equals_rendering_generated_rays(scene, rows, c, chunk_rays):
    """SURVEY 8f.1 -- ray generation fused into the renderer (NeRFRenderer.render_frames -> pn_render_frames): rendering
    target frames from (pose, focal, c, near, far) without a ray tensor must equal, BIT FOR BIT, rendering the tensor
    util.gen_rays builds (pinned to the live reference by test_gen_rays_kernel_vs_live_reference) through the ordinary
    path, with injected draws and with in-kernel draws of the same seed."""
    from pixelnerf_b200 import synthetic as syn
    from pixelnerf_b200.testing import build_net
    sc = syn.scene(scene)
    dev = torch.device("cuda")
    net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), dev, mlp_precision="fp16")
    w = h = 40 if scene == "sn64" else 48
    focal = sc["focal"] * w / sc["w"]
    poses = torch.stack([util.pose_spherical(a, -15.0, sc["radius"]) for a in (-70.0, 5.0, 140.0)])[:, :rows].contiguous().to(dev)
    n = poses.shape[0] * h * w
    rays = util.gen_rays(poses, w, h, focal, sc["near"], sc["far"], c=c).reshape(1, n, 8)
    if chunk_rays is not None:
        ren.max_rows_per_launch = chunk_rays * sc["ns"] * 96
    for injected in (True, False):
        draws = {k: v.to(dev) for k, v in syn.draws(9, n, 64, 16, 16).items()} if injected else None
        with torch.no_grad():
            torch.manual_seed(123)
            ren.injected_draws = dict(draws) if draws else None
            ref = ren(net, rays, want_weights=False)
            torch.manual_seed(123)
            got = ren.render_frames(net, poses, w, h, focal, sc["near"], sc["far"], c=c, draws=draws)
        for part in ("coarse", "fine"):
            assert got[part].rgb.shape == (poses.shape[0], h, w, 3) and got[part].depth.shape == (poses.shape[0], h, w)
            assert torch.equal(got[part].rgb.reshape(-1, 3), ref[part].rgb.reshape(-1, 3)), (scene, part, injected)
            assert torch.equal(got[part].depth.reshape(-1), ref[part].depth.reshape(-1)), (scene, part, injected)
    assert torch.isfinite(got.fine.rgb).all() and float(got.fine.rgb.std()) > 1e-3      # not a constant image


def test_render_frames_rejects_what_it_does_not_cover():
    from pixelnerf_b200 import _lib, synthetic as syn
    from pixelnerf_b200.testing import build_net
    sc = syn.scene("sn64")
    net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), torch.device("cuda"), mlp_precision="fp32")
    poses = util.pose_spherical(0.0, -10.0, 2.6)[None].cuda()
    with pytest.raises(NotImplementedError):
        ren.render_frames(net, poses, 8, 8, 10.0, 1.2, 4.0)
    net.mlp_precision = "fp16"
    with pytest.raises(RuntimeError, match="outside"):      # more rays than the frames have pixels
        fr = _lib.frames_struct(poses, 8, 8, 10.0, 1.2, 4.0)
        _, hl, wl, _ = net.latent_nhwc(torch.float32).shape
        mc = (net.mlp_coarse.packed("fp16"), net.latent_tables(True))
        _lib.render_rays(net._cam, hl, wl, mc, None, None, 64, 16, 16, 0.01, False, True, frames=fr, first_ray=10, n_rays=64)
