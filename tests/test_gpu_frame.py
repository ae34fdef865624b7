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
