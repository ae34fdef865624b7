"""Host-side helpers of the hot path, same names as the reference's util module.

Only what the render path and its callers need: ray
generation, spherical poses and the (SB, NS) row bookkeeping.  The image / bbox /
metric helpers of the reference's util.py are out of scope.
Citations are into the reference: src/util/util.py.
"""
import functools
import math

import numpy as np
import torch


def repeat_interleave(input, repeats, dim=0):
    """Repeat along axis 0, (a, b) -> (a, a, b, b) for repeats=2.  util.py:58-65."""
    assert dim == 0, "only dim 0 is used by the render path"
    return input.unsqueeze(1).expand(-1, repeats, *input.shape[1:]).reshape(-1, *input.shape[1:])


def combine_interleaved(t, inner_dims=(1,), agg_type="average"):
    """Reduce the NS axis of rows ordered (object, view, point).  util.py:461-471."""
    if len(inner_dims) == 1 and inner_dims[0] == 1:
        return t
    t = t.reshape(-1, *inner_dims, *t.shape[1:])
    if agg_type == "average":
        return torch.mean(t, dim=1)
    if agg_type == "max":
        return torch.max(t, dim=1)[0]
    raise NotImplementedError("Unsupported combine type " + agg_type)


def unproj_map(width, height, f, c=None, device="cpu"):
    """Unit camera-space ray direction per pixel, (H, W, 3), -z forward.  util.py:113-143."""
    if c is None:
        cx, cy = width * 0.5, height * 0.5
    else:
        c = torch.as_tensor(c).reshape(-1)
        cx, cy = float(c[0]), float(c[1])
    f = torch.as_tensor(f, dtype=torch.float32).reshape(-1)
    fx, fy = float(f[0]), float(f[-1])
    ys, xs = torch.meshgrid(torch.arange(height, dtype=torch.float32) - cy,
                            torch.arange(width, dtype=torch.float32) - cx, indexing="ij")
    xs, ys = xs.to(device) / fx, ys.to(device) / fy
    d = torch.stack((xs, -ys, -torch.ones_like(xs)), dim=-1)
    return d / torch.norm(d, dim=-1, keepdim=True)


def gen_rays(poses, width, height, focal, z_near, z_far, c=None, ndc=False):
    """Rays (B, H, W, 8) = [origin, unit direction, near, far].  util.py:238-276."""
    if ndc:
        raise NotImplementedError("ndc rays: the reference calls an undefined ndc_rays here (util.py:260)")
    n, dev = poses.shape[0], poses.device
    if poses.is_cuda:     # one kernel, no (H, W) host tables (csrc/pn_frame.cu)
        from .. import _lib
        return _lib.gen_rays(poses, width, height, focal, z_near, z_far, c=c)
    focal = torch.as_tensor(focal, dtype=torch.float32)
    cam = unproj_map(width, height, focal.squeeze(), c=c, device=dev).unsqueeze(0).repeat(n, 1, 1, 1)
    centers = poses[:, None, None, :3, 3].expand(-1, height, width, -1)
    dirs = torch.matmul(poses[:, None, None, :3, :3], cam.unsqueeze(-1))[..., 0]
    near = torch.tensor(z_near, device=dev, dtype=torch.float32).view(1, 1, 1, 1).expand(n, height, width, -1)
    far = torch.tensor(z_far, device=dev, dtype=torch.float32).view(1, 1, 1, 1).expand(n, height, width, -1)
    return torch.cat((centers, dirs, near, far), dim=-1)


def _homog(rows):
    return torch.tensor(rows, dtype=torch.float32)


def trans_t(t):
    return _homog([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, t], [0, 0, 0, 1]])


def rot_phi(phi):
    c, s = np.cos(phi), np.sin(phi)
    return _homog([[1, 0, 0, 0], [0, c, -s, 0], [0, s, c, 0], [0, 0, 0, 1]])


def rot_theta(th):
    c, s = np.cos(th), np.sin(th)
    return _homog([[c, 0, -s, 0], [0, 1, 0, 0], [s, 0, c, 0], [0, 0, 0, 1]])


def pose_spherical(theta, phi, radius):
    """Camera-to-world pose on a sphere around the origin (degrees).  util.py:309-323."""
    c2w = rot_theta(theta / 180.0 * np.pi) @ (rot_phi(phi / 180.0 * np.pi) @ trans_t(radius))
    return _homog([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1]]) @ c2w


def get_norm_layer(norm_type="instance", group_norm_groups=32):
    """Normalisation factory used by the encoder constructor.  util.py:326-341."""
    if norm_type == "batch":
        return functools.partial(torch.nn.BatchNorm2d, affine=True, track_running_stats=True)
    if norm_type == "instance":
        return functools.partial(torch.nn.InstanceNorm2d, affine=False, track_running_stats=False)
    if norm_type == "group":
        return functools.partial(torch.nn.GroupNorm, group_norm_groups)
    if norm_type == "none":
        return None
    raise NotImplementedError("normalization layer [%s] is not found" % norm_type)


def get_cuda(gpu_id):
    return torch.device("cuda:%d" % gpu_id) if torch.cuda.is_available() else torch.device("cpu")


def psnr(pred, target):
    mse = ((pred - target) ** 2).mean()
    return -10 * math.log10(float(mse))
