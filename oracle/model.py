"""Oracle: PixelNeRFNet.forward restated as flat functions (torch CPU).

Test infrastructure only (see oracle/__init__.py).  The encoder trunk
(SpatialEncoder.forward, src/model/encoder.py:111-164) is outside the hot path
and stays in torch; the oracle takes its output ``latent (SB*NS,C,Hl,Wl)`` as an
input.  Weights are passed as a plain dict keyed by the reference's state_dict
names without the ``mlp_coarse.`` / ``mlp_fine.`` prefix, e.g. ``lin_in.weight``,
``blocks.0.fc_0.weight``, ``lin_z.1.bias``, ``lin_out.weight``.

All ``path:line`` citations are into /root/reference/.
"""
import math

import torch
import torch.nn.functional as F


def camera_state(poses, focal, c, width, height):
    """Per-view camera buffers as cached by encode().  src/model/models.py:112-141.

    poses (V,4,4) camera-to-world.  Returns dict(poses (V,3,4) world-to-camera,
    focal (F,2) with fy negated, c (Cn,2), image_shape (2,) = [W,H]).
    """
    rot = poses[:, :3, :3].transpose(1, 2)
    trans = -torch.bmm(rot, poses[:, :3, 3:])
    w2c = torch.cat((rot, trans), dim=-1)
    image_shape = torch.tensor([float(width), float(height)]).to(poses.device)
    if focal.dim() == 0:
        focal = focal[None, None].repeat((1, 2))
    elif focal.dim() == 1:
        focal = focal.unsqueeze(-1).repeat((1, 2))
    else:
        focal = focal.clone()
    focal = focal.float()
    focal[..., 1] *= -1.0
    if c is None:
        c = (image_shape * 0.5).unsqueeze(0)
    elif c.dim() == 0:
        c = c[None, None].repeat((1, 2))
    elif c.dim() == 1:
        c = c.unsqueeze(-1).repeat((1, 2))
    return dict(poses=w2c, focal=focal, c=c, image_shape=image_shape)


def _repeat_interleave(x, n):
    # src/util/util.py:58-65
    return x.unsqueeze(1).expand(-1, n, *x.shape[1:]).reshape(-1, *x.shape[1:])


def positional_encoding(x, num_freqs=6, freq_factor=1.5):
    """[x, sin(f_k x), sin(f_k x + pi/2)] interleaved.  src/model/code.py:12-42."""
    freqs = (freq_factor * 2.0 ** torch.arange(0, num_freqs)).to(x.device)
    f = torch.repeat_interleave(freqs, 2).view(1, -1, 1)
    ph = torch.zeros(2 * num_freqs)
    ph[1::2] = math.pi * 0.5
    ph = ph.view(1, -1, 1).to(x.device)
    emb = x.unsqueeze(1).repeat(1, num_freqs * 2, 1)
    emb = torch.sin(torch.addcmul(ph, emb, f)).view(x.shape[0], -1)
    return torch.cat((x, emb), dim=-1)


def point_features(latent, cam, xyz, viewdirs, ns, num_freqs=6, freq_factor=1.5):
    """MLP input rows [latent(C), posenc(xyz_rot)(39), viewdir_rot(3)].

    src/model/models.py:157-227 with the shipped switches (use_xyz, normalize_z,
    use_code, use_viewdirs, not use_code_viewdirs) and src/model/encoder.py:80-109.
    xyz, viewdirs (SB,P,3) world space.  Returns (SB*NS*P, C+42) rows ordered
    (object, view, point), plus uv (SB*NS,P,2) for inspection.
    """
    sb, p, _ = xyz.shape
    poses = cam["poses"]
    xyz_r = _repeat_interleave(xyz, ns)
    xyz_rot = torch.matmul(poses[:, None, :3, :3], xyz_r.unsqueeze(-1))[..., 0]
    xyz_cam = xyz_rot + poses[:, None, :3, 3]
    z_feature = positional_encoding(xyz_rot.reshape(-1, 3), num_freqs, freq_factor)
    vd = _repeat_interleave(viewdirs.reshape(sb, p, 3, 1), ns)
    vd = torch.matmul(poses[:, None, :3, :3], vd).reshape(-1, 3)
    z_feature = torch.cat((z_feature, vd), dim=1)

    uv = -xyz_cam[:, :, :2] / xyz_cam[:, :, 2:]
    focal, c = cam["focal"], cam["c"]
    uv = uv * _repeat_interleave(focal.unsqueeze(1), ns if focal.shape[0] > 1 else 1)
    uv = uv + _repeat_interleave(c.unsqueeze(1), ns if c.shape[0] > 1 else 1)
    # encoder.index: src/model/encoder.py:94-109
    hl, wl = latent.shape[-2:]
    ls = torch.tensor([float(wl), float(hl)])
    ls = (ls / (ls - 1) * 2.0).to(latent.device)
    scale = ls / cam["image_shape"]
    uvn = (uv * scale - 1.0).unsqueeze(2)
    samples = F.grid_sample(latent, uvn, align_corners=True, mode="bilinear",
                            padding_mode="border")[:, :, :, 0]
    lat = samples.transpose(1, 2).reshape(-1, latent.shape[1])
    return torch.cat((lat, z_feature), dim=-1), uv


def resnetfc_forward(w, zx, ns, d_latent=512, n_blocks=5, combine_layer=3):
    """ResnetFC.forward with ReLU + average combine.  src/model/resnetfc.py:132-184,
    ResnetBlockFC.forward :53-62, combine_interleaved src/util/util.py:461-471.

    zx (N, d_latent + d_in), rows ordered (object, view, point); returns (N/ns, 4).
    """
    z, x = zx[..., :d_latent], zx[..., d_latent:]
    h = F.linear(x, w["lin_in.weight"], w["lin_in.bias"])
    for b in range(n_blocks):
        if b == combine_layer and ns > 1:
            # reshape(-1, NS, B, H).mean(1): the caller guarantees one object per
            # call or (object, view, point) ordering, both handled by this reshape.
            h = h.reshape(-1, ns, w["_points_per_obj"], h.shape[-1]).mean(dim=1)
            h = h.reshape(-1, h.shape[-1])
        if b < combine_layer:
            h = h + F.linear(z, w[f"lin_z.{b}.weight"], w[f"lin_z.{b}.bias"])
        net = F.linear(torch.relu(h), w[f"blocks.{b}.fc_0.weight"], w[f"blocks.{b}.fc_0.bias"])
        dx = F.linear(torch.relu(net), w[f"blocks.{b}.fc_1.weight"], w[f"blocks.{b}.fc_1.bias"])
        h = h + dx
    return F.linear(torch.relu(h), w["lin_out.weight"], w["lin_out.bias"])


def net_forward(weights, latent, cam, ns, xyz, viewdirs):
    """PixelNeRFNet.forward: (SB,P,3) -> (SB,P,4) = [sigmoid(rgb), relu(sigma)].

    src/model/models.py:146-266.  ``weights`` is the un-prefixed dict of the MLP
    selected by the caller (coarse or fine, models.py:242-255).
    """
    sb, p, _ = xyz.shape
    zx, _ = point_features(latent, cam, xyz, viewdirs, ns)
    w = dict(weights)
    w["_points_per_obj"] = p
    out = resnetfc_forward(w, zx, ns, d_latent=latent.shape[1]).reshape(-1, p, 4)
    return torch.cat([torch.sigmoid(out[..., :3]), torch.relu(out[..., 3:4])], dim=-1).reshape(sb, p, 4)


def split_state_dict(sd):
    """Reference state_dict -> (coarse weights, fine weights or None)."""
    def pick(prefix):
        d = {k[len(prefix):]: v.detach().float() for k, v in sd.items() if k.startswith(prefix)}
        return d or None
    return pick("mlp_coarse."), pick("mlp_fine.")
