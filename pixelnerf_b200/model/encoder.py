"""Spatial image encoder -- stays in torch by design (BASELINE.json north_star).

Same constructor keys, attributes (``latent``, ``latent_scaling``, ``latent_size``)
and state_dict names (``model.*`` = torchvision resnet) as the reference's
SpatialEncoder (src/model/encoder.py:13-177).  The feature lookup
(``SpatialEncoder.index``, encoder.py:80-109) is NOT here: it is part of the hot path
and lives in csrc/pn_features.cu.
"""
import torch
import torch.nn.functional as F
import torchvision
from torch import nn

from .. import util


def latent_scaling_host(wl, hl):
    """[W_l, H_l] / ([W_l, H_l] - 1) * 2 with fp32 rounding, as python floats (reference encoder.py:131-133)."""
    import numpy as np
    w, h = np.float32(wl), np.float32(hl)
    one, two = np.float32(1.0), np.float32(2.0)
    return float(w / (w - one) * two), float(h / (h - one) * two)


class SpatialEncoder(nn.Module):
    def __init__(self, backbone="resnet34", pretrained=True, num_layers=4, index_interp="bilinear",
                 index_padding="border", upsample_interp="bilinear", feature_scale=1.0,
                 use_first_pool=True, norm_type="batch"):
        super().__init__()
        if backbone == "custom":
            raise NotImplementedError("backbone=custom (experimental ConvEncoder) is outside the B200 hot path")
        if index_interp != "bilinear" or index_padding != "border":
            raise NotImplementedError("the gather kernel implements index_interp=bilinear, index_padding=border only")
        if norm_type != "batch":
            assert not pretrained
        self.feature_scale, self.use_first_pool = feature_scale, use_first_pool
        weights = "IMAGENET1K_V1" if pretrained else None
        self.model = getattr(torchvision.models, backbone)(weights=weights, norm_layer=util.get_norm_layer(norm_type))
        self.model.fc = nn.Sequential()
        self.model.avgpool = nn.Sequential()
        self.latent_size = [0, 64, 128, 256, 512, 1024][num_layers]
        self.num_layers = num_layers
        self.index_interp, self.index_padding, self.upsample_interp = index_interp, index_padding, upsample_interp
        self.register_buffer("latent", torch.empty(1, 1, 1, 1), persistent=False)
        self.register_buffer("latent_scaling", torch.empty(2, dtype=torch.float32), persistent=False)

    def index(self, uv, cam_z=None, image_size=(), z_bounds=None):
        """Bilinear latent lookup at pixel coordinates uv (B, N, 2) -> (B, L, N).  Reference encoder.py:80-109.
        The render path never calls this (the lookup is fused into pn_point_features / the chain's tables); direct
        callers get the same kernel through an identity camera: a point (u, v, -1) projects to uv = (u, v)."""
        from .. import _lib
        lat = self.latent
        if not lat.is_cuda:
            raise RuntimeError("pixelnerf_b200 kernels need CUDA tensors (there is no CPU path)")
        if uv.shape[0] == 1 and lat.shape[0] > 1:
            uv = uv.expand(lat.shape[0], -1, -1)
        b, n, _ = uv.shape
        dev = lat.device
        uv = uv.to(device=dev, dtype=torch.float32)
        if len(image_size) > 0:
            scale = (self.latent_scaling.to(dev) / torch.as_tensor(image_size, dtype=torch.float32, device=dev)).cpu()
        else:     # uv already normalised to [-1, 1] (encoder.py:95-101 skips the scaling): uv*1 - 1 after a shift
            scale, uv = torch.ones(2), uv + 1.0
        eye = torch.eye(3, 4, device=dev).unsqueeze(0).expand(b, 3, 4).contiguous()
        cam = dict(poses=eye, focal=torch.ones(b, 2, device=dev), c=torch.zeros(b, 2, device=dev),
                   scale_x=float(scale[0]), scale_y=float(scale[1]), n_objects=b, n_views=1)
        xyz = torch.cat((uv, -torch.ones(b, n, 1, device=dev)), dim=-1).contiguous()
        nhwc = lat.detach().permute(0, 2, 3, 1)
        if not nhwc.is_contiguous():
            nhwc = _lib.prepare_latent(lat.detach(), torch.float32)
        feat_z, _ = _lib.point_features(cam, nhwc, n, xyz=xyz, viewdirs=torch.zeros_like(xyz), ldx=44)
        return feat_z.reshape(b, n, -1).permute(0, 2, 1)

    def forward(self, x):
        """(B,3,H,W) image -> latent (B, latent_size, H/2, W/2), cached in ``self.latent``."""
        if self.feature_scale != 1.0:
            up = self.feature_scale > 1.0
            x = F.interpolate(x, scale_factor=self.feature_scale, mode="bilinear" if up else "area",
                              align_corners=True if up else None, recompute_scale_factor=True)
        x = x.to(device=self.latent.device)
        m = self.model
        x = m.relu(m.bn1(m.conv1(x)))
        latents = [x]
        if self.num_layers > 1:
            if self.use_first_pool:
                x = m.maxpool(x)
            x = m.layer1(x)
            latents.append(x)
        for i, layer in enumerate((m.layer2, m.layer3, m.layer4)):
            if self.num_layers > i + 2:
                x = layer(x)
                latents.append(x)
        self.latents = latents
        fused = (x.is_cuda and self.upsample_interp == "bilinear" and all(l.shape[1] % 32 == 0 for l in latents)
                 and not (torch.is_grad_enabled() and any(l.requires_grad for l in latents)))
        if fused:
            # inference hand-off: resize + concat + channels-last conversion in one kernel
            # (pn_encoder_latent); ``latent`` stays an NCHW-shaped tensor, as a view of channels-last memory,
            # which PixelNeRFNet.latent_nhwc recognises and uses without a transpose
            from .. import _lib
            self.latent = _lib.encoder_latent(latents, torch.float32).permute(0, 3, 1, 2)
        else:
            size = latents[0].shape[-2:]
            latents = [F.interpolate(l, size, mode=self.upsample_interp, align_corners=True) for l in latents]
            self.latent = torch.cat(latents, dim=1)
        # latent_scaling = [W_l, H_l] / ([W_l, H_l] - 1) * 2 in fp32 (encoder.py:131-133), evaluated on the host and
        # written with two fills: no host -> device copy, so a training step can be captured in a CUDA graph
        self._latent_scaling_host = latent_scaling_host(self.latent.shape[-1], self.latent.shape[-2])
        ls = torch.empty(2, dtype=torch.float32, device=self.latent.device)
        ls[0].fill_(self._latent_scaling_host[0])
        ls[1].fill_(self._latent_scaling_host[1])
        self.latent_scaling = ls
        return self.latent

    @classmethod
    def from_conf(cls, conf):
        return cls(conf.get_string("backbone"), pretrained=conf.get_bool("pretrained", True),
                   num_layers=conf.get_int("num_layers", 4),
                   index_interp=conf.get_string("index_interp", "bilinear"),
                   index_padding=conf.get_string("index_padding", "border"),
                   upsample_interp=conf.get_string("upsample_interp", "bilinear"),
                   feature_scale=conf.get_float("feature_scale", 1.0),
                   use_first_pool=conf.get_bool("use_first_pool", True))
