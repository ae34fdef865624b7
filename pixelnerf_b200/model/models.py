"""PixelNeRFNet -- same constructor keys, attributes, encode()/forward() contracts
and checkpoint format as the reference's src/model/models.py, with forward()
running on the sm_100a kernels (point features -> ResnetFC -> head).
"""
import os
import os.path as osp
import warnings

import torch

from .. import _lib
from .._prof import prof_range
from .code import PositionalEncoding
from .model_util import make_encoder, make_mlp

_PRECISIONS = ("fp32", "fp16", "fp16x2", "fp16x3", "bf16")


class PixelNeRFNet(torch.nn.Module):
    def __init__(self, conf, stop_encoder_grad=False):
        super().__init__()
        self.encoder = make_encoder(conf["encoder"])
        self.use_encoder = conf.get_bool("use_encoder", True)
        self.use_xyz = conf.get_bool("use_xyz", False)
        assert self.use_encoder or self.use_xyz
        self.normalize_z = conf.get_bool("normalize_z", True)
        self.stop_encoder_grad = stop_encoder_grad
        self.use_code = conf.get_bool("use_code", False)
        self.use_code_viewdirs = conf.get_bool("use_code_viewdirs", True)
        self.use_viewdirs = conf.get_bool("use_viewdirs", False)
        self.use_global_encoder = conf.get_bool("use_global_encoder", False)
        # The kernels implement the configuration every shipped conf resolves to
        # (conf/default.conf:5-29); other switch settings raise instead of falling back.
        bad = [n for n, ok in (("use_encoder=False", self.use_encoder), ("use_xyz=False", self.use_xyz),
                               ("normalize_z=False", self.normalize_z), ("use_code=False", self.use_code),
                               ("use_code_viewdirs=True", not self.use_code_viewdirs),
                               ("use_viewdirs=False", self.use_viewdirs),
                               ("use_global_encoder=True", not self.use_global_encoder)) if not ok]
        if bad:
            raise NotImplementedError("model switches outside the B200 hot path: " + ", ".join(bad))
        d_latent = self.encoder.latent_size
        self.code = PositionalEncoding.from_conf(conf["code"], d_in=3)
        if self.code.num_freqs != 6 or abs(self.code.freq_factor - 1.5) > 1e-12 or not self.code.include_input:
            raise NotImplementedError("positional encoding other than num_freqs=6, freq_factor=1.5, include_input")
        d_in = self.code.d_out + 3
        d_out = 4
        self.latent_size = self.encoder.latent_size
        self.mlp_coarse = make_mlp(conf["mlp_coarse"], d_in, d_latent, d_out=d_out)
        self.mlp_fine = make_mlp(conf["mlp_fine"], d_in, d_latent, d_out=d_out, allow_empty=True)
        self.register_buffer("poses", torch.empty(1, 3, 4), persistent=False)
        self.register_buffer("image_shape", torch.empty(2), persistent=False)
        self.d_in, self.d_out, self.d_latent = d_in, d_out, d_latent
        self.register_buffer("focal", torch.empty(1, 2), persistent=False)
        self.register_buffer("c", torch.empty(1, 2), persistent=False)
        self.num_objs = 0
        self.num_views_per_obj = 1
        # "fp16" (default): tcgen05 chain, 2e-3 parity -- the product path.  "fp16x3": the precise tcgen05 chain (three
        # hi/lo operand passes per GEMM, everything else fp32), 1e-4 parity; inference -- in training mode it runs the
        # fp32 kernels.  "fp16x2": the precise chain without its third pass (weights as fp16 pairs, activations single
        # fp16): 2e-3 parity on every scene class incl. DTU depth; inference -- training mode runs the fp16 path.
        # "fp32": the CUDA-core 1e-4 parity path.
        # Not a reference conf key: the reference's conf files never set it, so they get the tensor-core chain.
        self.mlp_precision = conf.get_string("mlp_precision", os.environ.get("PIXELNERF_B200_MLP", "fp16"))
        assert self.mlp_precision in _PRECISIONS, self.mlp_precision
        self._cam = None
        self._latent_nhwc = {}
        self._tables = {}

    # ------------------------------------------------------------------ encode
    def encode(self, images, poses, focal, z_bounds=None, c=None):
        """Cache the encoder latent and per-view cameras.  Reference models.py:89-144.

        images (NS,3,H,W) or (SB,NS,3,H,W); poses camera-to-world (...,4,4); focal (), (SB) or
        (SB,2); c None, (), (SB) or (SB,2).
        """
        self.num_objs = images.size(0)
        if len(images.shape) == 5:
            assert len(poses.shape) == 4
            assert poses.size(1) == images.size(1)
            self.num_views_per_obj = images.size(1)
            images = images.reshape(-1, *images.shape[2:])
            poses = poses.reshape(-1, 4, 4)
        else:
            self.num_views_per_obj = 1
        self.encoder(images)
        rot = poses[:, :3, :3].transpose(1, 2)
        trans = -torch.bmm(rot, poses[:, :3, 3:])
        self.poses = torch.cat((rot, trans), dim=-1)
        self.image_shape[0].fill_(images.shape[-1])     # fills, not scalar copies: capturable in a CUDA graph
        self.image_shape[1].fill_(images.shape[-2])
        if len(focal.shape) == 0:
            focal = focal[None, None].repeat((1, 2))
        elif len(focal.shape) == 1:
            focal = focal.unsqueeze(-1).repeat((1, 2))
        else:
            focal = focal.clone()
        self.focal = focal.float()
        self.focal[..., 1] *= -1.0
        if c is None:
            c = (self.image_shape * 0.5).unsqueeze(0)
        elif len(c.shape) == 0:
            c = c[None, None].repeat((1, 2))
        elif len(c.shape) == 1:
            c = c.unsqueeze(-1).repeat((1, 2))
        self.c = c
        # the uv -> texel scale from host-side shapes (no device read: the step stays capturable in a CUDA graph)
        import numpy as np
        ls = getattr(self.encoder, "_latent_scaling_host", None)
        host = None
        if ls is not None:
            host = (float(np.float32(ls[0]) / np.float32(images.shape[-1])),
                    float(np.float32(ls[1]) / np.float32(images.shape[-2])))
        self._refresh_camera(host)

    def _refresh_camera(self, host_scale=None):
        """Kernel-side camera block: per-view expanded focal / c and the uv -> texel scale (`host_scale`: that scale
        when the caller already has it on the host; otherwise it is read back from the device buffers)."""
        n_views_total = self.poses.shape[0]
        ns = self.num_views_per_obj
        dev = self.poses.device

        def per_view(t):  # (SB or 1, 2) -> (SB*NS, 2), the repeat_interleave of models.py:207-212
            t = t.to(device=dev, dtype=torch.float32)
            if t.shape[0] == 1:
                return t.expand(n_views_total, 2).contiguous()
            return t.unsqueeze(1).expand(-1, ns, 2).reshape(-1, 2).contiguous()

        if host_scale is not None:
            scale = host_scale
        else:
            ls = self.encoder.latent_scaling.float().cpu()
            img = self.image_shape.float().cpu()
            scale = ls / img  # encoder.py:98: latent_scaling / image_size, fp32 like the reference
        self._cam = dict(poses=self.poses.float().contiguous(), focal=per_view(self.focal), c=per_view(self.c),
                         scale_x=float(scale[0]), scale_y=float(scale[1]),
                         n_objects=n_views_total // ns, n_views=ns)
        self._latent_nhwc = {}
        self._tables = {}

    def latent_nhwc(self, dtype):
        """Channels-last copy of encoder.latent in `dtype`, made once per encode()."""
        lat = self.encoder.latent
        key = (lat.data_ptr(), lat._version, tuple(lat.shape))
        hit = self._latent_nhwc.get(dtype)
        if hit is None or hit[0] != key:
            src = lat.detach()
            nhwc = src.permute(0, 2, 3, 1)
            if lat.dim() == 4 and lat.dtype == torch.float32 and lat.is_cuda and nhwc.is_contiguous() \
                    and not src.is_contiguous():
                # the encoder already wrote channels-last memory (pn_encoder_latent): no transpose
                t = nhwc if dtype == torch.float32 else nhwc.to(dtype)
            else:
                t = _lib.prepare_latent(src, dtype)
            self._latent_nhwc[dtype] = hit = (key, t)
        return hit[1]

    def latent_tables(self, coarse, fp32=False):
        """Per-texel lin_z tables of one MLP for the tensor-core chain (lin_z of a bilinear blend = blend of
        lin_z of the texels, resnetfc.py:174-180 after encoder.py:102-108): rebuilt when the latent or any
        parameter of that MLP changed."""
        mlp = self._mlp(coarse)
        # training mode rebuilds the tables every optimiser step: tensor-core product from the fp16 latent
        # (hi + lo weights); eval mode builds them once per encode in fp32
        lat = self.latent_nhwc(torch.float16 if (self.training and not fp32) else torch.float32)
        key = (self._latent_nhwc[lat.dtype][0], lat.dtype, mlp.version_key())
        slot = ("coarse" if mlp is self.mlp_coarse else "fine") + ("32" if fp32 else "")
        hit = self._tables.get(slot)
        if hit is None or hit[0] != key:
            params = mlp.weight_dict()
            build = _lib.prepare_latent_tables_f32 if fp32 else _lib.prepare_latent_tables
            hit = (key, build({k: v.detach() for k, v in params.items()}, lat))
            self._tables[slot] = hit
        return hit[1]

    def invalidate(self):
        """Drop every kernel-side cache (packed weights, per-texel tables, channels-last latent).  Needed only after
        writes the version counters cannot see: ``p.data.copy_()`` / ``mul_()`` on a parameter, or a call of
        ``self.encoder(images)`` without ``encode()`` that re-uses the latent's storage."""
        self._latent_nhwc, self._tables = {}, {}
        for m in (self.mlp_coarse, self.mlp_fine):
            if m is not None:
                m.invalidate()

    def load_state_dict(self, *args, **kwargs):
        out = super().load_state_dict(*args, **kwargs)
        self.invalidate()
        return out

    # ----------------------------------------------------------------- forward
    def _mlp(self, coarse):
        return self.mlp_coarse if coarse or self.mlp_fine is None else self.mlp_fine

    def run_points(self, coarse, ppo, *, rays=None, z=None, xyz=None, viewdirs=None):
        """Feature gather + MLP + head for SB*ppo points -> (SB*ppo, 4) [rgb, sigma]."""
        if self._cam is None:
            raise RuntimeError("call encode() before forward()")
        with prof_range("model_inference"):
            return self._run_points(coarse, ppo, rays, z, xyz, viewdirs)

    def _run_points(self, coarse, ppo, rays, z, xyz, viewdirs):
        prec = self.mlp_precision
        if prec == "fp16x3" and (self.training or getattr(self, "_train_stash", None) is not None):
            prec = "fp32"      # the precise chain is an inference path; training uses the fp32 kernels (fwd and bwd)
        if prec == "fp16x2" and (self.training or getattr(self, "_train_stash", None) is not None):
            prec = "fp16"      # the two-pass chain is an inference path; training uses the tensor-core fp16 path
        mlp = self._mlp(coarse)
        cam = self._cam
        if prec == "fp32":
            with prof_range("encoder_index"):
                fz, fx = _lib.point_features(cam, self.latent_nhwc(torch.float32), ppo, rays=rays, z=z, xyz=xyz,
                                             viewdirs=viewdirs, ldx=44)
            with prof_range("resnetfc_infer"):
                return _lib.mlp_forward_fp32(mlp.packed("fp32"), fz, fx, cam["n_objects"], cam["n_views"], ppo)
        _, hl, wl, _ = self.latent_nhwc(torch.float32).shape
        if prec in ("fp16x3", "fp16x2"):     # fp16x2: the same chain without its third operand pass
            with prof_range("encoder_index"):
                fx, fx_lo = _lib.point_taps_f32(cam, hl, wl, ppo, rays=rays, z=z, xyz=xyz, viewdirs=viewdirs)
            with prof_range("resnetfc_infer"):
                return _lib.mlp_forward_tc_x3(mlp.packed("fp16x3"), self.latent_tables(coarse, fp32=True), fx, fx_lo,
                                              cam["n_objects"], cam["n_views"], ppo, passes=3 if prec == "fp16x3" else 2)
        tables = self.latent_tables(coarse)
        stash = getattr(self, "_train_stash", None)
        if cam["n_views"] > 1 and not self.training and not (stash is not None and rays is not None) and \
                os.environ.get("PIXELNERF_B200_TWO_LAUNCH") != "1":
            # several views, inference (eval mode): rows in grouped order, view mean on chip, one chain launch.
            # Training mode keeps the two-launch form, whose saved operands / recomputation the backward uses.
            with prof_range("encoder_index"):
                fx = _lib.point_taps_grouped(cam, hl, wl, ppo, rays=rays, z=z, xyz=xyz, viewdirs=viewdirs)
            with prof_range("resnetfc_infer"):
                return _lib.mlp_forward_tc_grouped(mlp.packed(prec), tables, fx, cam["n_objects"], cam["n_views"], ppo)
        with prof_range("encoder_index"):
            fx = _lib.point_taps(cam, hl, wl, ppo, rays=rays, z=z, xyz=xyz, viewdirs=viewdirs)
        if stash is not None and rays is not None:
            # training forward (render/autograd.py): keep what the tensor-core backward needs, so that it does
            # not have to re-run the chain
            out, acts = _lib.mlp_forward_tc_train(mlp.packed(prec), tables, fx, cam["n_objects"], cam["n_views"], ppo)
            stash.append(dict(coarse=coarse, ppo=ppo, rays=rays, z=z, fx=fx, out=out, acts=acts,
                              range=getattr(self, "_stash_range", None)))
            return out
        with prof_range("resnetfc_infer"):
            return _lib.mlp_forward_tc(mlp.packed(prec), tables, fx, cam["n_objects"], cam["n_views"], ppo)

    def forward(self, xyz, coarse=True, viewdirs=None, far=False):
        """(SB,B,3) world points -> (SB,B,4) [sigmoid rgb, relu sigma].  Reference models.py:146-266."""
        sb, b, _ = xyz.shape
        assert viewdirs is not None, "use_viewdirs=True needs viewdirs (reference models.py:186)"
        assert sb == self._cam["n_objects"], "xyz batch must match the encoded objects"
        out = self.run_points(coarse, b, xyz=xyz.reshape(sb, b, 3), viewdirs=viewdirs.reshape(sb, b, 3))
        return out.reshape(sb, b, 4)

    # -------------------------------------------------------------- checkpoints
    def load_weights(self, args, opt_init=False, strict=True, device=None):
        """Reference models.py:268-298: <checkpoints_path>/<name>/pixel_nerf_{init,latest}."""
        if opt_init and not args.resume:
            return
        ckpt_name = "pixel_nerf_init" if opt_init or not args.resume else "pixel_nerf_latest"
        model_path = "%s/%s/%s" % (args.checkpoints_path, args.name, ckpt_name)
        if device is None:
            device = self.poses.device
        if os.path.exists(model_path):
            print("Load", model_path)
            self.load_state_dict(torch.load(model_path, map_location=device), strict=strict)
        elif not opt_init:
            warnings.warn("WARNING: {} does not exist, not loaded!! Model will be re-initialized.".format(model_path))
        return self

    def save_weights(self, args, opt_init=False):
        """Reference models.py:300-316 (keeps one backup copy)."""
        from shutil import copyfile
        ckpt_name = "pixel_nerf_init" if opt_init else "pixel_nerf_latest"
        backup_name = "pixel_nerf_init_backup" if opt_init else "pixel_nerf_backup"
        ckpt_path = osp.join(args.checkpoints_path, args.name, ckpt_name)
        if osp.exists(ckpt_path):
            copyfile(ckpt_path, osp.join(args.checkpoints_path, args.name, backup_name))
        torch.save(self.state_dict(), ckpt_path)
        return self
