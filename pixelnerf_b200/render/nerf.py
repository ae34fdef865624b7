"""NeRFRenderer -- drop-in for the reference's src/render/nerf.py on sm_100a kernels.

Same constructor / from_conf keys, mutable ``n_coarse`` / ``n_fine`` / ``using_fine``
read at call time, persistent buffers ``iter_idx`` / ``last_sched``, ``sched_step``,
``bind_parallel`` and output structure.  Per ray chunk the forward is
    pn_sample_coarse -> pn_point_features -> MLP -> pn_composite_fwd
    -> pn_sample_fine (importance + depth + merge sort) -> features -> MLP -> composite
with no host synchronisation in between; on the tensor-core path without gradients the whole
chunk is one C call (pn_render_rays).
"""
import os

import torch

from .. import _lib
from .._prof import prof_range


class DotMap(dict):
    """Attribute dict with auto-vivification and toDict(): the slice of dotmap.DotMap the
    reference's callers rely on (``outputs.fine`` exists and is empty when n_fine == 0)."""

    def __getattr__(self, k):
        if k.startswith("__"):
            raise AttributeError(k)
        if k not in self:
            self[k] = DotMap()
        return self[k]

    def __setattr__(self, k, v):
        self[k] = v

    def toDict(self):
        return {k: (v.toDict() if isinstance(v, DotMap) else v) for k, v in self.items()}


class _RenderWrapper(torch.nn.Module):
    """Reference nerf.py:15-42."""

    def __init__(self, net, renderer, simple_output):
        super().__init__()
        self.net, self.renderer, self.simple_output = net, renderer, simple_output

    def forward(self, rays, want_weights=False):
        if rays.shape[0] == 0:
            return torch.zeros(0, 3, device=rays.device), torch.zeros(0, device=rays.device)
        outputs = self.renderer(self.net, rays, want_weights=want_weights and not self.simple_output)
        if self.simple_output:
            part = outputs.fine if self.renderer.using_fine else outputs.coarse
            return part.rgb, part.depth
        return outputs.toDict()


class NeRFRenderer(torch.nn.Module):
    def __init__(self, n_coarse=128, n_fine=0, n_fine_depth=0, noise_std=0.0, depth_std=0.01,
                 eval_batch_size=100000, white_bkgd=False, lindisp=False, sched=None):
        super().__init__()
        self.n_coarse, self.n_fine, self.n_fine_depth = n_coarse, n_fine, n_fine_depth
        self.noise_std, self.depth_std = noise_std, depth_std
        self.eval_batch_size = eval_batch_size
        self.white_bkgd, self.lindisp = white_bkgd, lindisp
        if lindisp:
            print("Using linear displacement rays")
        self.using_fine = n_fine > 0
        self.sched = sched if sched is not None and len(sched) > 0 else None
        self.register_buffer("iter_idx", torch.tensor(0, dtype=torch.long), persistent=True)
        self.register_buffer("last_sched", torch.tensor(0, dtype=torch.long), persistent=True)
        # Test hook: dict(u_coarse, u_bin, u_in, noise) consumed by the next forward() instead of
        # in-kernel draws (the reference's four torch.rand* calls, nerf.py:111,135,141,158).
        self.injected_draws = None
        self.injected_z = None  # teacher forcing: dict(z_coarse=..., z_fine=...)
        self.max_rows_per_launch = 1 << 22   # MLP rows per chunk: 0.5 GB of code rows (the two-launch NS > 1 form of the
                                             # training forward adds 4.3 GB of fp16 hand-off; inference has none)
        self.max_rows_train = 1 << 16  # rows whose 11 fp32 activation tensors are kept at once (1.4 GB)
        self.max_rows_train_tc = 1 << 20  # tensor-core training path: ~20 KB per row of activations + workspace
        # tensor-core training path: the forward keeps its GEMM operands (11.3 KB per MLP row) for the backward
        # when they fit in this many bytes; above it the backward re-runs the forward chunk by chunk
        self.train_stash_bytes = 64 << 30
        # training: the depth-guided samples are differentiable w.r.t. the coarse depth like in the reference
        # (nerf.py:150-161, 290-293); False treats every sample position as a constant
        self.depth_sample_grad = True
        self._keep_parts = False   # set by render/autograd.py: sample_fine also returns the unsorted depth samples
        # multi-GPU wrappers (pixelnerf_b200/parallel.py): every shard of a call derives its draws from one call
        # seed offset by the shard's first ray; consumed by the next forward()
        self.seed_offset = 0
        self.seed_override = None
        # CUDA-graph capture of a training step (use_device_seed): the call seed lives in device memory and is
        # advanced by a device-side add at the start of every forward, so a replayed graph draws new jitter
        self.device_seed = None

    _SEED_STRIDE = -7046029254386353131    # 0x9E3779B97F4A7C15 as int64: consecutive calls land far apart

    def use_device_seed(self, device=None, enable=True):
        """Keep the call seed of the in-kernel draws in device memory (one int64 word, initialised from torch's CPU
        generator like `_seed`).  Needed to capture a training step in a CUDA graph: the reference draws with
        torch.rand_like / randn_like (src/render/nerf.py:112,139,158), whose generator state torch advances across
        graph replays; the equivalent here is this word plus a captured `add_`."""
        if not enable:
            self.device_seed = None
            return self
        device = device if device is not None else self.last_sched.device
        self.device_seed = torch.full((1,), self._seed(), dtype=torch.int64, device=device)
        return self

    # ---------------------------------------------------------------- sampling
    @staticmethod
    def _seed():
        # one 63-bit seed per call from torch's CPU generator (honours torch.manual_seed)
        return int(torch.empty((), dtype=torch.int64).random_().item())

    def sample_coarse(self, rays, u=None, seed=None):
        """Stratified samples (B,8) -> (B,Kc).  Reference nerf.py:98-118."""
        return _lib.sample_coarse(rays, self.n_coarse, self.lindisp, u=u,
                                  seed=self._seed() if seed is None and u is None else (seed or 0))

    def sample_fine(self, rays, weights, u_bin=None, u_in=None, seed=None):
        """Importance samples (B, Kf-Kfd).  Reference nerf.py:120-148."""
        n_imp = self.n_fine - self.n_fine_depth
        z0 = torch.zeros(rays.shape[0], self.n_coarse, device=rays.device)
        _, z_imp, _ = _lib.sample_fine(rays, z0, weights, None, n_imp, 0, self.depth_std, self.lindisp, u_bin=u_bin,
                                       u_in=u_in, seed=self._seed() if seed is None else seed, want_parts=True)
        return z_imp

    def sample_fine_depth(self, rays, depth, noise=None, seed=None):
        """Samples around the coarse depth (B, Kfd).  Reference nerf.py:150-161."""
        z0 = torch.zeros(rays.shape[0], 1, device=rays.device)
        _, _, z_dep = _lib.sample_fine(rays, z0, None, depth, 0, self.n_fine_depth, self.depth_std, self.lindisp,
                                       noise=noise, seed=self._seed() if seed is None else seed, want_parts=True)
        return z_dep

    # --------------------------------------------------------------- composite
    def noisy_sigma(self, rgbs, sigma_noise):
        """sigmas + randn * noise_std in training mode (reference nerf.py:225-226); rgbs (B,K,4), draws (B,K)."""
        rgbs = rgbs.clone()
        rgbs[..., 3].add_(sigma_noise, alpha=float(self.noise_std))
        return rgbs

    def composite(self, model, rays, z_samp, coarse=True, sb=0, sigma_noise=None):
        """Model evaluation + alpha compositing.  Reference nerf.py:163-249.
        rays (B,8), z_samp (B,K) -> weights (B,K), rgb (B,3), depth (B).  sigma_noise (B,K): the standard normal
        draws of nerf.py:226 when training with noise_std > 0 (drawn here if not given)."""
        with prof_range("renderer_composite"):
            b, k = z_samp.shape
            n_obj = max(sb, 1)
            rgbs = model.run_points(coarse, (b // n_obj) * k, rays=rays, z=z_samp).reshape(b, k, 4)
            if self.training and self.noise_std > 0.0:
                rgbs = self.noisy_sigma(rgbs, torch.randn_like(z_samp) if sigma_noise is None else sigma_noise)
            return _lib.composite_fwd(rgbs, z_samp, rays, self.white_bkgd)

    # ----------------------------------------------------------------- forward
    def _fused_ok(self, model, forced):
        """The whole chunk can go through pn_render_rays: tensor-core MLP, nothing teacher-forced, and no
        training forward that has to keep intermediates for the backward."""
        return (model.mlp_precision == "fp16" and not forced and not self._keep_parts
                and not (self.training and self.noise_std > 0.0)
                and getattr(model, "_train_stash", None) is None and os.environ.get("PIXELNERF_B200_STAGED") != "1")

    def _render_fused(self, model, rays, draws, seed):
        if model._cam is None:
            raise RuntimeError("call encode() before forward()")
        prec = model.mlp_precision
        _, hl, wl, _ = model.latent_nhwc(torch.float32).shape
        mc = (model.mlp_coarse.packed(prec), model.latent_tables(True))
        fine_mlp = model._mlp(False)
        mf = None if fine_mlp is model.mlp_coarse else (fine_mlp.packed(prec), model.latent_tables(False))
        n_dep = self.n_fine_depth if self.using_fine else 0
        n_imp = self.n_fine - n_dep if self.using_fine else 0
        coarse, fine = _lib.render_rays(model._cam, hl, wl, mc, mf, rays, self.n_coarse, n_imp, n_dep, self.depth_std,
                                        self.lindisp, self.white_bkgd, seed=seed, draws=draws)
        if fine is not None:
            fine = fine + (fine[3].new_zeros(rays.shape[0], 0),)
        return coarse, fine

    def render_frames(self, model, poses, width, height, focal, z_near, z_far, c=None, draws=None):
        """SURVEY 8f.1 -- ray generation fused into the renderer: render every pixel of the target frames given by
        camera-to-world ``poses`` (NV, 3|4, 4) and the intrinsics -- the arguments of ``util.gen_rays``
        (src/util/util.py:238-276) -- without ever building the (NV, H, W, 8) ray tensor: the kernels compute each ray
        from the pose where they need it (pn_render_frames).  Same result, bit for bit, as
        ``self(model, util.gen_rays(poses, ...).reshape(1, -1, 8))`` with the same draws / seed and chunking.  Eval-mode
        tensor-core path only (one encoded object).  Returns DotMap{coarse:{rgb (NV,H,W,3), depth (NV,H,W)}, fine:{...}}."""
        if model.mlp_precision != "fp16" or (torch.is_grad_enabled() and self.training):
            raise NotImplementedError("render_frames: the no-grad tensor-core path only; build rays with util.gen_rays")
        if model._cam is None:
            raise RuntimeError("call encode() before render_frames()")
        if model._cam["n_objects"] != 1:
            raise NotImplementedError("render_frames renders target views of ONE encoded object")
        with prof_range("renderer_forward"), torch.no_grad():
            frames = _lib.frames_struct(poses.to(model.poses.device), width, height, focal, z_near, z_far, c=c)
            nv = frames[1].shape[0]
            n = nv * height * width
            ns = model.num_views_per_obj
            k_max = self.n_coarse + (self.n_fine if self.using_fine else 0)
            chunk = max(1, self.max_rows_per_launch // max(1, ns * k_max))
            seed = 0 if (draws and "u_coarse" in draws) else self._seed() + int(self.seed_offset)
            prec = model.mlp_precision
            _, hl, wl, _ = model.latent_nhwc(torch.float32).shape
            mc = (model.mlp_coarse.packed(prec), model.latent_tables(True))
            fine_mlp = model._mlp(False)
            mf = None if fine_mlp is model.mlp_coarse else (fine_mlp.packed(prec), model.latent_tables(False))
            n_dep = self.n_fine_depth if self.using_fine else 0
            n_imp = self.n_fine - n_dep if self.using_fine else 0
            parts = []
            for s in range(0, n, chunk):
                e = min(n, s + chunk)
                dr = {k: v[s:e].contiguous() for k, v in draws.items()} if draws else None
                parts.append(_lib.render_rays(model._cam, hl, wl, mc, mf, None, self.n_coarse, n_imp, n_dep,
                                              self.depth_std, self.lindisp, self.white_bkgd, seed=seed + s, draws=dr,
                                              frames=frames, first_ray=s, n_rays=e - s))

            def fmt(which):
                rgb = torch.cat([p[which][1] for p in parts]) if len(parts) > 1 else parts[0][which][1]
                dep = torch.cat([p[which][2] for p in parts]) if len(parts) > 1 else parts[0][which][2]
                return DotMap(rgb=rgb.reshape(nv, height, width, 3), depth=dep.reshape(nv, height, width))
            out = DotMap(coarse=fmt(0))
            if self.using_fine:
                out.fine = fmt(1)
            return out

    def _render_chunk(self, model, rays, sb, draws, forced, want_weights, seed):
        if self._fused_ok(model, forced):
            return self._render_fused(model, rays, draws, seed)
        d = draws or {}
        z_c = forced.get("z_coarse")
        if z_c is None:
            z_c = _lib.sample_coarse(rays, self.n_coarse, self.lindisp, u=d.get("u_coarse"), seed=seed)
        coarse = self.composite(model, rays, z_c, coarse=True, sb=sb, sigma_noise=d.get("sigma_noise_coarse"))
        fine = None
        if self.using_fine:
            n_dep = self.n_fine_depth
            n_imp = self.n_fine - n_dep
            z_f, z_dep = forced.get("z_fine"), forced.get("z_depth")
            if z_f is None:
                z_f = _lib.sample_fine(rays, z_c, forced.get("w_coarse", coarse[0]),
                                       forced.get("depth_coarse", coarse[2]), n_imp, n_dep, self.depth_std,
                                       self.lindisp, u_bin=d.get("u_bin"), u_in=d.get("u_in"), noise=d.get("noise"),
                                       seed=seed, want_parts=self._keep_parts)
                if self._keep_parts:
                    z_f, _, z_dep = z_f
            if z_dep is None:   # teacher-forced z_fine without its depth samples: nothing to differentiate
                z_dep = z_f.new_zeros(z_f.shape[0], 0)
            fine = self.composite(model, rays, z_f, coarse=False, sb=sb,
                                  sigma_noise=d.get("sigma_noise_fine")) + (z_f, z_dep)
        return coarse + (z_c,), fine

    def _render_all(self, model, rays, sb, draws, forced, seed):
        """rays (SB*B,8) -> coarse, fine tuples (weights, rgb, depth, z) with a flat (SB*B) leading axis."""
        b = rays.shape[0] // sb
        ns = getattr(model, "num_views_per_obj", 1)
        k_max = self.n_coarse + (self.n_fine if self.using_fine else 0)
        # chunk along the per-object ray axis so every launch keeps (object, view, point) row order
        chunk = max(1, self.max_rows_per_launch // max(1, sb * ns * k_max))
        if chunk >= b:
            model._stash_range = (0, b)
            return self._render_chunk(model, rays, sb, draws, forced, False, seed)

        def cut(t, s, e):  # (SB*B, ...) -> (SB*(e-s), ...), keeping object-major order
            return t.reshape(sb, b, *t.shape[1:])[:, s:e].reshape(sb * (e - s), *t.shape[1:]).contiguous()

        parts = []
        for s in range(0, b, chunk):
            e = min(b, s + chunk)
            dr = {k: cut(v, s, e) for k, v in draws.items()} if draws else None
            fz = {k: cut(v, s, e) for k, v in forced.items()}
            model._stash_range = (s, e)
            parts.append(self._render_chunk(model, cut(rays, s, e), sb, dr, fz, False, seed + s))

        def join(which):
            if parts[0][which] is None:
                return None
            return tuple(torch.cat([p[which][i].reshape(sb, p[which][i].shape[0] // sb, *p[which][i].shape[1:])
                                    for p in parts], dim=1)
                         .reshape(sb * b, *parts[0][which][i].shape[1:]) for i in range(len(parts[0][which])))
        return join(0), join(1)

    def forward(self, model, rays, want_weights=False):
        """rays (SB,B,8) -> DotMap{coarse:{rgb,depth[,weights]}, fine:{...}}.  Reference nerf.py:251-303."""
        with prof_range("renderer_forward"):
            return self._forward(model, rays, want_weights)

    def _forward(self, model, rays, want_weights):
        if self.sched is not None and self.last_sched.item() > 0:
            self.n_coarse = self.sched[1][self.last_sched.item() - 1]
            self.n_fine = self.sched[2][self.last_sched.item() - 1]
        assert len(rays.shape) == 3
        sb, b = rays.shape[0], rays.shape[1]
        if not rays.is_cuda:
            raise RuntimeError("pixelnerf_b200 renders on CUDA devices only (no CPU fallback)")
        rays = rays.reshape(-1, 8).float().contiguous()
        draws, forced = self.injected_draws, self.injected_z or {}
        self.injected_draws, self.injected_z = None, None
        dev_seed = None
        if draws and "u_coarse" in draws:
            seed = 0
        elif self.seed_override is not None:
            seed = int(self.seed_override)
        elif self.device_seed is not None:
            if self.device_seed.device != rays.device:
                raise RuntimeError("use_device_seed() was set up on another device than the rays")
            dev_seed = self.device_seed
            dev_seed.add_(self._SEED_STRIDE)       # a device op: captured with the step, replayed with it
            seed = int(self.seed_offset)
        else:
            seed = self._seed() + int(self.seed_offset)
        self.seed_override = None
        _lib.set_device_seed(dev_seed)
        try:
            return self._forward_seeded(model, rays, sb, b, draws, forced, seed, want_weights)
        finally:
            _lib.set_device_seed(None)

    def _forward_seeded(self, model, rays, sb, b, draws, forced, seed, want_weights):
        if self.training and self.noise_std > 0.0:
            # sigma noise (nerf.py:225-226): drawn once per call so that the backward sees the forward's draws
            draws = dict(draws or {})
            if "sigma_noise_coarse" not in draws:
                draws["sigma_noise_coarse"] = torch.randn(rays.shape[0], self.n_coarse, device=rays.device)
            if self.using_fine and "sigma_noise_fine" not in draws:
                draws["sigma_noise_fine"] = torch.randn(rays.shape[0], self.n_coarse + self.n_fine, device=rays.device)
        latent = model.encoder.latent
        if model.stop_encoder_grad:
            latent = latent.detach()
        needs_grad = torch.is_grad_enabled() and (latent.requires_grad or any(
            p.requires_grad for m in (model.mlp_coarse, model.mlp_fine) if m is not None for p in m.parameters()))
        if needs_grad:
            # training: hand-written backward behind one autograd node (render/autograd.py)
            from .autograd import RenderFunction, _flat_params
            coarse_p, fine_p = _flat_params(model)
            proxies = getattr(model, "_param_proxies", None)
            if proxies is not None:    # a replica of the in-process multi-GPU wrapper: gradients flow to the
                coarse_p, fine_p = proxies   # primary's parameters through differentiable peer copies
            res = RenderFunction.apply(self, model, rays, sb, draws, forced, seed, latent, *coarse_p, *fine_p)
            coarse, fine = res[:4], (res[4:9] if self.using_fine else None)
        else:
            coarse, fine = self._render_all(model, rays, sb, draws, forced, seed)

        def fmt(part):
            ret = DotMap(rgb=part[1].reshape(sb, b, 3), depth=part[2].reshape(sb, b))
            if want_weights:
                ret.weights = part[0].reshape(sb, b, -1)
            return ret

        outputs = DotMap(coarse=fmt(coarse))
        if self.using_fine:
            outputs.fine = fmt(fine)
        # sample positions of the last call, for the stage-wise parity tests
        self.last_z = (coarse[3].reshape(sb, b, -1), fine[3].reshape(sb, b, -1) if self.using_fine else None)
        self.last_z_depth = fine[4].reshape(sb, b, fine[4].shape[-1]) if self.using_fine else None
        return outputs

    def sched_step(self, steps=1):
        """Reference nerf.py:318-338."""
        if self.sched is None:
            return
        self.iter_idx += steps
        while self.last_sched.item() < len(self.sched[0]) and self.iter_idx.item() >= self.sched[0][self.last_sched.item()]:
            self.n_coarse = self.sched[1][self.last_sched.item()]
            self.n_fine = self.sched[2][self.last_sched.item()]
            print("INFO: NeRF sampling resolution changed on schedule ==> c", self.n_coarse, "f", self.n_fine)
            self.last_sched += 1

    @classmethod
    def from_conf(cls, conf, white_bkgd=False, lindisp=False, eval_batch_size=100000):
        """Reference nerf.py:340-352 (white_bkgd is read with get_float there; kept)."""
        return cls(conf.get_int("n_coarse", 128), conf.get_int("n_fine", 0),
                   n_fine_depth=conf.get_int("n_fine_depth", 0), noise_std=conf.get_float("noise_std", 0.0),
                   depth_std=conf.get_float("depth_std", 0.01), white_bkgd=conf.get_float("white_bkgd", white_bkgd),
                   lindisp=lindisp, eval_batch_size=conf.get_int("eval_batch_size", eval_batch_size),
                   sched=conf.get_list("sched", None))

    def bind_parallel(self, net, gpus=None, simple_output=False):
        """Reference nerf.py:354-371.  With more than one GPU the reference wraps the module in DataParallel
        (dim 1 = rays); here the rays of a call are sharded the same way over resident replicas, from one process
        (``DeviceSharded``) or, under torchrun, one process per GPU (``RaySharded``)."""
        wrapped = _RenderWrapper(net, self, simple_output=simple_output)
        if gpus is not None and len(gpus) > 1:
            import torch.distributed as dist
            from ..parallel import DeviceSharded, RaySharded
            if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
                wrapped = RaySharded(wrapped)           # torchrun: one process per GPU, NCCL gather
            else:
                wrapped = DeviceSharded(wrapped, gpus)  # one process driving all of `gpus`, like the reference
        return wrapped
