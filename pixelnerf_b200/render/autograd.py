"""Training path: one torch.autograd.Function around the render kernels.

The reference trains by autograd over its whole torch op graph (train/train.py:264-281:
``render_par(all_rays, want_weights=True)`` -> MSE on coarse/fine rgb -> ``loss.backward()``).
Here the forward is the ordinary kernel path and the backward is hand written:

    upstream (g_rgb, g_depth, g_weights) per pass
      -> pn_composite_bwd                    dL/d(rgb, sigma) per sample
      -> pn_mlp_backward_fp32 / _tc          head + 15 linears: weight/bias grads, dL/d feat_z
                                             (mlp_precision fp32: CUDA cores; fp16: tcgen05 dgrad/wgrad GEMMs
                                             with bf16 gradients, csrc/pn_train_tc.cu)
      -> pn_point_features_bwd               bilinear scatter-add into the latent gradient
      -> returned to torch autograd          the ResNet34 encoder's backward stays in torch

Activations are not kept from the forward: the backward re-runs the fp32 forward chunk by chunk
(pn_mlp_forward_fp32_train keeps the 11 activation tensors of one chunk), so memory is bounded for
any ray count (config 5's sweep goes to 2^22 rays).

Sample positions: the reference detaches all of them (``weights.detach()``, src/render/nerf.py:130, :288)
except the depth-guided samples of the fine pass, z_d = clamp(depth_coarse + noise * std, near, far)
(:150-161, 290-293), through which the fine loss reaches the coarse MLP.  That path is kept: the fine pass's
backward runs first and pn_depth_sample_bwd turns dL/dz at those samples (compositing + code rows + latent
rows) into an extra dL/d depth_coarse that joins the coarse pass's upstream gradient
(``renderer.depth_sample_grad = False`` drops it).
"""
import torch

from .. import _lib


def _flat_params(model):
    """(coarse tensors, fine tensors or []) in _lib.weight_names() order."""
    names = _lib.weight_names()
    wc = model.mlp_coarse.weight_dict()
    coarse = [wc[k] for k in names]
    fine = []
    if model.mlp_fine is not None:
        wf = model.mlp_fine.weight_dict()
        fine = [wf[k] for k in names]
    return coarse, fine


# precisions whose TRAINING path is the tensor-core one (fp16 forward keeping its operands, bf16 dgrad / wgrad GEMMs);
# the two-pass chain (fp16x2) is an inference form of the same path.  fp32 and fp16x3 train with the fp32 kernels.
_TC_TRAIN = ("fp16", "fp16x2")


class RenderFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, renderer, model, rays, sb, draws, forced, seed, latent, *params):
        ns = getattr(model, "num_views_per_obj", 1)
        k_all = renderer.n_coarse + ((renderer.n_coarse + renderer.n_fine) if renderer.using_fine else 0)
        stash_bytes = rays.shape[0] * ns * k_all * (11 * 1024 + 160)
        keep = model.mlp_precision in _TC_TRAIN and stash_bytes <= renderer.train_stash_bytes
        model._train_stash = [] if keep else None
        renderer._keep_parts = bool(renderer.depth_sample_grad and renderer.n_fine_depth > 0)
        try:
            with torch.no_grad():
                coarse, fine = renderer._render_all(model, rays, sb, draws, forced, seed)
        finally:
            ctx.stash, model._train_stash = model._train_stash, None
            renderer._keep_parts = False
        ctx.renderer, ctx.model, ctx.sb = renderer, model, sb
        # What the backward needs of the model is captured NOW (torch's autograd captures tensors the same way): a
        # second encode() or a parameter update between this forward and its backward must not change the camera,
        # latent or tables the gradients are computed against.
        tensor_core = model.mlp_precision in _TC_TRAIN
        ctx.snap = dict(cam=model._cam, ns=ns, latent32=model.latent_nhwc(torch.float32),
                        tables={c: model.latent_tables(c) for c in (True, False)} if tensor_core and not keep else None,
                        versions=tuple(p._version for p in params))
        noisy = renderer.training and renderer.noise_std > 0.0 and draws is not None
        ctx.sigma_noise = (draws.get("sigma_noise_coarse"), draws.get("sigma_noise_fine")) if noisy else (None, None)
        ctx.n_params = len(params)
        ctx.has_fine = fine is not None
        ctx.latent_shape = latent.shape
        z_f = fine[3] if fine is not None else coarse[3].new_zeros(0)
        z_dep = fine[4] if fine is not None else coarse[3].new_zeros(0)
        ctx.save_for_backward(rays, coarse[3], z_f, z_dep)
        outs = tuple(coarse) + (tuple(fine) if fine is not None else ())
        ctx.mark_non_differentiable(coarse[3], *((fine[3], fine[4]) if fine is not None else ()))
        return outs

    @staticmethod
    def backward(ctx, *g):
        renderer, model, sb = ctx.renderer, ctx.model, ctx.sb
        rays, z_c, z_f, z_dep = ctx.saved_tensors
        names = _lib.weight_names()
        coarse_p, fine_p = _flat_params(model)
        snap = ctx.snap
        if getattr(model, "_param_proxies", None) is None and \
                tuple(p._version for p in coarse_p + fine_p) != snap["versions"]:
            raise RuntimeError("one of the variables needed for gradient computation has been modified by an inplace "
                               "operation: ResnetFC parameters changed between the render's forward and its backward")
        dev = rays.device
        want_latent = ctx.needs_input_grad[7]
        g_lat = None
        if want_latent:
            v, ch, hl, wl = ctx.latent_shape
            g_lat = torch.zeros(v, hl, wl, ch, device=dev, dtype=torch.float32)

        def zero_grads(ps):
            # one zero-filled buffer per MLP, the 30 gradient tensors are views of it (one fill instead of 30)
            flat = torch.zeros(sum(p.numel() for p in ps), device=dev, dtype=torch.float32)
            out, off = {}, 0
            for k, p in zip(names, ps):
                out[k] = flat[off:off + p.numel()].view(p.shape)
                off += p.numel()
            return out

        grads_c = zero_grads(coarse_p)
        grads_f = zero_grads(fine_p) if fine_p else grads_c
        stash, ctx.stash = ctx.stash, None
        g_depth_c = g[2]
        if ctx.has_fine:    # fine pass first: its depth samples hand a gradient to the coarse depth
            g_dep = None
            if renderer.depth_sample_grad and z_dep.numel() > 0 and not all(x is None for x in (g[4], g[5], g[6])):
                g_dep = torch.zeros(rays.shape[0], device=dev, dtype=torch.float32)
            _backward_pass(renderer, model, snap, False, rays, z_f, sb, g[4], g[5], g[6], grads_f, g_lat, stash,
                           z_dep=z_dep, g_dep=g_dep, sigma_noise=ctx.sigma_noise[1])
            if g_dep is not None:
                g_depth_c = g_dep if g_depth_c is None else g_depth_c + g_dep
        _backward_pass(renderer, model, snap, True, rays, z_c, sb, g[0], g[1], g_depth_c, grads_c, g_lat, stash,
                       sigma_noise=ctx.sigma_noise[0])
        flat = [grads_c[k] for k in names] + ([grads_f[k] for k in names] if fine_p else [])
        g_latent = g_lat.permute(0, 3, 1, 2) if want_latent else None
        return (None, None, None, None, None, None, None, g_latent, *flat)


def _backward_pass(renderer, model, snap, coarse, rays, z, sb, g_w, g_rgb, g_depth, grads, g_lat, stash=None,
                   z_dep=None, g_dep=None, sigma_noise=None):
    """One pass's backward, chunk by chunk.  g_dep (R) / z_dep (R,Kd): fine pass only -- receives the gradient
    the depth-guided sample positions carry to the coarse depth.  ``snap``: the camera block, channels-last latent
    and lin_z tables the forward used (RenderFunction.forward)."""
    r_total, k = z.shape
    if g_w is None and g_rgb is None and g_depth is None:
        return
    b = r_total // sb
    ns = snap["ns"]
    cam = snap["cam"]
    mlp = model.mlp_coarse if coarse or model.mlp_fine is None else model.mlp_fine
    weights = {n: p.detach() for n, p in mlp.weight_dict().items()}
    packed = mlp.packed("fp32")
    latent = snap["latent32"]
    if g_rgb is None:
        g_rgb = torch.zeros(r_total, 3, device=z.device)
    chunk = max(1, renderer.max_rows_train // max(1, sb * ns * k))

    def cut(t, s, e):
        if t is None:
            return None
        return t.reshape(sb, b, *t.shape[1:])[:, s:e].reshape(sb * (e - s), *t.shape[1:]).contiguous()

    def composited(out, s, e):
        """What the forward composited: the model output, with the forward's sigma noise (nerf.py:225-226)."""
        rs = out.reshape(-1, k, 4)
        return rs if sigma_noise is None else renderer.noisy_sigma(rs, cut(sigma_noise, s, e))

    want_pos = g_dep is not None
    want_fz = g_lat is not None or want_pos
    w_in = weights["lin_in.weight"]

    def positions(rr, zz, s, e, g_z, g_h0, g_fz):
        """dL/dz of this chunk's depth samples -> g_dep[object, s:e]"""
        part = torch.zeros(rr.shape[0], device=rr.device, dtype=torch.float32)
        _lib.depth_sample_bwd(cam, latent, rr, zz, cut(z_dep, s, e), g_z, g_h0, g_fz, w_in, part)
        g_dep.view(sb, b)[:, s:e] += part.view(sb, e - s)

    tensor_core = model.mlp_precision in _TC_TRAIN
    if stash is not None:
        # the forward kept its operands: one backward per forward chunk, nothing is recomputed
        w16 = mlp.packed("bf16_train")
        for ent in stash:
            if ent.get("coarse") is None or ent["coarse"] != coarse:   # other pass, or already consumed
                continue
            s, e = ent["range"]
            out, ppo = ent["out"], ent["ppo"]
            g_rs, g_z = _lib.composite_bwd(composited(out, s, e), ent["z"], ent["rays"], renderer.white_bkgd,
                                           cut(g_rgb, s, e), cut(g_depth, s, e), cut(g_w, s, e), want_gz=want_pos)
            g_fz, g_h0 = _lib.mlp_backward_tc(packed, w16, ent["fx"], latent, sb, ns, ppo, ent["acts"], out,
                                              g_rs.reshape(-1, 4), grads, want_g_feat_z=want_fz, want_g_h0=True)
            if want_pos:
                positions(ent["rays"], ent["z"], s, e, g_z, g_h0, g_fz)
            if g_lat is not None:
                _lib.point_features_bwd(cam, g_lat, ppo, g_fz, rays=ent["rays"], z=ent["z"])
            ent.clear()     # release this chunk's activations
        return
    if tensor_core:
        packed_tc, w16 = mlp.packed("fp16"), mlp.packed("bf16_train")
        tables = snap["tables"][coarse] if snap["tables"] is not None else model.latent_tables(coarse)
        _, hl, wl, _ = latent.shape
        chunk = max(1, renderer.max_rows_train_tc // max(1, sb * ns * k))
    for s in range(0, b, chunk):
        e = min(b, s + chunk)
        rr, zz = cut(rays, s, e), cut(z, s, e)
        ppo = (e - s) * k
        if tensor_core:
            # forward chain on the tensor cores keeping its bf16 GEMM operands, then dgrad / wgrad GEMMs on them
            fx = _lib.point_taps(cam, hl, wl, ppo, rays=rr, z=zz)
            out, acts = _lib.mlp_forward_tc_train(packed_tc, tables, fx, sb, ns, ppo)
            g_rs, g_z = _lib.composite_bwd(composited(out, s, e), zz, rr, renderer.white_bkgd, cut(g_rgb, s, e),
                                           cut(g_depth, s, e), cut(g_w, s, e), want_gz=want_pos)
            g_fz, g_h0 = _lib.mlp_backward_tc(packed, w16, fx, latent, sb, ns, ppo, acts, out, g_rs.reshape(-1, 4),
                                              grads, want_g_feat_z=want_fz, want_g_h0=True)
            if want_pos:
                positions(rr, zz, s, e, g_z, g_h0, g_fz)
            if g_lat is not None:
                _lib.point_features_bwd(cam, g_lat, ppo, g_fz, rays=rr, z=zz)
            continue
        fz, fx = _lib.point_features(cam, latent, ppo, rays=rr, z=zz, ldx=44)
        out, acts = _lib.mlp_forward_fp32_train(packed, fz, fx, sb, ns, ppo)
        g_rs, g_z = _lib.composite_bwd(composited(out, s, e), zz, rr, renderer.white_bkgd, cut(g_rgb, s, e),
                                       cut(g_depth, s, e), cut(g_w, s, e), want_gz=want_pos)
        g_fz, g_h0 = _lib.mlp_backward_fp32(packed, weights, fz, fx, sb, ns, ppo, acts, out, g_rs.reshape(-1, 4),
                                            grads, want_g_feat_z=want_fz, want_g_h0=True)
        if want_pos:
            positions(rr, zz, s, e, g_z, g_h0, g_fz)
        if g_lat is not None:
            _lib.point_features_bwd(cam, g_lat, ppo, g_fz, rays=rr, z=zz)
