"""Generate tests/golden/*.npz from the LIVE reference (authoring container only).

Run:  python tests/golden/make_golden.py            (needs /root/reference)

The reference is imported unmodified from /root/reference/src with two small
stand-ins for its absent third-party imports (``dotmap.DotMap``,
``pyhocon.ConfigTree``); its renderer is fed pre-generated random draws by
temporarily replacing ``torch.rand`` / ``rand_like`` / ``randn_like`` (the
consumption order is src/render/nerf.py:111,135,141,158).  The encoder trunk is
outside the hot path: ``net.encode`` runs for real (so every camera buffer is
the reference's), then ``encoder.latent`` is overwritten with the synthetic
latent so the fixture does not depend on conv/batch-norm bit patterns.

Fixtures hold only seeds, small inputs and the reference's outputs; weights and
latents are regenerated from the seeds by pixelnerf_b200.synthetic.
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference/src"
sys.path.insert(0, ROOT)


def install_shims():
    class DotMap(dict):
        def __init__(self, *a, **k):
            super().__init__()
            for key, v in dict(*a, **k).items():
                self[key] = v

        def __setitem__(self, key, v):
            if isinstance(v, dict) and not isinstance(v, DotMap):
                v = DotMap(v)
            super().__setitem__(key, v)

        def __getattr__(self, key):
            if key.startswith("__"):
                raise AttributeError(key)
            if key not in self:
                self[key] = DotMap()
            return self[key]

        __setattr__ = __setitem__

        def toDict(self):
            return {k: (v.toDict() if isinstance(v, DotMap) else v) for k, v in self.items()}

    class ConfigTree(dict):
        def _get(self, key, default=KeyError):
            cur = self
            for part in key.split("."):
                if not isinstance(cur, dict) or part not in cur:
                    if default is KeyError:
                        raise KeyError(key)
                    return default
                cur = dict.__getitem__(cur, part)
            return ConfigTree(cur) if isinstance(cur, dict) and not isinstance(cur, ConfigTree) else cur

        def __getitem__(self, key):
            return self._get(key)

        def get_int(self, k, d=KeyError):
            return int(self._get(k, d)) if self._get(k, d) is not None else None

        def get_float(self, k, d=KeyError):
            return float(self._get(k, d))

        def get_bool(self, k, d=KeyError):
            return bool(self._get(k, d))

        def get_string(self, k, d=KeyError):
            return self._get(k, d)

        def get_list(self, k, d=KeyError):
            return self._get(k, d)

    m = types.ModuleType("dotmap")
    m.DotMap = DotMap
    sys.modules["dotmap"] = m
    p = types.ModuleType("pyhocon")
    p.ConfigTree = ConfigTree
    p.ConfigFactory = None
    sys.modules["pyhocon"] = p
    return ConfigTree


def reference_conf(ConfigTree, use_first_pool=True, white_bkgd=True):
    """conf/default_mv.conf merged over conf/default.conf (+ exp overrides), pretrained off."""
    mlp = dict(type="resnet", n_blocks=5, d_hidden=512, combine_layer=3, combine_type="average")
    return ConfigTree(
        model=dict(
            use_encoder=True, use_global_encoder=False, use_xyz=True, canon_xyz=False, use_code=True,
            code=dict(num_freqs=6, freq_factor=1.5, include_input=True),
            use_viewdirs=True, use_code_viewdirs=False, mlp_coarse=dict(mlp), mlp_fine=dict(mlp),
            encoder=dict(backbone="resnet34", pretrained=False, num_layers=4, use_first_pool=use_first_pool),
        ),
        renderer=dict(n_coarse=64, n_fine=32, n_fine_depth=16, depth_std=0.01, sched=[], white_bkgd=white_bkgd),
    )


class DrawInjector:
    """Feeds the renderer a fixed list of tensors instead of fresh random draws."""

    def __init__(self, tensors):
        self.q = list(tensors)

    def __enter__(self):
        self.saved = (torch.rand, torch.rand_like, torch.randn_like)
        pop = lambda *a, **k: self.q.pop(0).clone()
        torch.rand = pop
        torch.rand_like = pop
        torch.randn_like = pop
        return self

    def __exit__(self, *exc):
        torch.rand, torch.rand_like, torch.randn_like = self.saved


def load_reference():
    ConfigTree = install_shims()
    sys.path.insert(0, REF)
    import warnings
    warnings.filterwarnings("ignore")
    from model import make_model  # noqa: E402  (reference)
    from render import NeRFRenderer  # noqa: E402  (reference)
    return ConfigTree, make_model, NeRFRenderer


def build_reference_net(ConfigTree, make_model, conf, wc, wf):
    net = make_model(conf["model"])
    sd = net.state_dict()
    for k, v in wc.items():
        sd["mlp_coarse." + k].copy_(v)
    for k, v in wf.items():
        sd["mlp_fine." + k].copy_(v)
    return net.eval()


GRAD_SUB = 16   # stride of the subsample kept of every 512 x 512 gradient (32 x 32 entries)


def sub_grad(name, g):
    """What a grad_*.npz fixture keeps of one gradient tensor: small tensors whole, 512 x 512 matrices strided."""
    g = g.detach()
    if g.dim() == 2 and g.shape[0] == 512 and g.shape[1] == 512:
        return g[::GRAD_SUB, ::GRAD_SUB].contiguous().numpy()
    return g.numpy()


def case_render(name, ConfigTree, make_model, NeRFRenderer, *, h, w, ns, sb, n_rays, seed, focal, c,
                near_far, white, lindisp, hl=None, wl=None, per_ray_bounds=False, with_grad=False):
    """One full coarse+fine reference render on a synthetic scene; returns the fixture dict.  with_grad: also the
    reference's own autograd gradients of the training loss (train/train.py:264-281: MSE of the coarse and of the
    fine rgb against a target) w.r.t. both MLPs' parameters and the encoder latent."""
    from pixelnerf_b200 import synthetic as syn
    conf = reference_conf(ConfigTree, use_first_pool=not (h == 64 and w == 64), white_bkgd=white)
    wc, wf = syn.mlp_weights(seed + 1), syn.mlp_weights(seed + 2)
    net = build_reference_net(ConfigTree, make_model, conf, wc, wf)
    imgs = syn.images(seed + 3, sb, ns, h, w)
    g = np.random.Generator(np.random.PCG64(seed + 4))
    az = g.uniform(-60, 60, (sb, ns))
    src_poses = torch.stack([torch.stack([syn.pose_spherical(float(a), -10.0, 2.0) for a in row]) for row in az])
    focal_t = torch.tensor(focal, dtype=torch.float32)
    c_t = None if c is None else torch.tensor(c, dtype=torch.float32)
    with torch.no_grad():
        net.encode(imgs, src_poses, focal_t, c=c_t)
        enc_shape = tuple(net.encoder.latent.shape)
        lat = syn.latent_maps(seed + 5, sb * ns, enc_shape[2], enc_shape[3])
        assert tuple(lat.shape) == enc_shape, (lat.shape, enc_shape)
        net.encoder.latent = lat
    # target rays: random pixels of one target view per object
    rays = []
    for o in range(sb):
        tp = syn.pose_spherical(float(g.uniform(-80, 80)), -10.0, 2.0)
        f_arg = focal if np.isscalar(focal) else (focal[0] if np.ndim(focal) == 1 else focal[o])
        c_arg = None if c is None else (c if np.ndim(c) == 1 else c[o])
        full = syn.gen_rays(tp, w, h, f_arg, near_far[0], near_far[1], c_arg).reshape(-1, 8)
        pick = torch.from_numpy(g.choice(full.shape[0], n_rays, replace=False))
        rays.append(full[pick])
    rays = torch.stack(rays)  # (SB, B, 8)
    if per_ray_bounds:
        u = torch.from_numpy(g.random((sb, n_rays, 2), dtype=np.float32))
        rays[..., 6] = near_far[0] + u[..., 0] * 0.4
        rays[..., 7] = near_far[1] - u[..., 1] * 0.5
    renderer = NeRFRenderer.from_conf(conf["renderer"], lindisp=lindisp, eval_batch_size=50000).eval()
    renderer.white_bkgd = white
    par = renderer.bind_parallel(net, [0], simple_output=False).eval()
    R = sb * n_rays
    d = syn.draws(seed + 6, R, 64, 16, 16)
    with torch.no_grad(), DrawInjector([d["u_coarse"], d["u_bin"], d["u_in"], d["noise"]]):
        out = par(rays, want_weights=True)
    # extra: raw network outputs on the coarse samples of the first 8 rays of each object
    fx = dict(
        name=name, h=h, w=w, ns=ns, sb=sb, seed=seed, white=white, lindisp=lindisp,
        rays=rays.numpy(), src_poses=src_poses.numpy(),
        focal=np.asarray(focal, np.float32), c=np.zeros(0, np.float32) if c is None else np.asarray(c, np.float32),
        cam_poses=net.poses.numpy(), cam_focal=net.focal.numpy(), cam_c=net.c.numpy(),
        cam_image_shape=net.image_shape.numpy(), latent_shape=np.asarray(enc_shape),
        rgb_coarse=out["coarse"]["rgb"].numpy(), depth_coarse=out["coarse"]["depth"].numpy(),
        w_coarse=out["coarse"]["weights"].numpy(),
        rgb_fine=out["fine"]["rgb"].numpy(), depth_fine=out["fine"]["depth"].numpy(),
        w_fine=out["fine"]["weights"].numpy(),
    )
    # direct net.forward probe (models.py:146-266) on a few points
    pts = torch.from_numpy(g.uniform(-0.6, 0.6, (sb, 24, 3)).astype(np.float32))
    dirs = torch.nn.functional.normalize(torch.from_numpy(g.standard_normal((sb, 24, 3)).astype(np.float32)), dim=-1)
    with torch.no_grad():
        fx["probe_xyz"], fx["probe_dirs"] = pts.numpy(), dirs.numpy()
        fx["probe_coarse"] = net(pts, coarse=True, viewdirs=dirs).numpy()
        fx["probe_fine"] = net(pts, coarse=False, viewdirs=dirs).numpy()
    if with_grad:
        tgt = torch.from_numpy(g.random((sb, n_rays, 3), dtype=np.float32))
        lat_leaf = lat.clone().requires_grad_(True)
        net.encoder.latent = lat_leaf
        for q in net.parameters():
            q.requires_grad_(True)
            q.grad = None
        renderer.train()
        with DrawInjector([d["u_coarse"], d["u_bin"], d["u_in"], d["noise"]]):
            o2 = renderer.bind_parallel(net, [0], simple_output=False)(rays, want_weights=True)
        crit = torch.nn.MSELoss()
        loss = crit(o2["coarse"]["rgb"], tgt) + crit(o2["fine"]["rgb"], tgt)   # lambda_coarse = lambda_fine = 1
        loss.backward()
        fx["target_rgb"] = tgt.numpy()
        fx["loss"] = np.float64(loss.item())
        for prefix, mlp in (("mlp_coarse", net.mlp_coarse), ("mlp_fine", net.mlp_fine)):
            for k, q in mlp.named_parameters():
                fx[f"g.{prefix}.{k}"] = sub_grad(k, q.grad)
                fx[f"gmax.{prefix}.{k}"] = np.float64(q.grad.abs().max().item())
        fx["g.latent"] = lat_leaf.grad[:, ::16, ::4, ::4].contiguous().numpy()
        fx["gmax.latent"] = np.float64(lat_leaf.grad.abs().max().item())
        renderer.eval()
    sw = float(out["fine"]["weights"].sum(-1).mean())
    print(f"{name}: R={R} mean sum(w_fine)={sw:.3f} rgb_fine mean={float(out['fine']['rgb'].mean()):.3f} "
          f"sum(w) range [{float(out['fine']['weights'].sum(-1).min()):.3f}, {float(out['fine']['weights'].sum(-1).max()):.3f}]")
    return fx


def case_sampling(NeRFRenderer, n_coarse, lindisp, seed, n_rays=1024):
    """Sampling functions alone, fed synthetic coarse weights; bit-exact targets."""
    g = np.random.Generator(np.random.PCG64(seed))
    r = NeRFRenderer(n_coarse=n_coarse, n_fine=32, n_fine_depth=16, lindisp=lindisp)
    rays = torch.zeros(n_rays, 8)
    rays[:, 6] = torch.from_numpy((0.5 + g.random(n_rays)).astype(np.float32))
    rays[:, 7] = rays[:, 6] + torch.from_numpy((0.5 + 3 * g.random(n_rays)).astype(np.float32))
    u_c = torch.from_numpy(g.random((n_rays, n_coarse), dtype=np.float32))
    # peaky weights like a real coarse pass, some rows all-zero
    w = torch.from_numpy((g.random((n_rays, n_coarse)) ** 8).astype(np.float32))
    w[::17] = 0.0
    u_b = torch.from_numpy(g.random((n_rays, 16), dtype=np.float32))
    u_i = torch.from_numpy(g.random((n_rays, 16), dtype=np.float32))
    depth = rays[:, 6] + (rays[:, 7] - rays[:, 6]) * torch.from_numpy(g.random(n_rays).astype(np.float32)) * 1.1 - 0.02
    nz = torch.from_numpy(g.standard_normal((n_rays, 16)).astype(np.float32))
    with DrawInjector([u_c]):
        z_c = r.sample_coarse(rays)
    with DrawInjector([u_b, u_i]):
        z_i = r.sample_fine(rays, w)
    with DrawInjector([nz]):
        z_d = r.sample_fine_depth(rays, depth)
    z_sorted = torch.sort(torch.cat([z_c, z_i, z_d], -1), -1)[0]
    lin = torch.linspace(0, 1 - 1.0 / n_coarse, n_coarse)
    return dict(n_coarse=n_coarse, lindisp=lindisp, rays=rays.numpy(), u_coarse=u_c.numpy(), weights=w.numpy(),
                u_bin=u_b.numpy(), u_in=u_i.numpy(), depth=depth.numpy(), noise=nz.numpy(), lin=lin.numpy(),
                z_coarse=z_c.numpy(), z_imp=z_i.numpy(), z_depth=z_d.numpy(), z_sorted=z_sorted.numpy())


def rays_cases():
    """rays_*.npz: the LIVE reference's util.gen_rays / util.pose_spherical (src/util/util.py:238-276, 309-323) on the
    shapes of the BASELINE configs.  Poses are stored whole; of the (B, H, W, 8) ray tensor a pixel lattice (every
    ``step``-th row / column plus the last ones) is kept, which is enough to pin the camera model."""
    import util   # the reference's own
    cases = [
        dict(name="srn128", w=128, h=128, focal=131.25, c=None, rows=4, radius=1.3, elev=-10.0, near=0.8, far=1.8, step=9),
        dict(name="dtu400x300", w=400, h=300, focal=(360.0, 355.5), c=(201.3, 148.9), rows=4, radius=2.6, elev=-30.0,
             near=0.1, far=5.0, step=23),
        dict(name="sn64", w=64, h=64, focal=70.0, c=None, rows=4, radius=2.6, elev=-10.0, near=1.2, far=4.0, step=5),
        dict(name="tiny3x4", w=7, h=5, focal=3.0, c=None, rows=3, radius=1.0, elev=20.0, near=0.5, far=2.0, step=1),
    ]
    for cs in cases:
        angles = np.linspace(-180, 180, 7)[:-1]
        poses = torch.stack([util.pose_spherical(float(a), cs["elev"], cs["radius"]) for a in angles])
        poses_in = poses[:, :cs["rows"]].contiguous()
        focal = torch.tensor(cs["focal"], dtype=torch.float32)
        c = None if cs["c"] is None else torch.tensor(cs["c"], dtype=torch.float32)
        rays = util.gen_rays(poses_in, cs["w"], cs["h"], focal, cs["near"], cs["far"], c=c)
        ys = sorted(set(list(range(0, cs["h"], cs["step"])) + [cs["h"] - 1]))
        xs = sorted(set(list(range(0, cs["w"], cs["step"])) + [cs["w"] - 1]))
        sub = rays[:, ys][:, :, xs].contiguous()
        np.savez_compressed(os.path.join(HERE, f"rays_{cs['name']}.npz"), w=cs["w"], h=cs["h"],
                            focal=np.asarray(cs["focal"], np.float32),
                            c=np.zeros(0, np.float32) if cs["c"] is None else np.asarray(cs["c"], np.float32),
                            near=cs["near"], far=cs["far"], angles=angles.astype(np.float64), elev=cs["elev"],
                            radius=cs["radius"], poses=poses.numpy(), pose_rows=cs["rows"], ys=np.asarray(ys),
                            xs=np.asarray(xs), rays=sub.numpy())
        print("rays", cs["name"], tuple(sub.shape))


def case_dtu_full(ConfigTree, make_model, NeRFRenderer, n_rays=2048, seed=4100, wc_seed=4101, wf_seed=4901):
    """BASELINE config 4 at its REAL size: 300 x 400 images, three source views, the 150 x 200 latent, per-ray
    near / far columns, rays taken from a full frame built by the reference's own util.gen_rays / pose_spherical.
    The reference's composite() is wrapped (not modified) to record the sample positions of both passes."""
    import util
    from pixelnerf_b200 import synthetic as syn
    h, w, ns = 300, 400, 3
    conf = reference_conf(ConfigTree, use_first_pool=True, white_bkgd=False)
    # weight seeds picked (oracle scan) so that both passes span transparent to opaque rays on this latent
    wc, wf = syn.mlp_weights(wc_seed), syn.mlp_weights(wf_seed)
    net = build_reference_net(ConfigTree, make_model, conf, wc, wf)
    g = np.random.Generator(np.random.PCG64(seed + 4))
    imgs = syn.images(seed + 3, 1, ns, h, w)
    src_poses = torch.stack([util.pose_spherical(a, -10.0, 2.6) for a in (-20.0, 0.0, 25.0)])[None]
    focal_t = torch.tensor([[723.0, 723.0]])
    c_t = torch.tensor([[200.0, 150.0]])
    with torch.no_grad():
        net.encode(imgs, src_poses, focal_t, c=c_t)
        enc_shape = tuple(net.encoder.latent.shape)
        assert enc_shape == (3, 512, 150, 200), enc_shape
        net.encoder.latent = syn.latent_maps(seed + 5, ns, 150, 200)
    full = util.gen_rays(util.pose_spherical(8.0, -12.0, 2.6)[None], w, h, focal_t[0], 0.1, 5.0, c=c_t[0])[0]
    full = full.reshape(-1, 8).clone()
    u = torch.from_numpy(g.random((h * w, 2), dtype=np.float32))
    full[:, 6] = 0.1 + u[:, 0] * 0.4
    full[:, 7] = 5.0 - u[:, 1] * 0.5
    # a contiguous run of rays from the middle of the frame + a strided set over the whole frame
    start = (h // 2) * w + 37
    pick = torch.cat([torch.arange(start, start + n_rays // 2),
                      torch.linspace(0, h * w - 1, n_rays - n_rays // 2).long()])
    rays = full[pick][None]
    renderer = NeRFRenderer.from_conf(conf["renderer"], lindisp=False, eval_batch_size=50000).eval()
    renderer.white_bkgd = False
    seen = []
    orig = renderer.composite
    renderer.composite = lambda model, rr, z, coarse=True, sb=0: (seen.append(z.clone()), orig(model, rr, z, coarse=coarse, sb=sb))[1]
    par = renderer.bind_parallel(net, [0], simple_output=False).eval()
    d = syn.draws(seed + 6, n_rays, 64, 16, 16)
    with torch.no_grad(), DrawInjector([d["u_coarse"], d["u_bin"], d["u_in"], d["noise"]]):
        out = par(rays, want_weights=True)
    z_c, z_f = seen
    fx = dict(name="dtu_full_ns3", h=h, w=w, ns=ns, sb=1, seed=seed, wc_seed=wc_seed, wf_seed=wf_seed,
              white=False, lindisp=False,
              rays=rays.numpy(), pick=pick.numpy(), src_poses=src_poses.numpy(),
              focal=focal_t.numpy(), c=c_t.numpy(),
              cam_poses=net.poses.numpy(), cam_focal=net.focal.numpy(), cam_c=net.c.numpy(),
              cam_image_shape=net.image_shape.numpy(), latent_shape=np.asarray(enc_shape),
              rgb_coarse=out["coarse"]["rgb"].numpy(), depth_coarse=out["coarse"]["depth"].numpy(),
              w_coarse=out["coarse"]["weights"].numpy(),
              rgb_fine=out["fine"]["rgb"].numpy(), depth_fine=out["fine"]["depth"].numpy(),
              # every 4th ray's fine sample positions / every 8th ray's fine weights keep the fixture under 1 MB
              z_fine_sub=z_f[::4].numpy(), w_fine_sub=out["fine"]["weights"][0, ::8].numpy(),
              z_coarse_sub=z_c[::16].numpy())
    print(f"dtu_full_ns3: R={n_rays} mean sum(w_coarse)={float(out['coarse']['weights'].sum(-1).mean()):.3f} "
          f"mean sum(w_fine)={float(out['fine']['weights'].sum(-1).mean()):.3f} "
          f"depth range [{float(out['fine']['depth'].min()):.3f}, {float(out['fine']['depth'].max()):.3f}]")
    return fx


def main():
    torch.manual_seed(0)
    ConfigTree, make_model, NeRFRenderer = load_reference()
    args = (ConfigTree, make_model, NeRFRenderer)
    if "--rays-only" in sys.argv:   # the rays_*.npz fixtures alone
        rays_cases()
        return
    if "--dtu-full-only" in sys.argv:
        fx = case_dtu_full(*args)
        np.savez_compressed(os.path.join(HERE, f"full_{fx['name']}.npz"), **fx)
        return
    if "--grad-only" in sys.argv:   # the grad_*.npz fixtures alone (the render/sampling ones are left untouched)
        grad_cases(args)
        return
    for kc, lindisp in ((64, False), (64, True), (128, False), (48, False), (96, True)):
        fx = case_sampling(NeRFRenderer, kc, lindisp, seed=100 + kc + int(lindisp))
        np.savez_compressed(os.path.join(HERE, f"sampling_k{kc}_{'disp' if lindisp else 'lin'}.npz"), **fx)
        print("sampling", kc, lindisp, "ok")
    args = (ConfigTree, make_model, NeRFRenderer)
    cases = [
        case_render("sn64_ns1", *args, h=64, w=64, ns=1, sb=1, n_rays=96, seed=1000, focal=70.0, c=None,
                    near_far=(1.2, 2.8), white=True, lindisp=False),
        case_render("srn128_ns2", *args, h=128, w=128, ns=2, sb=1, n_rays=64, seed=2000, focal=131.25, c=None,
                    near_far=(1.25, 2.75), white=True, lindisp=False),
        case_render("dtu_ns3_sb2", *args, h=60, w=80, ns=3, sb=2, n_rays=40, seed=3000,
                    focal=[[140.0, 150.0], [150.0, 140.0]], c=[[41.0, 29.0], [39.5, 30.5]],
                    near_far=(1.0, 3.2), white=False, lindisp=True, per_ray_bounds=True),
    ]
    for fx in cases:
        np.savez_compressed(os.path.join(HERE, f"render_{fx['name']}.npz"), **fx)
    grad_cases(args)
    rays_cases()
    fx = case_dtu_full(*args)
    np.savez_compressed(os.path.join(HERE, f"full_{fx['name']}.npz"), **fx)


def grad_cases(args):
    """Forward fixture + the reference's autograd gradients: NS=2 (view-mean backward) and the
    two-object NS=3 lindisp case with per-ray bounds."""
    cases = [
        case_render("srn128_ns2", *args, h=128, w=128, ns=2, sb=1, n_rays=64, seed=2000, focal=131.25, c=None,
                    near_far=(1.25, 2.75), white=True, lindisp=False, with_grad=True),
        case_render("dtu_ns3_sb2", *args, h=60, w=80, ns=3, sb=2, n_rays=40, seed=3000,
                    focal=[[140.0, 150.0], [150.0, 140.0]], c=[[41.0, 29.0], [39.5, 30.5]],
                    near_far=(1.0, 3.2), white=False, lindisp=True, per_ray_bounds=True, with_grad=True),
    ]
    for fx in cases:
        np.savez_compressed(os.path.join(HERE, f"grad_{fx['name']}.npz"), **fx)


if __name__ == "__main__":
    main()
