"""Deterministic synthetic scenes, weights and random draws.

There is no dataset or checkpoint on the build or GPU machines, so parity tests,
the smoke test and the benchmark all run on synthetic inputs of the reference's
shapes.  Everything here is generated with numpy's PCG64 (bit-identical across
machines for one numpy version) and handed to torch afterwards, so the golden
fixtures under tests/golden/ can be regenerated from a few integers.

A random-init ResnetFC is degenerate for parity testing: every ``fc_1`` is
zero-initialised (reference src/model/resnetfc.py:38-39), so the five residual
blocks are identities and ten of fifteen GEMMs would go untested, and sigma is
so small that every ray is empty.  ``mlp_weights`` therefore conditions the
weights like a trained network: hidden activations O(1), sigma*delta O(0.1-1).
"""
import math

import numpy as np
import torch

D_LATENT = 512
D_HIDDEN = 512
D_IN = 42
N_BLOCKS = 5
COMBINE_LAYER = 3


def _rng(seed):
    return np.random.Generator(np.random.PCG64(seed))


def mlp_weights(seed, sigma_gain=10.0, sigma_bias=-1.0, fc1_gain=0.35, d_latent=D_LATENT):
    """Un-prefixed ResnetFC state_dict (reference key names, fp32 torch tensors)."""
    g = _rng(seed)

    def kaiming(out_f, in_f, gain=1.0):
        return (g.standard_normal((out_f, in_f)) * (gain * math.sqrt(2.0 / in_f))).astype(np.float32)

    def bias(n, s=0.1):
        return (g.standard_normal(n) * s).astype(np.float32)

    w = {"lin_in.weight": kaiming(D_HIDDEN, D_IN), "lin_in.bias": bias(D_HIDDEN)}
    for b in range(N_BLOCKS):
        w[f"blocks.{b}.fc_0.weight"] = kaiming(D_HIDDEN, D_HIDDEN)
        w[f"blocks.{b}.fc_0.bias"] = bias(D_HIDDEN)
        w[f"blocks.{b}.fc_1.weight"] = kaiming(D_HIDDEN, D_HIDDEN, fc1_gain)
        w[f"blocks.{b}.fc_1.bias"] = bias(D_HIDDEN)
    for b in range(COMBINE_LAYER):
        w[f"lin_z.{b}.weight"] = kaiming(D_HIDDEN, d_latent, 0.5)
        w[f"lin_z.{b}.bias"] = bias(D_HIDDEN)
    lo = kaiming(4, D_HIDDEN, 0.5)
    # zero-sum sigma row: cancels the common positive mean of relu(h), so the sign of
    # sigma's pre-activation varies from point to point instead of per weight seed
    lo[3] = (lo[3] - lo[3].mean()) * sigma_gain
    w["lin_out.weight"] = lo
    lb = bias(4)
    lb[3] = sigma_bias
    w["lin_out.bias"] = lb
    return {k: torch.from_numpy(v) for k, v in w.items()}


def latent_maps(seed, n_views, hl, wl, channels=D_LATENT):
    """Encoder-output stand-in (n_views, C, Hl, Wl): non-negative, spatially smooth-ish."""
    g = _rng(seed)
    lo = g.standard_normal((n_views, channels, (hl + 3) // 4 + 1, (wl + 3) // 4 + 1)).astype(np.float32)
    lo_t = torch.nn.functional.interpolate(torch.from_numpy(lo), size=(hl, wl), mode="bilinear",
                                           align_corners=True)
    hi = torch.from_numpy(g.standard_normal((n_views, channels, hl, wl)).astype(np.float32))
    return torch.relu(lo_t + 0.5 * hi).contiguous()


def images(seed, sb, ns, h, w):
    g = _rng(seed)
    return torch.from_numpy(g.uniform(-1, 1, (sb, ns, 3, h, w)).astype(np.float32))


def pose_spherical(theta, phi, radius):
    """Camera-to-world pose on a sphere looking at the origin.

    Same convention as the reference's util.pose_spherical (src/util/util.py:309-323):
    translate along z, rotate by phi about x, theta about y, then the fixed axis swap.
    """
    t, p = math.radians(theta), math.radians(phi)
    trans = np.eye(4)
    trans[2, 3] = radius
    rp = np.array([[1, 0, 0, 0], [0, math.cos(p), -math.sin(p), 0],
                   [0, math.sin(p), math.cos(p), 0], [0, 0, 0, 1]])
    rt = np.array([[math.cos(t), 0, -math.sin(t), 0], [0, 1, 0, 0],
                   [math.sin(t), 0, math.cos(t), 0], [0, 0, 0, 1]])
    swap = np.array([[-1, 0, 0, 0], [0, 0, 1, 0], [0, 1, 0, 0], [0, 0, 0, 1]])
    c2w = swap.astype(np.float32) @ (rt.astype(np.float32) @ (rp.astype(np.float32) @ trans.astype(np.float32)))
    return torch.from_numpy(c2w.astype(np.float32))


def gen_rays(pose, width, height, focal, z_near, z_far, c=None):
    """Rays (H,W,8) = [origin, unit dir, near, far] for one camera-to-world pose.

    Layout and camera convention of the reference's util.gen_rays / unproj_map
    (src/util/util.py:113-143, 238-276); host-side helper for synthetic inputs.
    """
    fx, fy = (focal, focal) if np.isscalar(focal) else (float(focal[0]), float(focal[1]))
    cx, cy = (width * 0.5, height * 0.5) if c is None else (float(c[0]), float(c[1]))
    ys, xs = torch.meshgrid(torch.arange(height, dtype=torch.float32) - cy,
                            torch.arange(width, dtype=torch.float32) - cx, indexing="ij")
    d = torch.stack((xs / fx, -ys / fy, -torch.ones_like(xs)), dim=-1)
    d = d / torch.norm(d, dim=-1, keepdim=True)
    d = torch.matmul(pose[None, None, :3, :3], d.unsqueeze(-1))[..., 0]
    o = pose[None, None, :3, 3].expand(height, width, -1)
    nf = torch.tensor([z_near, z_far], dtype=torch.float32).view(1, 1, 2).expand(height, width, -1)
    return torch.cat((o, d, nf), dim=-1)


def draws(seed, n_rays, n_coarse, n_imp, n_depth):
    """The four random tensors one NeRFRenderer.forward consumes, in order
    (reference src/render/nerf.py:111,135,141,158)."""
    g = _rng(seed)
    return dict(
        u_coarse=torch.from_numpy(g.random((n_rays, n_coarse), dtype=np.float32)),
        u_bin=torch.from_numpy(g.random((n_rays, n_imp), dtype=np.float32)),
        u_in=torch.from_numpy(g.random((n_rays, n_imp), dtype=np.float32)),
        noise=torch.from_numpy(g.standard_normal((n_rays, n_depth)).astype(np.float32)),
    )


SCENES = {
    # name: geometry of the BASELINE.json configs
    "sn64": dict(h=64, w=64, hl=32, wl=32, ns=1, focal=70.0, radius=2.6, near=1.2, far=4.0,
                 white=True, src_az=[30.0]),
    "srn_car": dict(h=128, w=128, hl=64, wl=64, ns=1, focal=131.25, radius=1.3, near=0.8, far=1.8,
                    white=True, src_az=[30.0]),
    "srn_chair": dict(h=128, w=128, hl=64, wl=64, ns=2, focal=131.25, radius=2.0, near=1.25, far=2.75,
                      white=True, src_az=[0.0, 40.0]),
    "dtu": dict(h=300, w=400, hl=150, wl=200, ns=3, focal=(723.0, 723.0), c=(200.0, 150.0), radius=2.6,
                near=0.1, far=5.0, white=False, src_az=[-20.0, 0.0, 25.0]),
}


def scene(name, seed=0, target_az=0.0, elev=-10.0):
    """Synthetic scene dict: latent, source poses, intrinsics, one frame of target rays."""
    s = dict(SCENES[name])
    s["name"] = name
    s["latent"] = latent_maps(seed + 11, s["ns"], s["hl"], s["wl"])
    s["src_poses"] = torch.stack([pose_spherical(a, elev, s["radius"]) for a in s["src_az"]])
    s["rays"] = gen_rays(pose_spherical(target_az, elev, s["radius"]), s["w"], s["h"], s["focal"],
                         s["near"], s["far"], s.get("c")).reshape(-1, 8).contiguous()
    f = s["focal"]
    s["focal_t"] = torch.tensor(f, dtype=torch.float32) if not np.isscalar(f) else torch.tensor(float(f))
    s["c_t"] = None if s.get("c") is None else torch.tensor(s["c"], dtype=torch.float32)
    return s
