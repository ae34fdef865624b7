"""Oracle: ray sampling and alpha compositing (torch-CPU restatement).

Test infrastructure only (see oracle/__init__.py).  Each function restates one
reference function as a flat, module-free function that takes the random draws
as explicit arguments, so that the reference, the oracle and the CUDA path can
be fed identical draws.  The arithmetic primitives are torch CPU ops because
that is where the reference's arithmetic lives; the
torch-free bit-exact emulation of the same recipes is in ``sampling_np.py``.

All ``path:line`` citations are into /root/reference/.
"""
import torch


def _lerp_depth(near, far, zs, lindisp):
    # src/render/nerf.py:112-115 (coarse) and :144-147 (fine): identical lerp.
    if not lindisp:
        return near * (1 - zs) + far * zs
    return 1 / (1 / near * (1 - zs) + 1 / far * zs)


def sample_coarse(rays, n_coarse, u, lindisp=False):
    """Stratified samples.  src/render/nerf.py:98-118.

    rays (R,8) = [o(3), d(3), near, far]; u (R,Kc) uniform draws that replace
    ``torch.rand_like`` at nerf.py:111.  Returns z (R,Kc).
    """
    near, far = rays[:, -2:-1], rays[:, -1:]
    step = 1.0 / n_coarse
    zs = torch.linspace(0, 1 - step, n_coarse).to(rays.device).unsqueeze(0).repeat(rays.shape[0], 1)
    zs = zs + u * step
    return _lerp_depth(near, far, zs, lindisp)


def sample_fine(rays, weights, n_coarse, u_bin, u_in, lindisp=False):
    """Inverse-CDF importance samples.  src/render/nerf.py:120-148.

    weights (R,Kc) coarse compositing weights; u_bin (R,Kf) replaces
    ``torch.rand`` at :135 and u_in (R,Kf) replaces ``torch.rand_like`` at :141.
    Returns (z (R,Kf), inds (R,Kf) float).
    """
    w = weights.detach() + 1e-5
    pdf = w / torch.sum(w, -1, keepdim=True)
    cdf = torch.cumsum(pdf, -1)
    cdf = torch.cat([torch.zeros_like(cdf[:, :1]), cdf], -1)
    inds = torch.searchsorted(cdf, u_bin.contiguous(), right=True).float() - 1.0
    inds = torch.clamp_min(inds, 0.0)
    zs = (inds + u_in) / n_coarse
    near, far = rays[:, -2:-1], rays[:, -1:]
    return _lerp_depth(near, far, zs, lindisp), inds


def sample_fine_depth(rays, depth, noise, depth_std):
    """Samples around the expected coarse depth.  src/render/nerf.py:150-161.

    noise (R,Kd) standard normal draws replacing ``torch.randn_like`` at :158.
    """
    z = depth.unsqueeze(1).repeat((1, noise.shape[1]))
    z = z + noise * depth_std
    return torch.max(torch.min(z, rays[:, -1:]), rays[:, -2:-1])


def merge_sorted(z_parts):
    """cat + sort of all sample sets.  src/render/nerf.py:294-295."""
    return torch.sort(torch.cat(z_parts, dim=-1), dim=-1)[0]


def composite(rgb_sigma, z, rays, white_bkgd, sigma_noise=None, noise_std=0.0):
    """Alpha compositing.  src/render/nerf.py:178-182, 221-249.

    rgb_sigma (R,K,4) is the model output already reshaped per ray; z (R,K).
    sigma_noise (R,K): standard normal draws replacing ``torch.randn_like`` at :226 (training
    with noise_std > 0 only; None = the branch at :225 is not taken).
    Returns (weights (R,K), rgb (R,3), depth (R,)).
    """
    deltas = z[:, 1:] - z[:, :-1]
    delta_inf = rays[:, -1:] - z[:, -1:]
    deltas = torch.cat([deltas, delta_inf], -1)
    rgbs = rgb_sigma[..., :3]
    sigmas = rgb_sigma[..., 3]
    if sigma_noise is not None and noise_std > 0.0:
        sigmas = sigmas + sigma_noise * noise_std
    alphas = 1 - torch.exp(-deltas * torch.relu(sigmas))
    shifted = torch.cat([torch.ones_like(alphas[:, :1]), 1 - alphas + 1e-10], -1)
    T = torch.cumprod(shifted, -1)
    weights = alphas * T[:, :-1]
    rgb = torch.sum(weights.unsqueeze(-1) * rgbs, -2)
    depth = torch.sum(weights * z, -1)
    if white_bkgd:
        rgb = rgb + 1 - weights.sum(dim=1).unsqueeze(-1)
    return weights, rgb, depth


def points_and_dirs(rays, z):
    """Sample positions o + z d and per-sample view directions.  nerf.py:185,203-208."""
    pts = rays[:, None, :3] + z.unsqueeze(2) * rays[:, None, 3:6]
    dirs = rays[:, None, 3:6].expand(-1, z.shape[1], -1)
    return pts, dirs


def render_rays(net_fn, rays, sb, n_coarse, n_fine, n_fine_depth, draws, *,
                depth_std=0.01, white_bkgd=False, lindisp=False, force=None, noise_std=0.0):
    """Whole renderer forward.  src/render/nerf.py:251-303.

    net_fn(xyz (SB,P,3), viewdirs (SB,P,3), coarse: bool) -> (SB,P,4)
    rays (SB*B,8) flattened; ``sb`` the number of objects.
    draws = dict(u_coarse (R,Kc), u_bin (R,Kf), u_in (R,Kf), noise (R,Kd)).
    force = optional dict for teacher forcing (SURVEY section 7 hard part 12):
      'z_coarse', 'w_coarse', 'depth_coarse', 'z_imp', 'z_fine' replace the values the
      oracle would compute itself at that stage ('z_imp' keeps the depth samples a
      differentiable function of the coarse depth, as in nerf.py:290-293).
    Returns a dict of every intermediate the parity tests compare.
    """
    force = force or {}
    R = rays.shape[0]
    out = {}

    def run(z, coarse):
        K = z.shape[1]
        pts, dirs = points_and_dirs(rays, z)
        rs = net_fn(pts.reshape(sb, -1, 3), dirs.reshape(sb, -1, 3), coarse)
        rs = rs.reshape(R, K, 4)
        # training with noise_std > 0 (nerf.py:225-226): draws["sigma_noise_coarse" | "sigma_noise_fine"]
        sn = draws.get("sigma_noise_coarse" if coarse else "sigma_noise_fine") if noise_std > 0.0 else None
        return rs, composite(rs, z, rays, white_bkgd, sn, noise_std)

    z_c = force.get("z_coarse")
    if z_c is None:
        z_c = sample_coarse(rays, n_coarse, draws["u_coarse"], lindisp)
    rs_c, (w_c, rgb_c, d_c) = run(z_c, True)
    out.update(z_coarse=z_c, rgbsigma_coarse=rs_c, w_coarse=w_c, rgb_coarse=rgb_c,
               depth_coarse=d_c)
    if n_fine > 0:
        parts = [z_c]
        w_in = force.get("w_coarse", w_c)
        d_in = force.get("depth_coarse", d_c)
        if n_fine - n_fine_depth > 0:
            z_i, inds = sample_fine(rays, w_in, n_coarse, draws["u_bin"], draws["u_in"], lindisp)
            z_i = force.get("z_imp", z_i)
            parts.append(z_i)
            out.update(z_imp=z_i, inds=inds)
        if n_fine_depth > 0:
            z_d = sample_fine_depth(rays, d_in, draws["noise"], depth_std)
            parts.append(z_d)
            out.update(z_depth=z_d)
        z_f = force.get("z_fine")
        if z_f is None:
            z_f = merge_sorted(parts)
        rs_f, (w_f, rgb_f, d_f) = run(z_f, False)
        out.update(z_fine=z_f, rgbsigma_fine=rs_f, w_fine=w_f, rgb_fine=rgb_f, depth_fine=d_f)
    return out
