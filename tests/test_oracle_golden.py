"""Pin the oracle against golden vectors produced by the live reference
(tests/golden/make_golden.py).  CPU only."""
import numpy as np
import pytest
import torch

import oracle
from oracle import model as omodel
from helpers import full_case_inputs, golden_files, load, render_case_inputs


@pytest.mark.parametrize("path", golden_files("sampling_"))
def test_sampling_bit_exact(path):
    fx = load(path)
    kc, lindisp = int(fx["n_coarse"]), bool(fx["lindisp"])
    rays = torch.from_numpy(fx["rays"])
    # torch restatement
    z_c = oracle.sample_coarse(rays, kc, torch.from_numpy(fx["u_coarse"]), lindisp)
    z_i, _ = oracle.sample_fine(rays, torch.from_numpy(fx["weights"]), kc, torch.from_numpy(fx["u_bin"]),
                                torch.from_numpy(fx["u_in"]), lindisp)
    z_d = oracle.sample_fine_depth(rays, torch.from_numpy(fx["depth"]), torch.from_numpy(fx["noise"]), 0.01)
    assert np.array_equal(z_c.numpy(), fx["z_coarse"])
    assert np.array_equal(z_i.numpy(), fx["z_imp"])
    assert np.array_equal(z_d.numpy(), fx["z_depth"])
    assert np.array_equal(oracle.merge_sorted([z_c, z_i, z_d]).numpy(), fx["z_sorted"])
    # numpy emulation of the exact operation order (what the CUDA kernels implement)
    assert np.array_equal(oracle.sample_coarse_np(fx["rays"], kc, fx["u_coarse"], lindisp, lin=fx["lin"]), fx["z_coarse"])
    z_i_np, _ = oracle.sample_fine_np(fx["rays"], fx["weights"], kc, fx["u_bin"], fx["u_in"], lindisp)
    assert np.array_equal(z_i_np, fx["z_imp"])
    assert np.array_equal(oracle.sample_fine_depth_np(fx["rays"], fx["depth"], fx["noise"], 0.01), fx["z_depth"])


def test_row_sum_matches_torch():
    g = np.random.Generator(np.random.PCG64(5))
    for k in (8, 24, 48, 64, 96, 100, 128):
        x = (g.random((4096, k)) ** 6).astype(np.float32)
        assert np.array_equal(oracle.aten_row_sum_f32(x), torch.sum(torch.from_numpy(x), -1).numpy())


@pytest.mark.parametrize("path", golden_files("render_"))
def test_render_matches_reference(path):
    fx = load(path)
    inp = render_case_inputs(fx)
    sb, ns = inp["sb"], inp["ns"]
    rays = inp["rays"].reshape(-1, 8)

    def net_fn(xyz, dirs, coarse):
        return omodel.net_forward(inp["wc"] if coarse else inp["wf"], inp["latent"], inp["cam"], ns, xyz, dirs)

    with torch.no_grad():
        probe_c = net_fn(torch.from_numpy(fx["probe_xyz"]), torch.from_numpy(fx["probe_dirs"]), True)
        probe_f = net_fn(torch.from_numpy(fx["probe_xyz"]), torch.from_numpy(fx["probe_dirs"]), False)
        out = oracle.render_rays(net_fn, rays, sb, 64, 32, 16, inp["draws"], depth_std=0.01,
                                 white_bkgd=inp["white"], lindisp=inp["lindisp"])
    # Same torch ops in the same order on the same machine class: expect (near) bit equality.
    tol = dict(rtol=0, atol=2e-6)
    np.testing.assert_allclose(probe_c.numpy(), fx["probe_coarse"], **tol)
    np.testing.assert_allclose(probe_f.numpy(), fx["probe_fine"], **tol)
    shp = fx["rgb_coarse"].shape[:2]
    np.testing.assert_allclose(out["rgb_coarse"].reshape(*shp, 3).numpy(), fx["rgb_coarse"], **tol)
    np.testing.assert_allclose(out["w_coarse"].reshape(*shp, -1).numpy(), fx["w_coarse"], **tol)
    np.testing.assert_allclose(out["depth_coarse"].reshape(shp).numpy(), fx["depth_coarse"], rtol=0, atol=1e-5)
    np.testing.assert_allclose(out["rgb_fine"].reshape(*shp, 3).numpy(), fx["rgb_fine"], rtol=0, atol=1e-5)
    np.testing.assert_allclose(out["w_fine"].reshape(*shp, -1).numpy(), fx["w_fine"], rtol=0, atol=1e-5)
    np.testing.assert_allclose(out["depth_fine"].reshape(shp).numpy(), fx["depth_fine"], rtol=0, atol=1e-5)


def grad_subsample(g):
    """The part of a gradient tensor the grad_*.npz fixtures keep (tests/golden/make_golden.py: sub_grad)."""
    if g.dim() == 2 and g.shape[0] == 512 and g.shape[1] == 512:
        return g[::16, ::16]
    return g


def grad_case_loss(fx, inp, out):
    """train/train.py:264-281 with lambda_coarse = lambda_fine = 1: MSE of both passes' rgb against the target."""
    tgt = torch.from_numpy(fx["target_rgb"]).reshape(-1, 3).to(out["rgb_coarse"].dtype)
    return ((out["rgb_coarse"] - tgt) ** 2).mean() + ((out["rgb_fine"] - tgt) ** 2).mean()


@pytest.mark.parametrize("path", golden_files("grad_"))
def test_gradients_match_reference_autograd(path):
    """autograd over the oracle reproduces the LIVE reference's own loss.backward() -- parameters of
    both MLPs and the encoder latent, through compositing, the view mean, the bilinear gather and the depth-guided
    sample positions.  Stratified and importance positions are the reference's (recomputed bit-exactly from its
    coarse weights, which the pinned sampler reproduces); the depth samples stay a function of the coarse depth."""
    fx = load(path)
    inp = render_case_inputs(fx)
    sb, ns = inp["sb"], inp["ns"]
    rays = inp["rays"].reshape(-1, 8)
    wc = {k: v.clone().requires_grad_(True) for k, v in inp["wc"].items()}
    wf = {k: v.clone().requires_grad_(True) for k, v in inp["wf"].items()}
    lat = inp["latent"].clone().requires_grad_(True)

    def net_fn(xyz, dirs, coarse):
        return omodel.net_forward(wc if coarse else wf, lat, inp["cam"], ns, xyz, dirs)

    d = inp["draws"]
    z_c = oracle.sample_coarse(rays, 64, d["u_coarse"], inp["lindisp"])
    z_i, _ = oracle.sample_fine(rays, torch.from_numpy(fx["w_coarse"]).reshape(-1, 64), 64, d["u_bin"], d["u_in"],
                                inp["lindisp"])
    out = oracle.render_rays(net_fn, rays, sb, 64, 32, 16, d, depth_std=0.01, white_bkgd=inp["white"],
                             lindisp=inp["lindisp"], force=dict(z_coarse=z_c, z_imp=z_i))
    loss = grad_case_loss(fx, inp, out)
    assert abs(loss.item() - float(fx["loss"])) < 1e-5 * max(1.0, float(fx["loss"]))
    loss.backward()
    worst = {}
    for prefix, w in (("mlp_coarse", wc), ("mlp_fine", wf)):
        for k, q in w.items():
            ref = fx[f"g.{prefix}.{k}"]
            got = grad_subsample(q.grad).numpy()
            worst[f"{prefix}.{k}"] = np.abs(got - ref).max() / max(float(fx[f"gmax.{prefix}.{k}"]), 1e-12)
    worst["latent"] = np.abs(lat.grad[:, ::16, ::4, ::4].numpy() - fx["g.latent"]).max() / float(fx["gmax.latent"])
    bad = {k: v for k, v in worst.items() if v > 2e-4}
    assert not bad, bad


def test_oracle_matches_reference_on_the_real_size_dtu_slice():
    """BASELINE config 4 at its real size (300 x 400, 150 x 200 latent, NS = 3, per-ray bounds): the oracle on 96 of
    the fixture's 2 048 rays -- half from the contiguous run, half from the frame-wide lattice -- against the live
    reference's outputs, sample positions bit for bit."""
    fx = load(golden_files("full_")[0])
    pick = np.concatenate([np.arange(0, 48), np.arange(1024, 2048, 21)[:48]])
    inp = full_case_inputs(fx, pick)
    rays = inp["rays"].reshape(-1, 8)

    def net_fn(xyz, dirs, coarse):
        return omodel.net_forward(inp["wc"] if coarse else inp["wf"], inp["latent"], inp["cam"], inp["ns"], xyz, dirs)
    with torch.no_grad():
        out = oracle.render_rays(net_fn, rays, 1, 64, 32, 16, inp["draws"], depth_std=0.01, white_bkgd=False,
                                 lindisp=False)
    tol = dict(rtol=0, atol=1e-5)
    np.testing.assert_allclose(out["rgb_coarse"].numpy(), fx["rgb_coarse"][0, pick], **tol)
    np.testing.assert_allclose(out["w_coarse"].numpy(), fx["w_coarse"][0, pick], **tol)
    np.testing.assert_allclose(out["depth_coarse"].numpy(), fx["depth_coarse"][0, pick], **tol)
    np.testing.assert_allclose(out["rgb_fine"].numpy(), fx["rgb_fine"][0, pick], **tol)
    np.testing.assert_allclose(out["depth_fine"].numpy(), fx["depth_fine"][0, pick], **tol)
    sub4 = [i for i, r in enumerate(pick) if r % 4 == 0]
    assert np.array_equal(out["z_fine"].numpy()[sub4], fx["z_fine_sub"][pick[sub4] // 4])
    sub16 = [i for i, r in enumerate(pick) if r % 16 == 0]
    assert np.array_equal(out["z_coarse"].numpy()[sub16], fx["z_coarse_sub"][pick[sub16] // 16])


@pytest.mark.parametrize("path", golden_files("rays_"))
def test_host_ray_generation_matches_the_live_reference(path):
    """util.gen_rays / pose_spherical (src/util/util.py:238-276, 309-323): the package's host code path and the
    synthetic-scene builder against rays produced by the reference's own functions."""
    from pixelnerf_b200 import synthetic as syn
    from pixelnerf_b200.util import util
    fx = load(path)
    w, h, rows = int(fx["w"]), int(fx["h"]), int(fx["pose_rows"])
    focal = fx["focal"].tolist() if fx["focal"].ndim else float(fx["focal"])
    c = fx["c"].tolist() if fx["c"].size else None
    near, far = float(fx["near"]), float(fx["far"])
    poses = torch.stack([util.pose_spherical(float(a), float(fx["elev"]), float(fx["radius"])) for a in fx["angles"]])
    assert np.array_equal(poses.numpy(), fx["poses"])                       # same op order: bit for bit
    rays = util.gen_rays(poses[:, :rows].contiguous(), w, h, focal, near, far, c=c)
    sub = rays[:, fx["ys"]][:, :, fx["xs"]].numpy()
    assert sub.shape == fx["rays"].shape
    assert np.array_equal(sub[..., :3], fx["rays"][..., :3]) and np.array_equal(sub[..., 6:], fx["rays"][..., 6:])
    assert np.abs(sub[..., 3:6] - fx["rays"][..., 3:6]).max() < 1e-6
    # the synthetic scene builder (numpy poses, one pose at a time) states the same camera model
    for i, a in enumerate(fx["angles"][:2]):
        p = syn.pose_spherical(float(a), float(fx["elev"]), float(fx["radius"]))
        assert np.abs(p.numpy() - fx["poses"][i]).max() < 1e-6
        r = syn.gen_rays(p, w, h, focal if np.ndim(focal) else float(focal), near, far, c)
        assert np.abs(r[fx["ys"]][:, fx["xs"]].numpy() - fx["rays"][i]).max() < 2e-6
