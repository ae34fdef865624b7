"""CUDA sampling kernels vs golden vectors of the live reference and vs the numpy oracle:
bit-exact.  Calls go through the C ABI (pixelnerf_b200._lib)."""
import numpy as np
import pytest
import torch

import oracle
from helpers import golden_files, load

pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("path", golden_files("sampling_"))
def test_sampling_golden_bit_exact(path):
    from pixelnerf_b200 import _lib
    fx = load(path)
    kc, lindisp = int(fx["n_coarse"]), bool(fx["lindisp"])
    rays = _dev(fx["rays"])
    z_c = _lib.sample_coarse(rays, kc, lindisp, u=_dev(fx["u_coarse"]))
    assert np.array_equal(z_c.cpu().numpy(), fx["z_coarse"])
    z_s, z_i, z_d = _lib.sample_fine(rays, z_c, _dev(fx["weights"]), _dev(fx["depth"]), 16, 16, 0.01, lindisp,
                                     u_bin=_dev(fx["u_bin"]), u_in=_dev(fx["u_in"]), noise=_dev(fx["noise"]),
                                     want_parts=True)
    assert np.array_equal(z_i.cpu().numpy(), fx["z_imp"])
    assert np.array_equal(z_d.cpu().numpy(), fx["z_depth"])
    assert np.array_equal(z_s.cpu().numpy(), fx["z_sorted"])


@pytest.mark.parametrize("kc,kf,kd,lindisp", [(64, 16, 16, False), (128, 32, 0, True), (44, 7, 3, False),
                                              (64, 0, 16, False), (200, 16, 16, False)])
def test_sampling_vs_numpy_oracle(kc, kf, kd, lindisp):
    from pixelnerf_b200 import _lib
    g = np.random.Generator(np.random.PCG64(kc * 7 + kf))
    r = 5000
    rays = np.zeros((r, 8), np.float32)
    rays[:, 6] = 0.3 + g.random(r)
    rays[:, 7] = rays[:, 6] + 0.2 + 4 * g.random(r)
    u_c = g.random((r, kc), dtype=np.float32)
    w = (g.random((r, kc)) ** 10).astype(np.float32)
    w[::13] = 0
    lin = torch.linspace(0, 1 - 1.0 / kc, kc).numpy()
    z_c_ref = oracle.sample_coarse_np(rays, kc, u_c, lindisp, lin=lin)
    z_c = _lib.sample_coarse(_dev(rays), kc, lindisp, u=_dev(u_c))
    assert np.array_equal(z_c.cpu().numpy(), z_c_ref)
    parts = [z_c_ref]
    kw = {}
    if kf:
        u_b, u_i = g.random((r, kf), dtype=np.float32), g.random((r, kf), dtype=np.float32)
        u_b[::50, 0] = 0.0  # exact-zero draw: searchsorted(right) lands past the leading 0
        parts.append(oracle.sample_fine_np(rays, w, kc, u_b, u_i, lindisp)[0])
        kw.update(u_bin=_dev(u_b), u_in=_dev(u_i))
    depth = (rays[:, 6] + (rays[:, 7] - rays[:, 6]) * (1.2 * g.random(r) - 0.1)).astype(np.float32)
    if kd:
        nz = g.standard_normal((r, kd)).astype(np.float32)
        parts.append(oracle.sample_fine_depth_np(rays, depth, nz, 0.01))
        kw.update(noise=_dev(nz))
    ref = np.sort(np.concatenate(parts, -1), -1)
    z_s = _lib.sample_fine(_dev(rays), z_c, _dev(w) if kf else None, _dev(depth) if kd else None, kf, kd, 0.01,
                           lindisp, **kw)
    assert np.array_equal(z_s.cpu().numpy(), ref)


def test_sampling_in_kernel_draws_properties():
    """Without injected draws the kernels draw themselves: check the size-independent properties
    at the full 128x128 frame size -- stratification, bounds, sortedness, determinism per seed."""
    from pixelnerf_b200 import _lib
    r, kc = 16384, 64
    rays = torch.zeros(r, 8, device="cuda")
    rays[:, 6], rays[:, 7] = 0.8, 1.8
    z = _lib.sample_coarse(rays, kc, False, seed=123)
    z2 = _lib.sample_coarse(rays, kc, False, seed=123)
    assert torch.equal(z, z2) and not torch.equal(z, _lib.sample_coarse(rays, kc, False, seed=124))
    edges = 0.8 + torch.arange(kc + 1, device="cuda") / kc
    assert bool(((z >= edges[:-1] - 1e-6) & (z <= edges[1:] + 1e-6)).all())
    u = (z - edges[:-1]) * kc
    assert abs(float(u.mean()) - 0.5) < 5e-3
    w = torch.rand(r, kc, device="cuda") ** 4
    zs = _lib.sample_fine(rays, z, w, torch.full((r,), 1.3, device="cuda"), 16, 16, 0.01, False, seed=5)
    assert zs.shape == (r, 96) and bool((zs[:, 1:] >= zs[:, :-1]).all())
    assert bool((zs >= 0.8).all()) and bool((zs <= 1.8).all())
    # multiset check: the coarse samples are all still present
    merged = torch.sort(torch.cat([z, zs], -1), -1)[0]
    assert merged.shape[1] == 160


def test_empty_and_bad_args():
    from pixelnerf_b200 import _lib
    rays = torch.zeros(0, 8, device="cuda")
    assert _lib.sample_coarse(rays, 64).shape == (0, 64)
    with pytest.raises(RuntimeError):
        _lib.sample_coarse(torch.zeros(4, 8), 64)  # CPU tensor: no fallback
    with pytest.raises(RuntimeError, match="512"):
        r4 = torch.ones(4, 8, device="cuda")
        _lib.sample_fine(r4, torch.ones(4, 600, device="cuda"), torch.ones(4, 600, device="cuda"), None, 4, 0, 0.01)
