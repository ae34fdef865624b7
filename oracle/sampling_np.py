"""Oracle: torch-free, bit-exact numpy emulation of the reference's sampling.

Test infrastructure only (see oracle/__init__.py).  This is the canonical
statement of the operation ORDER the CUDA sampling kernels reproduce: every
intermediate is rounded to fp32 exactly where torch-CPU rounds it.  It is pinned
against torch-CPU / the live reference in tests/test_oracle_golden.py.

Recipes (re-verified on torch 2.11 CPU under
both AVX2 and AVX512 dispatch):
  * torch.sum(float32, dim=-1) on a contiguous row: 8 lanes x 4 interleaved
    accumulators, leftover 8-vectors into accumulator 0, accumulators combined
    ((a0+a1)+a2)+a3 lane-wise; the scalar tail (K mod 8) is summed first, then the
    eight lane partials are added to it left to right.
  * torch.cumsum(float32): fp64 running sum, rounded to fp32 per element.
  * searchsorted(right=True) = number of cdf entries <= u.
"""
import numpy as np

f32 = np.float32


def aten_row_sum_f32(x):
    """torch.sum(x, -1) for contiguous float32 rows, bit-exact."""
    x = np.ascontiguousarray(x, dtype=f32)
    r, k = x.shape
    assert k < 512, "rows >= 512 elements enter ATen's cascade levels; not emulated"
    lanes, nacc = 8, 4
    nv = k // lanes
    v = x[:, : nv * lanes].reshape(r, nv, lanes)
    acc = [np.zeros((r, lanes), f32) for _ in range(nacc)]
    full = nv // nacc
    for i in range(full):
        for j in range(nacc):
            acc[j] = (acc[j] + v[:, i * nacc + j]).astype(f32)
    for i in range(full * nacc, nv):
        acc[0] = (acc[0] + v[:, i]).astype(f32)
    a = acc[0]
    for j in range(1, nacc):
        a = (a + acc[j]).astype(f32)
    # scalar tail (k % 8 elements): same 4-way ILP structure, summed FIRST, then
    # the 8 lane partials are added left to right (ATen SumKernel.cpp, vectorized_inner_sum).
    t = x[:, nv * lanes:]
    nt = t.shape[1]
    tacc = [np.zeros(r, f32) for _ in range(nacc)]
    tfull = nt // nacc
    for i in range(tfull):
        for j in range(nacc):
            tacc[j] = (tacc[j] + t[:, i * nacc + j]).astype(f32)
    for i in range(tfull * nacc, nt):
        tacc[0] = (tacc[0] + t[:, i]).astype(f32)
    s = tacc[0]
    for j in range(1, nacc):
        s = (s + tacc[j]).astype(f32)
    for l in range(lanes):
        s = (s + a[:, l]).astype(f32)
    return s


def _lerp(near, far, zs, lindisp):
    one = f32(1.0)
    a = (one - zs).astype(f32)
    if not lindisp:
        return ((near * a).astype(f32) + (far * zs).astype(f32)).astype(f32)
    inn = (one / near).astype(f32)
    inf_ = (one / far).astype(f32)
    den = ((inn * a).astype(f32) + (inf_ * zs).astype(f32)).astype(f32)
    return (one / den).astype(f32)


def linspace_table(n_coarse):
    """torch.linspace(0, 1-1/Kc, Kc): exactly k/Kc for power-of-two Kc; other Kc
    must use a table made by torch.linspace on the host."""
    assert n_coarse & (n_coarse - 1) == 0, "non power-of-two Kc: pass a torch.linspace table"
    return (np.arange(n_coarse, dtype=np.float64) / n_coarse).astype(f32)


def sample_coarse_np(rays, n_coarse, u, lindisp=False, lin=None):
    """src/render/nerf.py:98-118, every op rounded separately in fp32."""
    rays = np.asarray(rays, f32)
    u = np.asarray(u, f32)
    near, far = rays[:, 6:7], rays[:, 7:8]
    lin = linspace_table(n_coarse) if lin is None else np.asarray(lin, f32)
    step = f32(1.0 / n_coarse)
    zs = (lin[None, :] + (u * step).astype(f32)).astype(f32)
    return _lerp(near, far, zs, lindisp)


def sample_fine_np(rays, weights, n_coarse, u_bin, u_in, lindisp=False):
    """src/render/nerf.py:120-148.  Returns (z, inds)."""
    rays = np.asarray(rays, f32)
    w = (np.asarray(weights, f32) + f32(1e-5)).astype(f32)
    s = aten_row_sum_f32(w)
    pdf = (w / s[:, None]).astype(f32)
    cdf = np.cumsum(pdf.astype(np.float64), axis=-1).astype(f32)
    cdf = np.concatenate([np.zeros((w.shape[0], 1), f32), cdf], -1)
    u_bin = np.asarray(u_bin, f32)
    cnt = (cdf[:, None, :] <= u_bin[:, :, None]).sum(-1)
    inds = np.maximum(cnt.astype(f32) - f32(1.0), f32(0.0)).astype(f32)
    zs = ((inds + np.asarray(u_in, f32)).astype(f32) / f32(n_coarse)).astype(f32)
    return _lerp(rays[:, 6:7], rays[:, 7:8], zs, lindisp), inds


def sample_fine_depth_np(rays, depth, noise, depth_std):
    """src/render/nerf.py:150-161."""
    rays = np.asarray(rays, f32)
    z = (np.asarray(depth, f32)[:, None] + (np.asarray(noise, f32) * f32(depth_std)).astype(f32)).astype(f32)
    return np.maximum(np.minimum(z, rays[:, 7:8]), rays[:, 6:7]).astype(f32)
