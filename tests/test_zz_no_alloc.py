"""pn_mlp_forward_tc_static -- the no-allocation variant of the tensor-core ResnetFC chain (include/pixelnerf_b200.h;
no counterpart in the reference) against the existing entry point
pn_mlp_forward_tc, through the C ABI (ctypes), on two calls in a row.

The file name sorts last on purpose: a failure here must not hide other tests.
The scratch lives in a file-scope device array shared by every call and keeping the previous call's contents, so the
second call (different view count, different row count, different inputs) runs on top of the first one's leftovers.
Every buffer the entry point touches is allocated by the TEST before the call; between "before" and "after" the call
torch's caching allocator must not have been asked for anything (its counters are compared).

Tolerance: the chain has a fixed schedule and no atomics -- the existing determinism tests (tests/test_gpu_mlp_tc.py:
test_chain_is_deterministic) demand bit equality between launches, and so does this one."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

N_TEX = 8 * 24


def _case(sb, ns, ppo, seed):
    """Packed weights, per-texel tables and feat_x (positional code + taps record) of one call."""
    from helpers import make_taps
    from pixelnerf_b200 import _lib, synthetic as syn
    w = {k: v.cuda() for k, v in syn.mlp_weights(40 + seed).items()}
    g = torch.Generator().manual_seed(seed)
    latent = torch.rand(N_TEX, 512, generator=g)
    n = sb * ns * ppo
    idx = torch.randint(0, N_TEX, (n, 4), generator=g)
    wt = torch.rand(n, 4, generator=g)
    wt = wt / wt.sum(1, keepdim=True)
    fx = torch.zeros(n, 64)
    fx[:, :42] = torch.randn(n, 42, generator=g)
    packed = _lib.pack_weights_tc(w)
    tables = _lib.prepare_latent_tables(w, latent.reshape(1, 8, 24, 512).cuda())
    return packed, tables, make_taps(idx, wt, fx.half()).cuda()


def _static_call(packed, tables, feat_x, sb, ns, ppo, out, pre):
    """The new entry point through ctypes, raw pointers only; returns (rc, error text)."""
    from pixelnerf_b200 import _lib
    lib = _lib.lib()
    rc = lib.pn_mlp_forward_tc_static(packed.data_ptr(), tables.data_ptr(), tables.shape[1], feat_x.data_ptr(), sb, ns,
                                      ppo, out.data_ptr(), pre.data_ptr(), torch.cuda.current_stream().cuda_stream)
    return rc, lib.pn_last_error().decode()


def test_static_scratch_variant_matches_the_workspace_variant_twice_in_a_row():
    from pixelnerf_b200 import _lib
    lib = _lib.lib()
    limit = int(lib.pn_static_scratch_bytes())
    assert limit == 32 << 20
    torch.cuda.set_device(0)
    # call 1: two views (the scratch carries the fp16 hand-off of h, 1 KB per row) -- 2 x 2 x 1301 rows, ragged tiles
    # call 2: one view, more rows than call 1, different weights: runs over call 1's leftovers in the scratch
    for sb, ns, ppo, seed in ((2, 2, 1301, 1), (1, 1, 128 * 150 + 5, 2), (1, 3, 777, 3)):
        packed, tables, feat_x = _case(sb, ns, ppo, seed)
        ref_out, ref_pre = _lib.mlp_forward_tc(packed, tables, feat_x, sb, ns, ppo, want_pre=True)   # existing entry point
        out = torch.full((sb * ppo, 4), float("nan"), device="cuda")
        pre = torch.full((sb * ppo, 4), float("nan"), device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats()
        rc, err = _static_call(packed, tables, feat_x, sb, ns, ppo, out, pre)
        torch.cuda.synchronize()
        after = torch.cuda.memory_stats()
        assert rc == 0, err
        # nothing was requested from torch's allocator (and hence from cudaMalloc through it) during the call
        for key in ("allocation.all.allocated", "segment.all.allocated", "allocated_bytes.all.allocated"):
            assert after[key] == before[key], key
        assert torch.isfinite(out).all() and torch.isfinite(pre).all()
        assert torch.equal(out, ref_out) and torch.equal(pre, ref_pre), (
            float((out - ref_out).abs().max()), float((pre - ref_pre).abs().max()))
        # and once more on the same inputs: the array now holds this call's own leftovers
        out2, pre2 = torch.empty_like(out), torch.empty_like(pre)
        rc, err = _static_call(packed, tables, feat_x, sb, ns, ppo, out2, pre2)
        torch.cuda.synchronize()
        assert rc == 0, err
        assert torch.equal(out2, ref_out) and torch.equal(pre2, ref_pre)


def test_static_scratch_variant_rejects_sizes_beyond_its_array():
    from pixelnerf_b200 import _lib
    lib = _lib.lib()
    limit = int(lib.pn_static_scratch_bytes())
    packed, tables, feat_x = _case(1, 2, 64, 5)
    out = torch.empty(64, 4, device="cuda")
    # two views: 1 KB of scratch per row + the 256-byte header; the first row count that no longer fits
    too_many = (limit - 256) // 1024 + 1
    assert int(lib.pn_mlp_tc_workspace_bytes(too_many, 2)) > limit >= int(lib.pn_mlp_tc_workspace_bytes(too_many - 1, 2))
    assert too_many % 2 == 0
    rc = lib.pn_mlp_forward_tc_static(packed.data_ptr(), tables.data_ptr(), tables.shape[1], feat_x.data_ptr(), 1, 2,
                                      too_many // 2, out.data_ptr(), None, torch.cuda.current_stream().cuda_stream)
    msg = lib.pn_last_error().decode()
    assert rc != 0 and "static scratch" in msg and str(limit) in msg, (rc, msg)
    torch.cuda.synchronize()       # nothing was launched, nothing is pending
    # one view needs only the header: any row count is accepted (here simply a valid small call after the rejection)
    packed, tables, feat_x = _case(1, 1, 64, 6)
    pre = torch.empty(64, 4, device="cuda")
    rc, err = _static_call(packed, tables, feat_x, 1, 1, 64, out, pre)
    torch.cuda.synchronize()
    assert rc == 0, err
    ref = _lib.mlp_forward_tc(packed, tables, feat_x, 1, 1, 64)
    assert torch.equal(out, ref)
