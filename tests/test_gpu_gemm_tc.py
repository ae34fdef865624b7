"""Tensor-core GEMMs of the training path (csrc/pn_gemm_tc.cu) against torch fp32/fp64 matmuls of the same
16-bit-rounded operands: the only differences left are the fp32 accumulation order and the final bf16 rounding
of the 16-bit output copy (2^-9 relative)."""
import pytest
import torch

from pixelnerf_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rand(shape, dtype, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(dtype).to(DEV)


@pytest.mark.parametrize("rows", [1, 200, 256, 1000, 70000])
@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
def test_rows_gemm_plain(rows, dt):
    a, b = _rand((rows, 512), dt, 1), _rand((512, 512), dt, 2, 0.05)
    ref = a.double() @ b.double()
    acc = torch.empty(rows, 512, device=DEV)
    _, o16 = _lib.gemm_tc_rows(a, b, acc32=acc)
    torch.cuda.synchronize()
    scale = ref.abs().max().item()
    assert (acc.double() - ref).abs().max().item() < 2e-5 * scale
    assert (o16.double() - ref).abs().max().item() < 2.0 ** -8 * scale


def test_rows_gemm_mask_accumulate_scale():
    rows = 3001
    a, b = _rand((rows, 512), torch.bfloat16, 3), _rand((512, 512), torch.bfloat16, 4, 0.05)
    mask = torch.relu(_rand((rows, 512), torch.bfloat16, 5))
    prev = _rand((rows, 512), torch.float32, 6)
    ref = (a.double() @ b.double()) * 0.5 * (mask > 0) + prev.double()
    acc = prev.clone()
    _, o16 = _lib.gemm_tc_rows(a, b, mask=mask, acc32=acc, accumulate=True, scale=0.5)
    scale = ref.abs().max().item()
    assert (acc.double() - ref).abs().max().item() < 2e-5 * scale
    assert (o16.double() - ref).abs().max().item() < 2.0 ** -8 * scale
    # 16-bit output only
    _, only16 = _lib.gemm_tc_rows(a, b, mask=mask, scale=0.5)
    ref2 = (a.double() @ b.double()) * 0.5 * (mask > 0)
    assert (only16.double() - ref2).abs().max().item() < 2.0 ** -8 * ref2.abs().max().item()


@pytest.mark.parametrize("rows", [1, 63, 64, 777, 9473, 200000])
@pytest.mark.parametrize("dt", [torch.float16, torch.bfloat16])
def test_wgrad_gemm(rows, dt):
    g, x = _rand((rows, 512), dt, 7), _rand((rows, 512), dt, 8)
    out = torch.full((512, 512), 0.25, device=DEV)
    _lib.gemm_tc_wgrad(g, x, out)
    ref = g.double().t() @ x.double() + 0.25
    assert (out.double() - ref).abs().max().item() < 3e-5 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("rows", [5, 4097])
def test_wgrad_gemm_code_rows(rows):
    """lin_in: 64-wide code rows; columns past n_store may hold anything (NaN included)."""
    g = _rand((rows, 512), torch.bfloat16, 9)
    x = _rand((rows, 64), torch.bfloat16, 10)
    x[:, 48:] = float("nan")
    out = torch.zeros(512, 42, device=DEV)
    _lib.gemm_tc_wgrad(g, x, out, n_store=42)
    ref = g.double().t() @ x[:, :42].double()
    assert torch.isfinite(out).all()
    assert (out.double() - ref).abs().max().item() < 3e-5 * max(1.0, ref.abs().max().item())


def test_mixed_operand_formats_raise():
    """tcgen05 kind::f16 takes fp16 or bf16 operands but not one of each (illegal instruction on sm_100a)."""
    a, b = _rand((8, 512), torch.bfloat16, 1), _rand((512, 512), torch.float16, 2)
    with pytest.raises(RuntimeError, match="same format"):
        _lib.gemm_tc_rows(a, b)
    with pytest.raises(RuntimeError, match="same format"):
        _lib.gemm_tc_wgrad(a, _rand((8, 512), torch.float16, 3), torch.zeros(512, 512, device=DEV))


@pytest.mark.parametrize("rows", [77, 256 * 149 + 3])
def test_rows_gemm_residual_and_column_sums(rows):
    """dgrad of fc_0 as the backward uses it: masked product added to the bf16 residual-stream gradient in place,
    column sums of the result (bias gradients) added to two destinations."""
    a, b = _rand((rows, 512), torch.bfloat16, 11), _rand((512, 512), torch.bfloat16, 12, 0.05)
    mask = torch.relu(_rand((rows, 512), torch.bfloat16, 13))
    resid = _rand((rows, 512), torch.bfloat16, 14)
    ref = (a.double() @ b.double()) * (mask > 0) + resid.double()
    cs0, cs1 = torch.full((512,), 2.0, device=DEV), torch.zeros(512, device=DEV)
    out = resid.clone()
    lib = _lib.lib()
    _lib._check(lib.pn_gemm_tc_rows(a.data_ptr(), 1, b.data_ptr(), 1, rows, 512, 512, mask.data_ptr(), None, 0,
                                    out.data_ptr(), out.data_ptr(), 1.0, cs0.data_ptr(), cs1.data_ptr(), None,
                                    torch.cuda.current_stream().cuda_stream), "pn_gemm_tc_rows")
    scale = ref.abs().max().item()
    assert (out.double() - ref).abs().max().item() < 2.0 ** -8 * scale
    col = ref.sum(0)
    tol = 2.0 ** -9 * ref.abs().sum(0).max().item() / rows ** 0.5 * 4 + 1e-3   # bf16-exact values are not summed: fp32 sums of the unrounded result
    assert (cs1.double() - col).abs().max().item() < max(tol, 1e-4 * col.abs().max().item())
    assert torch.allclose(cs0 - 2.0, cs1, rtol=1e-4, atol=1e-3 * float(col.abs().max()))


@pytest.mark.parametrize("rows,slabs", [(1000, 3), (256, 1), (70000, 3), (1, 2)])
def test_rows_gemm_over_slabs(rows, slabs):
    """pn_gemm_tc_rows_slabs: K spread over several 512-wide operands stored slab_rows apart (dL/dz = sum_b dL/dh_b
    W_z[b] in one launch).  The padding rows of every slab hold NaN: they must never reach a stored row."""
    from pixelnerf_b200 import _lib
    g = torch.Generator().manual_seed(rows)
    pitch = (rows + 255) // 256 * 256
    a = torch.full((slabs, pitch, 512), float("nan"))
    a[:, :rows] = torch.randn(slabs, rows, 512, generator=g)
    a = a.bfloat16().cuda()
    b = (torch.randn(slabs * 512, 512, generator=g) * 0.05).bfloat16().cuda()
    out = _lib.gemm_tc_rows_slabs(a, b, rows)
    ref = sum(a[s, :rows].double() @ b[s * 512:(s + 1) * 512].double() for s in range(slabs))
    assert torch.isfinite(out).all()
    assert (out.double() - ref).abs().max().item() < 2e-5 * max(1.0, ref.abs().max().item())
    out2 = _lib.gemm_tc_rows_slabs(a, b, rows, acc32=out.clone(), accumulate=True, scale=0.5)
    assert (out2.double() - 1.5 * ref).abs().max().item() < 4e-5 * max(1.0, ref.abs().max().item())
