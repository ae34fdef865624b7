"""Diagnostic (not a test): the three training GEMM shapes at one training chunk's size, for ncu."""
import sys
import torch
sys.path[:0] = ["."]
from pixelnerf_b200 import _lib

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1310720
g = torch.Generator().manual_seed(0)
a = torch.randn(rows, 512, generator=g).bfloat16().cuda()
w = (torch.randn(512, 512, generator=g) * 0.05).bfloat16().cuda()
x = torch.relu(torch.randn(rows, 512, generator=g)).bfloat16().cuda()
gh = torch.randn(rows, 512, generator=g).bfloat16().cuda()
cs = torch.zeros(512, device="cuda")
dw = torch.zeros(512, 512, device="cuda")
lib = _lib.lib()
st = torch.cuda.current_stream().cuda_stream
out = torch.empty_like(a)
for rep in range(3 if len(sys.argv) < 3 else int(sys.argv[2])):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    ev[0].record()
    _lib._check(lib.pn_gemm_tc_rows(a.data_ptr(), 1, w.data_ptr(), 1, rows, 512, 512, x.data_ptr(), None, 0, None,
                                    out.data_ptr(), 1.0, cs.data_ptr(), None, None, st), "rows fc_1")
    ev[1].record()
    _lib._check(lib.pn_gemm_tc_rows(a.data_ptr(), 1, w.data_ptr(), 1, rows, 512, 512, x.data_ptr(), None, 0,
                                    gh.data_ptr(), gh.data_ptr(), 1.0, cs.data_ptr(), cs.data_ptr(), None, st), "rows fc_0")
    ev[2].record()
    _lib.gemm_tc_wgrad(a, x, dw)
    ev[3].record()
    torch.cuda.synchronize()
    t = [ev[i].elapsed_time(ev[i + 1]) for i in range(3)]
    fl = rows * 512 * 512 * 2 / 1e9
    print(f"rows {rows}: dgrad fc_1 {t[0]:.3f} ms ({fl / t[0]:.0f} TF/s, {rows * 3072 / t[0] / 1e6:.0f} GB/s)  "
          f"dgrad fc_0 {t[1]:.3f} ms ({fl / t[1]:.0f} TF/s, {rows * 4096 / t[1] / 1e6:.0f} GB/s)  wgrad {t[2]:.3f} ms ({fl / t[2]:.0f} TF/s)")
