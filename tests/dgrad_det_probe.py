"""Diagnostic: run-to-run determinism of the rows GEMM's dgrad forms (not a test)."""
import sys
import torch
sys.path[:0] = ["."]
from pixelnerf_b200 import _lib
lib = _lib.lib()
st = torch.cuda.current_stream().cuda_stream
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 76837
g = torch.Generator().manual_seed(1)
a = torch.randn(rows, 512, generator=g).bfloat16().cuda()
w = (torch.randn(512, 512, generator=g) * 0.05).bfloat16().cuda()
mask = torch.relu(torch.randn(rows, 512, generator=g)).bfloat16().cuda()
r0 = torch.randn(rows, 512, generator=g).bfloat16().cuda()
cs = torch.zeros(512, device="cuda")
ref1 = ((a.double() @ w.double()) * (mask.double() > 0))
first = None
for rep in range(int(sys.argv[2]) if len(sys.argv) > 2 else 12):
    out = torch.empty(rows, 512, device="cuda", dtype=torch.bfloat16)
    res = r0.clone()
    oop = torch.empty(rows, 512, device="cuda", dtype=torch.bfloat16)
    _lib._check(lib.pn_gemm_tc_rows(a.data_ptr(), 1, w.data_ptr(), 1, rows, 512, 512, mask.data_ptr(), None, 0, None,
                                    out.data_ptr(), 1.0, cs.data_ptr(), None, None, st), "fc_1 form")
    _lib._check(lib.pn_gemm_tc_rows(a.data_ptr(), 1, w.data_ptr(), 1, rows, 512, 512, mask.data_ptr(), None, 0,
                                    res.data_ptr(), res.data_ptr(), 1.0, cs.data_ptr(), cs.data_ptr(), None, st), "fc_0 in place")
    _lib._check(lib.pn_gemm_tc_rows(a.data_ptr(), 1, w.data_ptr(), 1, rows, 512, 512, mask.data_ptr(), None, 0,
                                    r0.data_ptr(), oop.data_ptr(), 1.0, cs.data_ptr(), cs.data_ptr(), None, st), "fc_0 out of place")
    torch.cuda.synchronize()
    e1 = (out.double() - ref1).abs().max().item()
    e2 = (res.double() - (ref1 + r0.double())).abs().max().item()
    e3 = (oop.double() - (ref1 + r0.double())).abs().max().item()
    if first is None:
        first = (out.clone(), res.clone(), oop.clone())
    d = [int((x != y).sum()) for x, y in zip((out, res, oop), first)]
    where = []
    for x, y in zip((out, res, oop), first):
        bad = (x != y).nonzero()
        if len(bad):
            where.append((bad[:, 0].min().item(), bad[:, 0].max().item(), bad[:, 1].min().item(), bad[:, 1].max().item()))
    print(f"rep {rep}: err vs fp64 fc_1 {e1:.3e} in-place {e2:.3e} out-of-place {e3:.3e}; elements differing from rep 0: {d} {where}")
