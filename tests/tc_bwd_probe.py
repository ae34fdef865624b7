"""Diagnostic (not a test): per-tensor gradient error of the tensor-core backward, in backward order."""
import sys
import torch
sys.path[:0] = ["tests", "."]
from oracle import model as omodel
from pixelnerf_b200 import _lib, synthetic as syn
from helpers import make_taps
from test_gpu_mlp_tc import _inputs, _tables

ns, ppo, sb = (int(v) for v in (sys.argv[1:4] if len(sys.argv) > 3 else (1, 300, 1)))
w = syn.mlp_weights(11)
n = sb * ns * ppo
idx, wt, latent, fx = _inputs(n, seed=5)
g_out = torch.randn(sb * ppo, 4, generator=torch.Generator().manual_seed(1))
torch.set_default_dtype(torch.float64)
wr = {k: v.double().requires_grad_(True) for k, v in w.items()}
zr = (wt.double()[:, :, None] * latent.double()[idx]).sum(1).requires_grad_(True)
outs, pres = [], []
for o in range(sb):
    sl = slice(o * ns * ppo, (o + 1) * ns * ppo)
    wd = dict(wr); wd["_points_per_obj"] = ppo
    pre = omodel.resnetfc_forward(wd, torch.cat([zr[sl], fx[sl, :42].double()], -1), ns)
    pres.append(pre)
    outs.append(torch.cat([torch.sigmoid(pre[:, :3]), torch.relu(pre[:, 3:4])], -1))
out_ref = torch.cat(outs)
MODE = sys.argv[4] if len(sys.argv) > 4 else "full"
if MODE == "full":
    (out_ref * g_out.double()).sum().backward()
torch.set_default_dtype(torch.float32)
wd = {k: v.cuda() for k, v in w.items()}
fx_t = make_taps(idx, wt, fx).cuda()
lat = latent.reshape(1, 8, 24, 512).cuda().contiguous()
out, acts = _lib.mlp_forward_tc_train(_lib.pack_weights_tc(wd), _tables(w, latent), fx_t, sb, ns, ppo)
if MODE != "full":   # reference backward from the kernel's own head gradient: isolates the MLP backward
    o = out.cpu().double()
    gp = torch.cat([o[:, :3] * (1 - o[:, :3]) * g_out[:, :3].double(), (o[:, 3:] > 0) * g_out[:, 3:].double()], -1)
    torch.cat(pres).backward(gp)
grads = {k: torch.zeros_like(v) for k, v in wd.items()}
g_fz = _lib.mlp_backward_tc(_lib.pack_weights_fp32(wd), _lib.pack_weights_bf16(wd), fx_t, lat, sb, ns, ppo, acts,
                            out, g_out.cuda(), grads, True)
torch.cuda.synchronize()
def rel(a, b): return (a.cpu().double() - b).abs().max().item() / max(b.abs().max().item(), 1e-12)
if MODE == "emu":
    from helpers import emulate_tc_backward
    m = sb * ppo
    a = _lib.tc_acts_view(acts, n)
    al = [a[i].cpu().double() if i < 6 or ns == 1 else a[i, :m].cpu().double() for i in range(11)]
    z64 = (wt.double()[:, :, None] * latent.double()[idx]).sum(1)
    ge, gz = emulate_tc_backward(w, al, z64, fx[:, :42].double(), out.cpu(), g_out, sb, ns, ppo)
    class _G:  # same interface as the autograd reference below
        def __init__(self, t): self.grad = t
    wr = {k: _G(v) for k, v in ge.items()}
    zr = _G(gz)
print("out", (out.cpu().double() - out_ref.detach()).abs().max().item())
order = ["lin_out"] + [f"blocks.{b}.fc_{j}" for b in (4, 3, 2, 1, 0) for j in (1, 0)] + ["lin_z.2", "lin_z.1", "lin_z.0", "lin_in"]
for k in order:
    print(f"{k:16s} W {rel(grads[k + '.weight'], wr[k + '.weight'].grad):.2e}  b {rel(grads[k + '.bias'], wr[k + '.bias'].grad):.2e}"
          f"   |gW|max {wr[k + '.weight'].grad.abs().max().item():.2e}")
print("g_feat_z", rel(g_fz, zr.grad))
# saved activations vs the float64 forward: slot 10 = relu(h5)
a = _lib.tc_acts_view(acts, n)
print("acts finite", torch.isfinite(a.float()).all().item(), "slot means", [round(a[i].float().mean().item(), 4) for i in range(11)])
