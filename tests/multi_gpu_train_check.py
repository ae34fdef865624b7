"""Run under torchrun (one rank per GPU): a ray-sharded TRAINING step -- every rank renders and
back-propagates its own ray range of the same batch through the hand-written backward, then
parallel.allreduce_gradients sums the weight gradients over NCCL -- must give the single-GPU gradients of the same
batch (same injected draws): fp32 path to 1e-4 of each tensor's largest entry (summation order differs), tensor-core
path to 3e-2 (bf16 gradients, relu decisions at the rounding level).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29512 tests/multi_gpu_train_check.py
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pixelnerf_b200 import parallel, synthetic as syn  # noqa: E402
from pixelnerf_b200.testing import build_net, scene_camera, set_encoded  # noqa: E402


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    sc = syn.scene("sn64")
    sb, n = 2, 301
    ok = True
    for prec, tol in (("fp32", 1e-4), ("fp16", 3e-2)):
        net, ren = build_net(sc, syn.mlp_weights(1), syn.mlp_weights(2), dev, mlp_precision=prec)
        cam = scene_camera(sc)
        lat = syn.latent_maps(21, sb, sc["hl"], sc["wl"])
        set_encoded(net, lat, cam["poses"].expand(sb, -1, -1).contiguous(), cam["focal"], cam["c"], cam["image_shape"], 1)
        net.train()
        ren.train()
        mlps = [net.mlp_coarse, net.mlp_fine]
        for q in net.parameters():
            q.requires_grad_(True)
        g = torch.Generator().manual_seed(5)
        idx = torch.randint(0, sc["rays"].shape[0], (sb, n), generator=g)
        rays = sc["rays"][idx].to(dev)
        target = torch.rand(sb, n, 3, generator=g).to(dev)
        draws = {k: v.to(dev) for k, v in syn.draws(9, sb * n, 64, 16, 16).items()}

        def run(par):
            for m in mlps:
                m.zero_grad(set_to_none=True)
            ren.injected_draws = dict(draws)
            out = par(rays, want_weights=True)
            loss = ((out["coarse"]["rgb"] - target) ** 2).mean() + ((out["fine"]["rgb"] - target) ** 2).mean()
            loss.backward()
            return float(loss)

        single = ren.bind_parallel(net, [local])
        sharded = ren.bind_parallel(net, list(range(world)))
        loss1 = run(single)
        ref = [q.grad.clone() for m in mlps for q in m.parameters()]
        lossn = run(sharded)
        nbytes = parallel.allreduce_gradients(mlps)
        got = [q.grad for m in mlps for q in m.parameters()]
        worst = max(float((a - b).abs().max() / b.abs().max().clamp_min(1e-12)) for a, b in zip(got, ref))
        same = worst < tol and abs(loss1 - lossn) < 1e-5 + tol * abs(loss1)
        print(f"rank {rank} {prec}: loss {loss1:.6f} vs {lossn:.6f}, worst gradient tensor error {worst:.2e} "
              f"(bar {tol}), all-reduced {nbytes / 1e6:.1f} MB: {same}", flush=True)
        ok = ok and same
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
