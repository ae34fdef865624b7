"""Run under torchrun (one rank per GPU): the ray-sharded render of one frame must equal rank 0's
single-GPU render of the same frame bit for bit (same injected draws, sliced by ray range), with the
encoded scene produced on rank 0 only and broadcast over NCCL.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/multi_gpu_check.py
"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pixelnerf_b200 import parallel, synthetic as syn  # noqa: E402
from pixelnerf_b200.testing import build_net  # noqa: E402


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rank, world = dist.get_rank(), dist.get_world_size()
    sc = syn.scene("srn_chair")          # NS = 2: exercises the two-segment chain as well
    wc, wf = syn.mlp_weights(1), syn.mlp_weights(2)
    ok = True
    for prec in ("fp32", "fp16"):
        net, ren = build_net(sc, wc, wf, dev, mlp_precision=prec)
        if rank != 0:   # only rank 0 "encoded" the scene; wipe the others, then broadcast
            net.encoder.latent = torch.zeros(1, 1, 1, 1, device=dev)
            net.poses = torch.zeros(1, 3, 4, device=dev)
        parallel.broadcast_encoded(net, src=0)
        n = 3001                          # not divisible by the world size
        rays = sc["rays"][:n].to(dev)[None]
        draws = {k: v.to(dev) for k, v in syn.draws(5, n, 64, 16, 16).items()}
        single = ren.bind_parallel(net, [local], simple_output=True).eval()
        sharded = ren.bind_parallel(net, list(range(world)), simple_output=True).eval()
        with torch.no_grad():
            ren.injected_draws = dict(draws)
            rgb1, d1 = single(rays)
            ren.injected_draws = dict(draws)
            rgbn, dn = sharded(rays)
        same = torch.equal(rgb1, rgbn) and torch.equal(d1, dn)
        # every rank must hold the same gathered frame
        chk = torch.stack([rgbn.double().sum(), dn.double().sum()])
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same = same and torch.equal(lo, hi)
        print(f"rank {rank} {prec}: sharded == single: {same}", flush=True)
        ok = ok and same
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
