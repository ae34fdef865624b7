# round-2: the training step ray-sharded over N GPUs (torchrun), strong scaling of one step
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 600 python bench.py --train --steps 5 > gpurun_out/r02w_train_n1.json 2> gpurun_out/r02w_train_n1.err; echo "n1 rc=$?"
for G in 2 4 8; do
  [ $G -le $N ] || continue
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 2954$G bench.py --train --gpus $G --steps 5 > gpurun_out/r02w_train_n$G.json 2> gpurun_out/r02w_train_n$G.err; echo "n$G rc=$?"
  tail -3 gpurun_out/r02w_train_n$G.err | cut -c1-300
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r02w_train_n*.json")):
    try:
        d = json.loads([l for l in open(f).read().splitlines() if l.startswith("{")][-1])
    except Exception as e:
        print("TRAINN", f, "unreadable", e); continue
    print("TRAINN", d.get("n_gpus"), [(s["rays_per_step"], round(s["ms_per_step"], 3), round(s["loss"], 5)) for s in d["sweep"]])
PY
