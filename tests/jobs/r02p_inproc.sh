# round-2 multi-GPU job: ONE process driving all visible GPUs (DeviceSharded), at 1 .. N devices
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
for G in 1 2 $N; do
  [ $G -le $N ] || continue
  timeout 600 python bench.py --in-process --gpus $G --steps 5 > gpurun_out/r02p_inproc_n$G.json 2> gpurun_out/r02p_inproc_n$G.err; echo "n$G rc=$?"
  tail -2 gpurun_out/r02p_inproc_n$G.err | cut -c1-300
  python - <<PY
import json
d = json.loads(open("gpurun_out/r02p_inproc_n$G.json").read().strip().splitlines()[-1])
print("INPROC n=$G", {k: (round(v["rays/s"]), round(v["ms_per_call"], 1)) for k, v in d["workloads"].items()})
PY
done
