# round-2 job: training step captured in a CUDA graph (device-resident call seed)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_graph.py -x -q -s > gpurun_out/r02r_graph_tests.log 2>&1; echo "tests rc=$?"
tail -30 gpurun_out/r02r_graph_tests.log | cut -c1-400
timeout 600 python bench.py --train --graph --steps 5 > gpurun_out/r02r_train_graph.json 2> gpurun_out/r02r_train_graph.err; echo "train rc=$?"
tail -3 gpurun_out/r02r_train_graph.err | cut -c1-400
python - <<PY
import json
d = json.loads(open("gpurun_out/r02r_train_graph.json").read().strip().splitlines()[-1])
for s in d["sweep"]:
    print("TRAIN", s["rays_per_step"], "eager", round(s["ms_per_step"], 3), "ms; graph", s.get("cuda_graph"))
PY
