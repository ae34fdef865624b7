# round-2 multi-GPU job (gpurun --gpus N): product wrappers on N GPUs -- tests, torchrun checks, sharded bench
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
echo "GPUs: $N"
timeout 1200 python -m pytest tests/test_gpu_multi.py -m gpu -q > gpurun_out/r02t_multi_tests_n$N.log 2>&1; echo "multi tests rc=$?" | tee -a gpurun_out/r02t_multi_tests_n$N.log
tail -40 gpurun_out/r02t_multi_tests_n$N.log | cut -c1-400
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29531 tests/multi_gpu_check.py > gpurun_out/r02t_multi_gpu_check_n$N.log 2>&1; echo "multi_gpu_check rc=$?" | tee -a gpurun_out/r02t_multi_gpu_check_n$N.log
grep -E "rank|rc=" gpurun_out/r02t_multi_gpu_check_n$N.log | tail -20
timeout 600 $TR --master-port 29532 tests/multi_gpu_train_check.py > gpurun_out/r02t_multi_gpu_train_check_n$N.log 2>&1; echo "multi_gpu_train_check rc=$?" | tee -a gpurun_out/r02t_multi_gpu_train_check_n$N.log
grep -E "rank|rc=" gpurun_out/r02t_multi_gpu_train_check_n$N.log | tail -20
timeout 900 $TR --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02t_bench_n$N.json 2> gpurun_out/r02t_bench_n$N.err; echo "bench rc=$?"
tail -3 gpurun_out/r02t_bench_n$N.err | cut -c1-600
python - <<PY
import json
lines = [l for l in open("gpurun_out/r02t_bench_n$N.json").read().splitlines() if l.startswith("{")]
d = json.loads(lines[-1])
print({k: d[k] for k in ("value", "n_gpus", "ms_per_step", "scaling", "gpu_launches")}, "e2e", d["e2e"]["value"])
print("roofline", {k: d["roofline"][k] for k in ("achieved", "executed_frac", "frac", "ms_per_step")}, "stage", d["stage_ms_per_step"])
print("sharded", json.dumps(d.get("sharded"), indent=0))
PY
timeout 600 $TR --master-port 29534 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/r02t_bench_reference_n$N.json 2> gpurun_out/r02t_bench_reference_n$N.err; echo "reference arm rc=$?"
tail -c 300 gpurun_out/r02t_bench_reference_n$N.json
