# round-2 GPU job: precise tensor-core chain (mlp_precision = fp16x3): its tests first, then the suite, parity, bench
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mlp_tc.py -m gpu -q -k "precise" > gpurun_out/r02j_x3_unit.log 2>&1; echo "x3 unit rc=$?"
tail -25 gpurun_out/r02j_x3_unit.log | cut -c1-300
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r02j_tests.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02j_tests.log
tail -25 gpurun_out/r02j_tests.log | cut -c1-300
timeout 600 python bench.py --precision fp16x3 --steps 5 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02j_bench_x3.json 2> gpurun_out/r02j_bench_x3.err; echo "bench x3 rc=$?"
tail -3 gpurun_out/r02j_bench_x3.err | cut -c1-300
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02j_bench_x3.json").read().strip().splitlines()[-1])
print("X3", round(d["value"]), "rays/s", round(d["ms_per_step"], 2), "ms", {k: round(d["roofline"][k], 3) for k in ("achieved", "executed", "executed_frac", "frac")}, d["stage_ms_per_step"], d["clocks"]["sm_mhz"])
PY
