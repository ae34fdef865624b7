# round-2: cooperative fp32 table gather in the precise / two-pass chains -- tests and bench
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mlp_tc.py tests/test_gpu_fullsize.py tests/test_gpu_render.py -m gpu -x -q > gpurun_out/r02y_tests.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/r02y_tests.log | cut -c1-300
for PR in fp16x3 fp16x2; do
  timeout 600 python bench.py --precision $PR --steps 5 --no-extras --no-cpu-baseline > gpurun_out/r02y_bench_$PR.json 2> gpurun_out/r02y_bench_$PR.err; echo "bench $PR rc=$?"
  python - <<PY
import json
d = json.loads(open("gpurun_out/r02y_bench_$PR.json").read().strip().splitlines()[-1])
print("PREC $PR", round(d["value"]), round(d["e2e"]["value"]), round(d["ms_per_step"], 2), {k: round(v, 3) for k, v in d["roofline"].items() if isinstance(v, float)}, d["clocks"]["sm_mhz"])
PY
done
