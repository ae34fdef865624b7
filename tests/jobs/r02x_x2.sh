# round-2: the two-pass chain (mlp_precision = fp16x2 = the precise chain without its a_lo pass) -- tests, bench, parity
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mlp_tc.py tests/test_gpu_fullsize.py -m gpu -x -q > gpurun_out/r02x_tests.log 2>&1; echo "tests rc=$?"
tail -15 gpurun_out/r02x_tests.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4
for SC in srn_car dtu srn_chair; do
  timeout 600 python bench.py --precision fp16x2 --scene $SC --steps 5 --no-extras --no-cpu-baseline > gpurun_out/r02x_bench_x2_$SC.json 2> gpurun_out/r02x_bench_x2_$SC.err; echo "bench $SC rc=$?"
  python - <<PY
import json
d = json.loads(open("gpurun_out/r02x_bench_x2_$SC.json").read().strip().splitlines()[-1])
print("X2 $SC", round(d["value"]), round(d["e2e"]["value"]), round(d["ms_per_step"], 2), {k: round(v, 3) for k, v in d["roofline"].items() if isinstance(v, float)}, d["clocks"]["sm_mhz"])
PY
done
timeout 600 python tests/parity_report.py --out gpurun_out/r02_parity.json > gpurun_out/r02x_parity.log 2>&1; echo "parity rc=$?"
grep "fp16x2" gpurun_out/r02x_parity.log | cut -c1-330
