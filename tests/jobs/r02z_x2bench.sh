# round-2: final kernels -- whole GPU suite, bench of the two-pass and precise chains on the three scene classes, parity report
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02z_gpu_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/r02z_gpu_tests.log
for PR in fp16x2 fp16x3; do
for SC in srn_car srn_chair dtu; do
  timeout 600 python bench.py --precision $PR --scene $SC --steps 5 --no-extras --no-cpu-baseline > gpurun_out/r02z_bench_${PR}_$SC.json 2> gpurun_out/r02z_bench_${PR}_$SC.err; echo "bench $PR $SC rc=$?"
  python - <<PY
import json
d = json.loads(open("gpurun_out/r02z_bench_${PR}_$SC.json").read().strip().splitlines()[-1])
print("PS $PR $SC", round(d["value"]), round(d["e2e"]["value"]), round(d["ms_per_step"], 2), round(d["roofline"]["executed_frac"], 3), d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
PY
done
done
timeout 600 python tests/parity_report.py --out gpurun_out/r02_parity.json > gpurun_out/r02z_parity.log 2>&1; echo "parity rc=$?"
grep "fp16x2\|fp16x3" gpurun_out/r02z_parity.log | cut -c1-230
