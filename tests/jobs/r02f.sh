# round-2 GPU job: full suite after the grouped (one-launch NS > 1) chain, then the multi-view benches
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r02f_tests.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02f_tests.log
tail -30 gpurun_out/r02f_tests.log | cut -c1-300
for SC in srn_chair dtu srn_car; do
  timeout 600 python bench.py --scene $SC --steps 5 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02f_bench_$SC.json 2> gpurun_out/r02f_bench_$SC.err; echo "bench $SC rc=$?"
  python - <<PY
import json
d = json.loads(open("gpurun_out/r02f_bench_$SC.json").read().strip().splitlines()[-1])
print("$SC", round(d["value"]), "rays/s", round(d["ms_per_step"], 2), "ms", {k: round(d["roofline"][k], 3) for k in ("achieved", "executed_frac", "frac", "ms_per_step", "launches_per_step")}, d["stage_ms_per_step"], d["clocks"]["sm_mhz"])
PY
done
