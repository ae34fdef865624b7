# round-2: bench of the two-pass and precise chains on the three scene classes (final kernels)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
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
