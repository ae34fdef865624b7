# timing experiment (wrong results by construction): how much of the chain is the table gather / the epilogues
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for EXP in none gather_loads_only gather_no_loads no_gather none gather_loads_only gather_no_loads; do
  if [ $EXP = none ]; then unset PN_TC_EXPERIMENT; else export PN_TC_EXPERIMENT=$EXP; fi
  PIXELNERF_B200_STAGED=1 timeout 300 python bench.py --steps 5 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02o_$EXP.json 2>/dev/null
  python - <<PY
import json
d = json.loads(open("gpurun_out/r02o_$EXP.json").read().strip().splitlines()[-1])
print("EXP $EXP", round(d["value"]), "rays/s; chain", round(d["stage_ms_per_step"]["mlp"], 2), "ms; clk", d["clocks"]["sm_mhz"])
PY
done
