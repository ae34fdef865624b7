# A/B in ONE job (same box, same clocks): n_views > 1 chain as one grouped launch vs the two-launch form
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for SC in srn_chair dtu; do
 for MODE in grouped two_launch grouped two_launch; do
  TL=0; [ $MODE = two_launch ] && TL=1
  PIXELNERF_B200_STAGED=1 PIXELNERF_B200_TWO_LAUNCH=$TL timeout 600 python bench.py --scene $SC --steps 5 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r02g_${SC}_$MODE.json 2> gpurun_out/r02g_${SC}_$MODE.err; echo "rc=$?"
  python - <<PY
import json
d = json.loads(open("gpurun_out/r02g_${SC}_$MODE.json").read().strip().splitlines()[-1])
print("AB $SC $MODE", round(d["value"]), "rays/s", round(d["ms_per_step"], 2), "ms; mlp", round(d["stage_ms_per_step"]["mlp"], 2), "ms in", d["roofline"]["launches_per_step"], "launches; features", round(d["stage_ms_per_step"]["features"], 2), "clk", d["clocks"]["sm_mhz"])
PY
 done
done
