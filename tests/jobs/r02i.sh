# round-2 GPU job: suite after the fused ray source (pn_render_frames), headline bench for regressions
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/r02i_tests.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02i_tests.log
tail -25 gpurun_out/r02i_tests.log | cut -c1-400
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02i_bench.json 2> gpurun_out/r02i_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02i_bench.json").read().strip().splitlines()[-1])
print("HEAD", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"], 3), d["stage_ms_per_step"], d["clocks"]["sm_mhz"], "traffic", d["roofline"]["traffic"])
print("KERN", {k: round(v["frac_of_measured_hbm"], 3) for k, v in d["roofline_kernels"]["kernels"].items()})
print("SHARDED", {k: round(v["rays/s"]) for k, v in d["sharded"].items()})
PY
