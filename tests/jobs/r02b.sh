# round-2 GPU job: full GPU test suite, measured parity table, headline bench (results land in gpurun_out/)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r02b_tests.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02b_tests.log
tail -25 gpurun_out/r02b_tests.log
timeout 900 python tests/parity_report.py --out gpurun_out/r02b_parity.json > gpurun_out/r02b_parity.log 2>&1; echo "parity rc=$?"
grep -E "fp16" gpurun_out/r02b_parity.log | cut -c1-420
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r02b_bench.json 2> gpurun_out/r02b_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02b_bench.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("value", "ms_per_step", "gpu_launches", "clocks")})
print("e2e", d["e2e"]["value"], "roofline", {k: d["roofline"][k] for k in ("achieved", "executed", "executed_frac", "frac", "ms_per_step")})
print("stage", d["stage_ms_per_step"]); print("parity", d.get("parity"))
print("sharded", {k: (v["rays/s"], v["ms_per_call"]) for k, v in d.get("sharded", {}).items()})
PY
tail -5 gpurun_out/r02b_bench.err
