# round-2 job: taps (SFU sine, 32-bit index split) and sample_coarse (template lindisp) -- tests, kernel table
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02q_gpu_tests.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/r02q_gpu_tests.log
timeout 300 python bench.py --kernels > gpurun_out/r02q_kernels.json 2> gpurun_out/r02q_kernels.err; echo "kernels rc=$?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r02q_kernels.json").read().strip().splitlines()[-1])
for k, v in d["kernels"].items():
    print("KERNEL", k, round(v["ms"], 4), "ms", round(v["frac_of_measured_hbm"], 3))
PY
timeout 300 python tests/parity_report.py > gpurun_out/r02q_parity.json 2> gpurun_out/r02q_parity.err; echo "parity rc=$?"
tail -c 1500 gpurun_out/r02q_parity.json
