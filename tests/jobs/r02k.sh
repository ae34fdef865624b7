# round-2 GPU job: suite (-x, as the driver runs it) + the measured parity table with all three MLP paths
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02k_tests.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02k_tests.log
tail -8 gpurun_out/r02k_tests.log | cut -c1-300
timeout 1200 python tests/parity_report.py --out gpurun_out/r02k_parity.json > gpurun_out/r02k_parity.log 2>&1; echo "parity rc=$?"
grep -E "fp16x3" gpurun_out/r02k_parity.log | cut -c1-330
