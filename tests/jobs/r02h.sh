# round-2 GPU job: suite + same-box A/B of the one-launch (per-view sub-tiles, running sum) chain vs the two-launch form
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/r02h_tests.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02h_tests.log
tail -12 gpurun_out/r02h_tests.log | cut -c1-300
bash tests/jobs/r02g_ab.sh 2>&1 | grep "^AB"
