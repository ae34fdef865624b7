# round-2 GPU job: loss-curve test (prints both curves)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_train_curve.py -m gpu -x -q -s > gpurun_out/r02l_curve.log 2>&1; echo "curve rc=$?" | tee -a gpurun_out/r02l_curve.log
tail -30 gpurun_out/r02l_curve.log | cut -c1-400
