# round-2 GPU job: the no-allocation entry point's test alone (fresh process), then at the end of the whole suite
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02c_noalloc_alone.log 2>&1; echo "alone rc=$?" | tee -a gpurun_out/r02c_noalloc_alone.log
tail -15 gpurun_out/r02c_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02c_tests.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/r02c_tests.log
tail -12 gpurun_out/r02c_tests.log
