# round-2 ncu job (ONE gpurun call, one GPU): --set full of the chain after the one-launch NS > 1 form and of the precise
# chain.  Every ncu run follows a plain run of the same command line that exited 0.
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for SPEC in "srn_car fp16" "srn_chair fp16" "dtu fp16" "srn_car fp16x3"; do
  set -- $SPEC
  CMD="python tests/ncu_frame.py --scene $1 --frames 2 --precision $2"
  $CMD > gpurun_out/r02m_plain_$1_$2.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:mlp_chain_kernel -s 2 -c 2 -f -o gpurun_out/r02m_chain_$1_$2 $CMD > gpurun_out/r02m_ncu_$1_$2.log 2>&1
  echo "$1 $2 rc=$?"; tail -1 gpurun_out/r02m_plain_$1_$2.log
  # gpurun returns at most 64 MiB: keep the raw page as csv, and the report itself only for the headline capture
  ncu -i gpurun_out/r02m_chain_$1_$2.ncu-rep --page raw --csv > gpurun_out/r02m_chain_$1_$2_raw.csv
  [ "$SPEC" = "srn_car fp16" ] || rm -f gpurun_out/r02m_chain_$1_$2.ncu-rep
done
ls -la gpurun_out/
