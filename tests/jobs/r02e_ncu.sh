# round-2 ncu job (ONE gpurun call, one GPU): launch list of a frame + --set full captures of the chain per view count.
# Every ncu run follows a plain run of the same command line that exited 0.
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD1="python tests/ncu_frame.py --scene srn_car --frames 3"
$CMD1 > gpurun_out/r02e_plain_car.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02e_launches_srn_car.csv $CMD1 > gpurun_out/r02e_ncu_launches.log 2>&1
echo "launch list rc=$?"; tail -2 gpurun_out/r02e_plain_car.log
for SC in srn_car srn_chair dtu; do
  L=4; [ $SC = srn_car ] && L=2
  CMD="python tests/ncu_frame.py --scene $SC --frames 2"
  $CMD > gpurun_out/r02e_plain_$SC.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:mlp_chain_kernel -s $L -c $L -f -o gpurun_out/r02e_chain_$SC $CMD > gpurun_out/r02e_ncu_$SC.log 2>&1
  echo "$SC full rc=$?"; tail -2 gpurun_out/r02e_ncu_$SC.log
done
ls -la gpurun_out/*.ncu-rep
