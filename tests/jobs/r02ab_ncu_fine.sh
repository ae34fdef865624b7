# round-2: ncu source-level profile of sample_fine_kernel (where its ~1100 warp instructions per ray go)
set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python bench.py --kernels"
$CMD > gpurun_out/r02ab_plain.json 2> gpurun_out/r02ab_plain.err; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sample_fine_kernel -s 4 -c 1 -f -o gpurun_out/r02ab_fine $CMD > gpurun_out/r02ab_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r02ab_fine.ncu-rep --page source --csv > gpurun_out/r02ab_fine_source.csv 2>/dev/null; echo "export rc=$?"
ncu -i gpurun_out/r02ab_fine.ncu-rep --page raw --csv > gpurun_out/r02ab_fine_raw.csv 2>/dev/null
rm -f gpurun_out/r02ab_fine.ncu-rep
head -3 gpurun_out/r02ab_fine_source.csv | cut -c1-600
wc -l gpurun_out/r02ab_fine_source.csv
