#!/bin/bash
# Tuning sweep of the composite kernels' block size / occupancy bounds on the GPU box (not a test):
#   gpurun -- 'bash tests/comp_sweep.sh'
cd pixelnerf_b200/csrc
for cfg in "4 6 8 4" "8 3 8 2" "4 8 4 6" "2 12 4 8"; do
  set -- $cfg
  rm -f build/pn_composite.o
  make NVCCFLAGS="-O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-fvisibility=hidden -Xptxas -v --expt-relaxed-constexpr -DPN_BUILDING -DPN_COMP_FWD_WARPS=$1 -DPN_COMP_FWD_MINB=$2 -DPN_COMP_BWD_WARPS=$3 -DPN_COMP_BWD_MINB=$4" > /dev/null 2>&1
  echo "fwd warps=$1 minb=$2, bwd warps=$3 minb=$4"
  (cd ../..; python bench.py --kernels 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin)
print('   ', ' '.join(f\"{k.replace('composite_','')}={v['frac_of_measured_hbm']:.3f}\" for k,v in d['kernels'].items() if 'composite' in k))")
done
