#!/bin/bash
# occupancy sweep of sample_fine_kernel on the GPU box (not a test)
cd pixelnerf_b200/csrc
for mb in 1 8 10 12 16; do
  rm -f build/pn_sampling.o
  make NVCCFLAGS="-O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-fvisibility=hidden -Xptxas -v --expt-relaxed-constexpr -DPN_BUILDING -DPN_FINE_MINB=$mb" > /dev/null 2>&1
  echo "minb=$mb $(grep -A2 sample_fine_kernel build/pn_sampling.ptxas.log | grep -o 'Used [0-9]* registers\|[0-9]* bytes spill stores' | paste -sd' ')"
  (cd ../..; python bench.py --kernels 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin); v=d['kernels']['sample_fine']; print('   sample_fine', round(v['ms'],4), round(v['frac_of_measured_hbm'],3))")
done
