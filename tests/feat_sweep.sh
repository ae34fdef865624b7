#!/bin/bash
# block-size sweep of the gather / taps kernels on the GPU box (not a test)
cd pixelnerf_b200/csrc
for wp in 8 4 2 16; do
  rm -f build/pn_features.o
  make NVCCFLAGS="-O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-fvisibility=hidden -Xptxas -v --expt-relaxed-constexpr -DPN_BUILDING -DPN_FEAT_WARPS=$wp" > /dev/null 2>&1
  (cd ../..; python bench.py --kernels 2>/dev/null | python -c "
import json,sys
d=json.load(sys.stdin)['kernels']; print('warps=$wp', 'features', round(d['point_features_fp16']['frac_of_measured_hbm'],3), 'taps', round(d['point_taps']['frac_of_measured_hbm'],3))")
done
