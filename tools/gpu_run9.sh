#!/bin/bash
# ncu captures of the three FP64-pipe kernels on C5 (one launch each), after a plain run of the same command
O=gpurun_out; mkdir -p $O
timeout 300 python tools/quick_time.py C5 > $O/r02_quick_c5.txt 2>&1 || exit 1
for k in k_apl_lanes k_glm_fit_lanes k_match_quantiles_warp; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 16 -c 1 -o $O/r02_prof_$k -f python tools/quick_time.py C5 > $O/r02_ncu_$k.log 2>&1
  echo "ncu $k rc=$?" >> $O/r02_ncu_$k.log
done
