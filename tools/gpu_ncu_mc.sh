#!/bin/bash
# ncu capture of the Monte-Carlo contraction kernel alone (tools/mc_time.py, one batch G=20000 n=100)
O=gpurun_out; mkdir -p $O
CMD="python tools/mc_time.py 20000 100 2:1"
$CMD > $O/mc_time_plain.txt 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_mc_gemm -s 1 -c 1 -f -o /tmp/prof_mc $CMD > $O/ncu_mc.log 2>&1
ncu -i /tmp/prof_mc.ncu-rep --page raw --csv > $O/mc_raw.csv 2>/dev/null
ncu -i /tmp/prof_mc.ncu-rep --page source --csv 2>/dev/null | gzip > $O/mc_source.csv.gz
exit 0
