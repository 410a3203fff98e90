#!/bin/bash
# 1 GPU: a short bench line, then the ncu launch list (gpu__time_duration.sum per launch) of the same command
O=gpurun_out; mkdir -p $O
CMD="python bench.py --steps 2 --warmup 1 --no-cpu"
timeout 600 $CMD > $O/r02_bench_c5_short_final.json 2> $O/r02_bench_c5_short_final.err; echo "rc=$?" >> $O/r02_bench_c5_short_final.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file $O/r02_ncu_launches_final.csv $CMD > $O/r02_ncu_launches_final.log 2>&1; echo "rc=$?" >> $O/r02_ncu_launches_final.log
exit 0
