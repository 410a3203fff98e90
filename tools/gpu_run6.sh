#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02_pytest_gpu_6.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_6.txt
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $O/r02_bench_c5_6.json 2> $O/r02_bench_c5_6.err; echo "rc=$?" >> $O/r02_bench_c5_6.err
timeout 300 python tools/mc_time.py 60000 100 2:1 > $O/r02_mc_time_60k_c.txt 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_scr_mma -s 1 -c 1 -o $O/r02_prof_scr3 -f python tools/mc_time.py 60000 100 2:1 > $O/r02_ncu_scr3.log 2>&1
echo "ncu rc=$?" >> $O/r02_ncu_scr3.log
