#!/bin/bash
# 1 GPU: tests, the default bench line, the ncu launch list of the same command
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -rs > $O/r02_pytest_gpu_16.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_16.txt
timeout 900 python bench.py > $O/r02_bench_c5_default_16.json 2> $O/r02_bench_c5_default_16.err; echo "rc=$?" >> $O/r02_bench_c5_default_16.err
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu > $O/r02_bench_c5_short_16.json 2> $O/r02_bench_c5_short_16.err; echo "rc=$?" >> $O/r02_bench_c5_short_16.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/r02_ncu_launches_16.csv python bench.py --steps 2 --warmup 1 --no-cpu > $O/r02_ncu_launches_16.log 2>&1; echo "rc=$?" >> $O/r02_ncu_launches_16.log
exit 0
