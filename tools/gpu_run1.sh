#!/bin/bash
# first GPU session of round 2: MC timing (band off / on), GPU parity tests, bench
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/gpus.txt 2>&1
timeout 600 python tools/mc_time.py 8000 100 2:0 2:1 > $O/r02_mc_time_8k.txt 2>&1; echo "rc=$?" >> $O/r02_mc_time_8k.txt
timeout 900 python tools/mc_time.py 60000 100 2:0 2:1 > $O/r02_mc_time_60k.txt 2>&1; echo "rc=$?" >> $O/r02_mc_time_60k.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/r02_pytest_gpu_1.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_1.txt
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $O/r02_bench_c5_1.json 2> $O/r02_bench_c5_1.err; echo "rc=$?" >> $O/r02_bench_c5_1.err
