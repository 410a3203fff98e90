#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -rs > $O/r02_pytest_gpu_23.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_23.txt
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu > $O/r02_bench_c5_23.json 2> $O/r02_bench_c5_23.err
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --tune mq_bins=0 > $O/r02_bench_c5_23_nobins.json 2> $O/r02_bench_c5_23_nobins.err
exit 0
