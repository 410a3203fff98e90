#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -rs > $O/r02_pytest_gpu_30.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_30.txt
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu > $O/r02_bench_c5_30.json 2> $O/r02_bench_c5_30.err
exit 0
