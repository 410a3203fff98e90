#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02_pytest_gpu_7.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_7.txt
timeout 900 python bench.py --steps 3 --warmup 3 > $O/r02_bench_c5_7.json 2> $O/r02_bench_c5_7.err; echo "rc=$?" >> $O/r02_bench_c5_7.err
