#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu7.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu7.txt
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $O/bench_c5_s7.json 2> $O/bench_c5_s7.err; echo "rc=$?" >> $O/bench_c5_s7.err
