#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu3.txt 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu3.txt
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $O/bench_c5_v3.json 2> $O/bench_c5_v3.err; echo "bench rc=$?" >> $O/bench_c5_v3.err
timeout 300 python tools/quick_time.py C3b > $O/quick_c3b.txt 2>&1
