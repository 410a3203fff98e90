#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_mc.py tests/test_gpu_pipeline.py -x -q > $O/pytest_gpu2.txt 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu2.txt
timeout 300 python tools/mc_time.py 20000 100 2 > $O/mc_time2.txt 2>&1
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $O/bench_c5_v2.json 2> $O/bench_c5_v2.err; echo "bench rc=$?" >> $O/bench_c5_v2.err
