#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu4.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu4.txt
timeout 600 python tools/ab_time.py C5 > $O/ab_c5.txt 2>&1; echo "rc=$?" >> $O/ab_c5.txt
timeout 300 python tools/ab_time.py C3b > $O/ab_c3b.txt 2>&1
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $O/bench_c5_s4.json 2> $O/bench_c5_s4.err; echo "rc=$?" >> $O/bench_c5_s4.err
