#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu6.txt 2>&1; echo "rc=$?" >> $O/pytest_gpu6.txt
timeout 300 python tools/mc_time.py 60000 100 2 > $O/mc_time_base.txt 2>&1
for v in nt4 nt4_noepi; do
  if [ -f tools/variants/libcombatseq_$v.so ]; then COMBATSEQ_LIB=$PWD/tools/variants/libcombatseq_$v.so timeout 300 python tools/mc_time.py 60000 100 2 > $O/mc_time_$v.txt 2>&1; fi
done
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $O/bench_c5_s6.json 2> $O/bench_c5_s6.err; echo "rc=$?" >> $O/bench_c5_s6.err
