#!/bin/bash
# 1 GPU: tests; APL staging variants A/B; C4 / C2 / C3 bench lines
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -rs > $O/r02_pytest_gpu_18.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_18.txt
for v in base aplint aplint5 aplint6 aplmb4; do
  lib=""; [ $v != base ] && lib=tools/variants/libcombatseq_$v.so
  COMBATSEQ_LIB=$lib timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu > $O/r02_ab18_${v}.json 2> $O/r02_ab18_${v}.err
done
for w in C4 C2 C3; do
  timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu > $O/r02_bench_${w}_18.json 2> $O/r02_bench_${w}_18.err
done
exit 0
