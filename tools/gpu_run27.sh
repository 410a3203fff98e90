#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -rs > $O/r02_pytest_gpu_27.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_27.txt
for v in base wide8 wide4; do
  lib=""; [ $v != base ] && lib=tools/variants/libcombatseq_$v.so
  COMBATSEQ_LIB=$lib timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu > $O/r02_ab27_${v}.json 2> $O/r02_ab27_${v}.err
done
exit 0
