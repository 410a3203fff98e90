#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 600 python tools/glm_iters.py C5 > $O/r02_glm_sweeps_c5.txt 2>&1; echo "rc=$?" >> $O/r02_glm_sweeps_c5.txt
for v in glm_u2b4 glm_u2b5 mq_b6 apl_l4; do
  COMBATSEQ_LIB=$PWD/tools/variants/libcombatseq_$v.so timeout 300 python tools/quick_time.py C5 > $O/r02_variant_$v.txt 2>&1
done
timeout 300 python tools/quick_time.py C5 > $O/r02_variant_base.txt 2>&1
