#!/bin/bash
# 1 GPU: 16-lane GLM variant A/B; one ncu --set full capture of the top kernels of one resident step
O=gpurun_out; mkdir -p $O
for v in base glm16; do
  lib=""; [ $v != base ] && lib=tools/variants/libcombatseq_$v.so
  COMBATSEQ_LIB=$lib timeout 600 python tools/quick_time.py C5 G=7500 > $O/r02_ab17_${v}_g7500.txt 2>&1
  COMBATSEQ_LIB=$lib timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu > $O/r02_ab17_${v}.json 2> $O/r02_ab17_${v}.err
done
CMD="python bench.py --steps 1 --warmup 1 --no-cpu"
timeout 600 $CMD > $O/r02_ncu17_plain.json 2> $O/r02_ncu17_plain.err || exit 0
K='regex:k_mc_gemm|k_apl_lanes|k_glm_fit_lanes|k_glm_fit_wide|k_match_quantiles_warp|k_scr_mma|k_mcg_prep_nodes'
timeout 1500 ncu --set full --clock-control none --import-source on -k "$K" -s 50 -c 50 -f -o /tmp/prof17 $CMD > $O/r02_ncu17.log 2>&1
ncu -i /tmp/prof17.ncu-rep --page raw --csv > $O/r02_ncu17_raw.csv 2>/dev/null
for k in k_mc_gemm k_apl_lanes k_glm_fit_lanes k_match_quantiles_warp k_scr_mma; do
  ncu -i /tmp/prof17.ncu-rep --page source --csv -k regex:$k -c 1 2>/dev/null | gzip > $O/r02_ncu17_source_$k.csv.gz
done
ls -la /tmp/prof17.ncu-rep >> $O/r02_ncu17.log
exit 0
