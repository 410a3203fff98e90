#!/bin/bash
# 1 GPU: the default bench line, then one ncu --set full capture of every launch of the top kernels in one resident C5 step
# (-> profiles/ncu_kernels.json via tools/ncu_kernels.py) and the launch list of the same command
O=gpurun_out; mkdir -p $O
timeout 900 python bench.py > $O/r02_bench_c5_default_final.json 2> $O/r02_bench_c5_default_final.err; echo "rc=$?" >> $O/r02_bench_c5_default_final.err
CMD="python bench.py --steps 1 --warmup 1 --no-cpu"
timeout 600 $CMD > $O/r02_ncu_step_plain.json 2> $O/r02_ncu_step_plain.err || exit 0
K='regex:k_mc_gemm|k_apl_lanes|k_glm_fit_lanes|k_glm_fit_wide|k_match_quantiles_binned|k_scr_mma|k_mcg_prep_nodes'
# the first (warm-up) step launches 57 of these kernels: skip them, capture the second step's
timeout 1500 ncu --set full --clock-control none --import-source on -k "$K" -s 57 -c 57 -f -o /tmp/prof_step $CMD > $O/r02_ncu_step.log 2>&1
ncu -i /tmp/prof_step.ncu-rep --page raw --csv > $O/r02_ncu_step_raw.csv 2>/dev/null
exit 0
