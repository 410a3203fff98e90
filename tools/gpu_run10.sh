#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/r02_pytest_gpu_10.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_10.txt
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu > $O/r02_bench_c5_10.json 2> $O/r02_bench_c5_10.err; echo "rc=$?" >> $O/r02_bench_c5_10.err
timeout 300 python tools/quick_time.py C5 > $O/r02_quick_c5.txt 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_glm_fit_lanes -s 4 -c 1 -o $O/r02_prof_k_glm_fit_lanes -f python tools/quick_time.py C5 > $O/r02_ncu_k_glm.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_match_quantiles_warp -s 1 -c 1 -o $O/r02_prof_k_match_quantiles_warp -f python tools/quick_time.py C5 > $O/r02_ncu_k_mq.log 2>&1
