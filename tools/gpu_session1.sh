#!/bin/bash
# GPU session: parity tests, bench, launch list, ncu captures of the top kernels (run under gpurun)
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.txt 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.txt
timeout 900 python bench.py --steps 3 --warmup 3 > $O/bench_c5.json 2> $O/bench_c5.err; echo "bench rc=$?" >> $O/bench_c5.err
timeout 300 python tools/mc_time.py 20000 100 2 > $O/mc_time.txt 2>&1
timeout 300 python __graft_entry__.py smoke > $O/smoke.txt 2>&1; echo "smoke rc=$?" >> $O/smoke.txt
timeout 300 python bench.py --impl reference --steps 1 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
timeout 300 python bench.py --workload C4 --steps 3 --warmup 3 --no-cpu > $O/bench_c4.json 2> $O/bench_c4.err
timeout 300 python bench.py --workload C2 --steps 3 --warmup 3 --no-cpu > $O/bench_c2.json 2> $O/bench_c2.err
CMD="python bench.py --steps 1 --warmup 1 --no-cpu"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1
for k in k_mc_gemm k_match_quantiles k_glm_fit k_apl; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -f -o /tmp/prof_$k $CMD > $O/ncu_$k.log 2>&1
  # the .ncu-rep files are 20-30 MB each and gpurun_out/ is capped at 64 MiB: export the pages on the box
  ncu -i /tmp/prof_$k.ncu-rep --page raw --csv > $O/prof_${k}_raw.csv 2>/dev/null
  ncu -i /tmp/prof_$k.ncu-rep --page details --csv > $O/prof_${k}_details.csv 2>/dev/null
  ncu -i /tmp/prof_$k.ncu-rep --page source --csv > $O/prof_${k}_source.csv 2>/dev/null
done
# gpurun_out/ is capped at 64 MiB: keep text summaries, gzip the big per-instruction pages, leave the .ncu-rep files on the box
for k in k_mc_gemm k_match_quantiles k_glm_fit k_apl; do
  { python tools/ncu_raw_summary.py $O/prof_${k}_raw.csv; echo; python tools/ncu_src_summary.py $O/prof_${k}_source.csv 18; } > $O/summary_$k.txt 2>&1
  gzip -f $O/prof_${k}_source.csv
done
du -sh $O
ls -la $O
