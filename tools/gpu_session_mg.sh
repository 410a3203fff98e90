#!/bin/bash
O=gpurun_out; mkdir -p $O
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tools/multi_gpu_check.py > $O/mg_check_$N.txt 2>&1; echo "rc=$?" >> $O/mg_check_$N.txt
timeout 600 $TR bench.py --gpus $N --steps 3 --warmup 3 > $O/bench_c5_g$N.json 2> $O/bench_c5_g$N.err; echo "rc=$?" >> $O/bench_c5_g$N.err
timeout 300 $TR bench.py --gpus $N --steps 1 --warmup 0 --impl reference --cpu-genes 120 > $O/bench_ref_g$N.json 2> $O/bench_ref_g$N.err; echo "rc=$?" >> $O/bench_ref_g$N.err
if [ "$N" = "2" ] && [ -x tools/dmma_bench ]; then timeout 120 tools/dmma_bench > $O/dmma_bench.txt 2>&1; fi
