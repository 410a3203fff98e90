#!/bin/bash
# 8-GPU session
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -rs > $O/r02_pytest_gpu_15.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_15.txt
for n in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 3 --warmup 3 --no-cpu > $O/r02_bench_c5_g${n}_b.json 2> $O/r02_bench_c5_g${n}_b.err; echo "rc=$?" >> $O/r02_bench_c5_g${n}_b.err
done
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $O/r02_bench_c5_g1_b.json 2> $O/r02_bench_c5_g1_b.err; echo "rc=$?" >> $O/r02_bench_c5_g1_b.err
timeout 600 python bench.py --inproc 8 --steps 3 --warmup 2 > $O/r02_bench_c5_inproc8_b.json 2> $O/r02_bench_c5_inproc8_b.err; echo "rc=$?" >> $O/r02_bench_c5_inproc8_b.err
