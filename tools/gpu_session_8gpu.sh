#!/bin/bash
# 8-GPU session: multi-GPU bit-identity test, C5 at 8 / 4 / 2 / 1 GPUs (one process per GPU) and 8 GPUs in one process
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q -rs -k "multi_gpu or static_scratch or pageable" > $O/r02_pytest_gpu_8gpu_final.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_8gpu_final.txt
for n in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 3 --warmup 3 --no-cpu > $O/r02_bench_c5_g${n}_final.json 2> $O/r02_bench_c5_g${n}_final.err; echo "rc=$?" >> $O/r02_bench_c5_g${n}_final.err
done
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $O/r02_bench_c5_g1_final.json 2> $O/r02_bench_c5_g1_final.err; echo "rc=$?" >> $O/r02_bench_c5_g1_final.err
timeout 600 python bench.py --inproc 8 --steps 3 --warmup 2 > $O/r02_bench_c5_inproc8_final.json 2> $O/r02_bench_c5_inproc8_final.err; echo "rc=$?" >> $O/r02_bench_c5_inproc8_final.err
exit 0
