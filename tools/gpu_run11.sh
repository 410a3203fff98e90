#!/bin/bash
# 8-GPU session: scaling of the one-process-per-GPU mode and of the single-process mode
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/gpus8.txt 2>&1
for n in 8 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 3 --warmup 3 --no-cpu > $O/r02_bench_c5_g${n}_a.json 2> $O/r02_bench_c5_g${n}_a.err; echo "rc=$?" >> $O/r02_bench_c5_g${n}_a.err
done
timeout 600 python bench.py --inproc 8 --steps 3 --warmup 2 > $O/r02_bench_c5_inproc8_a.json 2> $O/r02_bench_c5_inproc8_a.err; echo "rc=$?" >> $O/r02_bench_c5_inproc8_a.err
timeout 600 python -m pytest tests -m gpu -x -q -k "single_process_multi_gpu" > $O/r02_pytest_gpu_8gpu.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_8gpu.txt
