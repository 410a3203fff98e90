#!/bin/bash
# 2-GPU session: parity tests (incl. single-process multi-GPU), 1-GPU bench, 2-rank torchrun bench, in-process 2-GPU timing
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/gpus4.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q -rs > $O/r02_pytest_gpu_4.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_4.txt
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > $O/r02_bench_c5_4.json 2> $O/r02_bench_c5_4.err; echo "rc=$?" >> $O/r02_bench_c5_4.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu > $O/r02_bench_c5_4_g2.json 2> $O/r02_bench_c5_4_g2.err; echo "rc=$?" >> $O/r02_bench_c5_4_g2.err
