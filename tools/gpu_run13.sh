#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -rs > $O/r02_pytest_gpu_13.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_13.txt
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu > $O/r02_bench_c5_13.json 2> $O/r02_bench_c5_13.err; echo "rc=$?" >> $O/r02_bench_c5_13.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu > $O/r02_bench_c5_13_g2.json 2> $O/r02_bench_c5_13_g2.err; echo "rc=$?" >> $O/r02_bench_c5_13_g2.err
