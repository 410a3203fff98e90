#!/bin/bash
# 2-GPU box: the whole -m gpu tier in one process (multi-GPU bit-identity runs instead of skipping; torch is imported after NCCL)
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -rs > $O/r02_pytest_gpu_2gpu_final.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_2gpu_final.txt
exit 0
