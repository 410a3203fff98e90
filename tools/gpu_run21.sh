#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q -rs > $O/r02_pytest_no_alloc_alone.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_no_alloc_alone.txt
timeout 900 python -m pytest tests -m gpu -x -q -rs > $O/r02_pytest_gpu_21.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_21.txt
exit 0
