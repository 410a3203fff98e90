#!/bin/bash
# what the driver does at round end: build check, GPU tests, smoke(), default bench, reference arm
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -rs > $O/r02_pytest_gpu_final.txt 2>&1; echo "rc=$?" >> $O/r02_pytest_gpu_final.txt
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_smoke_final.txt 2>&1; echo "rc=$?" >> $O/r02_smoke_final.txt
timeout 900 python bench.py > $O/r02_bench_c5_default_final.json 2> $O/r02_bench_c5_default_final.err; echo "rc=$?" >> $O/r02_bench_c5_default_final.err
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > $O/r02_bench_c5_reference_arm.json 2> $O/r02_bench_c5_reference_arm.err; echo "rc=$?" >> $O/r02_bench_c5_reference_arm.err
exit 0
