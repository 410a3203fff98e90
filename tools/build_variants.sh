#!/bin/bash
# development builds of the library with a different Monte-Carlo ring geometry / diagnostics (loaded via COMBATSEQ_LIB)
cd "$(dirname "$0")/.."
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC -diag-suppress 177"
mkdir -p tools/variants
nvcc $F -DMCG_SUB_N=1 -DMCG_STAGES_N=8 -o tools/variants/libcombatseq_sub1x8.so combat-seq_b200/csrc/combatseq.cu -ldl &
nvcc $F -DMCG_DIAG_NO_EPILOGUE -o tools/variants/libcombatseq_noepi.so combat-seq_b200/csrc/combatseq.cu -ldl &
wait
ls -la tools/variants
