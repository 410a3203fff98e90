#!/bin/bash
# development builds with different compile-time knobs: tools/build_variant.sh NAME -DKNOB=... -> tools/variants/libcombatseq_NAME.so
# (select with COMBATSEQ_LIB=... ; tools/variants/ is git-ignored but travels to the GPU box)
set -e
name=$1; shift
cd "$(dirname "$0")/.."
out=tools/variants; obj=$out/obj_$name; mkdir -p $obj
for u in combatseq tu_apl tu_glm tu_mq tu_mc tu_screen tu_sim; do
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -diag-suppress 177 "$@" -c -o $obj/$u.o combat-seq_b200/csrc/$u.cu &
done
wait
/usr/local/cuda/bin/nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $out/libcombatseq_$name.so $obj/*.o -ldl
rm -rf $obj
echo $out/libcombatseq_$name.so
