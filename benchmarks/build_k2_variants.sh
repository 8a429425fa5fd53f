#!/bin/bash
# Builds lib/variants/libsplitq_<name>.so with extra -D flags for importance_fused.cu (A/B timing through SPLITQ_LIB).
# usage: build_k2_variants.sh name1:"flags" name2:"flags" ...
set -e
cd "$(dirname "$0")/../llm-inference-in-distributed-edge-networks_b200/csrc"
mkdir -p ../lib/variants ../build/variants
ARCH="-gencode arch=compute_100a,code=sm_100a"
OTHERS=$(ls ../build/*.o | grep -v importance_fused.o)
for v in "$@"; do
  n=${v%%:*}; f=${v#*:}
  nvcc $f -O3 -std=c++17 -lineinfo $ARCH -Xcompiler -fPIC,-Wall -I../../include --fmad=true -c importance_fused.cu -o ../build/variants/importance_fused_$n.o &
done
wait
for v in "$@"; do
  n=${v%%:*}
  nvcc -shared $ARCH -o ../lib/variants/libsplitq_$n.so $OTHERS ../build/variants/importance_fused_$n.o -lcudart
done
ls -la ../lib/variants
