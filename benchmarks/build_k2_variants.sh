#!/bin/bash
# Builds lib/variants/libsplitq_<name>.so with extra -D flags for importance_fused.cu (A/B timing through SPLITQ_LIB).
# Both builds of K2 (two / three epilogue groups) get the flags; SPLITQ_K2_GROUPS=2|3 pins one at run time.
# usage: build_k2_variants.sh name1:"flags" name2:"flags" ...
set -e
cd "$(dirname "$0")/../llm-inference-in-distributed-edge-networks_b200/csrc"
mkdir -p ../lib/variants ../build/variants
ARCH="-gencode arch=compute_100a,code=sm_100a"
OTHERS=$(ls ../build/*.o | grep -v "importance_fused_g[23].o")
for v in "$@"; do
  n=${v%%:*}; f=${v#*:}
  for g in 2 3; do
    nvcc $f -DSPLITQ_K2_EPI_GROUPS=$g -O3 -std=c++17 -lineinfo $ARCH -Xcompiler -fPIC,-Wall -I../../include --fmad=true -c importance_fused.cu -o ../build/variants/importance_fused_${n}_g$g.o &
  done
done
wait
for v in "$@"; do
  n=${v%%:*}
  nvcc -shared $ARCH -o ../lib/variants/libsplitq_$n.so $OTHERS ../build/variants/importance_fused_${n}_g2.o ../build/variants/importance_fused_${n}_g3.o -lcudart
done
ls -la ../lib/variants
