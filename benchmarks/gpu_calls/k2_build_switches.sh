# the build-time switches of K2 (benchmarks/build_k2_variants.sh n192:"-DSPLITQ_K2_NTILE=192" asmem:"-DSPLITQ_K2_A_SMEM=1"
# nofuse:"-DSPLITQ_K2_FUSE_QSPLIT=0") through the fused parity tests
mkdir -p gpurun_out
for v in n192 asmem nofuse; do
  echo "== $v"
  SPLITQ_LIB=$PWD/llm-inference-in-distributed-edge-networks_b200/lib/variants/libsplitq_$v.so timeout 900 python -m pytest tests/test_fused_gpu.py -q -x -m gpu 2>&1 | tail -4
  SPLITQ_LIB=$PWD/llm-inference-in-distributed-edge-networks_b200/lib/variants/libsplitq_$v.so SPLITQ_K2_GROUPS=2 timeout 300 python benchmarks/k2_experiments.py --short 2>&1 | head -1
done
