mkdir -p gpurun_out
timeout 120 python benchmarks/k2_one.py || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_importance_kernel --launch-skip 2 -c 2 -f -o gpurun_out/r2_k2_fast python benchmarks/k2_one.py > gpurun_out/ncu_k2_fast.log 2>&1
tail -3 gpurun_out/ncu_k2_fast.log; ls -la gpurun_out/*.ncu-rep
