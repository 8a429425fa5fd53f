mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none -k regex:"quant_pack_kernel|dequant_kernel|select_bits_kernel|split_f16|lse_combine|fused_finalize" --launch-skip 18 -c 6 -f -o gpurun_out/side_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 > gpurun_out/ncu_side.log 2>&1
timeout 300 ncu -i gpurun_out/side_full.ncu-rep --page raw --csv > gpurun_out/side_full_raw.csv 2>/dev/null
ls -la gpurun_out/side_full*; tail -2 gpurun_out/ncu_side.log
