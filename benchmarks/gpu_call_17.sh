mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/r2_g_bench_fused.json 2> gpurun_out/r2_g.err; tail -c 200 gpurun_out/r2_g.err
timeout 600 python bench.py --importance materialised --no-cpu-baseline > gpurun_out/r2_g_bench_materialised.json 2>> gpurun_out/r2_g.err
timeout 600 python bench.py --workload pythia-70m_split3_lastrow_s2048_int4 --no-cpu-baseline > gpurun_out/r2_g_bench_pythia_lastrow.json 2>> gpurun_out/r2_g.err
timeout 600 python bench.py --workload pythia-70m_split3_colmean_s512_int4group --no-cpu-baseline > gpurun_out/r2_g_bench_pythia_s512.json 2>> gpurun_out/r2_g.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_g_launches_fused.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 > gpurun_out/ncu_g.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_importance_kernel --launch-skip 6 -c 2 -f -o gpurun_out/r2_g_k2_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 > gpurun_out/ncu_g_full.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:"quant_pack_kernel|dequant_kernel|select_bits_kernel|split_f16|lse_combine|fused_finalize" --launch-skip 18 -c 6 -f -o gpurun_out/r2_g_side_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 > gpurun_out/ncu_g_side.log 2>&1
timeout 900 python benchmarks/sweep.py --out gpurun_out/r2_sweep.json --md gpurun_out/r2_sweep.md > gpurun_out/sweep.log 2>&1; tail -3 gpurun_out/sweep.log
timeout 600 python benchmarks/window_latency.py > gpurun_out/r2_window_latency.json 2> gpurun_out/wl.err; tail -2 gpurun_out/wl.err; head -c 600 gpurun_out/r2_window_latency.json
ls -la gpurun_out | tail -15
