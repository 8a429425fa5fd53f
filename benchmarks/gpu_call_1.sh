mkdir -p gpurun_out
timeout 120 benchmarks/probes/tmem_probe > gpurun_out/tmem_probe.txt 2>&1; tail -3 gpurun_out/tmem_probe.txt
timeout 600 python -m pytest tests -q -m gpu 2>&1 | tail -25
timeout 600 python bench.py > gpurun_out/bench_r2_base.json 2> gpurun_out/bench_r2_base.err; tail -c 300 gpurun_out/bench_r2_base.err
timeout 300 python benchmarks/k2_experiments.py > gpurun_out/k2_base.txt 2>&1; cat gpurun_out/k2_base.txt
