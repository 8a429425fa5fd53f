mkdir -p gpurun_out
timeout 120 benchmarks/probes/tmem_probe > gpurun_out/tmem_probe_chains.txt 2>&1
cat gpurun_out/tmem_probe_chains.txt
timeout 120 python benchmarks/k2_experiments.py --short 2>&1 | tail -6
