timeout 120 benchmarks/probes/tmem_probe 2>&1 | head -30
