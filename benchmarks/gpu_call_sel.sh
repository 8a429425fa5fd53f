mkdir -p gpurun_out
SPLITQ_LIB=$PWD/llm-inference-in-distributed-edge-networks_b200/lib/variants/libsplitq_seltiming.so timeout 120 python benchmarks/sel_timing.py 2>&1 | tail -6
timeout 120 benchmarks/probes/tmem_probe > gpurun_out/tmem_probe_chains.txt 2>&1; cat gpurun_out/tmem_probe_chains.txt
