mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_selection_analytics.py tests/test_kernels_gpu.py tests/test_property_gpu.py -q -x 2>&1 | tail -4
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/sl.csv python benchmarks/side_kernels.py > /dev/null 2>&1
for k in select_bits quant_pack_kernel dequant_kernel; do echo "$k $(grep $k gpurun_out/sl.csv | tail -5 | awk -F'","' '{print $NF}' | tr -d '"' | sort -n | head -3 | tr '\n' ' ')"; done
