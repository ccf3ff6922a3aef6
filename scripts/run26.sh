mkdir -p gpurun_out
NCU="ncu --clock-control none"
timeout 900 $NCU --set full --import-source on -k regex:flat_scan --launch-skip 4 --launch-count 1 -o gpurun_out/c1_scan_full python bench.py --workload c1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c1_full.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:adc_scan --launch-skip 6 --launch-count 1 -o gpurun_out/c3_adc_full python bench.py --workload c3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c3_full.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:flat_scan --launch-skip 6 --launch-count 1 -o gpurun_out/c3_probe_full python bench.py --workload c3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c3p_full.log 2>&1
