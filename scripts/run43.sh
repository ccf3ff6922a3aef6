mkdir -p gpurun_out
ncu --clock-control none --set full --import-source on -k regex:adc_scan --launch-skip 18 --launch-count 1 -o gpurun_out/c3b4096_adc_full python bench.py --workload c3 --batch 4096 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c3b_full.log 2>&1
tail -2 gpurun_out/ncu_c3b_full.log | cut -c 1-200
