set -x
mkdir -p gpurun_out
NCU="ncu --clock-control none"
timeout 900 $NCU --metrics gpu__time_duration.sum -k regex:'kernel' -c 300 --csv --log-file gpurun_out/c3_launches.csv python bench.py --workload c3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c3_list.log 2>&1
timeout 900 $NCU --metrics gpu__time_duration.sum -k regex:'kernel' -c 300 --csv --log-file gpurun_out/c1_launches.csv python bench.py --workload c1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c1_list.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:adc_scan --launch-skip 6 --launch-count 1 -o gpurun_out/c3_adc_full python bench.py --workload c3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c3_full.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:flat_scan --launch-skip 4 --launch-count 1 -o gpurun_out/c1_scan_full python bench.py --workload c1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c1_full.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:adc_scan --launch-skip 3 --launch-count 1 -o gpurun_out/c3b4096_adc_full python bench.py --workload c3 --batch 4096 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c3b_full.log 2>&1
ls -la gpurun_out | grep ncu-rep
