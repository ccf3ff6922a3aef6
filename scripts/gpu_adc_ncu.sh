# ncu --set full of the batched ADC scan (C3, B=4096)
set -x
mkdir -p gpurun_out/rot3
O=gpurun_out/rot3
timeout 1200 ncu --clock-control none --set full --import-source on -k regex:adc_scan --launch-skip 18 --launch-count 1 -o $O/c3b4096_adc_full python bench.py --workload c3 --batch 4096 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c3b4096.log 2>&1
tail -3 $O/ncu_c3b4096.log
ls -la $O
