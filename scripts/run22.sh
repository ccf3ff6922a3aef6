mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_flat_tc_gpu.py -x -q 2>&1 | tail -3
LANCE_B200_TIMER_DUMP=1 timeout 600 python bench.py --workload c4 --rows 12500000 --steps 2 --no-cpu-baseline > gpurun_out/dump_r12.json 2> gpurun_out/dump_r12.err
LANCE_B200_TIMER_DUMP=1 timeout 600 python bench.py --workload c4 --steps 2 --no-cpu-baseline > gpurun_out/dump_r100.json 2> gpurun_out/dump_r100.err
timeout 600 python bench.py --workload c4 --no-cpu-baseline > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err
timeout 600 python bench.py --workload c5 --batch 64 --no-cpu-baseline > gpurun_out/bench_c5b64.json 2> gpurun_out/bench_c5b64.err
