mkdir -p gpurun_out
LANCE_B200_TIMER_DUMP=1 timeout 600 python bench.py --workload c4 --rows 12500000 --steps 2 --no-cpu-baseline > gpurun_out/dump_r12.json 2> gpurun_out/dump_r12.err
LANCE_B200_TIMER_DUMP=1 timeout 600 python bench.py --workload c4 --steps 2 --no-cpu-baseline > gpurun_out/dump_r100.json 2> gpurun_out/dump_r100.err
LANCE_B200_TIMER_DUMP=1 timeout 600 python bench.py --workload c2 --steps 2 --no-cpu-baseline > gpurun_out/dump_c2.json 2> gpurun_out/dump_c2.err
