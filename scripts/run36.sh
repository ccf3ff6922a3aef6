mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_flat_tc_gpu.py -x -q 2>&1 | tail -2
run() { name=$1; shift; timeout 600 python bench.py "$@" --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; tail -c 300 gpurun_out/bench_$name.err; }
for j in 4 16 64; do
LANCE_B200_TC_JUMP=$j run k${j}_r12 --workload c4 --rows 12500000 --steps 10
done
LANCE_B200_TIMER_DUMP=1 run kd_r12 --workload c4 --rows 12500000 --steps 2
run k4_r100 --workload c4
run k4_c2 --workload c2
