mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run() { name=$1; shift; timeout 600 python bench.py "$@" --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; tail -c 300 gpurun_out/bench_$name.err; }
run c3 --workload c3 --steps 20
LANCE_B200_TIMER_DUMP=1 run c3d --workload c3 --steps 3
