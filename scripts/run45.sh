mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
run() { name=$1; shift; timeout 600 python bench.py "$@" --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; tail -c 300 gpurun_out/bench_$name.err; }
run c5 --workload c5 --steps 20
run c5s01 --workload c5 --selectivity 0.01 --steps 20
run c2b1 --workload c2 --batch 1 --steps 20
run c1 --workload c1 --steps 20
