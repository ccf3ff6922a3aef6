mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ivfpq_gpu.py tests/test_host_gpu.py -x -q 2>&1 | tail -2
run() { name=$1; shift; timeout 600 python bench.py "$@" --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; tail -c 300 gpurun_out/bench_$name.err; }
run c3 --workload c3 --steps 20
run c3b4 --workload c3 --batch 4 --steps 20
