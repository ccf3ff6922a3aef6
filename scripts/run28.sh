mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_flat_tc_gpu.py tests/test_flat_gpu.py -x -q 2>&1 | tail -3
run() { name=$1; shift; timeout 600 python bench.py "$@" --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; tail -c 300 gpurun_out/bench_$name.err; }
run c4 --workload c4
run c2b4096 --workload c2
run c2b64 --workload c2 --batch 64 --steps 20
run c4r12 --workload c4 --rows 12500000 --steps 10
