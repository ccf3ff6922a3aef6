mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run() { name=$1; shift; timeout 600 python bench.py "$@" --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; tail -c 300 gpurun_out/bench_$name.err; }
run c1 --workload c1 --steps 20
run c2b1 --workload c2 --batch 1 --steps 20
run c3 --workload c3 --steps 20
run c3b64 --workload c3 --batch 64 --steps 10
run c2b4096 --workload c2
run c2b64 --workload c2 --batch 64 --steps 20
run c4r12 --workload c4 --rows 12500000 --steps 10
