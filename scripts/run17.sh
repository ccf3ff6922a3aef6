# validation of: two-level fused final select (flat scan), fused ADC scan+merge+refine, warp-uniform GEMM emit path
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/pytest_gpu.txt; cat gpurun_out/pytest_gpu.txt
run() { name=$1; shift; timeout 600 python bench.py "$@" --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; tail -c 300 gpurun_out/bench_$name.err; }
run c1 --workload c1 --steps 20
run c2b1 --workload c2 --batch 1 --steps 20
run c2b64 --workload c2 --batch 64 --steps 20
run c2b4096 --workload c2
run c3 --workload c3 --steps 20
run c3b64 --workload c3 --batch 64 --steps 10
run c3b4096 --workload c3 --batch 4096 --steps 3
run c5 --workload c5 --steps 20
run c4 --workload c4
ls -la gpurun_out
