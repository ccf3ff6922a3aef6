set -x
mkdir -p gpurun_out/final2
O=gpurun_out/final2
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3 > $O/pytest_gpu.txt
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1 >> $O/pytest_gpu.txt
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/bench_$name.json 2> $O/bench_$name.err; tail -c 200 $O/bench_$name.err; }
run c4 --workload c4
run c4_ref --impl reference --workload c4
run c1 --workload c1 --steps 20
run c2b4096 --workload c2
run c2b64 --workload c2 --batch 64 --steps 20
run c2b1 --workload c2 --batch 1 --steps 20
run c3 --workload c3 --steps 20
run c3b64 --workload c3 --batch 64 --steps 10 --no-cpu-baseline
run c3b4096 --workload c3 --batch 4096 --steps 3 --no-cpu-baseline
run c5 --workload c5 --steps 20
run c5b64 --workload c5 --batch 64 --steps 10 --no-cpu-baseline
