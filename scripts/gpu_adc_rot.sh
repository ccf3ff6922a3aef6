# A/B of the lane-rotated (bank-conflict-free) ADC layout against the plain interleave, C3 (10M x 768, IVF4096 PQ96)
set -x
mkdir -p gpurun_out/rot2
O=gpurun_out/rot2
timeout 900 python -m pytest tests/test_ivfpq_gpu.py -m gpu -q -x 2>&1 | tail -5 > $O/pytest_ivfpq.txt
cat $O/pytest_ivfpq.txt
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/bench_$name.json 2> $O/bench_$name.err; tail -c 200 $O/bench_$name.err; head -c 200 $O/bench_$name.json; echo; }

run c3b4096 --workload c3 --batch 4096 --steps 3 --no-cpu-baseline
LANCE_B200_ADC_WARPS=8 run c3b4096_w8 --workload c3 --batch 4096 --steps 3 --no-cpu-baseline
run c3b64 --workload c3 --batch 64 --steps 10 --no-cpu-baseline
run c3 --workload c3 --steps 20 --no-cpu-baseline
ls $O
