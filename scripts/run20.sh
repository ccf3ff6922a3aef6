set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/pytest_gpu.txt; cat gpurun_out/pytest_gpu.txt
run() { name=$1; shift; timeout 600 python bench.py "$@" --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; tail -c 300 gpurun_out/bench_$name.err; }
run c2b4096 --workload c2
for g in 4 8 16; do for sd in 16 4; do
LANCE_B200_TC_GROWTH=$g LANCE_B200_TC_SEED=$sd run sw_r12_g${g}_s${sd} --workload c4 --rows 12500000 --steps 10
done; done
for g in 8 16; do
LANCE_B200_TC_GROWTH=$g run sw_r100_g${g} --workload c4
LANCE_B200_TC_GROWTH=$g run sw_c2_g${g} --workload c2
done
run c4 --workload c4
