mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_flat_tc_gpu.py tests/test_ivfpq_gpu.py -x -q 2>&1 | tail -3
run() { name=$1; shift; timeout 600 python bench.py "$@" --no-cpu-baseline > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; tail -c 300 gpurun_out/bench_$name.err; }
run c2b64 --workload c2 --batch 64 --steps 20
for c in 64 128 148 296; do LANCE_B200_ADC_CTAS=$c run c3_ctas$c --workload c3 --steps 20; done
for w in 8 16; do
LANCE_B200_ADC_WARPS=$w run c3b64_w$w --workload c3 --batch 64 --steps 10
LANCE_B200_ADC_WARPS=$w run c3b4096_w$w --workload c3 --batch 4096 --steps 3
done
