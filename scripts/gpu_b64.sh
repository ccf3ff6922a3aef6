# B=64 lines with the batch-size dependent segment growth + ADC CTA shape sweep at B=64
mkdir -p gpurun_out/b64
O=gpurun_out/b64
timeout 900 python -m pytest tests/test_flat_tc_gpu.py tests/test_host_gpu.py -m gpu -q -x 2>&1 | tail -3 > $O/pytest.txt; cat $O/pytest.txt
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/bench_$name.json 2> $O/bench_$name.err; python - "$O/bench_$name.json" <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); print(sys.argv[1].split("/")[-1], round(d["ms_per_step"],4), d["gpu_launches"]/d["steps"], d["roofline"]["frac"])
PY
}
run c2b64 --workload c2 --batch 64 --steps 20
run c5b64 --workload c5 --batch 64 --steps 10 --no-cpu-baseline
run c2b4096 --workload c2 --batch 4096 --steps 5 --no-cpu-baseline
for w in 8 16; do for c in 148 296 592 1184; do
LANCE_B200_ADC_WARPS=$w LANCE_B200_ADC_CTAS=$c run c3b64_w${w}_c$c --workload c3 --batch 64 --steps 10 --no-cpu-baseline
done; done
