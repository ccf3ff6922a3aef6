# self-restoring scan header / ADC tickets: whole GPU suite + the single-query lines
mkdir -p gpurun_out/hdr
O=gpurun_out/hdr
timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > $O/pytest_gpu.txt; cat $O/pytest_gpu.txt
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/bench_$name.json 2> $O/bench_$name.err; python - "$O/bench_$name.json" <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); print(sys.argv[1].split("/")[-1], round(d["ms_per_step"],4), d["gpu_launches"]/d["steps"], d["roofline"]["frac"], d["e2e"]["value"])
PY
}
run c1 --workload c1 --steps 20
run c3 --workload c3 --steps 20
run c2b1 --workload c2 --batch 1 --steps 20
run c5 --workload c5 --steps 20
run c5s01 --workload c5 --selectivity 0.01 --steps 20
