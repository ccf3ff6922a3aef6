# pinned staging of the host-buffer call: whole GPU suite + single-query lines (e2e column)
mkdir -p gpurun_out/e2e
O=gpurun_out/e2e
timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > $O/pytest_gpu.txt; cat $O/pytest_gpu.txt
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/bench_$name.json 2> $O/bench_$name.err; python - "$O/bench_$name.json" <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); print(sys.argv[1].split("/")[-1], round(d["ms_per_step"],4), d["gpu_launches"]/d["steps"], d["roofline"]["frac"], "e2e", d["e2e"]["value"], round(1e3/d["e2e"]["value"],4))
PY
}
run c1 --workload c1 --steps 20
run c3 --workload c3 --steps 20
run c2b1 --workload c2 --batch 1 --steps 20
run c2b64 --workload c2 --batch 64 --steps 20
run c5 --workload c5 --steps 20
