# final build with zero-copy result buffers (default mode 1): GPU suite, smoke, refreshed single-query lines (CPU leg included)
mkdir -p gpurun_out/zc2
O=gpurun_out/zc2
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > $O/pytest_gpu.txt; cat $O/pytest_gpu.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
run() { name=$1; shift; timeout 600 python bench.py "$@" > $O/bench_$name.json 2> $O/bench_$name.err; python - "$O/bench_$name.json" <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); print(sys.argv[1].split("/")[-1], "ms", round(d["ms_per_step"],4), "q/s", round(d["value"],1), "e2e q/s", round(d["e2e"]["value"],1), "frac", round(d["roofline"]["frac"],3), "cpu", d.get("cpu_baseline",{}).get("value"))
PY
}
run c1 --workload c1 --steps 20
run c3 --workload c3 --steps 20
run c5s01 --workload c5 --selectivity 0.01 --steps 20
run c2b1 --workload c2 --batch 1 --steps 20
run c5 --workload c5 --steps 20
