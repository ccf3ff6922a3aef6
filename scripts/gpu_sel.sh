# register-resident radix select in the standalone select kernel: GPU suite + C3 lines
mkdir -p gpurun_out/sel
O=gpurun_out/sel
timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > $O/pytest_gpu.txt; cat $O/pytest_gpu.txt
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/bench_$name.json 2> $O/bench_$name.err; python - "$O/bench_$name.json" <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); print(sys.argv[1].split("/")[-1], round(d["ms_per_step"],4), d["gpu_launches"]/d["steps"], d["roofline"]["frac"], d["e2e"]["value"])
PY
}
run c3 --workload c3 --steps 20
run c3b64 --workload c3 --batch 64 --steps 10 --no-cpu-baseline
run c3b4096 --workload c3 --batch 4096 --steps 3 --no-cpu-baseline
timeout 900 ncu --clock-control none --metrics gpu__time_duration.sum -k regex:'kernel' -c 400 --csv --log-file $O/c3_launches.csv python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c3_list.log 2>&1
timeout 900 ncu --clock-control none --metrics gpu__time_duration.sum -k regex:'kernel' -c 400 --csv --log-file $O/c3b4096_launches.csv python bench.py --workload c3 --batch 4096 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c3b4096_list.log 2>&1
