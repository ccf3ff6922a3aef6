# zero-copy result / query buffers of the small host-buffer call: GPU suite + e2e of the single-query lines per mode
mkdir -p gpurun_out/zc
O=gpurun_out/zc
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 > $O/pytest_gpu.txt; cat $O/pytest_gpu.txt
run() { name=$1; shift; timeout 600 python bench.py "$@" --no-cpu-baseline > $O/bench_$name.json 2> $O/bench_$name.err; python - "$O/bench_$name.json" <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); print(sys.argv[1].split("/")[-1], "ms", round(d["ms_per_step"],4), "e2e q/s", round(d["e2e"]["value"],1), "e2e ms", round(1e3*d["config"].get("batch",1)/d["e2e"]["value"],4))
PY
}
for m in 0 1 2; do
export LANCE_B200_ZEROCOPY=$m
run c1_zc$m --workload c1 --steps 20
run c3_zc$m --workload c3 --steps 20
run c2b1_zc$m --workload c2 --batch 1 --steps 20
run c5s01_zc$m --workload c5 --selectivity 0.01 --steps 20
done
export LANCE_B200_ZEROCOPY=1
run c2b64_zc1 --workload c2 --batch 64 --steps 20
export LANCE_B200_ZEROCOPY=0
run c2b64_zc0 --workload c2 --batch 64 --steps 20
