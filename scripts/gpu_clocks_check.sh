# clock sampler started before the warm-up + load extension of short timed regions: every line must carry samples
O=gpurun_out/clk
mkdir -p $O
run() { name=$1; shift; timeout 600 "$@" > $O/bench_$name.json 2> $O/bench_$name.err; python - "$O/bench_$name.json" <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); print(sys.argv[1].split("/")[-1], "ms", round(d["ms_per_step"],4), "q/s", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "frac", round(d["roofline"]["frac"],3), d["clocks"])
PY
}
run c4 python bench.py
run c1 python bench.py --workload c1 --steps 20
run c3 python bench.py --workload c3 --steps 20
run c5s01 python bench.py --workload c5 --selectivity 0.01 --steps 20
run c3b64 python bench.py --workload c3 --batch 64 --steps 20 --no-cpu-baseline
if [ "$(nvidia-smi -L | wc -l)" -ge 2 ]; then
run c4_n2 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 5 --warmup 3
fi
