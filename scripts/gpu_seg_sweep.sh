# segment schedule sweep of the batched flat path at B=64 (C2 1M x 768 cosine, C5 10M x 384 l2)
mkdir -p gpurun_out/seg
O=gpurun_out/seg
for cfg in "16 4" "16 8" "16 16" "64 8" "64 16" "32 32" "128 16"; do
  set -- $cfg
  for wl in c2 c5; do
    LANCE_B200_TC_SEED=$1 LANCE_B200_TC_GROWTH=$2 timeout 600 python bench.py --workload $wl --batch 64 --steps 10 --no-cpu-baseline > $O/bench_${wl}b64_s$1_g$2.json 2> $O/bench_${wl}b64_s$1_g$2.err
    python - "$O/bench_${wl}b64_s$1_g$2.json" <<'PY'
import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1]); print(sys.argv[1].split("/")[-1], round(d["ms_per_step"],4), d["gpu_launches"]/d["steps"])
PY
  done
done
