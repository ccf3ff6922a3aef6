timeout 600 python -m pytest tests/test_flat_tc_gpu.py -x -q 2>&1 | tail -3
for ks in 2 4 6; do
LANCE_B200_TSK_KS=$ks timeout 600 python bench.py --workload c2 --steps 5 --no-cpu-baseline > gpurun_out/t.json 2> gpurun_out/t.err; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/t.json")); print("c2 4096 KS=$ks", round(d["value"]), round(d["ms_per_step"],3), round(d["roofline"]["frac"],3), d["profile_ms"]["flat_gemm_ms"]/5)
except Exception as e: print(e, open("gpurun_out/t.err").read()[-1500:])
PY
done
