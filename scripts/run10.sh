timeout 600 python -m pytest tests/test_flat_tc_gpu.py -x -q 2>&1 | tail -8
timeout 600 python bench.py --workload c2 --steps 5 --no-cpu-baseline > gpurun_out/b_c2.json 2> gpurun_out/b_c2.err; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/b_c2.json")); print("c2 4096", round(d["value"]), d["ms_per_step"], d["roofline"]["frac"], d["profile_ms"], d["gpu_launches"], d["paths"])
except Exception as e: print(e, open("gpurun_out/b_c2.err").read()[-2000:])
PY
LANCE_B200_GEMM=ss timeout 600 python bench.py --workload c2 --steps 5 --no-cpu-baseline > gpurun_out/b_c2ss.json 2> gpurun_out/b_c2ss.err; python - <<PY
import json
d=json.load(open("gpurun_out/b_c2ss.json")); print("c2 4096 ss", round(d["value"]), d["ms_per_step"], d["roofline"]["frac"])
PY
timeout 600 python bench.py --workload c2 --batch 64 --steps 10 --no-cpu-baseline > gpurun_out/b_c2b64.json 2> gpurun_out/b_c2b64.err; python - <<PY
import json
d=json.load(open("gpurun_out/b_c2b64.json")); print("c2 64", round(d["value"]), d["ms_per_step"], d["roofline"])
PY
