timeout 600 python -m pytest tests/test_flat_tc_gpu.py tests/test_host_gpu.py -x -q 2>&1 | tail -8
LANCE_B200_GEMM=ss timeout 600 python -m pytest tests/test_flat_tc_gpu.py -x -q 2>&1 | tail -3
for v in ts ss; do
LANCE_B200_GEMM=$v timeout 300 python bench.py --workload c4 --rows 10000000 --steps 5 --no-cpu-baseline > gpurun_out/b_c4_10m_$v.json 2> gpurun_out/b_c4_10m_$v.err; python - <<PY
import json
d=json.load(open("gpurun_out/b_c4_10m_$v.json")); print("$v", round(d["value"]), d["ms_per_step"], d["roofline"]["achieved"], d["paths"])
PY
done
timeout 600 python bench.py --steps 5 > gpurun_out/b_c4.json 2> gpurun_out/b_c4.err; python - <<PY
import json
d=json.load(open("gpurun_out/b_c4.json")); print("c4 full", round(d["value"]), d["ms_per_step"], d["roofline"], d["e2e"])
PY
