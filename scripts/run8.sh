timeout 600 python -m pytest tests/test_flat_tc_gpu.py -x -q 2>&1 | tail -3
for g in 16 8 4; do for sd in 16 4; do
LANCE_B200_TC_GROWTH=$g LANCE_B200_TC_SEED=$sd timeout 300 python bench.py --workload c4 --rows 10000000 --steps 5 --no-cpu-baseline > gpurun_out/t.json 2> gpurun_out/t.err; python - <<PY
import json
d=json.load(open("gpurun_out/t.json")); print("growth $g seed $sd", round(d["value"]), round(d["ms_per_step"],3), d["gpu_launches"])
PY
done; done
timeout 600 python bench.py --steps 5 --no-cpu-baseline > gpurun_out/b_c4.json 2> gpurun_out/b_c4.err; python - <<PY
import json
d=json.load(open("gpurun_out/b_c4.json")); print("c4 full", round(d["value"]), d["ms_per_step"], d["roofline"]["frac"])
PY
