timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for w in "c3 --steps 20" "c3 --batch 1024 --steps 5 --no-cpu-baseline" "c2 --steps 5 --no-cpu-baseline" "c1 --steps 50 --no-cpu-baseline" "c5 --steps 10 --no-cpu-baseline" "c4 --rows 10000000 --steps 5 --no-cpu-baseline"; do
  name=$(echo $w | tr -d ' -' ); timeout 900 python bench.py --workload $w > gpurun_out/b_$name.json 2> gpurun_out/b_$name.err; echo "== $w"; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/b_$name.json")); r=d["roofline"] or {}
    print(round(d["value"],1),"q/s e2e",round(d["e2e"]["value"],1),"ms/step",round(d["ms_per_step"],3),r.get("bound"),r.get("achieved"),r.get("frac"),d["paths"],d["cpu_baseline"] and d["cpu_baseline"]["value"])
except Exception as e: print("fail",e); print(open("gpurun_out/b_$name.err").read()[-1500:])
PY
done
