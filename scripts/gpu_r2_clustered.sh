#!/bin/bash
# clustered (unit-norm mixture) data: parity tests, then C2 / C4 bench lines beside the Gaussian ones
O=gpurun_out/clu; mkdir -p $O
timeout 600 python -m pytest tests/test_flat_tc_gpu.py -m gpu -x -q 2>&1 | tail -15
for w in c2 c4; do
  timeout 600 python bench.py --workload $w --data clustered --no-configs --no-cpu-baseline > $O/bench_${w}_clustered.json 2> $O/bench_${w}_clustered.err; echo "$w rc=$?"
done
python scripts/sumbench.py $O/bench_c2_clustered.json $O/bench_c4_clustered.json
python - <<'PY'
import json
for w in ("c2", "c4"):
    d = json.loads([l for l in open(f"gpurun_out/clu/bench_{w}_clustered.json") if l.startswith("{")][-1])
    print(w, "paths", d.get("paths"), "ms", round(d["ms_per_step"], 3))
PY
