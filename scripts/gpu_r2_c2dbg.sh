#!/bin/bash
for dbg in 0 1 2 3 4 6; do
LANCE_B200_TSK_DBG=$dbg timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2s_c2dbg$dbg.json 2>/dev/null
done
python scripts/sumbench.py gpurun_out/r2s_c2dbg*.json
python - <<'PY'
import json
for dbg in (0,1,2,3,4,6):
    d=json.loads(open(f'gpurun_out/r2s_c2dbg{dbg}.json').read())
    pm=d['profile_ms']; print(dbg, round(pm['flat_gemm_ms']/10,3),'ms gemm/step', round(pm['rerank_ms']/10,3), 'rerank/step', d['paths'])
PY
