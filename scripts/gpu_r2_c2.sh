#!/bin/bash
( time timeout 600 python -m pytest tests/test_flat_tc_gpu.py tests/test_baseline_shapes_gpu.py -m gpu -x -q ) > gpurun_out/r2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest.log
tail -6 gpurun_out/r2_pytest.log
timeout 300 python bench.py --workload c2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2s_c2b4096.json 2> gpurun_out/r2s_c2b4096.err || tail -5 gpurun_out/r2s_c2b4096.err
LANCE_B200_TSK_MC=0 timeout 300 python bench.py --workload c2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2s_c2b4096_nomc.json 2>/dev/null
timeout 300 python bench.py --workload c2 --batch 64 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2s_c2b64.json 2>/dev/null
for ks in 2 3; do LANCE_B200_TSK_KS=$ks timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2s_c2b4096_ks$ks.json 2>/dev/null; done
LANCE_B200_TSK_DBG=1 timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2s_c2dbg1.json 2>/dev/null
python scripts/sumbench.py gpurun_out/r2s_c2b4096.json gpurun_out/r2s_c2b4096_nomc.json gpurun_out/r2s_c2b64.json gpurun_out/r2s_c2b4096_ks*.json gpurun_out/r2s_c2dbg1.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2s_c2b4096.json').read())
print(d['profile_ms'])
PY
