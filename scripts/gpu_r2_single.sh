#!/bin/bash
# round 2: single-query path (flat_scan1): GPU suite, then the single-query bench lines
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest.log
tail -12 gpurun_out/r2_pytest.log
for w in c1 c5; do
  python bench.py --workload $w --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2s_$w.json 2> gpurun_out/r2s_$w.err || tail -5 gpurun_out/r2s_$w.err
done
python bench.py --workload c5 --selectivity 0.01 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2s_c5s01.json 2> gpurun_out/r2s_c5s01.err
python bench.py --workload c2 --batch 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r2s_c2b1.json 2> gpurun_out/r2s_c2b1.err
LANCE_B200_POLL=0 python bench.py --workload c1 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2s_c1_nopoll.json 2>/dev/null
LANCE_B200_SCAN1=0 python bench.py --workload c1 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2s_c1_old.json 2>/dev/null
python scripts/sumbench.py gpurun_out/r2s_*.json
