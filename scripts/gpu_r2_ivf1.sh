#!/bin/bash
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest.log
tail -25 gpurun_out/r2_pytest.log
python bench.py --workload c3 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2s_c3b1.json 2> gpurun_out/r2s_c3b1.err || tail -5 gpurun_out/r2s_c3b1.err
LANCE_B200_IVF1=0 python bench.py --workload c3 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2s_c3b1_old.json 2>/dev/null
python bench.py --workload c1 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2s_c1.json 2>/dev/null
python bench.py --workload c5 --selectivity 0.01 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r2s_c5s01.json 2>/dev/null
python scripts/sumbench.py gpurun_out/r2s_c3b1.json gpurun_out/r2s_c3b1_old.json gpurun_out/r2s_c1.json gpurun_out/r2s_c5s01.json
