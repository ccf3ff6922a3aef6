#!/bin/bash
for eb in 4 8 12 16; do LANCE_B200_TSK_EB=$eb timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2s_c2eb$eb.json 2>/dev/null; done
LANCE_B200_TSK_EB=16 LANCE_B200_TSK_KS=3 timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2s_c2eb16ks3.json 2>/dev/null
LANCE_B200_TC_GROWTH=2 timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2s_c2g2.json 2>/dev/null
LANCE_B200_TC_GROWTH=3 timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2s_c2g3.json 2>/dev/null
LANCE_B200_TC_GROWTH=8 timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2s_c2g8.json 2>/dev/null
python scripts/sumbench.py gpurun_out/r2s_c2eb*.json gpurun_out/r2s_c2g*.json
