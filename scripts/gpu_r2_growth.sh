#!/bin/bash
for g in 8 12 16; do LANCE_B200_TSK_EB=12 LANCE_B200_TC_GROWTH=$g timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2s_c2g$g.json 2>/dev/null; done
LANCE_B200_TSK_EB=12 LANCE_B200_TC_GROWTH=8 LANCE_B200_TC_SEED=32 timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2s_c2g8s32.json 2>/dev/null
for g in 4 8 16; do LANCE_B200_TC_GROWTH=$g timeout 300 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/r2s_c4g$g.json 2>/dev/null; done
for g in 8 16; do LANCE_B200_TC_GROWTH=$g timeout 300 python bench.py --workload c4 --rows 12500000 --steps 20 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/r2s_c4shard_g$g.json 2>/dev/null; done
LANCE_B200_TC_GROWTH=4 timeout 300 python bench.py --workload c4 --rows 12500000 --steps 20 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/r2s_c4shard_g4.json 2>/dev/null
python scripts/sumbench.py gpurun_out/r2s_c2g*.json gpurun_out/r2s_c4g*.json gpurun_out/r2s_c4shard*.json
