#!/bin/bash
# threshold warm-up of the tcgen05 path on an 8-GPU-sized shard (12.5 M x 128, B=1024): seed tiles x growth x jump
for seed in 1 2 4 8 16; do for growth in 4 8; do for jump in 0 16; do
  if [ $jump = 0 ]; then J=""; else J="LANCE_B200_TC_JUMP=$jump"; fi
  r=$(env LANCE_B200_TC_SEED=$seed LANCE_B200_TC_GROWTH=$growth $J timeout 120 python bench.py --workload c4 --rows ${ROWS:-12500000} --steps 10 --warmup 3 --no-configs --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('%.4f ms  L=%d' % (d['ms_per_step'], d['gpu_launches']))")
  echo "seed=$seed growth=$growth jump=$jump : $r"
done; done; done
