#!/bin/bash
# growth factor of the emitting segments once the tile-max seeding pass has set the thresholds (12.5 M x 128 shard, B=1024)
for sm in 1024 2048; do for g in 2 3 4 6 8; do
  r=$(env LANCE_B200_TC_SEEDMAX=$sm LANCE_B200_TC_GROWTH=$g timeout 300 python bench.py --workload c4 --rows 12500000 --steps 20 --warmup 3 --no-configs --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('%.4f ms  L=%d' % (d['ms_per_step'], d['gpu_launches']))")
  echo "seedmax=$sm growth=$g : $r"
done; done
