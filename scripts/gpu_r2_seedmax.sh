#!/bin/bash
# tile-max seeding pass of the TMEM-resident kernel: parity test, then the 8-GPU-sized shard (12.5 M x 128, B=1024) and
# the full C4 table with the pass off / on at several seed sizes
timeout 900 python -m pytest tests/test_flat_tc_gpu.py -m gpu -x -q -k "seeding or random_parity or prefilter" 2>&1 | tail -3
for rows in 12500000 100000000; do for sm in 0 256 512 1024 2048 4096; do
  r=$(env LANCE_B200_TC_SEEDMAX=$sm timeout 300 python bench.py --workload c4 --rows $rows --steps 10 --warmup 3 --no-configs --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('%.4f ms  L=%d fb=%s' % (d['ms_per_step'], d['gpu_launches'], d['config'].get('tc_fallback_queries')))")
  echo "rows=$rows seedmax=$sm : $r"
done; done
