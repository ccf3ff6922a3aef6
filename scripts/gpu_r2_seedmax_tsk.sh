#!/bin/bash
# seeding pass on the K-streamed kernel: parity, then C2 (1 M x 768 cosine k=100) at B=64 / 4096 and C5-shape B=64 with the pass off / on
timeout 900 python -m pytest tests/test_flat_tc_gpu.py -m gpu -x -q 2>&1 | tail -3
for wl in "c2 --batch 64" "c2 --batch 4096" "c5 --batch 64"; do for sm in 0 128 256 512; do
  r=$(env LANCE_B200_TC_SEEDMAX=$sm timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-configs --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('%.4f ms  L=%d step_frac=%.3f' % (d['ms_per_step'], d['gpu_launches'], d['roofline']['step_frac']))")
  echo "$wl seedmax=$sm : $r"
done; done
