#!/bin/bash
# final ncu evidence of round 2: launch list of one C4 step (seeding pass + segments), full captures of the seeding pass and
# the largest segment (25 M-row table: the capture replays each kernel ~40 times), and of the single-query scan on C2 / C5 / C5 1 %
O=gpurun_out/final; mkdir -p $O
ok() { timeout 600 python bench.py "$@" --steps 2 --warmup 3 --no-cpu-baseline --no-configs > $O/plain.log 2>&1 || { tail -3 $O/plain.log; return 1; }; }
ok --workload c4 && timeout 900 ncu --clock-control none --metrics gpu__time_duration.sum -c 400 --csv --log-file $O/c4_launches.csv python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu-baseline --no-configs > $O/ncu_list_c4.log 2>&1
full() { name=$1; regex=$2; skip=$3; cnt=$4; shift 4; ok "$@" || return
  timeout 900 ncu --clock-control none --set full --import-source on -k regex:$regex --launch-skip $skip --launch-count $cnt -o $O/$name python bench.py "$@" --steps 1 --warmup 3 --no-cpu-baseline --no-configs > $O/ncu_full_$name.log 2>&1; tail -1 $O/ncu_full_$name.log
  python scripts/ncu_summary.py $O/$name.ncu-rep $O/${name}_ncu_full.txt; rm -f $O/$name.ncu-rep; }
# 25 M rows: seeding pass + 4 segments per step; skip the first step (shadow build interleaved), take one whole step
full c4_ts flat_gemm_ts_kernel 5 5 --workload c4 --rows 25000000
full c2b1_scan1 flat_scan1_kernel 3 1 --workload c2 --batch 1
full c5_scan1 flat_scan1_kernel 3 1 --workload c5
full c5s01_scan1 flat_scan1_kernel 3 1 --workload c5 --selectivity 0.01
ls -la $O
