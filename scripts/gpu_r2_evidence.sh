#!/bin/bash
# round 2 evidence on one B200: GPU suite, smoke, default bench line (all configs), launch lists and
# ncu --set full captures of the dominant kernels (each only after its plain command exited 0)
set -x
O=gpurun_out/r2ev2
mkdir -p $O
nproc; nvidia-smi --query-gpu=name,memory.total --format=csv
( time timeout 1500 python -m pytest tests -m gpu -x -q --durations=10 ) > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -18 $O/pytest.log
( time timeout 300 python __graft_entry__.py smoke ) > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
tail -3 $O/smoke.log
( time timeout 1200 python bench.py ) > $O/bench_default.json 2> $O/bench_default.err; echo "bench rc=$?"
tail -c 300 $O/bench_default.err
python scripts/sumbench.py $O/bench_default.json
ncu_list() { name=$1; shift; timeout 600 python bench.py "$@" --steps 2 --warmup 3 --no-cpu-baseline --no-configs > $O/plain_$name.log 2>&1 || { tail -3 $O/plain_$name.log; return; }
  timeout 900 ncu --clock-control none --metrics gpu__time_duration.sum -c 600 --csv --log-file $O/${name}_launches.csv python bench.py "$@" --steps 2 --warmup 3 --no-cpu-baseline --no-configs > $O/ncu_list_$name.log 2>&1; }
ncu_list c1 --workload c1
ncu_list c3 --workload c3
ncu_list c2 --workload c2
ncu_list c4 --workload c4
ncu_full() { name=$1; regex=$2; skip=$3; shift 3; timeout 900 ncu --clock-control none --set full --import-source on -k regex:$regex --launch-skip $skip --launch-count ${CNT:-1} -o $O/$name python bench.py "$@" --steps 1 --warmup 3 --no-cpu-baseline --no-configs > $O/ncu_full_$name.log 2>&1; tail -1 $O/ncu_full_$name.log
  # the report itself stays on the box (gpurun_out is capped at 64 MiB): summarise it here
  python scripts/ncu_summary.py $O/$name.ncu-rep $O/${name}_ncu_full.txt; rm -f $O/$name.ncu-rep; }
ncu_full c1_flat_scan1 flat_scan1_kernel 3 --workload c1
ncu_full c3_ivfpq1 ivfpq1_kernel 3 --workload c3
CNT=10 ncu_full c2_tsk flat_gemm_tsk_kernel 24 --workload c2
CNT=1 ncu_full c5_predicate eval_predicate_kernel 0 --workload c5 --selectivity 0.01
bash scripts/gpu_r2_sql.sh > $O/sql.log 2>&1; tail -10 $O/sql.log
ls -la $O
