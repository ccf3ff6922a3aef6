# last refresh of the round: C4 line + ncu of the ts kernel with the final build
set -x
mkdir -p gpurun_out/final3
O=gpurun_out/final3
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3 > $O/pytest_gpu.txt
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/bench_$name.json 2> $O/bench_$name.err; tail -c 200 $O/bench_$name.err; }
run c4 --workload c4
run c4_ref --impl reference --workload c4
run c1 --workload c1 --steps 20
run c2b1 --workload c2 --batch 1 --steps 20
run c3 --workload c3 --steps 20
run c5 --workload c5 --steps 20
run c5s01 --workload c5 --selectivity 0.01 --steps 20
NCU="ncu --clock-control none"
timeout 900 $NCU --metrics gpu__time_duration.sum -k regex:'kernel' -c 400 --csv --log-file $O/c4_launches.csv python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c4_list.log 2>&1
timeout 900 $NCU --metrics gpu__time_duration.sum -k regex:'kernel' -c 400 --csv --log-file $O/c3_launches.csv python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c3_list.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:flat_gemm --launch-skip 31 --launch-count 1 -o $O/c4_gemm_full python bench.py --rows 25000000 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c4_full.log 2>&1
ls $O
