# round-end check of the final build: whole GPU suite, smoke, C3 lines with the CPU leg, C3 launch list, default bench line
set -x
mkdir -p gpurun_out/final4
O=gpurun_out/final4
timeout 1200 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > $O/pytest_gpu.txt; cat $O/pytest_gpu.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2 > $O/smoke.txt; cat $O/smoke.txt
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/bench_$name.json 2> $O/bench_$name.err; tail -c 200 $O/bench_$name.err; head -c 150 $O/bench_$name.json; echo; }
run c3 --workload c3 --steps 20
timeout 900 ncu --clock-control none --metrics gpu__time_duration.sum -k regex:'kernel' -c 400 --csv --log-file $O/c3_launches.csv python bench.py --workload c3 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c3_list.log 2>&1
LANCE_B200_ADC_CTAS=148 run c3_ctas148 --workload c3 --steps 20 --no-cpu-baseline
LANCE_B200_ADC_CTAS=74 run c3_ctas74 --workload c3 --steps 20 --no-cpu-baseline
run c1 --workload c1 --steps 20
run c4 --workload c4
ls $O
