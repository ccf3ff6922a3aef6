# final round-1 evidence: bench lines of every config (+ CPU baseline), launch lists, full captures of the dominant kernels
set -x
mkdir -p gpurun_out/final
O=gpurun_out/final
nproc > $O/nproc.txt
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3 > $O/pytest_gpu.txt
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1 >> $O/pytest_gpu.txt
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/bench_$name.json 2> $O/bench_$name.err; tail -c 200 $O/bench_$name.err; }
run c4 --workload c4
run c4_ref --impl reference --workload c4
run c1 --workload c1 --steps 20
run c2b4096 --workload c2
run c2b64 --workload c2 --batch 64 --steps 20
run c2b1 --workload c2 --batch 1 --steps 20
run c3 --workload c3 --steps 20
run c3b64 --workload c3 --batch 64 --steps 10 --no-cpu-baseline
run c3b4096 --workload c3 --batch 4096 --steps 3 --no-cpu-baseline
run c5 --workload c5 --steps 20
run c5b64 --workload c5 --batch 64 --steps 10 --no-cpu-baseline
NCU="ncu --clock-control none"
for w in c4 c2 c3 c1; do
timeout 900 $NCU --metrics gpu__time_duration.sum -k regex:'kernel' -c 400 --csv --log-file $O/${w}_launches.csv python bench.py --workload $w --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_${w}_list.log 2>&1
done
timeout 900 $NCU --set full --import-source on -k regex:flat_gemm --launch-skip 31 --launch-count 1 -o $O/c4_gemm_full python bench.py --rows 25000000 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c4_full.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:flat_gemm --launch-skip 34 --launch-count 1 -o $O/c2_gemm_full python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c2_full.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:flat_scan --launch-skip 3 --launch-count 1 -o $O/c2b1_scan_full python bench.py --workload c2 --batch 1 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c2b1_full.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:flat_scan --launch-skip 3 --launch-count 1 -o $O/c5_scan_full python bench.py --workload c5 --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu_c5_full.log 2>&1
ls -la $O
