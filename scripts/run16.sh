# fused-merge validation, launch lists, full captures of the largest GEMM segments (c4 ts kernel, c2 tsk kernel)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_gpu.txt; cat gpurun_out/pytest_gpu.txt
for w in c1 c5; do timeout 600 python bench.py --workload $w --steps 20 --no-cpu-baseline > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; done
timeout 600 python bench.py --workload c2 --batch 1 --steps 20 --no-cpu-baseline > gpurun_out/bench_c2b1.json 2> gpurun_out/bench_c2b1.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'kernel' -c 400 --csv --log-file gpurun_out/c4_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c4_list.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'kernel' -c 400 --csv --log-file gpurun_out/c2_launches.csv python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c2_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:flat_gemm --launch-skip 31 --launch-count 1 -o gpurun_out/c4_gemm_full python bench.py --rows 25000000 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c4_full.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:flat_gemm --launch-skip 34 --launch-count 1 -o gpurun_out/c2_gemm_full python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c2_full.log 2>&1
ls -la gpurun_out
