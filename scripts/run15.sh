# round-1 re-entry: full GPU suite, every workload's bench line, launch list + one full ncu capture of the GEMM kernel
set -x
mkdir -p gpurun_out
nproc > gpurun_out/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_gpu.txt; cat gpurun_out/pytest_gpu.txt
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; tail -c 600 gpurun_out/bench_c4.err
timeout 400 python bench.py --impl reference > gpurun_out/bench_c4_ref.json 2> gpurun_out/bench_c4_ref.err
for w in c1 c3 c5; do timeout 600 python bench.py --workload $w --steps 20 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; done
timeout 600 python bench.py --workload c2 > gpurun_out/bench_c2b4096.json 2> gpurun_out/bench_c2b4096.err
timeout 600 python bench.py --workload c2 --batch 64 --steps 20 --no-cpu-baseline > gpurun_out/bench_c2b64.json 2> gpurun_out/bench_c2b64.err
timeout 600 python bench.py --workload c2 --batch 1 --steps 20 --no-cpu-baseline > gpurun_out/bench_c2b1.json 2> gpurun_out/bench_c2b1.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'lb::' -c 400 --csv --log-file gpurun_out/c4_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c4_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:flat_gemm --launch-skip 19 --launch-count 1 -o gpurun_out/c4_gemm_full python bench.py --rows 25000000 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c4_full.log 2>&1
tail -3 gpurun_out/ncu_c4_full.log
ls -la gpurun_out
