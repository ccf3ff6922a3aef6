mkdir -p gpurun_out
NCU="ncu --clock-control none"
timeout 900 python -m pytest tests/test_flat_tc_gpu.py tests/test_ivfpq_gpu.py -x -q 2>&1 | tail -3
timeout 900 $NCU --set full --import-source on -k regex:rerank_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/c2_rerank_full python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c2_rerank.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:refresh_threshold --launch-skip 30 --launch-count 2 -o gpurun_out/c2_refresh_full python bench.py --workload c2 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c2_refresh.log 2>&1
