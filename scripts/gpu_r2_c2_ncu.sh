# ncu --set full of the C2 (B=4096) step's three kernel classes
set -x
O=gpurun_out/r2c2
mkdir -p $O
CMD="python bench.py --workload c2 --steps 1 --warmup 3 --no-cpu-baseline"
$CMD > $O/plain.log 2>&1 || { tail -5 $O/plain.log; exit 1; }
timeout 600 ncu --clock-control none --set full --import-source on -k regex:rerank_kernel --launch-skip 0 --launch-count 1 -o $O/c2_rerank $CMD > $O/ncu1.log 2>&1
timeout 600 ncu --clock-control none --set full --import-source on -k regex:refresh_threshold --launch-skip 0 --launch-count 5 -o $O/c2_refresh $CMD > $O/ncu2.log 2>&1
timeout 600 ncu --clock-control none --set full --import-source on -k regex:flat_gemm_tsk --launch-skip 4 --launch-count 1 -o $O/c2_tsk $CMD > $O/ncu3.log 2>&1
tail -2 $O/ncu1.log $O/ncu2.log $O/ncu3.log
ls -la $O
