#!/bin/bash
# 8 GPUs of one box: sharded-handle tests, C4 (flat, rows sharded), C3 B=4096 (IVF_PQ, lists sharded), the in-process handle
N=${1:-8}; O=gpurun_out/n$N; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -$N
timeout 300 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q 2>&1 | tail -3
run() { name=$1; shift; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N "$@" --no-cpu-baseline --no-configs > $O/bench_${name}_n$N.json 2> $O/bench_${name}_n$N.err; echo "$name rc=$?"; }
run c4 --workload c4 --steps 20 --warmup 5
run c3b4096 --workload c3 --batch 4096 --steps 10 --warmup 3
python scripts/sumbench.py $O/bench_c4_n$N.json $O/bench_c3b4096_n$N.json
timeout 600 python scripts/bench_sharded_inproc.py $((8000000 * N)) 1024 | tee $O/inproc_n$N.json
