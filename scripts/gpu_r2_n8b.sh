#!/bin/bash
# 8 GPUs of one box after the seeding pass / parallel shard merge: C4 (rows sharded, one process per GPU) and the in-process handle
N=${1:-8}; O=gpurun_out/n${N}b; mkdir -p $O
timeout 300 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q 2>&1 | tail -2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --workload c4 --steps 20 --warmup 5 --no-cpu-baseline --no-configs > $O/bench_c4_n$N.json 2> $O/bench_c4_n$N.err; echo "c4 rc=$?"
python scripts/sumbench.py $O/bench_c4_n$N.json
LANCE_B200_SHARD_TRACE=1 timeout 600 python scripts/bench_sharded_inproc.py $((8000000 * N)) 1024 2> $O/inproc_trace.txt | tee $O/inproc_n$N.json
tail -$N $O/inproc_trace.txt
