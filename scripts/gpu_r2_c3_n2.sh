#!/bin/bash
# C3 (IVF_PQ) sharded by partition on N GPUs: single query and B=4096
N=${1:-2}; O=gpurun_out/c3n$N; mkdir -p $O
for B in 1 4096; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --workload c3 --batch $B --steps 10 --warmup 3 --no-cpu-baseline --no-configs > $O/bench_c3b${B}_n$N.json 2> $O/bench_c3b${B}_n$N.err; echo "B=$B rc=$?"; tail -c 600 $O/bench_c3b${B}_n$N.err
done
python scripts/sumbench.py $O/bench_c3b1_n$N.json $O/bench_c3b4096_n$N.json
