# final build on 8 ranks: the sharded C4 step under torchrun
O=gpurun_out/reentry
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512"
timeout 400 $TR bench.py --gpus 8 --steps 5 --warmup 3 > $O/bench_c4_n8.json 2> $O/bench_c4_n8.err; tail -c 300 $O/bench_c4_n8.err; head -c 600 $O/bench_c4_n8.json; echo
