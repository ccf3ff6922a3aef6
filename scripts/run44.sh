mkdir -p gpurun_out/final2
O=gpurun_out/final2
run() { name=$1; shift; timeout 900 python bench.py "$@" > $O/bench_$name.json 2> $O/bench_$name.err; tail -c 200 $O/bench_$name.err; }
run c3 --workload c3 --steps 20
run c5s01 --workload c5 --selectivity 0.01 --steps 20
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29531 bench.py --impl reference --gpus 1 --steps 2 --warmup 1 > $O/bench_c4_ref_torchrun.json 2> $O/bench_c4_ref_torchrun.err
