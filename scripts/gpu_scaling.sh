mkdir -p gpurun_out/final3
O=gpurun_out/final3
for N in 2 4 8; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 10 --warmup 3 > $O/bench_c4_n$N.json 2> $O/bench_c4_n$N.err
head -c 300 $O/bench_c4_n$N.json; echo
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --impl reference --gpus 8 --steps 2 --warmup 1 > $O/bench_c4_n8_ref.json 2> $O/bench_c4_n8_ref.err
head -c 200 $O/bench_c4_n8_ref.json
