# re-entry check of the torchrun path (2 ranks): reference arm first, then the sharded C4 step
set -x
O=gpurun_out/reentry
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 400 $TR bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $O/bench_c4_n2_ref.json 2> $O/bench_c4_n2_ref.err; tail -c 300 $O/bench_c4_n2_ref.err; head -c 300 $O/bench_c4_n2_ref.json; echo
timeout 500 $TR bench.py --gpus 2 --steps 5 --warmup 3 > $O/bench_c4_n2.json 2> $O/bench_c4_n2.err; tail -c 300 $O/bench_c4_n2.err; head -c 300 $O/bench_c4_n2.json; echo
