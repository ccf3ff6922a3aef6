set -x
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_c4_n2.json 2> gpurun_out/bench_c4_n2.err; tail -c 1500 gpurun_out/bench_c4_n2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 10 --warmup 3 --rows 25000000 > gpurun_out/bench_c4_n2_r25.json 2> gpurun_out/bench_c4_n2_r25.err; tail -c 500 gpurun_out/bench_c4_n2_r25.err
timeout 600 python bench.py --rows 12500000 --steps 10 --no-cpu-baseline > gpurun_out/bench_c4_n1_r12.json 2> gpurun_out/bench_c4_n1_r12.err
