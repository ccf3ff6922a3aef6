timeout 600 python -m pytest tests/test_host_gpu.py -x -q 2>&1 | tail -3
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c4_10m.csv python bench.py --workload c4 --rows 10000000 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l.log 2>&1; tail -2 gpurun_out/ncu_l.log
