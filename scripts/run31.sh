timeout 600 python bench.py --workload c3 --steps 2 --no-cpu-baseline 2>&1 | grep "ADC last" | tail -4
