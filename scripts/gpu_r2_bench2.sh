#!/bin/bash
( time python bench.py --steps 20 --warmup 5 ) > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo "bench rc=$?"
tail -c 400 gpurun_out/r2_bench.err
python scripts/sumbench.py gpurun_out/r2_bench.json
