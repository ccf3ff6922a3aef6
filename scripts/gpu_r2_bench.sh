#!/bin/bash
# round 2: GPU suite, then the default bench line (headline c4 + configs) and the reference arm
set -x
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest.log
tail -15 gpurun_out/r2_pytest.log
( time python bench.py --steps 20 --warmup 5 ) > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/r2_bench.err
( time python bench.py --impl reference --steps 20 --warmup 5 ) > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err; echo "ref rc=$?"
python scripts/sumbench.py gpurun_out/r2_bench.json || true
