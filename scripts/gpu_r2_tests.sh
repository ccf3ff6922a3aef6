#!/bin/bash
# round 2: the whole GPU suite (with the full-shape parity tests), then smoke()
set -x
nproc; free -g | head -2; nvidia-smi --query-gpu=name,memory.total --format=csv
( time python -m pytest tests -m gpu -x -q --durations=15 ) > gpurun_out/r2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest.log
tail -40 gpurun_out/r2_pytest.log
( time python __graft_entry__.py smoke ) > gpurun_out/r2_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r2_smoke.log
tail -5 gpurun_out/r2_smoke.log
