# re-entry check of the restored build: GPU suite, smoke, default bench line
set -x
O=gpurun_out/reentry
mkdir -p $O
timeout 700 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > $O/pytest_gpu.txt; cat $O/pytest_gpu.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2 > $O/smoke.txt; cat $O/smoke.txt
timeout 500 python bench.py > $O/bench_c4.json 2> $O/bench_c4.err; tail -c 300 $O/bench_c4.err; head -c 400 $O/bench_c4.json; echo
