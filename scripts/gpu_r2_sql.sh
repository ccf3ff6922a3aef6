#!/bin/bash
# Runs the reference's own SQL tests (+ integration/test/sql/lance_b200_rewrite.test) through the vendored DuckDB's
# unittest with the patched extension over liblance_b200 (integration/build_duckdb_harness.sh staged the binaries).
O=gpurun_out/sql; mkdir -p $O
B=integration/_build
[ -x $B/unittest ] || { echo "no staged harness under $B"; exit 1; }
mkdir -p /tmp/lance_ext/test && rm -rf /tmp/lance_ext/test/sql && cp -r $B/test/sql /tmp/lance_ext/test/
export LD_LIBRARY_PATH=$PWD/$B:$LD_LIBRARY_PATH
rc=0
for t in /tmp/lance_ext/test/sql/*.test; do
  timeout 300 $PWD/$B/unittest --test-dir /tmp/lance_ext "$t" > $O/$(basename $t).log 2>&1; r=$?
  echo "$(basename $t) rc=$r $(grep -E 'All tests passed|test cases:' $O/$(basename $t).log | tail -1)" | tee -a $O/summary.txt
  [ $r -ne 0 ] && rc=1
done
exit $rc
