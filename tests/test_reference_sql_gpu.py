"""The reference's own SQL tests through the vendored DuckDB with the patched extension over liblance_b200.

Runs only where integration/build_duckdb_harness.sh staged its binaries (integration/_build/, git-ignored: needs
/root/reference and ~25 min of build); skipped elsewhere.  Evidence of the last run: profiles/r2_sql_reference_tests.txt."""
import os
import shutil
import subprocess

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "integration", "_build")
# lance_optimizer_filter.test pins the order of an exact tie (the tie-neutral copy is the one expected to pass);
# lance_hnsw.test: the HNSW request is accepted and answered by the exact scan (include/lance_b200.h)
EXPECTED_TO_PASS = ["lance_b200_rewrite", "lance_basic", "lance_delete", "lance_hnsw", "lance_optimizer", "lance_optimizer_filter_tie_neutral",
                    "lance_persist", "lance_text_columns"]


@pytest.mark.parametrize("name", EXPECTED_TO_PASS)
def test_reference_sql_file(gpu, name, tmp_path):
    exe = os.path.join(BUILD, "unittest")
    if not os.path.exists(exe):
        pytest.skip("DuckDB harness not staged (integration/build_duckdb_harness.sh)")
    # the test paths are compiled into the binary: /tmp/lance_ext/test/sql
    dst = "/tmp/lance_ext/test/sql"
    os.makedirs(dst, exist_ok=True)
    shutil.copy(os.path.join(BUILD, "test", "sql", name + ".test"), dst)
    env = dict(os.environ, LD_LIBRARY_PATH=BUILD + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    r = subprocess.run([exe, "--test-dir", "/tmp/lance_ext", f"{dst}/{name}.test"], env=env, capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0 and "All tests passed" in r.stdout, r.stdout[-3000:] + r.stderr[-1000:]
