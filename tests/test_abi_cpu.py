"""CPU-side checks of the drop-in boundary: the library loads, exports every symbol the header declares,
the harness binds every one of them, and without a GPU the product fails loudly (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest


def declared_symbols(header_path):
    text = open(header_path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lance_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(pkg):
    names = declared_symbols(pkg.HEADER)
    assert len(names) >= 30
    L = ctypes.CDLL(pkg.LIB_PATH)
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    # the harness binds exactly the declared set
    assert sorted(pkg.SIGNATURES) == names


def test_reference_symbols_are_all_there(pkg):
    """The 20 `lance_*` symbols of /root/reference/src/rust_ffi.cpp:9-41 (names frozen here: the GPU box has
    no /root/reference)."""
    ref = """lance_create_detached lance_create_detached_from_arrow lance_open_detached lance_free_detached
    lance_detached_has_extra_columns lance_detached_dimension lance_detached_add lance_detached_add_batch
    lance_detached_add_batch_arrow lance_detached_merge lance_detached_search lance_detached_count
    lance_detached_delete lance_detached_delete_batch lance_detached_create_index
    lance_detached_create_hnsw_index lance_detached_compact lance_detached_get_vector
    lance_detached_get_all_vectors""".split()
    L = ctypes.CDLL(pkg.LIB_PATH)
    for n in ref:
        assert hasattr(L, n), n
    if os.path.isdir("/root/reference/src"):
        text = open("/root/reference/src/rust_ffi.cpp").read()
        block = text[text.index('extern "C" {'):text.index("namespace duckdb")]
        in_ref = sorted(set(re.findall(r"\b(lance_[a-z0-9_]+)\s*\(", block)))
        assert sorted(ref) == in_ref


def test_no_cpu_fallback(pkg):
    """With no CUDA device, create fails with a message through err_buf; nothing is computed on the CPU."""
    if pkg.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.LanceError, match="no CUDA device"):
        pkg.LanceHandle(8, "l2")
    # NULL handle convention (rust_lib/src/ffi.rs: "null handle")
    err = ctypes.create_string_buffer(64)
    q = np.zeros(8, np.float32)
    rc = pkg.lib().lance_detached_search(None, q.ctypes.data_as(ctypes.c_void_p), 8, 1, 1, 1, None, None, err, 64)
    assert rc == -1 and err.value == b"null handle"
    assert pkg.lib().lance_detached_count(None, err, 64) == -1
    assert pkg.lib().lance_detached_dimension(None) == 0
    pkg.lib().lance_free_detached(None)  # no-op


def test_err_buf_truncation(pkg):
    err = ctypes.create_string_buffer(8)
    rc = pkg.lib().lance_detached_count(None, err, 8)
    assert rc == -1 and err.value == b"null ha"  # len-1 bytes + NUL
    assert pkg.lib().lance_detached_count(None, None, 0) == -1  # err_buf may be NULL


def test_bitmap_layout(pkg, oracle_mod):
    rng = np.random.default_rng(0)
    m = rng.random(1000) < 0.3
    a, b = pkg.make_bitmap(m), oracle_mod.make_bitmap(m)
    assert a.dtype == np.uint32 and (a == b).all()
    for i in (0, 31, 32, 999):
        assert bool((a[i >> 5] >> (i & 31)) & 1) == bool(m[i])


def test_product_does_not_import_oracle():
    """The oracle is test infrastructure: nothing under the product package may reference it."""
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "duckdb-lancedb_b200")
    for dp, _, files in os.walk(root):
        if "build" in dp:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")):
                text = open(os.path.join(dp, f)).read()
                assert "import oracle" not in text and "liboracle" not in text and '#include "oracle' not in text, f
