"""ctypes harness over the C++ host mirror (csrc/host/, include/lance_b200_host.h) — tests only.

`HostIndex` drives lance_host::LanceIndex (the mirror of src/lance_index.cpp's class), `lance_search()` the
bind/init/scan triple of src/lance_search.cpp, `lance_index_scan()` the optimizer's replacement scan of
src/lance_optimizer.cpp:35-109, `Columns.evaluate()` the predicate push-down of :204-344.  Nothing is computed here.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import ERR_LEN, LIB_PATH, LanceError

HOST_HEADER = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "lance_b200_host.h")

_vp, _cp, _i32, _i64 = C.c_void_p, C.c_char_p, C.c_int32, C.c_int64
_ERR = [C.c_char_p, C.c_int]
HOST_SIGNATURES = {
    "lbh_columns_create": (_vp, []),
    "lbh_columns_free": (None, [_vp]),
    "lbh_columns_append_i64": (_i32, [_vp, _cp, _vp, _vp, _i64]),
    "lbh_columns_append_f64": (_i32, [_vp, _cp, _vp, _vp, _i64]),
    "lbh_columns_append_bool": (_i32, [_vp, _cp, _vp, _vp, _i64]),
    "lbh_columns_append_str": (_i32, [_vp, _cp, C.POINTER(C.c_char_p), _i64]),
    "lbh_eval_predicate": (_i32, [_vp, _cp, _i64, _vp] + _ERR),
    "lbh_distance_function_metric": (_i32, [_cp, C.c_char_p, C.c_int]),
    "lbh_metric_matches_index": (_i32, [_cp, _cp]),
    "lbh_index_create": (_vp, [_cp, _i32, _cp, _i32, _i32] + _ERR),
    "lbh_index_free": (None, [_vp]),
    "lbh_index_columns": (_vp, [_vp]),
    "lbh_index_append": (_i32, [_vp, _vp, _i64, _vp] + _ERR),
    "lbh_index_delete": (_i32, [_vp, _vp, _i64] + _ERR),
    "lbh_index_vacuum": (_i32, [_vp] + _ERR),
    "lbh_index_merge": (_i32, [_vp, _vp] + _ERR),
    "lbh_index_count": (_i64, [_vp] + _ERR),
    "lbh_index_create_ann": (_i32, [_vp, _i32, _i32] + _ERR),
    "lbh_index_search": (_i32, [_vp, _vp, _i32, _i32, _cp, _vp, _vp] + _ERR),
    "lbh_index_search_batch": (_i32, [_vp, _vp, _i32, _i32, _i32, _cp, _vp, _vp, _vp] + _ERR),
    "lbh_lance_search": (_i32, [_vp, _vp, _i32, _i32, _vp, _vp, _vp] + _ERR),
    "lbh_lance_index_scan": (_i32, [_vp, _vp, _i32, _i32, _cp, _vp] + _ERR),
}

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise LanceError(f"{LIB_PATH} is missing: run __graft_entry__.build()")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in HOST_SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Columns:
    """A lance_host::ColumnSet (standalone, or the one an index owns)."""

    def __init__(self, _raw=None):
        self._own = _raw is None
        self._c = lib().lbh_columns_create() if _raw is None else _raw

    def close(self):
        if self._own and self._c:
            lib().lbh_columns_free(self._c)
        self._c = None

    __del__ = close

    def append(self, name: str, values):
        """values: list of str/None, int/None, float/None or bool/None (type fixed by the first non-null)."""
        vals = list(values)
        first = next((v for v in vals if v is not None), None)
        valid = np.array([v is not None for v in vals], np.uint8)
        nm = name.encode()
        if isinstance(first, str) or first is None:
            arr = (C.c_char_p * len(vals))(*[None if v is None else v.encode() for v in vals])
            lib().lbh_columns_append_str(self._c, nm, arr, len(vals))
        elif isinstance(first, (bool, np.bool_)):
            a = np.array([bool(v) if v is not None else False for v in vals], np.uint8)
            lib().lbh_columns_append_bool(self._c, nm, _p(a), _p(valid), len(vals))
        elif isinstance(first, (int, np.integer)):
            a = np.array([int(v) if v is not None else 0 for v in vals], np.int64)
            lib().lbh_columns_append_i64(self._c, nm, _p(a), _p(valid), len(vals))
        else:
            a = np.array([float(v) if v is not None else 0.0 for v in vals], np.float64)
            lib().lbh_columns_append_f64(self._c, nm, _p(a), _p(valid), len(vals))

    def evaluate(self, predicate: str, n: int) -> np.ndarray:
        """-> bool[n]: rows for which the predicate is TRUE."""
        bm = np.zeros((n + 31) // 32 + 1, np.uint32)
        err = C.create_string_buffer(ERR_LEN)
        if lib().lbh_eval_predicate(self._c, predicate.encode(), n, _p(bm), err, ERR_LEN) != 0:
            raise LanceError(err.value.decode())
        bits = np.unpackbits(bm.view(np.uint8), bitorder="little")[:n]
        return bits.astype(bool)


def distance_function_metric(name: str):
    out = C.create_string_buffer(32)
    return out.value.decode() if lib().lbh_distance_function_metric(name.encode(), out, 32) else None


def metric_matches_index(fm: str, im: str) -> bool:
    return bool(lib().lbh_metric_matches_index(fm.encode(), im.encode()))


class HostIndex:
    """lance_host::LanceIndex: CREATE INDEX .. USING LANCE (v) WITH (metric, nprobes, refine_factor)."""

    def __init__(self, dimension: int, metric: str = "", nprobes: int = 0, refine_factor: int = -1, name: str = "idx"):
        self._err = C.create_string_buffer(ERR_LEN)
        self.dim = dimension
        self._h = lib().lbh_index_create(name.encode(), dimension, metric.encode(), nprobes, refine_factor, self._err, ERR_LEN)
        if not self._h:
            raise LanceError(self._err.value.decode())
        self.columns = Columns(_raw=lib().lbh_index_columns(self._h))

    def close(self):
        if getattr(self, "_h", None):
            lib().lbh_index_free(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc):
        if rc is None or rc < 0:
            raise LanceError(self._err.value.decode())
        return rc

    def append(self, vectors, row_ids):
        v = np.ascontiguousarray(vectors, np.float32)
        r = np.ascontiguousarray(row_ids, np.int64)
        self._check(lib().lbh_index_append(self._h, _p(v), v.shape[0], _p(r), self._err, ERR_LEN))

    def delete(self, row_ids):
        r = np.ascontiguousarray(row_ids, np.int64)
        self._check(lib().lbh_index_delete(self._h, _p(r), r.shape[0], self._err, ERR_LEN))

    def vacuum(self):
        self._check(lib().lbh_index_vacuum(self._h, self._err, ERR_LEN))

    def merge_indexes(self, other) -> int:
        """LanceIndex::MergeIndexes (src/lance_index.cpp:624-668): the live rows of `other` move into this index."""
        return int(self._check(lib().lbh_index_merge(self._h, other._h, self._err, ERR_LEN)))

    def count(self) -> int:
        return int(self._check(lib().lbh_index_count(self._h, self._err, ERR_LEN)))

    def create_ann_index(self, num_partitions: int, num_sub_vectors: int):
        self._check(lib().lbh_index_create_ann(self._h, num_partitions, num_sub_vectors, self._err, ERR_LEN))

    def search(self, query, k: int, predicate: str = ""):
        q = np.ascontiguousarray(query, np.float32)
        ids = np.empty(max(k, 1), np.int64)
        d = np.empty(max(k, 1), np.float32)
        n = self._check(lib().lbh_index_search(self._h, _p(q), q.shape[0], k, predicate.encode(), _p(ids), _p(d), self._err,
                                               ERR_LEN))
        return ids[:n].copy(), d[:n].copy()

    def search_batch(self, queries, k: int, predicate: str = ""):
        Q = np.ascontiguousarray(queries, np.float32)
        nq = Q.shape[0]
        ids = np.full((nq, k), -1, np.int64)
        d = np.full((nq, k), np.inf, np.float32)
        c = np.zeros(nq, np.int32)
        self._check(lib().lbh_index_search_batch(self._h, _p(Q), nq, Q.shape[1], k, predicate.encode(), _p(ids), _p(d), _p(c),
                                                 self._err, ERR_LEN))
        return ids, d, c

    def lance_search(self, query, k: int):
        """SELECT row_id, distance FROM lance_search(table, index, query, k) -> (row_ids, distances, scan_chunks)."""
        q = np.ascontiguousarray(query, np.float32)
        ids = np.empty(max(k, 1), np.int64)
        d = np.empty(max(k, 1), np.float32)
        chunks = C.c_int32(0)
        n = self._check(lib().lbh_lance_search(self._h, _p(q), q.shape[0], k, _p(ids), _p(d), C.addressof(chunks), self._err,
                                               ERR_LEN))
        return ids[:n].copy(), d[:n].copy(), chunks.value

    def lance_index_scan(self, query, limit: int, predicate: str = ""):
        """The rewritten `.. [WHERE p] ORDER BY array_distance(v, q) LIMIT k`: row ids in result order."""
        q = np.ascontiguousarray(query, np.float32)
        ids = np.empty(max(limit, 1), np.int64)
        n = self._check(lib().lbh_lance_index_scan(self._h, _p(q), q.shape[0], limit, predicate.encode(), _p(ids), self._err,
                                                   ERR_LEN))
        return ids[:n].copy()
