"""ctypes harness over liblance_b200.so — the C ABI declared in include/lance_b200.h.

The product is the shared library (CUDA kernels + C++ host code under csrc/); this module only
exists so that tests/, bench.py and __graft_entry__.py can drive the very entry points the DuckDB
extension's C++ side binds (/root/reference/src/rust_ffi.cpp:9-41).  It never computes anything
itself and never imports oracle/: if the library or a CUDA device is missing, calls raise.

The directory name carries a hyphen, so load it with `__graft_entry__.load_package()` (importlib),
which registers it as module `duckdb_lancedb_b200`.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liblance_b200.so")
CSRC = os.path.join(_HERE, "csrc")
HEADER = os.path.join(os.path.dirname(_HERE), "include", "lance_b200.h")

ERR_LEN = 2048  # src/rust_ffi.cpp:46 uses a 2048-byte stack buffer


class LanceError(IOError):
    """Mirrors duckdb::IOException("Lance <op>: ...") thrown by src/rust_ffi.cpp:44-206."""


def build(force: bool = False) -> str:
    """Compile the CUDA library for sm_100a (nvcc cross-compiles without a GPU)."""
    if force:
        subprocess.run(["make", "-C", CSRC, "clean"], check=True, capture_output=True)
    r = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True, env={**os.environ, "MAKEFLAGS": ""})
    if r.returncode != 0:
        raise RuntimeError("building liblance_b200.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    return LIB_PATH


class IvfPqInfo(C.Structure):
    _fields_ = [("metric", C.c_int32), ("dim", C.c_int32), ("P", C.c_int32), ("M", C.c_int32),
                ("n_indexed", C.c_int64)]


class Stats(C.Structure):
    _fields_ = [("kernel_launches", C.c_int64), ("searches", C.c_int64), ("flat_simt_batches", C.c_int64),
                ("flat_tc_batches", C.c_int64), ("tc_fallback_queries", C.c_int64), ("ivfpq_batches", C.c_int64),
                ("rows", C.c_int64), ("live_rows", C.c_int64), ("device", C.c_int32), ("has_index", C.c_int32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class Profile(C.Structure):
    _fields_ = [("flat_scan_ms", C.c_double), ("flat_scan_launches", C.c_int64), ("flat_gemm_ms", C.c_double),
                ("flat_gemm_launches", C.c_int64), ("adc_scan_ms", C.c_double), ("adc_scan_launches", C.c_int64),
                ("rerank_ms", C.c_double), ("rerank_launches", C.c_int64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


_lib = None
_vp, _cp, _i32, _i64 = C.c_void_p, C.c_char_p, C.c_int32, C.c_int64
_ERR = [C.c_char_p, C.c_int]

# name -> (restype, argtypes): one entry per declaration in include/lance_b200.h
SIGNATURES = {
    "lance_create_detached": (_vp, [_cp, _i32, _cp, _cp] + _ERR),
    "lance_create_detached_from_arrow": (_vp, [_cp, _vp, _cp, _cp] + _ERR),
    "lance_open_detached": (_vp, [_cp, _cp, _cp] + _ERR),
    "lance_free_detached": (None, [_vp]),
    "lance_detached_has_extra_columns": (_i32, [_vp]),
    "lance_detached_dimension": (_i32, [_vp]),
    "lance_detached_add": (_i64, [_vp, _vp, _i32] + _ERR),
    "lance_detached_add_batch": (_i32, [_vp, _vp, _i32, _i32, _vp] + _ERR),
    "lance_detached_add_batch_arrow": (_i32, [_vp, _vp, _vp, _vp] + _ERR),
    "lance_detached_merge": (_i32, [_vp, _vp, _vp, _i32, _vp, _vp] + _ERR),
    "lance_detached_search": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp] + _ERR),
    "lance_detached_search_predicate": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _cp, _vp, _vp] + _ERR),
    "lance_detached_search_batch_predicate": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _i32, _cp, _vp, _vp, _vp] + _ERR),
    "lance_detached_create_scalar_index": (_i32, [_vp, _cp, _cp] + _ERR),
    "lance_b200_set_column": (_i32, [_vp, _cp, _i32, _i64, _i64, _vp, _vp, _vp] + _ERR),
    "lance_b200_predicate_bitmap": (_i64, [_vp, _cp, _vp, _i64, _vp, _vp] + _ERR),
    "lance_b200_flat_search_noalloc": (_i32, [_i32, _vp, _i64, _i32, _i32, _vp, _i32, _i64, _vp, _i64, _vp, _vp, _vp, _vp] + _ERR),
    "lance_b200_add_batch_labeled": (_i32, [_vp, _vp, _i64, _i32, _i64] + _ERR),
    "lance_b200_add_batch_device_labels": (_i32, [_vp, _vp, _i64, _i32, _vp, _vp] + _ERR),
    "lance_b200_create_sharded": (_vp, [_vp, _i32, _i32, _cp] + _ERR),
    "lance_b200_create_sharded_at": (_vp, [_cp, _cp, _vp, _i32, _i32, _cp] + _ERR),
    "lance_b200_open_sharded": (_vp, [_cp, _cp, _cp, _vp, _i32] + _ERR),
    "lance_b200_sharded_rows": (_i32, [_vp, _vp, _i32]),
    "lance_detached_count": (_i64, [_vp] + _ERR),
    "lance_detached_delete": (_i32, [_vp, _i64] + _ERR),
    "lance_detached_delete_batch": (_i32, [_vp, _vp, _i32] + _ERR),
    "lance_detached_create_index": (_i32, [_vp, _i32, _i32] + _ERR),
    "lance_detached_create_hnsw_index": (_i32, [_vp, _i32, _i32] + _ERR),
    "lance_detached_compact": (_i32, [_vp] + _ERR),
    "lance_detached_get_vector": (_i32, [_vp, _i64, _vp, _i32] + _ERR),
    "lance_detached_get_all_vectors": (_i32, [_vp, _vp, _vp, _vp] + _ERR),
    "lance_detached_search_filtered": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _i64, _vp, _vp] + _ERR),
    "lance_detached_search_batch": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _i32, _vp, _i64, _vp, _vp, _vp] + _ERR),
    "lance_b200_search_device": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _i64, _vp, _vp, _vp, _vp] + _ERR),
    "lance_b200_search_device_async": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _i64, _vp, _vp, _vp, _vp, _vp] + _ERR),
    "lance_b200_search_finish": (_i32, [_vp, _vp] + _ERR),
    "lance_b200_merge_topk_device": (_i32, [_i32, _vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp] + _ERR),
    "lance_b200_merge_topk_strided_device": (_i32, [_i32, _vp, _vp, _vp, _i64, _i64, _i64, _i32, _i32, _i32, _vp, _vp, _vp, _vp] + _ERR),
    "lance_b200_set_label_base": (_i32, [_vp, _i64] + _ERR),
    "lance_b200_reserve": (_i32, [_vp, _i64] + _ERR),
    "lance_b200_add_batch_device": (_i64, [_vp, _vp, _i64, _i32, _vp] + _ERR),
    "lance_b200_set_ivfpq": (_i32, [_vp, _i32, _i32, _i64, _vp, _vp, _vp, _vp, _vp] + _ERR),
    "lance_b200_get_ivfpq": (_i32, [_vp, C.POINTER(IvfPqInfo), _vp, _vp, _vp, _vp, _vp] + _ERR),
    "lance_b200_drop_index": (_i32, [_vp] + _ERR),
    "lance_b200_ivfpq_search_debug": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp] + _ERR),
    "lance_b200_save": (_i32, [_vp] + _ERR),
    "lance_b200_stats": (_i32, [_vp, C.POINTER(Stats)]),
    "lance_b200_profile_enable": (_i32, [_vp, _i32]),
    "lance_b200_profile_read": (_i32, [_vp, C.POINTER(Profile)]),
    "lance_b200_time_search_device": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _i64, _vp, _vp, _vp, _vp, _i32, _vp, _i64,
                                             _vp] + _ERR),
    "lance_b200_set_flat_path": (_i32, [_vp, _i32]),
    "lance_b200_set_default_device": (_i32, [_i32]),
    "lance_b200_device_count": (_i32, []),
    "lance_b200_version": (_cp, []),
}


def lib():
    """dlopen the library (never builds implicitly on the GPU box: a missing .so is an error)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise LanceError(f"{LIB_PATH} is missing: run __graft_entry__.build() (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    return C.c_void_p(int(a))  # a raw device/host address (e.g. torch.Tensor.data_ptr())


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def make_bitmap(mask) -> np.ndarray:
    """bool[n] -> uint32 words, bit (i & 31) of word (i >> 5): the layout of the allow_bitmap arguments."""
    mask = np.asarray(mask, dtype=bool)
    n = mask.shape[0]
    padded = np.zeros(((n + 31) // 32) * 32, dtype=bool)
    padded[:n] = mask
    if padded.size == 0:
        return np.zeros(1, np.uint32)
    return np.packbits(padded.reshape(-1, 32), axis=1, bitorder="little").view(np.uint32).reshape(-1).copy()


class LanceHandle:
    """One `lance_*` handle.  Method names follow the wrappers of src/include/rust_ffi.hpp:9-58."""

    def __init__(self, dimension: int, metric: str = "l2", db_path: str = "", table_name: str = "vectors", _raw=None):
        self._err = C.create_string_buffer(ERR_LEN)
        self.metric = metric
        if _raw is not None:
            self._h = _raw
        else:
            self._h = lib().lance_create_detached(db_path.encode(), int(dimension), metric.encode(), table_name.encode(),
                                                  self._err, ERR_LEN)
            if not self._h:
                raise LanceError("Lance create: " + self._err.value.decode())
        self.dim = int(lib().lance_detached_dimension(self._h))

    @classmethod
    def open(cls, db_path: str, table_name: str = "vectors", metric: str = "l2"):
        err = C.create_string_buffer(ERR_LEN)
        h = lib().lance_open_detached(db_path.encode(), table_name.encode(), metric.encode(), err, ERR_LEN)
        if not h:
            raise LanceError("Lance open: " + err.value.decode())
        return cls(0, metric, _raw=h)

    @classmethod
    def create_sharded(cls, devices, dimension: int, metric: str = "l2", db_path: str | None = None, table: str = "vectors"):
        """lance_b200_create_sharded[_at]: one handle over several devices of the box, same calls as a plain handle
        (db_path given: it can be saved and reopened with open_sharded)."""
        err = C.create_string_buffer(ERR_LEN)
        dv = np.ascontiguousarray(devices, np.int32)
        if db_path is None:
            h = lib().lance_b200_create_sharded(_ptr(dv), dv.shape[0], int(dimension), metric.encode(), err, ERR_LEN)
        else:
            h = lib().lance_b200_create_sharded_at(str(db_path).encode(), table.encode(), _ptr(dv), dv.shape[0], int(dimension),
                                                   metric.encode(), err, ERR_LEN)
        if not h:
            raise LanceError("Lance create_sharded: " + err.value.decode())
        return cls(0, metric, _raw=h)

    @classmethod
    def open_sharded(cls, db_path: str, devices, metric: str = "l2", table: str = "vectors"):
        """lance_b200_open_sharded: reopens what save() of a create_sharded(..., db_path=...) handle wrote."""
        err = C.create_string_buffer(ERR_LEN)
        dv = np.ascontiguousarray(devices, np.int32)
        h = lib().lance_b200_open_sharded(str(db_path).encode(), table.encode(), metric.encode(), _ptr(dv), dv.shape[0], err, ERR_LEN)
        if not h:
            raise LanceError("Lance open_sharded: " + err.value.decode())
        return cls(0, metric, _raw=h)

    def shard_rows(self) -> np.ndarray:
        out = np.zeros(64, np.int64)
        n = lib().lance_b200_sharded_rows(self._h, _ptr(out), 64)
        if n < 0:
            raise LanceError("not a sharded handle")
        return out[:n].copy()

    def close(self):
        if getattr(self, "_h", None):
            lib().lance_free_detached(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc, op):
        if rc is None or rc < 0:
            raise LanceError(f"Lance {op}: " + self._err.value.decode())
        return rc

    # ---- ingest / lifecycle
    def add_batch(self, vectors) -> np.ndarray:
        v = _f32(vectors)
        if v.ndim == 1:
            v = v.reshape(1, -1)
        labels = np.empty(v.shape[0], np.int64)
        self._check(lib().lance_detached_add_batch(self._h, _ptr(v), v.shape[0], v.shape[1], _ptr(labels), self._err,
                                                   ERR_LEN), "add_batch")
        return labels

    def add_batch_labeled(self, vectors, first_label: int):
        """lance_b200_add_batch_labeled: rows with the caller's labels first_label .. first_label + n - 1."""
        v = _f32(vectors)
        self._check(lib().lance_b200_add_batch_labeled(self._h, _ptr(v), v.shape[0], v.shape[1], int(first_label), self._err,
                                                       ERR_LEN), "add_batch")

    def add_batch_device_labels(self, dev_ptr: int, num: int, dim: int, labels, stream: int = 0):
        """lance_b200_add_batch_device_labels: device rows with one (ascending) label each."""
        l = np.ascontiguousarray(labels, np.int64)
        assert l.shape[0] == num
        self._check(lib().lance_b200_add_batch_device_labels(self._h, _ptr(dev_ptr), num, dim, _ptr(l),
                                                             _ptr(stream) if stream else None, self._err, ERR_LEN), "add_batch")

    def add(self, vector) -> int:
        v = _f32(vector)
        return int(self._check(lib().lance_detached_add(self._h, _ptr(v), v.shape[0], self._err, ERR_LEN), "add"))

    def add_batch_device(self, dev_ptr: int, num: int, dim: int, stream: int = 0) -> int:
        return int(self._check(lib().lance_b200_add_batch_device(self._h, _ptr(dev_ptr), num, dim, _ptr(stream),
                                                                 self._err, ERR_LEN), "add_batch"))

    def reserve(self, rows: int):
        self._check(lib().lance_b200_reserve(self._h, rows, self._err, ERR_LEN), "reserve")

    def set_label_base(self, base: int):
        self._check(lib().lance_b200_set_label_base(self._h, base, self._err, ERR_LEN), "set_label_base")

    def count(self) -> int:
        return int(self._check(lib().lance_detached_count(self._h, self._err, ERR_LEN), "count"))

    def delete_batch(self, labels):
        l = np.ascontiguousarray(labels, np.int64)
        self._check(lib().lance_detached_delete_batch(self._h, _ptr(l), l.shape[0], self._err, ERR_LEN), "delete_batch")

    def delete(self, label: int):
        self._check(lib().lance_detached_delete(self._h, int(label), self._err, ERR_LEN), "delete")

    def merge(self, source: "LanceHandle", live_source_labels):
        """lance_detached_merge (src/rust_ffi.cpp:119-128): append the listed live rows of `source` under fresh labels;
        returns (old_labels, new_labels)."""
        l = np.ascontiguousarray(live_source_labels, np.int64)
        old = np.empty(max(l.shape[0], 1), np.int64)
        new = np.empty(max(l.shape[0], 1), np.int64)
        n = self._check(lib().lance_detached_merge(self._h, source._h, _ptr(l), l.shape[0], _ptr(old), _ptr(new),
                                                   self._err, ERR_LEN), "merge")
        return old[:n], new[:n]

    def get_vector(self, label: int) -> np.ndarray:
        out = np.empty(self.dim, np.float32)
        self._check(lib().lance_detached_get_vector(self._h, int(label), _ptr(out), self.dim, self._err, ERR_LEN),
                    "get_vector")
        return out

    def get_all_vectors(self):
        cnt = C.c_int64(0)
        self._check(lib().lance_detached_get_all_vectors(self._h, None, None, C.addressof(cnt), self._err, ERR_LEN),
                    "get_all_vectors")
        labels = np.empty(cnt.value, np.int64)
        vecs = np.empty((cnt.value, self.dim), np.float32)
        self._check(lib().lance_detached_get_all_vectors(self._h, _ptr(labels), _ptr(vecs), C.addressof(cnt), self._err,
                                                         ERR_LEN), "get_all_vectors")
        return labels, vecs

    def save(self):
        self._check(lib().lance_b200_save(self._h, self._err, ERR_LEN), "save")

    def compact(self):
        self._check(lib().lance_detached_compact(self._h, self._err, ERR_LEN), "compact")

    # ---- extra columns + predicate pushdown (include/lance_b200.h, "extra columns")
    def set_column(self, name: str, values, first_label: int = 0, valid=None, type: str | None = None):
        """lance_b200_set_column: values for the rows of labels [first_label, first_label + len(values))."""
        vb = None if valid is None else np.ascontiguousarray(np.asarray(valid, bool).astype(np.uint8))
        if type == "str":
            enc = [b"" if v is None else (v.encode() if isinstance(v, str) else bytes(v)) for v in values]
            if valid is None and any(v is None for v in values):
                vb = np.array([v is not None for v in values], np.uint8)
            off = np.zeros(len(enc) + 1, np.int64)
            if enc:
                off[1:] = np.cumsum([len(e) for e in enc])
            data = np.frombuffer(b"".join(enc) + b"\0", np.uint8).copy()
            rc = lib().lance_b200_set_column(self._h, name.encode(), 3, int(first_label), len(enc), _ptr(data), _ptr(off),
                                             _ptr(vb), self._err, ERR_LEN)
            return self._check(rc, "set_column")
        a = np.asarray(values)
        if type == "bool" or (type is None and a.dtype == np.bool_):
            t, a = 2, np.ascontiguousarray(a.astype(np.uint8))
        elif type == "double" or (type is None and a.dtype.kind == "f"):
            t, a = 1, np.ascontiguousarray(a, np.float64)
        else:
            t, a = 0, np.ascontiguousarray(a, np.int64)
        rc = lib().lance_b200_set_column(self._h, name.encode(), t, int(first_label), a.shape[0], _ptr(a), None, _ptr(vb),
                                         self._err, ERR_LEN)
        return self._check(rc, "set_column")

    def has_extra_columns(self) -> bool:
        return bool(lib().lance_detached_has_extra_columns(self._h))

    def create_scalar_index(self, column: str, index_type: str = "BTREE"):
        self._check(lib().lance_detached_create_scalar_index(self._h, column.encode(), index_type.encode(), self._err,
                                                             ERR_LEN), "create_scalar_index")

    def add_batch_arrow(self, schema_ptr: int, array_ptr: int, n: int) -> np.ndarray:
        """lance_detached_add_batch_arrow over exported ArrowSchema / ArrowArray structs (consumes the array)."""
        labels = np.empty(max(n, 1), np.int64)
        c = self._check(lib().lance_detached_add_batch_arrow(self._h, C.c_void_p(schema_ptr), C.c_void_p(array_ptr),
                                                             _ptr(labels), self._err, ERR_LEN), "add_batch_arrow")
        return labels[:c].copy()

    @classmethod
    def create_from_arrow(cls, schema_ptr: int, metric: str = "l2", db_path: str = "", table_name: str = "vectors"):
        err = C.create_string_buffer(ERR_LEN)
        h = lib().lance_create_detached_from_arrow(db_path.encode(), C.c_void_p(schema_ptr), metric.encode(),
                                                   table_name.encode(), err, ERR_LEN)
        if not h:
            raise LanceError("Lance create_from_arrow: " + err.value.decode())
        return cls(0, metric, _raw=h)

    def predicate_bitmap(self, predicate: str, words: int | None = None):
        """Returns (host copy of the allow bitmap uint32[words], rows covered, device pointer, was_cached)."""
        w = int(words) if words is not None else (int(self.stats()["rows"]) + 31) // 32 + 1
        out = np.zeros(max(w, 1), np.uint32)
        dptr = C.c_void_p(0)
        cached = C.c_int32(0)
        bits = lib().lance_b200_predicate_bitmap(self._h, predicate.encode(), _ptr(out), out.shape[0], C.byref(dptr),
                                                 C.byref(cached), self._err, ERR_LEN)
        self._check(bits, "predicate")
        return out, int(bits), int(dptr.value or 0), bool(cached.value)

    def search_predicate(self, query, k: int, predicate: str | None, nprobes: int = 20, refine_factor: int = 1):
        """The newer revision's LanceDetachedSearch(.., predicate, ..) (src/lance_index.cpp:452-453)."""
        q = _f32(query)
        kk = max(int(k), 1)
        labels = np.empty(kk, np.int64)
        dist = np.empty(kk, np.float32)
        c = lib().lance_detached_search_predicate(self._h, _ptr(q), q.shape[0], int(k), nprobes, refine_factor,
                                                  None if predicate is None else predicate.encode(), _ptr(labels),
                                                  _ptr(dist), self._err, ERR_LEN)
        self._check(c, "search")
        return labels[:c].copy(), dist[:c].copy()

    def search_batch_predicate(self, queries, k: int, predicate: str | None, nprobes: int = 20, refine_factor: int = 1):
        Q = _f32(queries)
        nq, d = Q.shape
        kk = max(int(k), 1)
        labels = np.empty((nq, kk), np.int64)
        dist = np.empty((nq, kk), np.float32)
        counts = np.empty(nq, np.int32)
        rc = lib().lance_detached_search_batch_predicate(self._h, _ptr(Q), nq, d, int(k), nprobes, refine_factor,
                                                         None if predicate is None else predicate.encode(), _ptr(labels),
                                                         _ptr(dist), _ptr(counts), self._err, ERR_LEN)
        self._check(rc, "search")
        return labels, dist, counts

    # ---- search
    def search(self, query, k: int, nprobes: int = 20, refine_factor: int = 1, allow=None):
        """LanceDetachedSearch (src/rust_ffi.cpp:130-139): returns (labels int64[c], distances float32[c])."""
        q = _f32(query)
        kk = max(int(k), 1)
        labels = np.empty(kk, np.int64)
        dist = np.empty(kk, np.float32)
        if allow is None:
            c = lib().lance_detached_search(self._h, _ptr(q), q.shape[0], int(k), nprobes, refine_factor, _ptr(labels),
                                            _ptr(dist), self._err, ERR_LEN)
        else:
            bm = np.ascontiguousarray(allow, np.uint32)
            c = lib().lance_detached_search_filtered(self._h, _ptr(q), q.shape[0], int(k), nprobes, refine_factor,
                                                     _ptr(bm), bm.shape[0] * 32, _ptr(labels), _ptr(dist), self._err,
                                                     ERR_LEN)
        self._check(c, "search")
        return labels[:c].copy(), dist[:c].copy()

    def search_batch(self, queries, k: int, nprobes: int = 20, refine_factor: int = 1, allow=None, out=None):
        """Returns (labels[nq,k], distances[nq,k], counts[nq]).  `out` = preallocated (labels, dist, counts)."""
        Q = queries if (isinstance(queries, np.ndarray) and queries.dtype == np.float32 and queries.flags.c_contiguous) \
            else _f32(queries)
        nq, d = Q.shape
        kk = max(int(k), 1)
        if out is None:
            labels = np.empty((nq, kk), np.int64)
            dist = np.empty((nq, kk), np.float32)
            counts = np.empty(nq, np.int32)
        else:
            labels, dist, counts = out
        bm = None if allow is None else np.ascontiguousarray(allow, np.uint32)
        rc = lib().lance_detached_search_batch(self._h, _ptr(Q), nq, d, int(k), nprobes, refine_factor, _ptr(bm),
                                               0 if bm is None else bm.shape[0] * 32, _ptr(labels), _ptr(dist),
                                               _ptr(counts), self._err, ERR_LEN)
        self._check(rc, "search")
        return labels, dist, counts

    def bind_search_batch(self, Q, k: int, out, nprobes: int = 20, refine_factor: int = 1, allow=None, predicate=None):
        """lance_detached_search_batch with every argument converted ONCE: returns a zero-argument callable that makes
        exactly the C call (bench.py's end-to-end loop: the per-call cost of building ctypes arguments in Python is
        harness time, not the library's).  Q / out / allow must stay alive and unchanged in place."""
        assert isinstance(Q, np.ndarray) and Q.dtype == np.float32 and Q.flags.c_contiguous
        labels, dist, counts = out
        bm = None if allow is None else np.ascontiguousarray(allow, np.uint32)
        fn = lib().lance_detached_search_batch
        args = (self._h, _ptr(Q), C.c_int32(Q.shape[0]), C.c_int32(Q.shape[1]), C.c_int32(int(k)), C.c_int32(nprobes),
                C.c_int32(refine_factor), _ptr(bm), C.c_int64(0 if bm is None else bm.shape[0] * 32), _ptr(labels), _ptr(dist),
                _ptr(counts), self._err, C.c_int(ERR_LEN))
        if predicate is not None:  # lance_detached_search_batch_predicate: the WHERE as a string, evaluated on the device
            fn = lib().lance_detached_search_batch_predicate
            pred = C.c_char_p(predicate.encode())
            args = (self._h, _ptr(Q), C.c_int32(Q.shape[0]), C.c_int32(Q.shape[1]), C.c_int32(int(k)), C.c_int32(nprobes),
                    C.c_int32(refine_factor), pred, _ptr(labels), _ptr(dist), _ptr(counts), self._err, C.c_int(ERR_LEN))
        keep = (Q, labels, dist, counts, bm)

        def call():
            if fn(*args) < 0:
                raise LanceError("Lance search: " + self._err.value.decode())
            return keep[1], keep[2], keep[3]

        return call

    def bind_search_device(self, d_queries: int, nq: int, k: int, d_labels: int, d_dist: int, d_counts: int,
                           nprobes: int = 20, refine_factor: int = 1, d_allow: int = 0, allow_bits: int = 0, stream: int = 0):
        """lance_b200_search_device with the arguments converted once (see bind_search_batch)."""
        fn = lib().lance_b200_search_device
        args = (self._h, _ptr(d_queries), C.c_int32(nq), C.c_int32(k), C.c_int32(nprobes), C.c_int32(refine_factor),
                _ptr(d_allow) if d_allow else None, C.c_int64(allow_bits), _ptr(d_labels), _ptr(d_dist), _ptr(d_counts),
                _ptr(stream) if stream else None, self._err, C.c_int(ERR_LEN))

        def call():
            if fn(*args) < 0:
                raise LanceError("Lance search: " + self._err.value.decode())

        return call

    def time_search_device(self, d_queries: int, nq: int, k: int, d_labels: int, d_dist: int, d_counts: int, iters: int,
                           nprobes: int = 20, refine_factor: int = 1, d_allow: int = 0, allow_bits: int = 0, stream: int = 0,
                           d_flush: int = 0, flush_bytes: int = 0) -> np.ndarray:
        """lance_b200_time_search_device: per-call device milliseconds of `iters` searches, events recorded in C."""
        out = np.zeros(iters, np.float32)
        rc = lib().lance_b200_time_search_device(self._h, _ptr(d_queries), nq, k, nprobes, refine_factor,
                                                 _ptr(d_allow) if d_allow else None, allow_bits, _ptr(d_labels), _ptr(d_dist),
                                                 _ptr(d_counts), _ptr(stream) if stream else None, iters,
                                                 _ptr(d_flush) if d_flush else None, flush_bytes, _ptr(out), self._err, ERR_LEN)
        self._check(rc, "timed search")
        return out

    def search_device(self, d_queries: int, nq: int, k: int, d_labels: int, d_dist: int, d_counts: int,
                      nprobes: int = 20, refine_factor: int = 1, d_allow: int = 0, allow_bits: int = 0, stream: int = 0):
        rc = lib().lance_b200_search_device(self._h, _ptr(d_queries), nq, k, nprobes, refine_factor,
                                            _ptr(d_allow) if d_allow else None, allow_bits, _ptr(d_labels), _ptr(d_dist),
                                            _ptr(d_counts), _ptr(stream) if stream else None, self._err, ERR_LEN)
        self._check(rc, "search")

    def search_device_async(self, d_queries: int, nq: int, k: int, d_labels: int, d_dist: int, d_counts: int,
                            nprobes: int = 20, refine_factor: int = 1, d_allow: int = 0, allow_bits: int = 0, stream: int = 0,
                            d_overflow: int = 0):
        """search_device without its host synchronisation; pair with search_finish (include/lance_b200.h)."""
        rc = lib().lance_b200_search_device_async(self._h, _ptr(d_queries), nq, k, nprobes, refine_factor,
                                                  _ptr(d_allow) if d_allow else None, allow_bits, _ptr(d_labels),
                                                  _ptr(d_dist), _ptr(d_counts), _ptr(d_overflow) if d_overflow else None,
                                                  _ptr(stream) if stream else None, self._err, ERR_LEN)
        self._check(rc, "search")

    def search_finish(self, stream: int = 0) -> int:
        """Waits for the stream and settles the pending overflow check; returns the number of re-run queries."""
        rc = lib().lance_b200_search_finish(self._h, _ptr(stream) if stream else None, self._err, ERR_LEN)
        if rc < 0:
            self._check(rc, "search")
        return int(rc)

    # ---- IVF_PQ
    def create_index(self, num_partitions: int = 0, num_sub_vectors: int = 0):
        self._check(lib().lance_detached_create_index(self._h, num_partitions, num_sub_vectors, self._err, ERR_LEN),
                    "create_index")

    def create_hnsw_index(self, m: int = 16, ef: int = 100):
        self._check(lib().lance_detached_create_hnsw_index(self._h, m, ef, self._err, ERR_LEN), "create_hnsw_index")

    def drop_index(self):
        self._check(lib().lance_b200_drop_index(self._h, self._err, ERR_LEN), "drop_index")

    def set_ivfpq(self, centroids, codebook, part_offsets, row_ids, codes):
        cen = _f32(centroids)
        P = cen.shape[0]
        codes = np.ascontiguousarray(codes, np.uint8)
        M = codes.shape[1]
        cb = _f32(codebook)
        off = np.ascontiguousarray(part_offsets, np.int64)
        ids = np.ascontiguousarray(row_ids, np.int64)
        self._check(lib().lance_b200_set_ivfpq(self._h, P, M, ids.shape[0], _ptr(cen), _ptr(cb), _ptr(off), _ptr(ids),
                                               _ptr(codes), self._err, ERR_LEN), "set_ivfpq")

    def get_ivfpq(self):
        info = IvfPqInfo()
        self._check(lib().lance_b200_get_ivfpq(self._h, C.byref(info), None, None, None, None, None, self._err, ERR_LEN),
                    "get_ivfpq")
        cen = np.empty((info.P, info.dim), np.float32)
        cb = np.empty((info.M, 256, info.dim // info.M), np.float32)
        off = np.empty(info.P + 1, np.int64)
        ids = np.empty(info.n_indexed, np.int64)
        codes = np.empty((info.n_indexed, info.M), np.uint8)
        self._check(lib().lance_b200_get_ivfpq(self._h, C.byref(info), _ptr(cen), _ptr(cb), _ptr(off), _ptr(ids),
                                               _ptr(codes), self._err, ERR_LEN), "get_ivfpq")
        return dict(metric=info.metric, dim=info.dim, P=info.P, M=info.M, n_indexed=info.n_indexed, centroids=cen,
                    codebook=cb, part_offsets=off, row_ids=ids, codes=codes)

    def ivfpq_part_offsets(self) -> np.ndarray:
        """part_offsets[P+1] only (no bulk export)."""
        info = IvfPqInfo()
        self._check(lib().lance_b200_get_ivfpq(self._h, C.byref(info), None, None, None, None, None, self._err, ERR_LEN),
                    "get_ivfpq")
        off = np.empty(info.P + 1, np.int64)
        self._check(lib().lance_b200_get_ivfpq(self._h, C.byref(info), None, None, _ptr(off), None, None, self._err,
                                               ERR_LEN), "get_ivfpq")
        return off

    def ivfpq_search_debug(self, query, k, nprobes, refine_factor=1, allow=None):
        q = _f32(query)
        rf = max(int(refine_factor), 1)
        parts = np.full(max(nprobes, 1), -1, np.int32)
        cl = np.empty(k * rf, np.int64)
        ca = np.empty(k * rf, np.float32)
        nc = C.c_int32(0)
        labels = np.empty(k, np.int64)
        dist = np.empty(k, np.float32)
        bm = None if allow is None else np.ascontiguousarray(allow, np.uint32)
        c = lib().lance_b200_ivfpq_search_debug(self._h, _ptr(q), q.shape[0], k, nprobes, refine_factor, _ptr(bm),
                                                0 if bm is None else bm.shape[0] * 32, _ptr(parts), _ptr(cl), _ptr(ca),
                                                C.addressof(nc), _ptr(labels), _ptr(dist), self._err, ERR_LEN)
        self._check(c, "search")
        return dict(parts=parts[parts >= 0], cand_labels=cl[:nc.value].copy(), cand_adc=ca[:nc.value].copy(),
                    labels=labels[:c].copy(), dist=dist[:c].copy())

    # ---- introspection / hooks
    def stats(self) -> dict:
        s = Stats()
        lib().lance_b200_stats(self._h, C.byref(s))
        return s.as_dict()

    def profile_enable(self, on: bool = True):
        lib().lance_b200_profile_enable(self._h, 1 if on else 0)

    def profile_read(self) -> dict:
        p = Profile()
        lib().lance_b200_profile_read(self._h, C.byref(p))
        return p.as_dict()

    def set_flat_path(self, mode: int):
        """0 auto, 1 fp32 streaming scan, 2 tcgen05 candidate GEMM + re-rank."""
        if lib().lance_b200_set_flat_path(self._h, mode) != 0:
            raise LanceError("Lance set_flat_path: bad mode")


def merge_topk_device(device: int, d_labels: int, d_dists: int, d_counts: int, G: int, nq: int, k: int,
                      d_out_labels: int, d_out_dists: int, d_out_counts: int, stream: int = 0):
    err = C.create_string_buffer(ERR_LEN)
    rc = lib().lance_b200_merge_topk_device(device, _ptr(d_labels), _ptr(d_dists), _ptr(d_counts), G, nq, k,
                                            _ptr(d_out_labels), _ptr(d_out_dists), _ptr(d_out_counts),
                                            _ptr(stream) if stream else None, err, ERR_LEN)
    if rc != 0:
        raise LanceError("Lance merge: " + err.value.decode())


def merge_topk_strided_device(device: int, d_labels: int, d_dists: int, d_counts: int, ls: int, ds: int, cs: int, G: int,
                              nq: int, k: int, d_out_labels: int, d_out_dists: int, d_out_counts: int, stream: int = 0):
    """Final select over per-rank packed records (element strides ls/ds/cs between ranks); one launch."""
    err = C.create_string_buffer(ERR_LEN)
    rc = lib().lance_b200_merge_topk_strided_device(device, _ptr(d_labels), _ptr(d_dists), _ptr(d_counts), ls, ds, cs, G,
                                                    nq, k, _ptr(d_out_labels), _ptr(d_out_dists), _ptr(d_out_counts),
                                                    _ptr(stream) if stream else None, err, ERR_LEN)
    if rc != 0:
        raise LanceError("Lance merge: " + err.value.decode())


def flat_search_noalloc(device: int, d_vectors: int, rows: int, dim: int, metric: int, d_query: int, k: int, d_labels: int,
                        d_dist: int, d_count: int, label_base: int = 0, d_allow: int = 0, allow_bits: int = 0, stream: int = 0):
    """lance_b200_flat_search_noalloc: device pointers in and out, enqueued on `stream`, no allocation (static scratch)."""
    err = C.create_string_buffer(ERR_LEN)
    rc = lib().lance_b200_flat_search_noalloc(device, _ptr(d_vectors) if d_vectors else None, rows, dim, metric, _ptr(d_query), k,
                                              label_base, _ptr(d_allow) if d_allow else None, allow_bits, _ptr(d_labels),
                                              _ptr(d_dist), _ptr(d_count), _ptr(stream) if stream else None, err, ERR_LEN)
    if rc < 0:
        raise LanceError("Lance " + err.value.decode())


def device_count() -> int:
    return int(lib().lance_b200_device_count())


def set_default_device(device: int):
    lib().lance_b200_set_default_device(int(device))
