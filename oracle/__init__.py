"""ctypes front-end of oracle/liboracle.so (and oracle/_ref when built).

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this.
See oracle/oracle.h for the parity-pinning status of each function.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_REF_PATH = os.path.join(_HERE, "_ref", "libref_array_kernels.so")

L2, COSINE, DOT = 0, 1, 2
METRICS = {"l2": L2, "cosine": COSINE, "dot": DOT, "ip": DOT}


def build(force: bool = False) -> None:
    """Compile the checker (and oracle/_ref when /root/reference is present)."""
    src_newer = (
        not os.path.exists(_LIB_PATH)
        or os.path.getmtime(_LIB_PATH) < max(os.path.getmtime(os.path.join(_HERE, f)) for f in ("oracle.c", "oracle.h"))
    )
    if force or src_newer or (os.path.isdir("/root/reference") and not os.path.exists(_REF_PATH)):
        subprocess.run(["make", "-C", _HERE, "-s"], check=True, env={**os.environ, "MAKEFLAGS": ""})


_lib = None
_ref = None

_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


class _IvfPq(C.Structure):
    _fields_ = [
        ("metric", C.c_int32),
        ("dim", C.c_int32),
        ("P", C.c_int32),
        ("M", C.c_int32),
        ("n_indexed", C.c_int64),
        ("centroids", C.c_void_p),
        ("codebook", C.c_void_p),
        ("part_offsets", C.c_void_p),
        ("row_ids", C.c_void_p),
        ("codes", C.c_void_p),
    ]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.oracle_distance_f32.restype = C.c_float
        _lib.oracle_distance_f32.argtypes = [C.c_int, _f32p, _f32p, C.c_int]
        _lib.oracle_distance_f64.restype = C.c_double
        _lib.oracle_distance_f64.argtypes = [C.c_int, _f32p, _f32p, C.c_int]
        _lib.oracle_flat_search.restype = C.c_int
        _lib.oracle_flat_search_batch_f32.restype = C.c_int
        _lib.oracle_merge_topk.restype = C.c_int
        _lib.oracle_num_threads.restype = C.c_int
        _lib.oracle_ivf_probe.restype = C.c_int
        _lib.oracle_ivfpq_search.restype = C.c_int
        _lib.oracle_adc_distance.restype = C.c_float
        _lib.oracle_metric_from_string.restype = C.c_int
        _lib.oracle_ivfpq_search_parts.restype = C.c_int
        _lib.oracle_flat_search_batch_blocked.restype = C.c_int
        _lib.oracle_acc_new.restype = C.c_void_p
        _lib.oracle_acc_add.restype = None
        _lib.oracle_acc_finish.restype = None
    return _lib


def ref():
    """The reference's own DuckDB distance kernels (oracle/_ref), or None when not built."""
    global _ref
    if _ref is None and os.path.exists(_REF_PATH):
        _ref = C.CDLL(_REF_PATH)
        for name in ("ref_array_distance", "ref_array_distance_squared", "ref_array_cosine_distance",
                     "ref_array_inner_product"):
            fn = getattr(_ref, name)
            fn.restype = C.c_float
            fn.argtypes = [_f32p, _f32p, C.c_int64]
    return _ref


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _bitmap(bm):
    return None if bm is None else np.ascontiguousarray(bm, dtype=np.uint32)


def metric_id(metric) -> int:
    return metric if isinstance(metric, int) else lib().oracle_metric_from_string(str(metric).encode())


def make_bitmap(mask: np.ndarray) -> np.ndarray:
    """bool[n] -> uint32 words, bit (i&31) of word i>>5 (the layout the C ABI takes)."""
    mask = np.asarray(mask, dtype=bool)
    n = mask.shape[0]
    padded = np.zeros(((n + 31) // 32) * 32, dtype=bool)
    padded[:n] = mask
    return np.packbits(padded.reshape(-1, 32), axis=1, bitorder="little").view(np.uint32).reshape(-1).copy()


def distance_f32(metric, q, x) -> float:
    q, x = _f32(q), _f32(x)
    return float(lib().oracle_distance_f32(metric_id(metric), q, x, q.shape[0]))


def distance_f64(metric, q, x) -> float:
    q, x = _f32(q), _f32(x)
    return float(lib().oracle_distance_f64(metric_id(metric), q, x, q.shape[0]))


def flat_search(X, q, k, metric="l2", allow=None, deleted=None, label_base=0):
    """Truth top-k of one query.  Returns (labels int64[c], dist32 float32[c], dist64 float64[c])."""
    X, q = _f32(X), _f32(q)
    n, d = X.shape if X.ndim == 2 else (0, q.shape[0])
    allow, deleted = _bitmap(allow), _bitmap(deleted)
    k = int(k)
    labels = np.empty(max(k, 1), np.int64)
    d32 = np.empty(max(k, 1), np.float32)
    d64 = np.empty(max(k, 1), np.float64)
    c = lib().oracle_flat_search(_ptr(X), C.c_int64(n), C.c_int(d), C.c_int(metric_id(metric)), _ptr(q),
                                 C.c_int(k), _ptr(allow), _ptr(deleted), C.c_int64(label_base),
                                 _ptr(labels), _ptr(d32), _ptr(d64))
    return labels[:c].copy(), d32[:c].copy(), d64[:c].copy()


def flat_search_batch_f32(X, Q, k, metric="l2", allow=None, deleted=None, nthreads=0):
    """OpenMP fp32 batch search (the timed CPU baseline).  Returns (labels[nq,k], dist[nq,k], counts[nq])."""
    X, Q = _f32(X), _f32(Q)
    n, d = X.shape
    nq = Q.shape[0]
    allow, deleted = _bitmap(allow), _bitmap(deleted)
    labels = np.full((nq, max(k, 1)), -1, np.int64)
    dist = np.full((nq, max(k, 1)), np.inf, np.float32)
    counts = np.zeros(nq, np.int32)
    lib().oracle_flat_search_batch_f32(_ptr(X), C.c_int64(n), C.c_int(d), C.c_int(metric_id(metric)), _ptr(Q),
                                       C.c_int(nq), C.c_int(k), _ptr(allow), _ptr(deleted), _ptr(labels),
                                       _ptr(dist), _ptr(counts), C.c_int(nthreads))
    return labels, dist, counts


def flat_search_batch_blocked(X, Q, k, metric="l2", allow=None, deleted=None, nthreads=0):
    """Cache-blocked SGEMM-style fp32 batch search (the timed CPU baseline)."""
    X, Q = _f32(X), _f32(Q)
    n, d = X.shape
    nq = Q.shape[0]
    allow, deleted = _bitmap(allow), _bitmap(deleted)
    labels = np.full((nq, max(k, 1)), -1, np.int64)
    dist = np.full((nq, max(k, 1)), np.inf, np.float32)
    counts = np.zeros(nq, np.int32)
    lib().oracle_flat_search_batch_blocked(_ptr(X), C.c_int64(n), C.c_int(d), C.c_int(metric_id(metric)), _ptr(Q),
                                           C.c_int(nq), C.c_int(k), _ptr(allow), _ptr(deleted), _ptr(labels),
                                           _ptr(dist), _ptr(counts), C.c_int(nthreads))
    return labels, dist, counts


class FlatAccumulator:
    """oracle_flat_search for several queries over a table that arrives in row chunks (fp64 truth, OpenMP)."""

    def __init__(self, Q, k, metric="l2", label_base=0):
        Q = _f32(Q)
        if Q.ndim == 1:
            Q = Q.reshape(1, -1)
        self.nq, self.d = Q.shape
        self.k, self.label_base = int(k), int(label_base)
        lib().oracle_acc_new.restype = C.c_void_p
        self._a = lib().oracle_acc_new(C.c_int(metric_id(metric)), _ptr(Q), C.c_int(self.nq), C.c_int(self.d),
                                       C.c_int(self.k))

    def add(self, X, row0=0, allow=None, deleted=None, allow_bits=None, nthreads=0):
        X = _f32(X)
        assert X.ndim == 2 and X.shape[1] == self.d and self._a
        allow, deleted = _bitmap(allow), _bitmap(deleted)
        if allow_bits is None:
            allow_bits = 0 if allow is None else allow.shape[0] * 32
        lib().oracle_acc_add(C.c_void_p(self._a), _ptr(X), C.c_int64(X.shape[0]), C.c_int64(int(row0)), _ptr(allow),
                             C.c_int64(int(allow_bits)), _ptr(deleted), C.c_int64(self.label_base), C.c_int(nthreads))

    def finish(self):
        """-> list of (labels int64[c], dist64 float64[c]) per query."""
        kk = max(self.k, 1)
        labels = np.empty((self.nq, kk), np.int64)
        d64 = np.empty((self.nq, kk), np.float64)
        counts = np.zeros(self.nq, np.int32)
        lib().oracle_acc_finish(C.c_void_p(self._a), _ptr(labels), _ptr(d64), _ptr(counts))
        self._a = None
        return [(labels[j, :counts[j]].copy(), d64[j, :counts[j]].copy()) for j in range(self.nq)]


def num_threads() -> int:
    return int(lib().oracle_num_threads())


def merge_topk(labels, dists, counts, k):
    labels = np.ascontiguousarray(labels, np.int64)
    dists = _f32(dists)
    counts = np.ascontiguousarray(counts, np.int32)
    G = labels.shape[0]
    ol = np.empty(max(k, 1), np.int64)
    od = np.empty(max(k, 1), np.float32)
    c = lib().oracle_merge_topk(_ptr(labels), _ptr(dists), _ptr(counts), C.c_int(G), C.c_int(k), _ptr(ol), _ptr(od))
    return ol[:c].copy(), od[:c].copy()


class IvfPqIndex:
    """Host copy of an IVF_PQ index in the oracle's layout (oracle.h: oracle_ivfpq_t)."""

    def __init__(self, metric, dim, P, M, centroids, codebook, part_offsets, row_ids, codes, n_indexed=None):
        self.metric = metric_id(metric)
        self.dim, self.P, self.M = int(dim), int(P), int(M)
        self.dsub = self.dim // self.M
        self.centroids = _f32(centroids).reshape(self.P, self.dim)
        self.codebook = _f32(codebook).reshape(self.M, 256, self.dsub)
        self.part_offsets = np.ascontiguousarray(part_offsets, np.int64)
        self.row_ids = np.ascontiguousarray(row_ids, np.int64)
        self.codes = np.ascontiguousarray(codes, np.uint8).reshape(-1, self.M)
        self.n_indexed = int(self.row_ids.shape[0] if n_indexed is None else n_indexed)

    def _c(self):
        s = _IvfPq(self.metric, self.dim, self.P, self.M, self.n_indexed, _ptr(self.centroids),
                   _ptr(self.codebook), _ptr(self.part_offsets), _ptr(self.row_ids), _ptr(self.codes))
        return s

    @staticmethod
    def train(X, metric, P, M, iters=10):
        X = _f32(X)
        n, d = X.shape
        assert d % M == 0
        cen = np.empty((P, d), np.float32)
        cb = np.empty((M, 256, d // M), np.float32)
        off = np.empty(P + 1, np.int64)
        ids = np.empty(n, np.int64)
        codes = np.empty((n, M), np.uint8)
        lib().oracle_ivfpq_train(_ptr(X), C.c_int64(n), C.c_int(d), C.c_int(metric_id(metric)), C.c_int(P),
                                 C.c_int(M), C.c_int(iters), _ptr(cen), _ptr(cb), _ptr(off), _ptr(ids), _ptr(codes))
        return IvfPqIndex(metric, d, P, M, cen, cb, off, ids, codes)

    def probe(self, q, nprobes):
        q = _f32(q)
        nprobes = min(int(nprobes), self.P)
        parts = np.empty(nprobes, np.int32)
        dists = np.empty(nprobes, np.float32)
        s = self._c()
        c = lib().oracle_ivf_probe(C.byref(s), _ptr(q), C.c_int(nprobes), _ptr(parts), _ptr(dists))
        return parts[:c].copy(), dists[:c].copy()

    def lut(self, q, part):
        q = _f32(q)
        out = np.empty((self.M, 256), np.float32)
        s = self._c()
        lib().oracle_pq_lut(C.byref(s), _ptr(q), C.c_int(int(part)), _ptr(out))
        return out

    def search(self, X, q, k, nprobes, refine_factor=1, allow=None, deleted=None):
        """Returns dict(parts, cand_labels, cand_adc, labels, dist)."""
        X, q = _f32(X), _f32(q)
        n = X.shape[0]
        allow, deleted = _bitmap(allow), _bitmap(deleted)
        nprobes = max(1, min(int(nprobes), self.P))
        rf = max(int(refine_factor), 1)
        parts = np.empty(nprobes, np.int32)
        cl = np.empty(k * rf, np.int64)
        ca = np.empty(k * rf, np.float32)
        nc = C.c_int32(0)
        ol = np.empty(k, np.int64)
        od = np.empty(k, np.float32)
        s = self._c()
        c = lib().oracle_ivfpq_search(C.byref(s), _ptr(X), C.c_int64(n), _ptr(q), C.c_int(k), C.c_int(nprobes),
                                      C.c_int(int(refine_factor)), _ptr(allow), _ptr(deleted), _ptr(parts),
                                      _ptr(cl), _ptr(ca), C.byref(nc), _ptr(ol), _ptr(od))
        return dict(parts=parts, cand_labels=cl[: nc.value].copy(), cand_adc=ca[: nc.value].copy(),
                    labels=ol[:c].copy(), dist=od[:c].copy())

    def search_parts(self, X, q, k, parts, refine_factor=1, allow=None, deleted=None):
        """The same pipeline over a GIVEN list of partitions (oracle_ivfpq_search_parts).  X may be None when
        refine_factor == 0 and the index covers every row (no row is then read)."""
        q = _f32(q)
        Xc = None if X is None else _f32(X)
        n = self.n_indexed if Xc is None else Xc.shape[0]
        allow, deleted = _bitmap(allow), _bitmap(deleted)
        parts = np.ascontiguousarray(parts, np.int32)
        rf = max(int(refine_factor), 1)
        cl = np.empty(k * rf, np.int64)
        ca = np.empty(k * rf, np.float32)
        nc = C.c_int32(0)
        ol = np.empty(k, np.int64)
        od = np.empty(k, np.float32)
        s = self._c()
        c = lib().oracle_ivfpq_search_parts(C.byref(s), _ptr(Xc), C.c_int64(n), _ptr(q), C.c_int(k), _ptr(parts),
                                            C.c_int(parts.shape[0]), C.c_int(int(refine_factor)), _ptr(allow),
                                            _ptr(deleted), _ptr(cl), _ptr(ca), C.byref(nc), _ptr(ol), _ptr(od))
        assert c >= 0
        return dict(parts=parts, cand_labels=cl[: nc.value].copy(), cand_adc=ca[: nc.value].copy(),
                    labels=ol[:c].copy(), dist=od[:c].copy())


def kmeans(X, k, iters=10, metric="l2"):
    X = _f32(X)
    n, d = X.shape
    cen = np.empty((k, d), np.float32)
    assign = np.empty(n, np.int32)
    lib().oracle_kmeans(_ptr(X), C.c_int64(n), C.c_int(d), C.c_int(k), C.c_int(iters), C.c_int(metric_id(metric)),
                        _ptr(cen), _ptr(assign))
    return cen, assign
