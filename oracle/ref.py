"""ctypes binding of the compiled, unmodified reference (oracle/_ref/libref_usearch.so). Test infrastructure only."""
import ctypes as C
import numpy as np
from . import REF_SO

_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(REF_SO)
        vp, u32, u64, i64p, f32p = C.c_void_p, C.c_uint32, C.c_uint64, C.POINTER(C.c_int64), C.POINTER(C.c_float)
        u64p, u32p, u8p = C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.POINTER(C.c_uint8)
        L.ref_create.restype = vp
        L.ref_create.argtypes = [u32, C.c_char, u32, u32, u32, u32]
        L.ref_destroy.argtypes = [vp]
        L.ref_reserve.argtypes = [vp, u64, u32]
        L.ref_add.argtypes = [vp, C.c_int64, f32p, u32, u64p, u64p]
        L.ref_add_bulk.restype = u64
        L.ref_add_bulk.argtypes = [vp, i64p, f32p, u64, u32, u64p, u64p, C.POINTER(C.c_double)]
        L.ref_search.restype = u32
        L.ref_search.argtypes = [vp, f32p, u32, u32, i64p, f32p, u64p, u64p]
        L.ref_search_batch.restype = C.c_double
        L.ref_search_batch.argtypes = [vp, f32p, u64, u32, u32, u32, i64p, f32p, u32p, u64p, u64p]
        L.ref_remove.restype = u64
        L.ref_remove.argtypes = [vp, C.c_int64]
        for n in ("ref_size", "ref_capacity", "ref_max_level", "ref_memory_usage", "ref_serialized_length"):
            getattr(L, n).restype = u64
            getattr(L, n).argtypes = [vp]
        L.ref_level_stats.argtypes = [vp, u32, u64p]
        L.ref_save.restype = u64
        L.ref_save.argtypes = [vp, u8p, u64]
        L.ref_load.argtypes = [vp, u8p, u64]
        L.ref_metric.restype = C.c_float
        L.ref_metric.argtypes = [C.c_char, f32p, f32p, u32]
        L.ref_duckdb_distance.restype = C.c_float
        L.ref_duckdb_distance.argtypes = [C.c_char, f32p, f32p, u32]
        L.ref_duckdb_topk.argtypes = [C.c_char, f32p, u64, f32p, u64, u32, u32, u32, i64p, f32p]
        L.ref_fast_topk.argtypes = [C.c_char, f32p, u64, f32p, u64, u32, u32, u32, i64p, f32p]
        L.ref_compact.argtypes = [vp]
        L.ref_hardware_threads.restype = u32
        _lib = L
    return _lib


def _f32p(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _i64p(a):
    return a.ctypes.data_as(C.POINTER(C.c_int64))


METRIC_CHAR = {"l2sq": b"e", "cosine": b"c", "cos": b"c", "ip": b"i"}


def hardware_threads() -> int:
    return int(lib().ref_hardware_threads())


class RefIndex:
    """The reference's ``index_dense_gt<row_t>`` configured as in hnsw_index.cpp:162-189."""

    def __init__(self, dims, metric="l2sq", M=16, M0=0, ef_construction=128, ef_search=64):
        self.dims, self.metric = int(dims), metric
        self.h = lib().ref_create(dims, METRIC_CHAR[metric], M, M0, ef_construction, ef_search)

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().ref_destroy(self.h)
                self.h = None
        except Exception:  # interpreter shutdown
            pass

    def reserve(self, n, threads=1):
        if lib().ref_reserve(self.h, n, threads):
            raise MemoryError("reserve failed")

    def add(self, key, vec, thread=0):
        vec = np.ascontiguousarray(vec, dtype=np.float32)
        c, v = C.c_uint64(), C.c_uint64()
        if lib().ref_add(self.h, int(key), _f32p(vec), thread, C.byref(c), C.byref(v)):
            raise RuntimeError("reference add failed")
        return c.value, v.value

    def add_bulk(self, vecs, keys=None, threads=1):
        vecs = np.ascontiguousarray(vecs, dtype=np.float32)
        n = vecs.shape[0]
        kp = _i64p(np.ascontiguousarray(keys, dtype=np.int64)) if keys is not None else None
        c, v, s = C.c_uint64(), C.c_uint64(), C.c_double()
        failed = lib().ref_add_bulk(self.h, kp, _f32p(vecs), n, threads, C.byref(c), C.byref(v), C.byref(s))
        if failed:
            raise RuntimeError(f"{failed} reference adds failed")
        return {"seconds": s.value, "computed_distances": c.value, "visited_members": v.value}

    def search(self, q, k, ef):
        q = np.ascontiguousarray(q, dtype=np.float32)
        keys = np.full(k, -1, dtype=np.int64)
        dist = np.full(k, np.inf, dtype=np.float32)
        c, v = C.c_uint64(), C.c_uint64()
        n = lib().ref_search(self.h, _f32p(q), k, ef, _i64p(keys), _f32p(dist), C.byref(c), C.byref(v))
        return keys[:n], dist[:n], c.value, v.value

    def search_batch(self, qs, k, ef, threads=1):
        qs = np.ascontiguousarray(qs, dtype=np.float32)
        nq = qs.shape[0]
        keys = np.empty((nq, k), dtype=np.int64)
        dist = np.empty((nq, k), dtype=np.float32)
        counts = np.empty(nq, dtype=np.uint32)
        comp = np.empty(nq, dtype=np.uint64)
        vis = np.empty(nq, dtype=np.uint64)
        secs = lib().ref_search_batch(self.h, _f32p(qs), nq, k, ef, threads, _i64p(keys), _f32p(dist),
                                      counts.ctypes.data_as(C.POINTER(C.c_uint32)),
                                      comp.ctypes.data_as(C.POINTER(C.c_uint64)),
                                      vis.ctypes.data_as(C.POINTER(C.c_uint64)))
        return {"keys": keys, "dist": dist, "counts": counts, "computed": comp, "visited": vis, "seconds": secs}

    def remove(self, key):
        return int(lib().ref_remove(self.h, int(key)))

    def size(self):
        return int(lib().ref_size(self.h))

    def capacity(self):
        return int(lib().ref_capacity(self.h))

    def max_level(self):
        return int(lib().ref_max_level(self.h))

    def level_stats(self, level):
        out = (C.c_uint64 * 4)()
        lib().ref_level_stats(self.h, level, out)
        return {"nodes": out[0], "edges": out[1], "max_edges": out[2], "allocated_bytes": out[3]}

    def save(self) -> bytes:
        n = int(lib().ref_serialized_length(self.h))
        buf = np.empty(n, dtype=np.uint8)
        w = lib().ref_save(self.h, buf.ctypes.data_as(C.POINTER(C.c_uint8)), n)
        if w != n:
            raise RuntimeError(f"reference save wrote {w} of {n} bytes")
        return buf.tobytes()

    def load(self, blob: bytes):
        buf = np.frombuffer(blob, dtype=np.uint8)
        if lib().ref_load(self.h, buf.ctypes.data_as(C.POINTER(C.c_uint8)), buf.size):
            raise RuntimeError("reference load failed")

    def compact(self):
        """index.compact() (hnsw_index.cpp:458): re-orders the nodes, keeps removed members."""
        if lib().ref_compact(self.h):
            raise RuntimeError("reference compact failed")


def metric(kind, a, b):
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    return float(lib().ref_metric(METRIC_CHAR[kind], _f32p(a), _f32p(b), a.size))


def duckdb_distance(kind, a, b):
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    return float(lib().ref_duckdb_distance(METRIC_CHAR[kind], _f32p(a), _f32p(b), a.size))


def duckdb_topk(kind, base, qs, k, threads=1):
    base = np.ascontiguousarray(base, dtype=np.float32)
    qs = np.ascontiguousarray(qs, dtype=np.float32)
    ids = np.empty((qs.shape[0], k), dtype=np.int64)
    dist = np.empty((qs.shape[0], k), dtype=np.float32)
    lib().ref_duckdb_topk(METRIC_CHAR[kind], _f32p(base), base.shape[0], _f32p(qs), qs.shape[0], base.shape[1], k,
                          threads, _i64p(ids), _f32p(dist))
    return ids, dist


def fast_topk(kind, base, qs, k, threads=1):
    """Ground truth at full size: vectorised candidate ranking + exact re-score with DuckDB's folds, (distance, rowid) order."""
    base = np.ascontiguousarray(base, dtype=np.float32)
    qs = np.ascontiguousarray(qs, dtype=np.float32)
    ids = np.empty((qs.shape[0], k), dtype=np.int64)
    dist = np.empty((qs.shape[0], k), dtype=np.float32)
    lib().ref_fast_topk(METRIC_CHAR[kind], _f32p(base), base.shape[0], _f32p(qs), qs.shape[0], base.shape[1], k,
                        threads, _i64p(ids), _f32p(dist))
    return ids, dist
