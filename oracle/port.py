"""ctypes binding of the plain-C restatement (oracle/libhnsw_oracle.so). Test infrastructure only."""
import ctypes as C
import numpy as np
from . import PORT_SO

_lib = None
METRIC_CHAR = {"l2sq": b"e", "cosine": b"c", "cos": b"c", "ip": b"i"}


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(PORT_SO)
        vp, u32, u64, i64p, f32p = C.c_void_p, C.c_uint32, C.c_uint64, C.POINTER(C.c_int64), C.POINTER(C.c_float)
        u64p, u32p, u8p = C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.POINTER(C.c_uint8)
        L.orc_create.restype = vp
        L.orc_create.argtypes = [u32, C.c_char, u32, u32, u32, u32]
        L.orc_destroy.argtypes = [vp]
        L.orc_reserve.argtypes = [vp, u64]
        L.orc_add.argtypes = [vp, C.c_int64, f32p, u64p, u64p]
        L.orc_search.restype = u32
        L.orc_search.argtypes = [vp, f32p, u32, u32, i64p, f32p, u64p, u64p, u64p, u32p]
        L.orc_remove.restype = u64
        L.orc_remove.argtypes = [vp, C.c_int64]
        for n in ("orc_size", "orc_node_count", "orc_entry_slot", "orc_serialized_length"):
            getattr(L, n).restype = u64
            getattr(L, n).argtypes = [vp]
        L.orc_max_level.restype = C.c_int32
        L.orc_max_level.argtypes = [vp]
        L.orc_save.restype = u64
        L.orc_save.argtypes = [vp, u8p, u64]
        L.orc_load.argtypes = [vp, u8p, u64]
        L.orc_level_stats.argtypes = [vp, u32, u64p]
        L.orc_metric.restype = C.c_float
        L.orc_metric.argtypes = [C.c_char, f32p, f32p, u32]
        L.orc_duckdb_distance.restype = C.c_float
        L.orc_duckdb_distance.argtypes = [C.c_char, f32p, f32p, u32]
        L.orc_exact_topk.argtypes = [C.c_char, f32p, u64, f32p, u64, u32, u32, i64p, f32p]
        L.orc_level_from_state.restype = C.c_int16
        L.orc_level_from_state.argtypes = [u64p, u32]
        _lib = L
    return _lib


def _f32p(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _i64p(a):
    return a.ctypes.data_as(C.POINTER(C.c_int64))


class PortIndex:
    def __init__(self, dims, metric="l2sq", M=16, M0=0, ef_construction=128, ef_search=64):
        self.dims, self.metric = int(dims), metric
        self.h = lib().orc_create(dims, METRIC_CHAR[metric], M, M0, ef_construction, ef_search)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_destroy(self.h)
            self.h = None

    def reserve(self, n, threads=1):
        lib().orc_reserve(self.h, n)

    def add(self, key, vec):
        vec = np.ascontiguousarray(vec, dtype=np.float32)
        c, v = C.c_uint64(), C.c_uint64()
        if lib().orc_add(self.h, int(key), _f32p(vec), C.byref(c), C.byref(v)):
            raise RuntimeError("Reserve capacity ahead of insertions!")
        return c.value, v.value

    def add_bulk(self, vecs, keys=None, threads=1):
        vecs = np.ascontiguousarray(vecs, dtype=np.float32)
        comp = vis = 0
        for i in range(vecs.shape[0]):
            c, v = self.add(i if keys is None else keys[i], vecs[i])
            comp += c
            vis += v
        return {"computed_distances": comp, "visited_members": vis}

    def search(self, q, k, ef, detail=False):
        q = np.ascontiguousarray(q, dtype=np.float32)
        keys = np.full(k, -1, dtype=np.int64)
        dist = np.full(k, np.inf, dtype=np.float32)
        c, v, h, pk = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_uint32()
        n = lib().orc_search(self.h, _f32p(q), k, ef, _i64p(keys), _f32p(dist), C.byref(c), C.byref(v), C.byref(h),
                             C.byref(pk))
        if detail:
            return keys[:n], dist[:n], c.value, v.value, h.value, pk.value
        return keys[:n], dist[:n], c.value, v.value

    def search_batch(self, qs, k, ef, threads=1):
        qs = np.ascontiguousarray(qs, dtype=np.float32)
        nq = qs.shape[0]
        keys = np.full((nq, k), -1, dtype=np.int64)
        dist = np.full((nq, k), np.inf, dtype=np.float32)
        counts = np.zeros(nq, dtype=np.uint32)
        comp = np.zeros(nq, dtype=np.uint64)
        vis = np.zeros(nq, dtype=np.uint64)
        hops = np.zeros(nq, dtype=np.uint64)
        for i in range(nq):
            kk, dd, c, v, h, _ = self.search(qs[i], k, ef, detail=True)
            counts[i] = len(kk)
            keys[i, :len(kk)] = kk
            dist[i, :len(kk)] = dd
            comp[i], vis[i], hops[i] = c, v, h
        return {"keys": keys, "dist": dist, "counts": counts, "computed": comp, "visited": vis, "base_hops": hops}

    def remove(self, key):
        return int(lib().orc_remove(self.h, int(key)))

    def size(self):
        return int(lib().orc_size(self.h))

    def max_level(self):
        return int(lib().orc_max_level(self.h))

    def entry_slot(self):
        return int(lib().orc_entry_slot(self.h))

    def level_stats(self, level):
        out = (C.c_uint64 * 4)()
        lib().orc_level_stats(self.h, level, out)
        return {"nodes": out[0], "edges": out[1], "max_edges": out[2], "allocated_bytes": out[3]}

    def save(self) -> bytes:
        n = int(lib().orc_serialized_length(self.h))
        buf = np.empty(n, dtype=np.uint8)
        w = lib().orc_save(self.h, buf.ctypes.data_as(C.POINTER(C.c_uint8)), n)
        if w != n:
            raise RuntimeError(f"oracle save wrote {w} of {n} bytes")
        return buf.tobytes()

    def load(self, blob: bytes):
        buf = np.frombuffer(blob, dtype=np.uint8)
        rc = lib().orc_load(self.h, buf.ctypes.data_as(C.POINTER(C.c_uint8)), buf.size)
        if rc:
            raise RuntimeError(f"oracle load failed ({rc})")


def metric(kind, a, b):
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    return float(lib().orc_metric(METRIC_CHAR[kind], _f32p(a), _f32p(b), a.size))


def duckdb_distance(kind, a, b):
    a = np.ascontiguousarray(a, dtype=np.float32)
    b = np.ascontiguousarray(b, dtype=np.float32)
    return float(lib().orc_duckdb_distance(METRIC_CHAR[kind], _f32p(a), _f32p(b), a.size))


def exact_topk(kind, base, qs, k):
    base = np.ascontiguousarray(base, dtype=np.float32)
    qs = np.ascontiguousarray(qs, dtype=np.float32)
    ids = np.empty((qs.shape[0], k), dtype=np.int64)
    dist = np.empty((qs.shape[0], k), dtype=np.float32)
    lib().orc_exact_topk(METRIC_CHAR[kind], _f32p(base), base.shape[0], _f32p(qs), qs.shape[0], base.shape[1], k,
                         _i64p(ids), _f32p(dist))
    return ids, dist
