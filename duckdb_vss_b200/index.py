"""Python face of the C ABI, with the usearch method names the extension calls on ``HNSWIndex::index``
(reference src/hnsw/hnsw_index.cpp): reserve / add / ef_search / remove / compact /
save_to_stream / load_from_stream / size / capacity / max_level / stats. Host buffers are numpy arrays;
the ``*_device`` calls take raw device pointers (e.g. ``torch.Tensor.data_ptr()``)."""
import ctypes as C
import numpy as np
from . import capi

METRIC_CHAR = {"l2sq": b"e", "cosine": b"c", "cos": b"c", "ip": b"i"}
METRIC_NAME = {b"e": "l2sq", b"c": "cosine", b"i": "ip"}


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Index:
    """``index_dense_gt<row_t>::make(metric, config)`` (hnsw_index.cpp:162-189) on a B200."""

    def __init__(self, dims, metric="l2sq", M=0, M0=0, ef_construction=0, ef_search=0, device=-1, devices=None):
        """``devices=[0, 1, ...]``: one logical index partitioned over those GPUs (``gvss_create_sharded``)."""
        self._h = C.c_void_p()
        cfg = capi.Config(int(dims), METRIC_CHAR[metric], int(M), int(M0), int(ef_construction), int(ef_search),
                          int(device))
        if devices is None:
            capi.check(capi.lib().gvss_create(C.byref(cfg), C.byref(self._h)))
        else:
            arr = (C.c_int32 * len(devices))(*[int(v) for v in devices])
            capi.check(capi.lib().gvss_create_sharded(C.byref(cfg), arr, len(devices), C.byref(self._h)))
        self.dims = int(dims)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            capi.lib().gvss_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # interpreter shutdown: the module globals are already gone
            pass

    # --- construction ---
    def reserve(self, members, threads=1):
        capi.check(capi.lib().gvss_reserve(self._h, int(members), int(threads)))

    def add(self, key, vec, thread=0):
        """``index.add(key, vec, thread)``; ``thread=capi.ANY_THREAD`` is ``index_dense_t::any_thread()``."""
        vec = np.ascontiguousarray(vec, dtype=np.float32)
        assert vec.size == self.dims
        capi.check(capi.lib().gvss_add(self._h, int(key), _ptr(vec), int(thread)))

    def add_batch(self, vecs, keys=None):
        vecs = np.ascontiguousarray(vecs, dtype=np.float32).reshape(-1, self.dims)
        k = None if keys is None else np.ascontiguousarray(keys, dtype=np.int64)
        capi.check(capi.lib().gvss_add_batch(self._h, _ptr(k), _ptr(vecs), vecs.shape[0]))

    def add_batch_device(self, vecs_ptr, n, keys=None):
        k = None if keys is None else np.ascontiguousarray(keys, dtype=np.int64)
        capi.check(capi.lib().gvss_add_batch_device(self._h, _ptr(k), C.c_void_p(vecs_ptr), int(n)))

    def flush(self):
        capi.check(capi.lib().gvss_flush(self._h))

    def set_build_concurrency(self, n):
        capi.check(capi.lib().gvss_set_build_concurrency(self._h, int(n)))

    def set_rows_only(self, on=True):
        """Rows, keys and removals only: no graph is built and exact_search is the one search answered (gvss_set_rows_only)."""
        capi.check(capi.lib().gvss_set_rows_only(self._h, int(bool(on))))

    def set_background_build(self, min_rows):
        """Rows staged by add() are inserted by a library thread as soon as ``min_rows`` wait (0 = off)."""
        capi.check(capi.lib().gvss_set_background_build(self._h, int(min_rows)))

    def remove(self, key):
        n = C.c_uint64()
        capi.check(capi.lib().gvss_remove(self._h, int(key), C.byref(n)))
        return n.value

    def compact(self):
        capi.check(capi.lib().gvss_compact(self._h))

    def compact_rebuild(self):
        """Deleting compaction: rebuilds the graph over the live members; returns how many removed members were dropped."""
        n = C.c_uint64(0)
        capi.check(capi.lib().gvss_compact_rebuild(self._h, C.byref(n)))
        return n.value

    def reset(self):
        capi.check(capi.lib().gvss_reset(self._h))

    # --- search ---
    def ef_search(self, queries, k, ef, stats=False):
        """Batched ``index.ef_search(q, k, ef)``; returns (keys[nq,k], dist[nq,k], counts[nq]) (+ stats)."""
        q = np.ascontiguousarray(queries, dtype=np.float32).reshape(-1, self.dims)
        nq = q.shape[0]
        keys = np.full((nq, max(k, 1)), -1, dtype=np.int64)
        dist = np.full((nq, max(k, 1)), np.inf, dtype=np.float32)
        counts = np.zeros(nq, dtype=np.uint32)
        st = np.zeros((nq, 4), dtype=np.uint32)
        capi.check(capi.lib().gvss_search(self._h, _ptr(q), nq, int(k), int(ef), _ptr(keys), _ptr(dist),
                                          _ptr(counts), _ptr(st)))
        if stats:
            return keys[:, :k], dist[:, :k], counts, {"computed": st[:, 0], "visited": st[:, 1],
                                                      "base_hops": st[:, 2], "flags": st[:, 3]}
        return keys[:, :k], dist[:, :k], counts

    def exact_search(self, queries, k):
        """Brute-force top-k over the rows the index holds in HBM (``gvss_exact_search``): (keys, DuckDB-fold distances, counts)."""
        q = np.ascontiguousarray(queries, dtype=np.float32).reshape(-1, self.dims)
        nq = q.shape[0]
        keys = np.full((nq, k), -1, dtype=np.int64)
        dist = np.full((nq, k), np.inf, dtype=np.float32)
        counts = np.zeros(nq, dtype=np.uint32)
        capi.check(capi.lib().gvss_exact_search(self._h, _ptr(q), nq, int(k), _ptr(keys), _ptr(dist), _ptr(counts)))
        return keys, dist, counts

    def search_device(self, q_ptr, nq, k, ef, keys_ptr, dist_ptr=0, counts_ptr=0, stats_ptr=0, stream=0):
        capi.check(capi.lib().gvss_search_device(self._h, C.c_void_p(q_ptr), int(nq), int(k), int(ef),
                                                 C.c_void_p(keys_ptr), C.c_void_p(dist_ptr or None),
                                                 C.c_void_p(counts_ptr or None), C.c_void_p(stats_ptr or None),
                                                 C.c_void_p(stream or None)))

    def search_device_noalloc(self, q_ptr, nq, k, ef, keys_ptr, dist_ptr=0, counts_ptr=0, stats_ptr=0, stream=0):
        """``gvss_search_device_noalloc``: same results as :meth:`search_device`, scratch in the library's built-in array."""
        capi.check(capi.lib().gvss_search_device_noalloc(self._h, C.c_void_p(q_ptr), int(nq), int(k), int(ef),
                                                         C.c_void_p(keys_ptr), C.c_void_p(dist_ptr or None),
                                                         C.c_void_p(counts_ptr or None), C.c_void_p(stats_ptr or None),
                                                         C.c_void_p(stream or None)))

    # --- persistence ---
    def save_to_stream(self, write):
        """``write(bytes) -> bool`` receives the usearch v2.12 stream in pieces (hnsw_index.cpp:526-529)."""
        def cb(_user, data, size):
            return 1 if write(C.string_at(data, size)) else 0
        fn = capi.WRITE_CB(cb)
        capi.check(capi.lib().gvss_save(self._h, fn, None))

    def save(self) -> bytes:
        parts = []
        self.save_to_stream(lambda b: parts.append(b) or True)
        return b"".join(parts)

    def load_from_stream(self, read):
        """``read(n) -> bytes`` must return exactly n bytes (hnsw_index.cpp:204-205)."""
        def cb(_user, data, size):
            b = read(size)
            if b is None or len(b) != size:
                return 0
            C.memmove(data, b, size)
            return 1
        fn = capi.READ_CB(cb)
        capi.check(capi.lib().gvss_load(self._h, fn, None))

    def load(self, blob: bytes):
        view = memoryview(blob)
        pos = [0]

        def read(n):
            b = view[pos[0]:pos[0] + n]
            pos[0] += n
            return bytes(b)
        self.load_from_stream(read)

    def serialized_length(self):
        n = C.c_uint64()
        capi.check(capi.lib().gvss_serialized_length(self._h, C.byref(n)))
        return n.value

    # --- introspection ---
    def info(self):
        i = capi.Info()
        capi.check(capi.lib().gvss_get_info(self._h, C.byref(i)))
        return {f: getattr(i, f) for f, _ in capi.Info._fields_}

    def size(self):
        return self.info()["size"]

    def capacity(self):
        return self.info()["capacity"]

    def max_level(self):
        return self.info()["max_level"]

    def stats(self, level):
        s = capi.LevelStats()
        capi.check(capi.lib().gvss_get_level_stats(self._h, int(level), C.byref(s)))
        return {f: getattr(s, f) for f, _ in capi.LevelStats._fields_}

    def build_stats(self):
        """Traversal counters summed over every insert (add_result_t::computed_distances / visited_members in bulk)."""
        s = capi.BuildStats()
        capi.check(capi.lib().gvss_get_build_stats(self._h, C.byref(s)))
        return {f: getattr(s, f) for f, _ in capi.BuildStats._fields_}


def exact_topk(metric, base, queries, k, keys=None):
    """Brute-force top-k with DuckDB's folds (array_kernels.hpp:11-105), ties by row index."""
    base = np.ascontiguousarray(base, dtype=np.float32)
    queries = np.ascontiguousarray(queries, dtype=np.float32).reshape(-1, base.shape[1])
    ids = np.empty((queries.shape[0], k), dtype=np.int64)
    dist = np.empty((queries.shape[0], k), dtype=np.float32)
    kk = None if keys is None else np.ascontiguousarray(keys, dtype=np.int64)
    capi.check(capi.lib().gvss_exact_topk(METRIC_CHAR[metric], _ptr(base), base.shape[0], _ptr(kk), _ptr(queries),
                                          queries.shape[0], base.shape[1], int(k), _ptr(ids), _ptr(dist)))
    return ids, dist


def exact_topk_device(metric, base_ptr, n, q_ptr, nq, d, k, ids_ptr, dist_ptr, keys_ptr=0, stream=0):
    capi.check(capi.lib().gvss_exact_topk_device(METRIC_CHAR[metric], C.c_void_p(base_ptr), int(n),
                                                 C.c_void_p(keys_ptr or None), C.c_void_p(q_ptr), int(nq), int(d),
                                                 int(k), C.c_void_p(ids_ptr), C.c_void_p(dist_ptr or None),
                                                 C.c_void_p(stream or None)))


def merge_topk_device(keys_ptr, dist_ptr, shards, nq, k, out_keys_ptr, out_dist_ptr, stream=0):
    capi.check(capi.lib().gvss_merge_topk_device(C.c_void_p(keys_ptr), C.c_void_p(dist_ptr), int(shards), int(nq),
                                                 int(k), C.c_void_p(out_keys_ptr), C.c_void_p(out_dist_ptr),
                                                 C.c_void_p(stream or None)))


def merge_topk_peers_device(peer_keys, peer_dist, peer_flags, rank, step, nq, k, out_keys_ptr, out_dist_ptr, stream=0):
    """Merge that reads the peers' lists over NVLink (gvss_merge_topk_peers_device). peer_*: sequences of device
    pointers (ints), one per shard, as mapped into this process."""
    shards = len(peer_keys)
    arr = C.c_void_p * shards
    capi.check(capi.lib().gvss_merge_topk_peers_device(arr(*peer_keys), arr(*peer_dist), arr(*peer_flags), int(rank), shards,
                                                       int(step), int(nq), int(k), C.c_void_p(out_keys_ptr),
                                                       C.c_void_p(out_dist_ptr or None), C.c_void_p(stream or None)))
