"""GPU: host-side behaviour of the C ABI that the extension relies on — locking (adds, scans and deletes from several
threads the way HNSWIndex calls them), error paths that must leave the index usable, validation of loaded streams, and
index.compact() against the reference's."""
import threading
import numpy as np
import pytest
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth, capi
from conftest import load_golden

pytestmark = pytest.mark.gpu


def _built(n=6000, d=32, M=16, metric="l2sq", seed=5):
    base, q = synth.clustered(n, d, max(16, n // 100), 6, seed, 200)
    ix = gv.Index(d, metric, M=M)
    ix.reserve(n, 1)
    ix.add_batch(base)
    ix.flush()
    return ix, base, q


def test_compact_reorders_like_the_reference(checker):
    """index.compact() (hnsw_index.cpp:458 -> index.hpp:3007-3096): nodes re-ordered by (level desc, nearest upper-level node),
    links renamed, removed members kept. Same input graph -> the stream after compact() equals the reference's byte for byte,
    and searches (ids, distance bits, counters) are the same on both."""
    Checker, kind = checker
    if kind != "reference":
        pytest.skip("compact() exists only in the compiled reference")
    for name, removes in (("clustered16_l2sq", 0), ("clustered24_cos", 40), ("readme_l2sq", 5)):
        z, blob = load_golden(name)
        d = z["base"].shape[1]
        ref = Checker(d, str(z["metric"]), M=int(z["M"]), M0=int(z["M0"]), ef_construction=int(z["efc"]))
        ref.load(blob)
        ix = gv.Index(d, str(z["metric"]), M=int(z["M"]), M0=int(z["M0"]), ef_construction=int(z["efc"]))
        ix.load(blob)
        ref.compact()
        ix.compact()
        got, want = ix.save(), ref.save()
        assert got == want, f"{name}: stream after compact() differs from the reference's"
        # removed members survive compact() in this usearch version. (Removing BEFORE the reference's compact would hit its
        # stale free-slot ring — SURVEY a17 latent bug — so the removal is replayed on the compacted graphs.)
        keys = z["keys_ef64"][:, 0]
        for key in keys[:removes]:
            ix.remove(int(key))
        iy = gv.Index(d, str(z["metric"]), M=int(z["M"]), M0=int(z["M0"]), ef_construction=int(z["efc"]))
        iy.load(got)
        for key in keys[:removes]:
            iy.remove(int(key))
        a = ix.ef_search(z["queries"], 10, 64, stats=True)
        b = iy.ef_search(z["queries"], 10, 64, stats=True)
        w = ref.search_batch(z["queries"], 10, 64)
        if not removes:
            np.testing.assert_array_equal(a[0], w["keys"])
            np.testing.assert_array_equal(a[1].view(np.uint32), w["dist"].view(np.uint32))
        np.testing.assert_array_equal(a[0], b[0])
        np.testing.assert_array_equal(a[3]["computed"], b[3]["computed"])
        ix.compact()   # compacting twice, with removed members present, still works
        assert ix.size() == z["base"].shape[0] - len(set(int(k) for k in keys[:removes]))


def test_any_thread_adds_while_other_threads_search():
    """HNSWIndex::Construct adds with any_thread() under the SHARED rwlock (hnsw_index.cpp:433-447): INSERTs and index scans
    of other connections overlap. Every row must arrive exactly once and every search must succeed."""
    d, n0, n1 = 16, 3000, 2000
    base, q = synth.clustered(n0 + n1, d, 50, 4, 11, 64)
    ix = gv.Index(d, "l2sq")
    ix.reserve(n0 + n1, 2)
    ix.add_batch(base[:n0], keys=np.arange(n0))
    ix.flush()
    errs, stop = [], threading.Event()

    def scan():
        try:
            while not stop.is_set():
                keys, _, counts = ix.ef_search(q, 5, 32)
                assert (counts == 5).all()
        except Exception as e:  # pragma: no cover
            errs.append(e)

    def insert(lo, hi):
        try:
            for i in range(lo, hi):
                ix.add(i, base[i], thread=capi.ANY_THREAD)
        except Exception as e:  # pragma: no cover
            errs.append(e)
    scanners = [threading.Thread(target=scan) for _ in range(3)]
    writers = [threading.Thread(target=insert, args=(n0 + j * (n1 // 4), n0 + (j + 1) * (n1 // 4))) for j in range(4)]
    [t.start() for t in scanners + writers]
    [t.join() for t in writers]
    stop.set()
    [t.join() for t in scanners]
    assert not errs, errs
    assert ix.size() == n0 + n1 and ix.info()["node_count"] == n0 + n1
    keys, dist, _ = ix.ef_search(base[n0:], 1, 64)
    assert (dist[:, 0] == 0).all()                                  # every inserted row is findable ...
    assert sorted(keys[:, 0]) == list(range(n0, n0 + n1))          # ... under its own key, once
    with pytest.raises(gv.GvssError, match="Reserve capacity"):
        ix.add(1, base[0], thread=capi.ANY_THREAD)


def test_thread_id_beyond_reserve_is_rejected():
    ix = gv.Index(4, "l2sq")
    ix.reserve(16, 2)
    ix.add(0, np.zeros(4, np.float32), thread=1)
    with pytest.raises(gv.GvssError, match="thread id 2"):
        ix.add(1, np.zeros(4, np.float32), thread=2)
    assert ix.size() == 1


def test_concurrent_host_searches_run_at_the_same_time_with_identical_results():
    """Scans hold only the shared side of HNSWIndex::rwlock (hnsw_index.cpp:285-311). gvss_search calls from several host
    threads each get a stream and scratch of their own: they overlap on the GPU and return the single-thread answer."""
    ix, base, q = _built(20000, 48, metric="cosine", seed=77)
    want = [ix.ef_search(q[i * 25:(i + 1) * 25], 10, 64) for i in range(8)]
    got, errs = [None] * 8, []
    gate = threading.Barrier(8)

    def work(i):
        try:
            gate.wait()
            for _ in range(200):
                got[i] = ix.ef_search(q[i * 25:(i + 1) * 25], 10, 64)
        except Exception as e:  # pragma: no cover
            errs.append(e)
    ts = [threading.Thread(target=work, args=(i,)) for i in range(8)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs
    for i in range(8):
        np.testing.assert_array_equal(got[i][0], want[i][0])
        np.testing.assert_array_equal(got[i][1].view(np.uint32), want[i][1].view(np.uint32))
    assert ix.info()["concurrent_searches_peak"] >= 2   # more than one search was in flight at once


def test_corrupt_stream_is_refused_and_the_index_keeps_its_contents():
    z, blob = load_golden("clustered16_l2sq")
    d = z["base"].shape[1]
    ix = gv.Index(d, "l2sq", M=int(z["M"]), M0=int(z["M0"]), ef_construction=int(z["efc"]))
    ix.load(blob)
    want = ix.ef_search(z["queries"], 10, 64)[0]
    n = z["base"].shape[0]
    hdr = 8 + n * d * 4                     # offset of the 64-byte dense header
    gh = hdr + 64                           # graph header: size, M, M0, max_level, entry
    tapes = gh + 40 + 2 * n                 # first node tape: key(8) level(2) count(4) slots...

    def corrupt(off, data):
        b = bytearray(blob)
        b[off:off + len(data)] = data
        return bytes(b)
    bad = {
        "magic": corrupt(hdr, b"x"),
        "metric byte": corrupt(hdr + 13, b"z"),
        "entry slot out of range": corrupt(gh + 32, (10 ** 9).to_bytes(8, "little")),
        "max level": corrupt(gh + 24, (500).to_bytes(8, "little")),
        "connectivity": corrupt(gh + 8, (2 ** 40).to_bytes(8, "little")),
        "negative level": corrupt(gh + 40, (-3).to_bytes(2, "little", signed=True)),
        "neighbour slot out of range": corrupt(tapes + 14, (n + 7).to_bytes(4, "little")),
        "truncated": blob[:len(blob) // 2],
    }
    for what, b in bad.items():
        with pytest.raises(gv.GvssError):
            ix.load(b)
        np.testing.assert_array_equal(ix.ef_search(z["queries"], 10, 64)[0], want, err_msg=f"after refusing: {what}")
    assert ix.save() == blob


def test_device_batches_may_reuse_removed_slots():
    """gvss_add_batch_device after gvss_remove: the rows land in recycled slots first (index_dense.hpp:1766-1793), then in
    fresh ones — not an error, and nothing is inserted twice."""
    import torch
    ix, base, q = _built(4000, 16, seed=3)
    for key in (5, 6, 7, 900):
        assert ix.remove(key) == 1
    extra, _ = synth.clustered(10, 16, 4, 4, 99, 0)
    dev = torch.from_numpy(extra).cuda()
    ix.reserve(4010, 1)
    ix.add_batch_device(dev.data_ptr(), 10, keys=np.arange(10_000, 10_010))
    ix.flush()
    info = ix.info()
    assert info["size"] == 4006 and info["node_count"] == 4006
    keys, dist, _ = ix.ef_search(extra, 1, 64)
    assert list(keys[:, 0]) == list(range(10_000, 10_010)) and (dist[:, 0] == 0).all()


def test_flush_that_cannot_allocate_keeps_the_rows_and_the_index():
    """An impossible reservation fails cleanly and the index still answers; staged rows are not lost by a failed call."""
    ix, base, q = _built(2000, 8, seed=4)
    want = ix.ef_search(q, 5, 32)[0]
    with pytest.raises(gv.GvssError):
        ix.reserve(2 ** 31 - 100, 1)       # beyond the addressable slots (and beyond any HBM)
    np.testing.assert_array_equal(ix.ef_search(q, 5, 32)[0], want)
    assert ix.capacity() == 2000


def _host_merge(parts, k):
    """(keys, dist) per shard -> merged top-k by (distance, key), the order k_merge / k_merge_peers define"""
    keys = np.concatenate([p[0] for p in parts], axis=1)
    dist = np.concatenate([p[1] for p in parts], axis=1)
    out_k = np.empty((keys.shape[0], k), np.int64)
    out_d = np.empty((keys.shape[0], k), np.float32)
    for i in range(keys.shape[0]):
        valid = keys[i] >= 0
        order = np.lexsort((keys[i][valid], dist[i][valid]))[:k]
        out_k[i], out_d[i] = -1, np.inf
        out_k[i, :len(order)] = keys[i][valid][order]
        out_d[i, :len(order)] = dist[i][valid][order]
    return out_k, out_d


def test_sharded_handle_equals_separate_indexes_plus_merge():
    """gvss_create_sharded (here: three shards on ONE GPU, which exercises all of the routing): rows are dealt out in chunks of
    2048, every shard is searched, lists are merged by (distance, key). The same thing done by hand with three plain indexes
    gives identical ids and distance bits; counts, size, removal and the refusals behave as documented."""
    n, d, k, ef, G = 9000, 24, 10, 96, 3
    base, q = synth.clustered(n, d, 90, 6, 21, 150)
    keys = np.arange(n, dtype=np.int64) + 10_000
    ix = gv.Index(d, "l2sq", M=16, devices=[0] * G)
    ix.reserve(n, 1)
    ix.set_build_concurrency(1)          # sequential builds are deterministic: the hand-made copy below gets the same graphs
    ix.add_batch(base, keys=keys)
    ix.flush()
    info = ix.info()
    assert info["size"] == n and info["node_count"] == n and info["capacity"] == n
    got_k, got_d, got_c, st = ix.ef_search(q, k, ef, stats=True)
    parts = []
    for s in range(G):
        rows = np.concatenate([np.arange(c, min(n, c + 2048)) for c in range(s * 2048, n, G * 2048)])
        iy = gv.Index(d, "l2sq", M=16)
        iy.reserve(len(rows) + 2048, 1)
        iy.set_build_concurrency(1)
        iy.add_batch(base[rows], keys=keys[rows])
        iy.flush()
        parts.append(iy.ef_search(q, k, ef, stats=True))
    want_k, want_d = _host_merge(parts, k)
    np.testing.assert_array_equal(got_k, want_k)
    np.testing.assert_array_equal(got_d.view(np.uint32), want_d.view(np.uint32))
    assert (got_c == k).all()
    np.testing.assert_array_equal(st["computed"], sum(p[3]["computed"] for p in parts))
    truth, _ = gv.exact_topk("l2sq", base, q, k, keys=keys)
    assert synth.recall_at_k(got_k, truth) > 0.97
    # removal reaches whichever shard holds the row
    victim = int(got_k[0, 0])
    assert ix.remove(victim) == 1 and ix.size() == n - 1
    assert victim not in ix.ef_search(q[:1], k, ef)[0]
    with pytest.raises(gv.GvssError, match="not one usearch stream"):
        ix.save()
    with pytest.raises(gv.GvssError, match="Reserve capacity"):
        ix.add_batch(base[:5], keys=np.arange(5))     # n - 1 live rows + 5 > capacity n


def test_sharded_handle_takes_adds_from_host_threads():
    """T construct tasks call add(key, vec, thread) on a sharded handle exactly as on a plain one."""
    n, d, T = 5000, 8, 4
    base, q = synth.clustered(n, d, 40, 4, 23, 0)
    ix = gv.Index(d, "l2sq", devices=[0, 0])
    ix.reserve(n, T)
    errs = []

    def work(t):
        try:
            for i in range(t, n, T):
                ix.add(i, base[i], thread=t)
        except Exception as e:  # pragma: no cover
            errs.append(e)
    ts = [threading.Thread(target=work, args=(t,)) for t in range(T)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs, errs
    assert ix.size() == n
    keys, dist, _ = ix.ef_search(base[:500], 1, 64)
    assert (dist[:, 0] == 0).all() and list(keys[:, 0]) == list(range(500))
