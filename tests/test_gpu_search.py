"""GPU parity #1: searching a reference-built (deserialised) graph through the C ABI returns the reference's row ids,
distances (bit for bit) and traversal counters. Checker = compiled reference when present, else the pinned C port."""
import numpy as np
import pytest
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth
from conftest import golden_cases, load_golden

pytestmark = pytest.mark.gpu


def _load(z, blob):
    ix = gv.Index(z["base"].shape[1], str(z["metric"]), M=int(z["M"]), M0=int(z["M0"]), ef_construction=int(z["efc"]))
    ix.load(blob)
    return ix


@pytest.mark.parametrize("name", golden_cases())
def test_golden_search_identical_ids_distances_counters(name):
    z, blob = load_golden(name)
    ix = _load(z, blob)
    assert ix.size() == z["base"].shape[0]
    for ef in (3, 10, 64, 128):
        keys, dist, counts, st = ix.ef_search(z["queries"], 10, ef, stats=True)
        np.testing.assert_array_equal(keys, z[f"keys_ef{ef}"], err_msg=f"{name} ef={ef}")
        np.testing.assert_array_equal(dist.view(np.uint32), z[f"dist_ef{ef}"].view(np.uint32))
        np.testing.assert_array_equal(st["computed"], z[f"computed_ef{ef}"].astype(np.uint32))
        np.testing.assert_array_equal(st["visited"], z[f"visited_ef{ef}"].astype(np.uint32))
        assert (st["flags"] & 3 == 0).all()  # no queue overflow, no bitmap fallback
        if name.startswith("readme"):  # the grid is all ties: every query must have been replayed with usearch's heap (flag bit2)
            assert (st["flags"] & 4 != 0).any()


@pytest.mark.parametrize("name", golden_cases())
def test_stream_round_trip_is_byte_identical(name):
    z, blob = load_golden(name)
    ix = _load(z, blob)
    assert ix.serialized_length() == len(blob)
    assert ix.save() == blob


def test_readme_known_answer_through_c_abi():
    z, blob = load_golden("readme_l2sq")
    ix = _load(z, blob)
    keys, dist, counts = ix.ef_search(np.array([[1, 2, 3]], np.float32), 3, 64)
    assert list(keys[0]) == [11, 92, 12]          # SURVEY §8c (compiled reference)
    assert list(np.sqrt(dist[0])) == [0.0, 1.0, 1.0]  # hnsw_result.test:20-25
    assert counts[0] == 3


@pytest.mark.parametrize("metric,d,n,M", [("l2sq", 128, 20000, 16), ("cosine", 200, 6000, 16), ("ip", 96, 8000, 8), ("l2sq", 5, 3000, 3)])
def test_reference_built_graph_search_parity(metric, d, n, M, checker):
    Checker, kind = checker
    base, q = synth.clustered(n, d, max(16, n // 100), 8, 100 + d, 300)
    if metric == "ip":
        base /= np.linalg.norm(base, axis=1, keepdims=True)
    ref = Checker(d, metric, M=M)
    ref.reserve(n, 8)
    ref.add_bulk(base, threads=8)       # multi-threaded reference build; the saved graph is the shared input
    blob = ref.save()
    ix = gv.Index(d, metric, M=M)
    ix.load(blob)
    assert ix.save() == blob
    for ef in (10, 64, 128):
        want = ref.search_batch(q, 10, ef)
        keys, dist, counts, st = ix.ef_search(q, 10, ef, stats=True)
        np.testing.assert_array_equal(keys, want["keys"], err_msg=f"{metric} ef={ef} ({kind})")
        np.testing.assert_array_equal(dist.view(np.uint32), want["dist"].view(np.uint32))
        np.testing.assert_array_equal(counts, want["counts"])
        np.testing.assert_array_equal(st["computed"], want["computed"].astype(np.uint32))
        np.testing.assert_array_equal(st["visited"], want["visited"].astype(np.uint32))


@pytest.mark.parametrize("metric,d,n,levels", [("l2sq", 32, 12000, 4), ("ip", 24, 8000, 3), ("cosine", 16, 6000, 3)])
def test_tie_heavy_integer_data_matches_reference_pop_order(metric, d, n, levels, checker):
    """SIFT-like integer-valued vectors: distances are small integers, so live candidates tie for the head of the queue
    all the time. The list-as-queue fast path must notice every such tie and continue with usearch's heap (flag bit2);
    ids, order among ties, distance bits and traversal counters still equal the reference's on its own graph."""
    Checker, kind = checker
    rng = np.random.default_rng(7 + d)
    base = rng.integers(0, levels, size=(n, d)).astype(np.float32)
    q = rng.integers(0, levels, size=(200, d)).astype(np.float32)
    base[:, 0] += 1.0  # no all-zero rows (cosine)
    q[:, 0] += 1.0
    ref = Checker(d, metric, M=8)
    ref.reserve(n, 4)
    ref.add_bulk(base, threads=4)
    ix = gv.Index(d, metric, M=8)
    ix.load(ref.save())
    tied = 0
    for ef in (10, 64, 200):
        want = ref.search_batch(q, 10, ef)
        keys, dist, counts, st = ix.ef_search(q, 10, ef, stats=True)
        np.testing.assert_array_equal(keys, want["keys"], err_msg=f"{metric} ef={ef} ({kind})")
        np.testing.assert_array_equal(dist.view(np.uint32), want["dist"].view(np.uint32))
        np.testing.assert_array_equal(st["computed"], want["computed"].astype(np.uint32))
        np.testing.assert_array_equal(st["visited"], want["visited"].astype(np.uint32))
        assert (st["flags"] & 3 == 0).all()
        tied += int((st["flags"] & 4 != 0).sum())
    assert tied > 300  # the tie path really ran (3 x 200 queries)


@pytest.mark.parametrize("n,d", [(40, 40), (7, 128), (1, 32), (33, 8)])
def test_graph_smaller_than_the_beam(n, d, checker):
    """ef larger than the whole graph: the result list never fills, the search ends when every member is expanded."""
    Checker, kind = checker
    base, q = synth.clustered(n, d, 4, 3, 900 + n, 25)
    ref = Checker(d, "l2sq", M=4)
    ref.reserve(n, 1)
    ref.add_bulk(base, threads=1)
    ix = gv.Index(d, "l2sq", M=4)
    ix.load(ref.save())
    for ef in (3, 64):
        want = ref.search_batch(q, 10, ef)
        keys, dist, counts, st = ix.ef_search(q, 10, ef, stats=True)
        np.testing.assert_array_equal(keys, want["keys"], err_msg=f"n={n} d={d} ef={ef} ({kind})")
        np.testing.assert_array_equal(dist.view(np.uint32), want["dist"].view(np.uint32))
        np.testing.assert_array_equal(counts, want["counts"])
        np.testing.assert_array_equal(st["computed"], want["computed"].astype(np.uint32))
        np.testing.assert_array_equal(st["visited"], want["visited"].astype(np.uint32))


def test_removed_members_are_traversed_but_never_returned(checker):
    Checker, _ = checker
    n, d = 3000, 16
    base, q = synth.clustered(n, d, 30, 6, 5, 100)
    ref = Checker(d, "l2sq")
    ref.reserve(n, 1)
    ref.add_bulk(base)
    ix = gv.Index(d, "l2sq")
    ix.load(ref.save())
    gone = list(range(0, n, 3))
    for key in gone:
        assert ref.remove(key) == 1
        assert ix.remove(key) == 1
    assert ix.remove(10 ** 9) == 0                      # unknown key: silent no-op (hnsw_index.cpp:478)
    assert ix.size() == n - len(gone)
    want = ref.search_batch(q, 10, 64)
    keys, dist, counts = ix.ef_search(q, 10, 64)
    np.testing.assert_array_equal(keys, want["keys"])
    np.testing.assert_array_equal(dist.view(np.uint32), want["dist"].view(np.uint32))
    assert not np.isin(keys, gone).any()
    # removed slots survive a save/load cycle (hnsw_crud.test:12-47)
    blob = ix.save()
    want_blob = ref.save()
    # the only bytes allowed to differ are count_present / count_deleted in the 64-byte header: the reference derives
    # them from its free-slot ring, whose size() wraps to 0 whenever the ring is exactly full (1000 removes -> it
    # reports 2960 present / 40 deleted); gvss writes the true counts. Loaders ignore both fields.
    off = 8 + n * d * 4 + 17
    assert blob[:off] == want_blob[:off] and blob[off + 16:] == want_blob[off + 16:]
    assert int.from_bytes(blob[off:off + 8], "little") == n - len(gone)
    assert int.from_bytes(blob[off + 8:off + 16], "little") == len(gone)
    iy = gv.Index(d, "l2sq")
    iy.load(blob)
    assert iy.size() == n - len(gone)
    k2, _, _ = iy.ef_search(q, 10, 64)
    np.testing.assert_array_equal(k2, keys)


def test_edge_cases():
    z, blob = load_golden("readme_l2sq")
    ix = _load(z, blob)
    q = np.array([[1, 2, 3]], np.float32)
    keys, dist, counts = ix.ef_search(q, 0, 64)           # k == 0 -> empty (index.hpp:2489-2490)
    assert counts[0] == 0
    keys, dist, counts = ix.ef_search(q, 1000, 64)        # k > N: beam = max(ef, k), returns <= N
    assert counts[0] == 729 and len(set(keys[0][:729])) == 729 and (keys[0][729:] == -1).all()
    assert (np.diff(dist[0][:729]) >= 0).all()
    keys, dist, counts = ix.ef_search(q, 3, 0)            # ef == 0 -> 64 (index.hpp:2493-2494)
    assert list(keys[0]) == [11, 92, 12]
    nanq = np.full((2, 3), np.nan, np.float32)            # NULL outer row of the index join (hnsw_lateral_join.test:57-67)
    keys, dist, counts = ix.ef_search(nanq, 3, 64)
    assert (counts <= 3).all()
    empty = gv.Index(3, "l2sq")                           # empty index -> 0 rows (hnsw_insert.test:18-21)
    empty.reserve(32, 1)
    keys, dist, counts = empty.ef_search(q, 3, 64)
    assert counts[0] == 0 and (keys == -1).all()
    assert empty.size() == 0 and len(empty.save()) == 8 + 64 + 40


def test_device_resident_search_matches_host_path():
    import torch
    z, blob = load_golden("clustered16_l2sq")
    ix = _load(z, blob)
    q = torch.from_numpy(z["queries"]).cuda()
    nq, k = q.shape[0], 10
    keys = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    dist = torch.empty((nq, k), dtype=torch.float32, device="cuda")
    counts = torch.empty(nq, dtype=torch.int32, device="cuda")
    ix.search_device(q.data_ptr(), nq, k, 64, keys.data_ptr(), dist.data_ptr(), counts.data_ptr(), 0,
                     torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(keys.cpu().numpy(), z["keys_ef64"])
    np.testing.assert_array_equal(dist.cpu().numpy().view(np.uint32), z["dist_ef64"].view(np.uint32))


def test_tiny_connectivity_options_build_and_search_like_the_reference(checker):
    # hnsw_options.test:49-50 accepts ef_construction=100, ef_search=100, M=3, M0=3
    Checker, _ = checker
    base, q = synth.clustered(700, 5, 10, 3, 21, 50)
    ref = Checker(5, "l2sq", M=3, M0=3, ef_construction=100, ef_search=100)
    ref.reserve(700, 1)
    ref.add_bulk(base, threads=1)
    ix = gv.Index(5, "l2sq", M=3, M0=3, ef_construction=100, ef_search=100)
    ix.reserve(700, 1)
    ix.set_build_concurrency(1)
    ix.add_batch(base)
    ix.flush()
    assert ix.save() == ref.save()
    want = ref.search_batch(q, 3, 100)
    keys, dist, counts = ix.ef_search(q, 3, 100)
    np.testing.assert_array_equal(keys, want["keys"])
    np.testing.assert_array_equal(dist.view(np.uint32), want["dist"].view(np.uint32))


def test_m0_smaller_than_m_is_clamped_not_overrun():
    # the reference writes M links into a level-0 list sized M0 (out of bounds); gvss clamps to M0
    base, q = synth.clustered(400, 6, 8, 3, 22, 20)
    ix = gv.Index(6, "l2sq", M=6, M0=3, ef_construction=40)
    ix.reserve(400, 1)
    ix.add_batch(base)
    ix.flush()
    st = ix.stats(0)
    assert st["nodes"] == 400 and st["edges"] <= 400 * 3
    keys, dist, counts = ix.ef_search(base[:20], 1, 64)
    assert (counts == 1).all() and (dist[:, 0] >= 0).all()


def test_duplicate_vectors_and_removed_entry_point(checker):
    Checker, _ = checker
    base, q = synth.clustered(600, 8, 6, 3, 23, 30)
    base[100:200] = base[0:100]                      # exact duplicates: distance-0 ties between nodes
    ref = Checker(8, "l2sq")
    ref.reserve(600, 1)
    ref.add_bulk(base, threads=1)
    ix = gv.Index(8, "l2sq")
    ix.load(ref.save())
    for ef in (10, 64):
        want = ref.search_batch(q, 10, ef)
        keys, dist, _ = ix.ef_search(q, 10, ef)
        np.testing.assert_array_equal(keys, want["keys"])
        np.testing.assert_array_equal(dist.view(np.uint32), want["dist"].view(np.uint32))
    # remove the entry point itself: it is still traversed, only withheld from the results (index.hpp:3529-3533)
    blob = ref.save()
    off = 8 + 600 * 8 * 4 + 64
    entry = int.from_bytes(blob[off + 32:off + 40], "little")
    assert ref.remove(entry) == 1 and ix.remove(entry) == 1
    want = ref.search_batch(q, 10, 64)
    keys, dist, _ = ix.ef_search(q, 10, 64)
    np.testing.assert_array_equal(keys, want["keys"])
    assert not (keys == entry).any()


def test_searches_issued_on_different_streams_do_not_share_scratch_at_the_same_time():
    """gvss_search_device is asynchronous: two batches issued back to back on two streams use the same visited tables and
    work counter, so the library orders them on the device (event wait). Results must equal the one-at-a-time results."""
    import torch
    n, d, nq, k, ef = 30000, 64, 3000, 10, 96
    base, q = synth.clustered(n, d, 300, 8, 321, 2 * nq)
    ix = gv.Index(d, "l2sq", M=16)
    ix.reserve(n, 1)
    ix.add_batch(base)
    ix.flush()
    dev = torch.device("cuda", 0)
    qa, qb = torch.from_numpy(q[:nq]).to(dev), torch.from_numpy(q[nq:]).to(dev)
    outs = [torch.empty((nq, k), dtype=torch.int64, device=dev) for _ in range(4)]
    dists = [torch.empty((nq, k), dtype=torch.float32, device=dev) for _ in range(4)]
    cnts = [torch.empty(nq, dtype=torch.int32, device=dev) for _ in range(4)]
    torch.cuda.synchronize()
    # one at a time
    for i, qq in enumerate((qa, qb)):
        ix.search_device(qq.data_ptr(), nq, k, ef, outs[i].data_ptr(), dists[i].data_ptr(), cnts[i].data_ptr(), 0, 0)
        torch.cuda.synchronize()
    # back to back on two streams, several rounds
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    for _ in range(5):
        ix.search_device(qa.data_ptr(), nq, k, ef, outs[2].data_ptr(), dists[2].data_ptr(), cnts[2].data_ptr(), 0, s1.cuda_stream)
        ix.search_device(qb.data_ptr(), nq, k, ef, outs[3].data_ptr(), dists[3].data_ptr(), cnts[3].data_ptr(), 0, s2.cuda_stream)
    torch.cuda.synchronize()
    assert torch.equal(outs[0], outs[2]) and torch.equal(outs[1], outs[3])
    assert torch.equal(dists[0], dists[2]) and torch.equal(dists[1], dists[3])


def test_concurrent_host_threads_searching_one_index():
    """InitializeScan runs under the extension's SHARED lock (hnsw_index.cpp:301-303): several host threads may call
    gvss_search on one index at the same time. The library serialises them; every thread gets the single-thread answer."""
    import threading
    n, d = 20000, 48
    base, q = synth.clustered(n, d, 200, 8, 77, 800)
    ix = gv.Index(d, "cosine", M=16)
    ix.reserve(n, 1)
    ix.add_batch(base)
    ix.flush()
    want = [ix.ef_search(q[i * 100:(i + 1) * 100], 10, 64)[0] for i in range(8)]
    got, errs = [None] * 8, []

    def work(i):
        try:
            for _ in range(3):
                got[i] = ix.ef_search(q[i * 100:(i + 1) * 100], 10, 64)[0]
        except Exception as e:  # pragma: no cover
            errs.append(e)
    ts = [threading.Thread(target=work, args=(i,)) for i in range(8)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errs
    for i in range(8):
        np.testing.assert_array_equal(got[i], want[i])
