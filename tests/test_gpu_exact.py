"""GPU parity #2: exact brute-force top-k ids and distances are bit-exact against DuckDB's array_distance /
array_cosine_distance / array_negative_inner_product folds ordered by (distance, rowid); plus the top-k merge."""
import numpy as np
import pytest
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth
from conftest import golden_cases, load_golden

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", golden_cases())
def test_exact_topk_matches_golden_duckdb_folds(name):
    z, _ = load_golden(name)
    ids, dist = gv.exact_topk(str(z["metric"]), z["base"], z["queries"], 10)
    np.testing.assert_array_equal(ids, z["exact_ids"])
    np.testing.assert_array_equal(dist.view(np.uint32), z["exact_dist"].view(np.uint32))


def test_duckdb_known_answers():
    f = lambda *v: np.array([v], np.float32)
    base = np.array([[1, 2, 3], [2, 2, 3], [7, 8, 9], [-1, -2, -3], [0, 0, 0]], np.float32)
    ids, dist = gv.exact_topk("l2sq", base, f(1, 2, 3), 3)     # array_distance.test:20-27
    assert list(ids[0]) == [0, 1, 4] and dist[0][0] == 0.0 and dist[0][1] == 1.0
    ids, dist = gv.exact_topk("cosine", base, f(1, 2, 3), 5)   # array_cosine_distance.test:26-33 (+ zero vector -> 2.0)
    assert ids[0][0] == 0 and abs(dist[0][0]) < 1e-7 and dist[0][-1] == 2.0 and set(ids[0][-2:]) == {3, 4}
    ids, dist = gv.exact_topk("ip", base, f(1, 2, 3), 2)       # array_inner_product.test:36-43 (negated)
    assert list(ids[0]) == [2, 1] and list(dist[0]) == [-50.0, -15.0]


@pytest.mark.parametrize("metric,d,n,nq,k", [("l2sq", 128, 50000, 333, 10), ("cosine", 768, 3000, 70, 10), ("ip", 96, 20000, 17, 100), ("l2sq", 3, 729, 5, 7)])
def test_exact_topk_parity_against_checker(metric, d, n, nq, k):
    import oracle
    base, q = synth.clustered(n, d, 50, 8, 400 + d, nq) if d > 3 else (synth.readme_grid(), np.array([[1, 2, 3], [5, 5, 5], [9, 9, 9], [0, 0, 0], [4.5, 4.5, 4.5]], np.float32))
    if oracle.have_ref():
        from oracle import ref
        want_ids, want_d = ref.duckdb_topk(metric, base, q, k, threads=8)
    else:
        from oracle import port
        want_ids, want_d = port.exact_topk(metric, base, q, k)
    ids, dist = gv.exact_topk(metric, base, q, k)
    np.testing.assert_array_equal(ids, want_ids)
    np.testing.assert_array_equal(dist.view(np.uint32), want_d.view(np.uint32))


def test_exact_topk_key_mapping_and_small_inputs():
    base = synth.readme_grid()
    keys = np.arange(729, dtype=np.int64) * 10 + 5
    ids, dist = gv.exact_topk("l2sq", base, np.array([[1, 2, 3]], np.float32), 3, keys=keys)
    assert list(ids[0]) == [115, 25, 105]
    ids, dist = gv.exact_topk("l2sq", base[:2], np.array([[1, 1, 1]], np.float32), 5)   # k > n: tail = -1 / inf
    assert list(ids[0]) == [0, 1, -1, -1, -1] and np.isinf(dist[0][2:]).all()


def test_merge_topk_orders_by_distance_then_key():
    import torch
    rng = np.random.default_rng(3)
    shards, nq, k = 4, 50, 10
    d = np.sort(rng.integers(0, 30, size=(shards, nq, k)).astype(np.float32), axis=2)   # many ties across shards
    keys = rng.permutation(shards * nq * k).reshape(shards, nq, k).astype(np.int64)
    d[3, :, 7:] = np.inf
    keys[3, :, 7:] = -1                                                                  # a short list
    gk, gd = torch.from_numpy(keys).cuda(), torch.from_numpy(d).cuda()
    ok = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    od = torch.empty((nq, k), dtype=torch.float32, device="cuda")
    gv.merge_topk_device(gk.data_ptr(), gd.data_ptr(), shards, nq, k, ok.data_ptr(), od.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    for qi in range(nq):
        cand = sorted((float(d[s, qi, i]), int(keys[s, qi, i])) for s in range(shards) for i in range(k) if keys[s, qi, i] >= 0)[:k]
        assert [c[1] for c in cand] == ok[qi].tolist()
        assert [c[0] for c in cand] == od[qi].tolist()


def _tc_exact(metric, base, q, k, want_dots=False):
    """Drives the tensor-core path directly (test hook gvssdbg_tc_exact, not part of the public ABI)."""
    import ctypes as C
    import torch
    from duckdb_vss_b200.index import METRIC_CHAR
    fn = gv.lib().gvssdbg_tc_exact
    fn.restype = C.c_int
    fn.argtypes = [C.c_char, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p,
                   C.c_void_p, C.POINTER(C.c_uint64)]
    b, qq = torch.from_numpy(base).cuda(), torch.from_numpy(q).cuda()
    n, nq = base.shape[0], q.shape[0]
    ids = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    dist = torch.empty((nq, k), dtype=torch.float32, device="cuda")
    dots = torch.zeros((nq, n), dtype=torch.float32, device="cuda") if want_dots else None
    flagged = C.c_uint64(0)
    rc = fn(METRIC_CHAR[metric], b.data_ptr(), n, qq.data_ptr(), nq, base.shape[1], k, ids.data_ptr(), dist.data_ptr(),
            dots.data_ptr() if want_dots else None, C.byref(flagged))
    assert rc == 0, gv.lib().gvss_last_error().decode()
    return ids.cpu().numpy(), dist.cpu().numpy(), (dots.cpu().numpy() if want_dots else None), flagged.value


def _checker_topk(metric, base, q, k):
    import oracle
    if oracle.have_ref():
        from oracle import ref
        return ref.duckdb_topk(metric, base, q, k, threads=8)
    from oracle import port
    return port.exact_topk(metric, base, q, k)


@pytest.mark.parametrize("metric,d,n,nq", [("l2sq", 128, 30000, 257), ("cosine", 768, 4000, 130), ("ip", 96, 9000, 64), ("l2sq", 20, 1500, 40)])
def test_tensor_core_candidates_plus_rerank_are_bit_exact(metric, d, n, nq):
    base, q = synth.clustered(n, d, 40, 8, 900 + d, nq)
    if metric == "ip":
        base /= np.linalg.norm(base, axis=1, keepdims=True)
    ids, dist, dots, flagged = _tc_exact(metric, base, q, 10, want_dots=n <= 9000)
    want_ids, want_d = _checker_topk(metric, base, q, 10)
    np.testing.assert_array_equal(ids, want_ids)
    np.testing.assert_array_equal(dist.view(np.uint32), want_d.view(np.uint32))
    if dots is not None:   # 3xTF32 holds fp32 tolerance: error relative to |q||b| ~1e-6..1e-5 (grows with d through the
        # tensor core's fp32 accumulation), far below a single TF32 pass (~5e-4)
        ref_dots = q.astype(np.float64) @ base.astype(np.float64).T
        scale = np.linalg.norm(q, axis=1)[:, None] * np.linalg.norm(base, axis=1)[None, :]
        assert (np.abs(dots - ref_dots) / scale).max() < 3e-5
    assert flagged <= nq // 20


def test_tensor_core_path_falls_back_on_ties_it_cannot_resolve():
    # integer grid {0,1,2}^6 padded to 8 dims, k = 16: from the centre point 12 rows sit at distance 1 and 60 at sqrt(2),
    # so ranks 14..16 are taken from 60 tied rows — more than the 32 candidate slots. The candidate set cannot be proven
    # sufficient -> the query is recomputed by the all-pairs kernel; the result is still exact
    g = np.array(np.meshgrid(*[np.arange(3)] * 6, indexing="ij")).reshape(6, -1).T.astype(np.float32)
    base = np.concatenate([g, np.zeros((g.shape[0], 2), np.float32)], axis=1)
    q = base[[0, 364, 728]].copy()
    ids, dist, _, flagged = _tc_exact("l2sq", base, q, 16)
    want_ids, want_d = _checker_topk("l2sq", base, q, 16)
    np.testing.assert_array_equal(ids, want_ids)
    np.testing.assert_array_equal(dist.view(np.uint32), want_d.view(np.uint32))
    assert flagged >= 1


def test_peer_merge_kernel_equals_gathered_merge_single_process():
    """gvss_merge_topk_peers_device with both 'peers' in this process (flags pre-set, so nobody waits): same output as
    gvss_merge_topk_device on the gathered copy. The cross-process use (symmetric memory, one flag per pair of ranks)
    is exercised by bench.py --gpus N."""
    import torch
    from duckdb_vss_b200 import index as gvi
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(3)
    shards, nq, k = 3, 500, 10
    d = np.sort(rng.random((shards, nq, k), dtype=np.float32), axis=2)
    d[1, :, 7:] = np.inf                                    # short lists: padding (key -1, +inf)
    keys = rng.permutation(shards * nq * k).astype(np.int64).reshape(shards, nq, k)
    keys[1, :, 7:] = -1
    d[2, :, 0] = d[0, :, 0]                                 # exact distance ties across shards: ordered by key
    gk, gd = torch.from_numpy(keys).to(dev), torch.from_numpy(d).to(dev)
    want_k = torch.empty((nq, k), dtype=torch.int64, device=dev)
    want_d = torch.empty((nq, k), dtype=torch.float32, device=dev)
    gvi.merge_topk_device(gk.data_ptr(), gd.data_ptr(), shards, nq, k, want_k.data_ptr(), want_d.data_ptr())
    flags = [torch.full((16,), 5, dtype=torch.int32, device=dev) for _ in range(shards)]
    got_k, got_d = torch.empty_like(want_k), torch.empty_like(want_d)
    for rank in range(shards):
        got_k.fill_(-7)
        gvi.merge_topk_peers_device([gk[s].data_ptr() for s in range(shards)], [gd[s].data_ptr() for s in range(shards)],
                                    [f.data_ptr() for f in flags], rank, 5, nq, k, got_k.data_ptr(), got_d.data_ptr())
        torch.cuda.synchronize()
        assert torch.equal(got_k, want_k) and torch.equal(got_d, want_d)
    with pytest.raises(gv.GvssError, match="256"):
        gvi.merge_topk_peers_device([gk.data_ptr()] * 16, [gd.data_ptr()] * 16, [flags[0].data_ptr()] * 16, 0, 6, nq, 32,
                                    got_k.data_ptr(), got_d.data_ptr())


def test_exact_search_over_the_rows_an_index_holds(checker):
    """gvss_exact_search: brute force over the index's own HBM-resident rows (zero-padded rows, d = 12 -> 16) equals
    gvss_exact_topk over the host copy, for all three metrics, and skips removed members."""
    n, d, k = 7000, 12, 10
    base, q = synth.clustered(n, d, 70, 4, 61, 120)
    keys = np.arange(n, dtype=np.int64) * 3 + 5
    for metric in ("l2sq", "cosine", "ip"):
        ix = gv.Index(d, metric, M=8)
        ix.reserve(n, 1)
        ix.add_batch(base, keys=keys)
        ix.flush()
        got_k, got_d, got_c = ix.exact_search(q, k)
        want_k, want_d = gv.exact_topk(metric, base, q, k, keys=keys)
        np.testing.assert_array_equal(got_k, want_k)
        np.testing.assert_array_equal(got_d.view(np.uint32), want_d.view(np.uint32))
        assert (got_c == k).all()
        gone = want_k[:, 0][:20]
        for key in set(int(v) for v in gone):
            assert ix.remove(key) == 1
        got2, _, c2 = ix.exact_search(q, k)
        assert not np.isin(got2, gone).any() and (c2 == k).all()
        live = ~np.isin(keys, gone)
        want2, _ = gv.exact_topk(metric, base[live], q, k, keys=keys[live])
        np.testing.assert_array_equal(got2, want2)


def test_rows_only_index_answers_exact_search_without_a_graph():
    """gvss_set_rows_only (un-indexed tables): the flush uploads and stops; exact_search equals gvss_exact_topk over the
    host copy; the graph calls refuse instead of walking empty lists."""
    n, d, k = 20000, 24, 7
    base, q = synth.clustered(n, d, 200, 4, 62, 64)
    keys = np.arange(n, dtype=np.int64) * 2 + 1
    ix = gv.Index(d, "cosine", M=16)
    ix.set_rows_only()
    ix.reserve(n, 2)
    ix.add_batch(base[:n // 2], keys=keys[:n // 2])
    ix.flush()
    ix.add_batch(base[n // 2:], keys=keys[n // 2:])      # second flush appends to the resident rows
    got_k, got_d, got_c = ix.exact_search(q, k)
    want_k, want_d = gv.exact_topk("cosine", base, q, k, keys=keys)
    np.testing.assert_array_equal(got_k, want_k)
    np.testing.assert_array_equal(got_d.view(np.uint32), want_d.view(np.uint32))
    assert (got_c == k).all() and ix.size() == n
    with pytest.raises(gv.GvssError, match="rows only"):
        ix.ef_search(q, k, 64)
    with pytest.raises(gv.GvssError, match="rows only"):
        ix.save()
    assert ix.remove(int(want_k[0, 0])) == 1
    got2, _, _ = ix.exact_search(q[:1], k)
    assert int(want_k[0, 0]) not in got2
    iy = gv.Index(d, "cosine", M=16)
    iy.reserve(10, 1)
    iy.add_batch(base[:10])
    with pytest.raises(gv.GvssError, match="already has members"):
        iy.set_rows_only()
