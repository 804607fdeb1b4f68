"""CPU: pins the plain-C oracle (oracle/hnsw_oracle.c) against (a) the golden fixtures written by the compiled
reference, (b) the compiled reference itself when it is present, (c) known answers of the reference's / DuckDB's
own tests."""
import numpy as np
import pytest
import oracle
from oracle import port
from duckdb_vss_b200 import synth
from conftest import golden_cases, load_golden


@pytest.mark.parametrize("name", golden_cases())
def test_port_loads_golden_stream_and_searches_identically(name):
    z, blob = load_golden(name)
    ix = port.PortIndex(z["base"].shape[1], str(z["metric"]), M=int(z["M"]), M0=int(z["M0"]), ef_construction=int(z["efc"]))
    ix.load(blob)
    assert ix.save() == blob  # parse -> re-serialise is the identity
    for ef in (3, 10, 64, 128):
        r = ix.search_batch(z["queries"], 10, ef)
        np.testing.assert_array_equal(r["keys"], z[f"keys_ef{ef}"])
        np.testing.assert_array_equal(r["dist"].view(np.uint32), z[f"dist_ef{ef}"].view(np.uint32))
        np.testing.assert_array_equal(r["computed"], z[f"computed_ef{ef}"])
        np.testing.assert_array_equal(r["visited"], z[f"visited_ef{ef}"])


@pytest.mark.parametrize("name", golden_cases())
def test_port_single_thread_build_is_byte_identical_to_reference_stream(name):
    z, blob = load_golden(name)
    ix = port.PortIndex(z["base"].shape[1], str(z["metric"]), M=int(z["M"]), M0=int(z["M0"]), ef_construction=int(z["efc"]))
    ix.reserve(z["base"].shape[0])
    ix.add_bulk(z["base"])
    assert ix.save() == blob


@pytest.mark.parametrize("name", golden_cases())
def test_port_exact_topk_matches_duckdb_folds(name):
    z, _ = load_golden(name)
    ids, dist = port.exact_topk(str(z["metric"]), z["base"], z["queries"], 10)
    np.testing.assert_array_equal(ids, z["exact_ids"])
    np.testing.assert_array_equal(dist.view(np.uint32), z["exact_dist"].view(np.uint32))


def test_readme_known_answer():
    # hnsw_result.test:20-25 — distances 0.0, 1.0, 1.0 for query [1,2,3] (array_distance = sqrt of l2sq)
    g = synth.readme_grid()
    ix = port.PortIndex(3)
    ix.reserve(729)
    ix.add_bulk(g)
    keys, dist, _, _ = ix.search(np.array([1, 2, 3], np.float32), 3, 64)
    assert list(np.sqrt(dist)) == [0.0, 1.0, 1.0]
    assert list(keys) == [11, 92, 12]  # SURVEY.md §8c probe of the compiled reference
    assert len(ix.save()) == 117372    # SURVEY.md Appendix B worked example


def test_duckdb_fold_known_answers():
    # duckdb/test/sql/function/array/array_distance.test:20-27
    f = lambda *v: np.array(v, np.float32)
    assert port.duckdb_distance("l2sq", f(1, 2, 3), f(1, 2, 3)) == 0.0
    assert port.duckdb_distance("l2sq", f(1, 2, 3), f(2, 2, 3)) == 1.0
    assert abs(port.duckdb_distance("l2sq", f(1, 2, 3), f(7, 8, 9)) - 10.392304) < 1e-6
    # array_cosine_distance.test:26-33: identical direction -> 0, opposite -> 2
    assert port.duckdb_distance("cosine", f(1, 2, 3), f(1, 2, 3)) == pytest.approx(0.0, abs=1e-7)
    assert port.duckdb_distance("cosine", f(1, 2, 3), f(-1, -2, -3)) == pytest.approx(2.0, abs=1e-6)
    # array_inner_product.test:36-43 (negative inner product = -dot)
    assert port.duckdb_distance("ip", f(1, 2, 3), f(1, 2, 3)) == -14.0
    assert port.duckdb_distance("ip", f(1, 2, 3), f(2, 3, 4)) == -20.0
    # zero vector: DuckDB's clamp turns the NaN similarity into -1 -> distance 2 (array_kernels.hpp:62-63)
    assert port.duckdb_distance("cosine", f(0, 0, 0), f(1, 2, 3)) == 2.0


def test_usearch_metric_special_cases():
    f = lambda *v: np.array(v, np.float32)
    assert port.metric("cosine", f(0, 0), f(0, 0)) == 0.0   # index_plugins.hpp:1021-1025
    assert port.metric("cosine", f(0, 0), f(1, 0)) == 1.0
    assert port.metric("ip", f(1, 2), f(3, 4)) == 1 - 11.0  # 1 - ab, not -ab (index_plugins.hpp:992)


def test_level_law_matches_reference_levels():
    # the levels stored in the README golden stream came from std::default_random_engine in the reference
    z, blob = load_golden("readme_l2sq")
    n = 729
    off = 8 + n * 12 + 64 + 40
    levels = np.frombuffer(blob[off:off + 2 * n], dtype=np.int16)
    import ctypes as C
    st = C.c_uint64(1)
    mine = [port.lib().orc_level_from_state(C.byref(st), 16) for _ in range(n)]
    assert list(levels) == mine


@pytest.mark.skipif(not oracle.have_ref(), reason="compiled reference not present")
@pytest.mark.parametrize("metric,d,n,M,efc", [("l2sq", 8, 1500, 16, 128), ("cosine", 20, 800, 5, 50), ("ip", 6, 800, 3, 100)])
def test_port_vs_compiled_reference_random(metric, d, n, M, efc):
    from oracle import ref
    rng = np.random.default_rng(d)
    x = rng.standard_normal((n, d)).astype(np.float32)
    r = ref.RefIndex(d, metric, M=M, ef_construction=efc)
    r.reserve(n, 1)
    r.add_bulk(x)
    p = port.PortIndex(d, metric, M=M, ef_construction=efc)
    p.reserve(n)
    p.add_bulk(x)
    assert r.save() == p.save()
    # lazy deletes survive in both and are never returned (hnsw_crud.test:12-47)
    for key in (3, 77, 500):
        assert r.remove(key) == 1 and p.remove(key) == 1
    p2 = port.PortIndex(d, metric, M=M, ef_construction=efc)
    p2.load(p.save())
    q = rng.standard_normal((60, d)).astype(np.float32)
    for ef in (10, 64):
        a, b = r.search_batch(q, 10, ef), p2.search_batch(q, 10, ef)
        np.testing.assert_array_equal(a["keys"], b["keys"])
        np.testing.assert_array_equal(a["dist"].view(np.uint32), b["dist"].view(np.uint32))
        assert not np.isin(a["keys"], [3, 77, 500]).any()
    ids_r, d_r = ref.duckdb_topk(metric, x, q, 7)
    ids_p, d_p = port.exact_topk(metric, x, q, 7)
    np.testing.assert_array_equal(ids_r, ids_p)
    np.testing.assert_array_equal(d_r.view(np.uint32), d_p.view(np.uint32))
