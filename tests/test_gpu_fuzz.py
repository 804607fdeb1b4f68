"""GPU parity fuzz: random shapes through every staging / queue path of k_search (rows shorter and longer than one
128-byte line, partial last lines, batches of <= 8 / <= 16 / 32 rows, tie-free and tie-heavy data, k above and below
ef) against the compiled reference on the reference's own graph: ids, distance bits and traversal counters."""
import numpy as np
import pytest
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth

pytestmark = pytest.mark.gpu

CASES = []
_rng = np.random.default_rng(20261107)
for i in range(14):
    CASES.append(dict(seed=int(_rng.integers(1, 1 << 30)), d=int(_rng.choice([8, 31, 32, 33, 40, 64, 96, 100, 128, 160, 257])),
                      metric=str(_rng.choice(["l2sq", "cosine", "ip"])), M=int(_rng.choice([3, 4, 8, 16])),
                      M0=int(_rng.choice([0, 0, 7, 24])), n=int(_rng.integers(150, 2500)), integer=bool(i % 2)))


@pytest.mark.parametrize("case", CASES, ids=lambda c: f"{c['metric']}-d{c['d']}-M{c['M']}-{c['M0']}-n{c['n']}-{'int' if c['integer'] else 'real'}")
def test_random_shape_matches_reference(case, checker):
    Checker, kind = checker
    rng = np.random.default_rng(case["seed"])
    n, d = case["n"], case["d"]
    if case["integer"]:
        base = rng.integers(0, 3, size=(n, d)).astype(np.float32)
        q = rng.integers(0, 3, size=(60, d)).astype(np.float32)
        base[:, 0] += 1.0
        q[:, 0] += 1.0
    else:
        base, q = synth.clustered(n, d, max(4, n // 60), min(6, d), case["seed"] % 1000, 60)
    M0 = max(case["M0"], case["M"]) if case["M0"] else 0   # M0 < M is out of bounds in the reference (DESIGN.md deviation 2)
    ref = Checker(d, case["metric"], M=case["M"], M0=M0, ef_construction=40)
    ref.reserve(n, 4)
    ref.add_bulk(base, threads=4)
    ix = gv.Index(d, case["metric"], M=case["M"], M0=M0, ef_construction=40)
    ix.load(ref.save())
    for k, ef in ((1, 1), (3, 5), (10, 17), (25, 64), (10, 200)):
        want = ref.search_batch(q, k, ef)
        keys, dist, counts, st = ix.ef_search(q, k, ef, stats=True)
        msg = f"{case} k={k} ef={ef} ({kind})"
        np.testing.assert_array_equal(keys, want["keys"], err_msg=msg)
        np.testing.assert_array_equal(dist.view(np.uint32), want["dist"].view(np.uint32), err_msg=msg)
        np.testing.assert_array_equal(counts, want["counts"], err_msg=msg)
        np.testing.assert_array_equal(st["computed"], want["computed"].astype(np.uint32), err_msg=msg)
        np.testing.assert_array_equal(st["visited"], want["visited"].astype(np.uint32), err_msg=msg)
        assert (st["flags"] & 3 == 0).all()


BUILD_CASES = []
for i in range(8):
    BUILD_CASES.append(dict(seed=int(_rng.integers(1, 1 << 30)), d=int(_rng.choice([5, 24, 32, 48, 100, 128, 200])),
                            metric=str(_rng.choice(["l2sq", "cosine", "ip"])), M=int(_rng.choice([3, 4, 8, 16])),
                            M0=int(_rng.choice([0, 0, 9, 40])), n=int(_rng.integers(200, 900)), efc=int(_rng.choice([8, 40, 128])),
                            integer=bool(i % 2)))


@pytest.mark.parametrize("case", BUILD_CASES, ids=lambda c: f"{c['metric']}-d{c['d']}-M{c['M']}-{c['M0']}-efc{c['efc']}-n{c['n']}-{'int' if c['integer'] else 'real'}")
def test_random_shape_sequential_build_is_the_reference_stream(case, checker):
    """One insert in flight: the GPU build (list-as-queue beam search with heap replay on ties, batched neighbour selection,
    reverse links) writes the reference's single-thread stream byte for byte — also on tie-heavy integer data."""
    Checker, kind = checker
    if kind != "reference":
        pytest.skip("needs the compiled reference")
    rng = np.random.default_rng(case["seed"])
    n, d = case["n"], case["d"]
    if case["integer"]:
        base = rng.integers(0, 3, size=(n, d)).astype(np.float32)
        base[:, 0] += 1.0
    else:
        base, _ = synth.clustered(n, d, max(4, n // 60), min(6, d), case["seed"] % 1000, 0)
    M0 = max(case["M0"], case["M"]) if case["M0"] else 0
    ref = Checker(d, case["metric"], M=case["M"], M0=M0, ef_construction=case["efc"])
    ref.reserve(n, 1)
    ref.add_bulk(base, threads=1)
    ix = gv.Index(d, case["metric"], M=case["M"], M0=M0, ef_construction=case["efc"])
    ix.reserve(n, 1)
    ix.set_build_concurrency(1)
    ix.add_batch(base)
    ix.flush()
    assert ix.save() == ref.save(), str(case)


@pytest.mark.parametrize("case", CASES[:8], ids=lambda c: f"{c['metric']}-d{c['d']}-M{c['M']}-{c['M0']}-n{c['n']}-{'int' if c['integer'] else 'real'}")
def test_random_shape_with_removed_members_matches_reference(case, checker):
    """A third of the members removed (lazy delete): they are still traversed — in list-as-queue mode from a side heap —
    and never returned; ids, distance bits and traversal counters equal the reference's, entry point removed or not."""
    Checker, kind = checker
    rng = np.random.default_rng(case["seed"] + 1)
    n, d = case["n"], case["d"]
    if case["integer"]:
        base = rng.integers(0, 3, size=(n, d)).astype(np.float32)
        q = rng.integers(0, 3, size=(60, d)).astype(np.float32)
        base[:, 0] += 1.0
        q[:, 0] += 1.0
    else:
        base, q = synth.clustered(n, d, max(4, n // 60), min(6, d), case["seed"] % 1000 + 1, 60)
    M0 = max(case["M0"], case["M"]) if case["M0"] else 0
    ref = Checker(d, case["metric"], M=case["M"], M0=M0, ef_construction=40)
    ref.reserve(n, 1)
    ref.add_bulk(base, threads=1)
    ix = gv.Index(d, case["metric"], M=case["M"], M0=M0, ef_construction=40)
    ix.load(ref.save())
    gone = rng.choice(n, size=n // 3, replace=False)
    for key in gone:
        assert ref.remove(int(key)) == 1
        assert ix.remove(int(key)) == 1
    for k, ef in ((1, 1), (3, 5), (10, 17), (10, 64), (10, 200)):
        want = ref.search_batch(q, k, ef)
        keys, dist, counts, st = ix.ef_search(q, k, ef, stats=True)
        msg = f"{case} k={k} ef={ef} ({kind})"
        np.testing.assert_array_equal(keys, want["keys"], err_msg=msg)
        np.testing.assert_array_equal(dist.view(np.uint32), want["dist"].view(np.uint32), err_msg=msg)
        np.testing.assert_array_equal(counts, want["counts"], err_msg=msg)
        np.testing.assert_array_equal(st["computed"], want["computed"].astype(np.uint32), err_msg=msg)
        np.testing.assert_array_equal(st["visited"], want["visited"].astype(np.uint32), err_msg=msg)
        assert not np.isin(keys, gone).any()
