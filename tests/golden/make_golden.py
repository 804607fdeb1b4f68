"""Regenerates the golden fixtures in this directory from the COMPILED REFERENCE (oracle/_ref, i.e. the unmodified
vendored usearch + DuckDB array kernels under /root/reference). Run in the build container only:

    python tests/golden/make_golden.py [case ...]

Fixtures (all small):
  <name>.usearch   — stream written by the reference's save_to_stream after a single-thread build
  <name>.npz       — inputs (base, queries) and the reference's answers (keys, dist, computed, visited) per ef,
                     plus DuckDB-fold brute-force top-k (ids, dist) ordered by (distance, rowid)
"""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import ref  # noqa: E402
from duckdb_vss_b200 import synth  # noqa: E402

CASES = {
    # name: (metric, base, queries, M, M0, efc)
    "readme_l2sq": ("l2sq", synth.readme_grid(), np.array([[1, 2, 3], [5, 5, 5], [9.5, 0.5, 4.25]], np.float32), 16, 0, 128),
}
rng = np.random.default_rng(7)
b, q = synth.clustered(1200, 16, 24, 6, 11, 40)
CASES["clustered16_l2sq"] = ("l2sq", b, q, 16, 0, 128)
b, q = synth.clustered(900, 24, 16, 6, 12, 40)
b[5] = 0.0  # a zero vector: cosine special cases (index_plugins.hpp:1021-1025)
CASES["clustered24_cos"] = ("cosine", b, q, 8, 0, 64)
b, q = synth.clustered(900, 12, 16, 4, 13, 40)
b /= np.linalg.norm(b, axis=1, keepdims=True)
CASES["clustered12_ip"] = ("ip", b, q, 4, 6, 40)
b, q = synth.clustered(1000, 40, 20, 8, 14, 40)  # rows longer than one 128-byte line (second line partial): the GPU's gather4 staging path
CASES["clustered40_l2sq"] = ("l2sq", b, q, 16, 0, 128)
only = set(sys.argv[1:])  # optional: regenerate only the named cases

for name, (metric, base, queries, M, M0, efc) in CASES.items():
    if only and name not in only:
        continue
    ix = ref.RefIndex(base.shape[1], metric, M=M, M0=M0, ef_construction=efc)
    ix.reserve(base.shape[0], 1)
    ix.add_bulk(base, threads=1)
    blob = ix.save()
    open(os.path.join(HERE, name + ".usearch"), "wb").write(blob)
    out = {"base": base, "queries": queries, "M": M, "M0": M0, "efc": efc, "metric": metric}
    for ef in (3, 10, 64, 128):
        r = ix.search_batch(queries, 10, ef)
        out[f"keys_ef{ef}"] = r["keys"]
        out[f"dist_ef{ef}"] = r["dist"]
        out[f"computed_ef{ef}"] = r["computed"]
        out[f"visited_ef{ef}"] = r["visited"]
    ids, dist = ref.duckdb_topk(metric, base, queries, 10)
    out["exact_ids"], out["exact_dist"] = ids, dist
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, len(blob), "bytes; max_level", ix.max_level())

# known answers of DuckDB's own tests for the exact folds (duckdb/test/sql/function/array/array_distance.test:20-27,
# array_cosine_distance.test:26-33, array_inner_product.test:36-43) are hard-coded in tests/test_oracle.py.
