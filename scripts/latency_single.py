"""Single-query latency of the HNSW_INDEX_SCAN path (one gvss_search call per query, host pointers) next to the compiled
reference's single-thread ef_search on the same GPU-built graph: python scripts/latency_single.py [n] [ef]"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
ef = int(sys.argv[2]) if len(sys.argv) > 2 else 64
base, q = synth.clustered(n, 128, max(1024, n // 100), 16, 42, 400)
dev = torch.from_numpy(base).cuda()
for M in (16, 32):
    ix = gv.Index(128, "l2sq", M=M, ef_construction=128 if M == 16 else 200)
    ix.reserve(n, 1)
    ix.add_batch_device(dev.data_ptr(), n)
    ix.flush()
    for i in range(20):
        ix.ef_search(q[i:i + 1], 10, ef)
    lat = []
    for i in range(20, 400):
        t = time.perf_counter()
        ix.ef_search(q[i:i + 1], 10, ef)
        lat.append(time.perf_counter() - t)
    lat = np.array(lat) * 1e3
    line = f"n={n} M={M} ef={ef}: gvss_search nq=1 median {np.median(lat):.3f} ms  p95 {np.percentile(lat, 95):.3f} ms"
    try:
        from oracle import ref
        r = ref.RefIndex(128, "l2sq", M=M, ef_construction=128 if M == 16 else 200)
        r.load(ix.save())
        rl = []
        for i in range(20, 400):
            t = time.perf_counter()
            r.search_batch(q[i:i + 1], 10, ef, threads=1)
            rl.append(time.perf_counter() - t)
        rl = np.array(rl) * 1e3
        line += f" | reference (1 thread, same graph) median {np.median(rl):.3f} ms  p95 {np.percentile(rl, 95):.3f} ms"
    except Exception as e:
        line += f" | no reference here ({e})"
    print(line, flush=True)
    ix.close()
