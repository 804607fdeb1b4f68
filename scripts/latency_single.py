"""Single-query latency of the HNSW_INDEX_SCAN path (one gvss_search call per query, host pointers) next to the compiled
reference's single-thread ef_search on the same GPU-built graph, and the rate T host threads reach when each of them issues
single queries (concurrent index scans under the extension's shared lock): python scripts/latency_single.py [n] [ef] [threads]"""
import os, sys, time, threading
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
ef = int(sys.argv[2]) if len(sys.argv) > 2 else 64
T = int(sys.argv[3]) if len(sys.argv) > 3 else 16
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
    # T host threads, one query per call each
    per = 150
    def work(t):
        for i in range(per):
            ix.ef_search(q[(t * 7 + i) % 400:(t * 7 + i) % 400 + 1], 10, ef)
    ts = [threading.Thread(target=work, args=(t,)) for t in range(T)]
    t0 = time.perf_counter()
    [t.start() for t in ts]
    [t.join() for t in ts]
    dt = time.perf_counter() - t0
    line += f" | {T} host threads x single queries: {T * per / dt:.0f} QPS (peak concurrent searches {ix.info()['concurrent_searches_peak']})"
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
        qq = np.concatenate([q] * 6)
        t0 = time.perf_counter()
        r.search_batch(qq, 10, ef, threads=T)
        line += f"; reference {T} threads: {len(qq) / (time.perf_counter() - t0):.0f} QPS"
    except Exception as e:
        line += f" | no reference here ({e})"
    print(line, flush=True)
    ix.close()
