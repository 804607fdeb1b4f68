"""Config check (SURVEY.md App. D last paragraph): is recall@10 >= 0.95 reachable at 1M with M=16/efc=128, on the
reference-built graph and on the GPU-built graph? Otherwise which (M, efc) is."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth
from oracle import ref
n, d, nq = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000, 128, 2000
base, _ = synth.clustered(n, d, max(1024, n // 100), 16, 42, 0)
_, q = synth.clustered(1, d, max(1024, n // 100), 16, 42, nq)
truth, _ = gv.exact_topk("l2sq", base, q, 10)
T = ref.hardware_threads()
for (M, efc) in ((16, 128), (32, 200)):
    ix = gv.Index(d, "l2sq", M=M, ef_construction=efc); ix.reserve(n, 1); ix.add_batch(base)
    t = time.time(); ix.flush(); tg = time.time() - t
    rg = {ef: round(synth.recall_at_k(ix.ef_search(q, 10, ef)[0], truth), 4) for ef in (64, 128, 256, 512)}
    print(f"M={M} efc={efc} GPU build {tg:.2f}s ({n/tg:.0f} vec/s) recall {rg}", flush=True)
    if M == 16 or len(sys.argv) > 2:
        r = ref.RefIndex(d, "l2sq", M=M, ef_construction=efc); r.reserve(n, T)
        t = time.time(); r.add_bulk(base, threads=T); tr = time.time() - t
        rr = {ef: round(synth.recall_at_k(r.search_batch(q, 10, ef, threads=T)["keys"], truth), 4) for ef in (64, 128, 256, 512)}
        print(f"M={M} efc={efc} REF build {tr:.2f}s ({n/tr:.0f} vec/s, {T} threads) recall {rr}", flush=True)
