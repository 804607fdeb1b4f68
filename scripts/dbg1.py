import os, sys, time, subprocess
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth
from oracle import ref

def search_dbg():
    n, d, M = 20000, 128, 16
    base, q = synth.clustered(n, d, max(16, n // 100), 8, 100 + d, 300)
    for trial in range(4):
        r = ref.RefIndex(d, "l2sq", M=M); r.reserve(n, 8); r.add_bulk(base, threads=8)
        blob = r.save()
        ix = gv.Index(d, "l2sq", M=M); ix.load(blob)
        for ef in (10, 64, 128):
            want = r.search_batch(q, 10, ef)
            for rep in range(2):
                keys, dist, counts, st = ix.ef_search(q, 10, ef, stats=True)
                bad = np.where((keys != want["keys"]).any(axis=1))[0]
                badc = np.where(st["computed"] != want["computed"].astype(np.uint32))[0]
                badv = np.where(st["visited"] != want["visited"].astype(np.uint32))[0]
                if len(bad) or len(badc) or len(badv):
                    print("trial", trial, "ef", ef, "rep", rep, "bad keys", bad[:5], "bad computed", badc[:5], "bad visited", badv[:5])
                    for qi in list(badc[:2]):
                        print("  q", qi, "gpu", keys[qi], st["computed"][qi], st["visited"][qi], "ref", want["keys"][qi], want["computed"][qi], want["visited"][qi])
                        print("  dist gpu", dist[qi], "ref", want["dist"][qi])
        # check for duplicate neighbours in the reference graph
    print("search dbg done")

def build_dbg():
    n, d = 30000, 128
    base, q = synth.clustered(n, d, max(64, n // 100), 16, 300 + d, 500)
    truth, _ = gv.exact_topk("l2sq", base, q, 10)
    r = ref.RefIndex(d, "l2sq"); r.reserve(n, 8); r.add_bulk(base, threads=8)
    rr = {ef: synth.recall_at_k(r.search_batch(q, 10, ef)["keys"], truth) for ef in (64, 128)}
    print("ref recall", rr)
    for ramp in (sys.argv[2:] or ["16", "64", "256", "1024"]):
        os.environ["GVSS_BUILD_RAMP"] = ramp
        ix = gv.Index(d, "l2sq"); ix.reserve(n, 1); ix.add_batch(base)
        t = time.time(); ix.flush(); t = time.time() - t
        rec = {ef: synth.recall_at_k(ix.ef_search(q, 10, ef)[0], truth) for ef in (64, 128)}
        st = ix.stats(0)
        print("ramp", ramp, "build s %.3f" % t, "vec/s %.0f" % (n / t), "recall", rec, "edges/node %.2f" % (st["edges"] / st["nodes"]))

if sys.argv[1] == "search": search_dbg()
else: build_dbg()
