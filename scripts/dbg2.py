import os, sys, subprocess
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth
from oracle import ref
n, d, M = 20000, 128, 16
base, q = synth.clustered(n, d, max(16, n // 100), 8, 100 + d, 300)
os.makedirs("gpurun_out", exist_ok=True)
for trial in range(8):
    r = ref.RefIndex(d, "l2sq", M=M); r.reserve(n, 8); r.add_bulk(base, threads=8)
    blob = r.save()
    ix = gv.Index(d, "l2sq", M=M); ix.load(blob)
    for ef in (64, 128):
        want = r.search_batch(q, 10, ef)
        keys, dist, counts, st = ix.ef_search(q, 10, ef, stats=True)
        bad = np.where((st["computed"] != want["computed"].astype(np.uint32)) | (st["visited"] != want["visited"].astype(np.uint32)))[0]
        if len(bad):
            qi = int(bad[0])
            print("mismatch trial", trial, "ef", ef, "query", qi, st["computed"][qi], want["computed"][qi], st["visited"][qi], want["visited"][qi], "flags", st["flags"][qi])
            open("gpurun_out/bad.usearch", "wb").write(blob)
            np.save("gpurun_out/bad_q.npy", q[qi]); open("gpurun_out/bad_ef.txt", "w").write(str(ef))
            os.environ["GVSS_TRACE_QUERY"] = "0"
            sys.stdout.flush()
            ix.ef_search(q[qi:qi+1], 10, ef)
            import ctypes; 
            sys.exit(0)
print("no mismatch")
