"""Recall of GPU-built graphs vs build concurrency (is the residual gap to the reference due to in-flight inserts?)"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth
n, M, efc = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
base, q = synth.clustered(n, 128, max(1024, n // 100), 16, 42, 2000)
truth, _ = gv.exact_topk("l2sq", base, q, 10)
dev = torch.from_numpy(base).cuda()
for conc in [int(c) for c in sys.argv[4:]]:
    ix = gv.Index(128, "l2sq", M=M, ef_construction=efc); ix.reserve(n, 1)
    ix.set_build_concurrency(conc)
    ix.add_batch_device(dev.data_ptr(), n)
    torch.cuda.synchronize(); t = time.time(); ix.flush(); torch.cuda.synchronize(); t = time.time() - t
    rec = {ef: round(synth.recall_at_k(ix.ef_search(q, 10, ef)[0], truth), 4) for ef in (64, 128, 256)}
    st = ix.stats(0)
    print(f"n={n} M={M} efc={efc} max_concurrency={conc or 'all'} build {t:.2f}s recall {rec} edges/node {st['edges']/st['nodes']:.2f}", flush=True)
