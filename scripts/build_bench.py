"""Times gvss_flush (bulk build) alone: python scripts/build_bench.py [n] [M] [efc]"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
M = int(sys.argv[2]) if len(sys.argv) > 2 else 32
efc = int(sys.argv[3]) if len(sys.argv) > 3 else 200
base, q = synth.clustered(n, 128, max(1024, n // 100), 16, 42, 2000)
dev = torch.from_numpy(base).cuda()
ix = gv.Index(128, "l2sq", M=M, ef_construction=efc); ix.reserve(n, 1)
ix.add_batch_device(dev.data_ptr(), n)
torch.cuda.synchronize(); t = time.time(); ix.flush(); torch.cuda.synchronize(); t = time.time() - t
truth, _ = gv.exact_topk("l2sq", base, q, 10)
rec = {ef: round(synth.recall_at_k(ix.ef_search(q, 10, ef)[0], truth), 4) for ef in (128, 384)}
print(f"n={n} M={M} efc={efc} ramp={os.environ.get('GVSS_BUILD_RAMP','256')} build {t:.2f}s {n/t:.0f} vec/s recall {rec} launches {gv.lib().gvss_kernel_launches()}")
