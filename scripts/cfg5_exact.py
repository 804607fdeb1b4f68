"""Config 5 (BASELINE.json): index join / exact top-k, 100k queries x 10M x FLOAT[128] l2sq, k=10 — timed on one B200
with everything resident in HBM, plus a parity spot check of a query sub-sample against the CPU reference folds."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth, index as gvi
n, nq, d, k = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000, int(sys.argv[2]) if len(sys.argv) > 2 else 100_000, 128, 10
t = time.time()
base, q = synth.clustered(n, d, n // 100, 16, 48, nq)
print(f"generated {n} x {d} + {nq} queries in {time.time()-t:.1f}s", flush=True)
b, qq = torch.from_numpy(base).cuda(), torch.from_numpy(q).cuda()
ids = torch.empty((nq, k), dtype=torch.int64, device="cuda"); dist = torch.empty((nq, k), dtype=torch.float32, device="cuda")
for rep in range(2):
    torch.cuda.synchronize(); t = time.time()
    gvi.exact_topk_device("l2sq", b.data_ptr(), n, qq.data_ptr(), nq, d, k, ids.data_ptr(), dist.data_ptr())
    torch.cuda.synchronize(); t = time.time() - t
    print(f"exact top-{k}: {nq} x {n} x {d}: {t*1e3:.0f} ms ({nq/t:.0f} queries/s, {2*nq*n*d/t/1e12:.1f} TFLOP/s fp32-accurate)", flush=True)
try:
    from oracle import ref
    sub = np.arange(0, nq, max(1, nq // 16))[:16]
    t = time.time()
    wi, wd = ref.duckdb_topk("l2sq", base, q[sub], k, threads=ref.hardware_threads())
    print(f"CPU reference folds on {len(sub)} queries: {time.time()-t:.1f}s")
    gi, gd = ids.cpu().numpy()[sub], dist.cpu().numpy()[sub]
    print("ids identical", np.array_equal(gi, wi), "distance bits identical", np.array_equal(gd.view(np.uint32), wd.view(np.uint32)))
except Exception as e:
    print("no CPU reference here:", e)
