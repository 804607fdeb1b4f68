import time, numpy as np, torch, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth, index as gvi
n, nq, d = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000, int(sys.argv[2]) if len(sys.argv) > 2 else 10_000, int(os.environ.get('EXACT_D', '128'))
base, q = synth.clustered(n, d, max(64, n // 100), 16, 42, nq)
b, qq = torch.from_numpy(base).cuda(), torch.from_numpy(q).cuda()
ids = torch.empty((nq, 10), dtype=torch.int64, device="cuda"); dist = torch.empty((nq, 10), dtype=torch.float32, device="cuda")
for rep in range(int(sys.argv[3]) if len(sys.argv) > 3 else 3):
    torch.cuda.synchronize(); t = time.time()
    gvi.exact_topk_device("l2sq", b.data_ptr(), n, qq.data_ptr(), nq, d, 10, ids.data_ptr(), dist.data_ptr())
    torch.cuda.synchronize(); t = time.time() - t
print("GVSS_EXACT_TC=", os.environ.get("GVSS_EXACT_TC", "1"), f"{nq} x {n} x {d} exact top-10: {t*1e3:.1f} ms", "checksum", int(ids.sum().item()))
