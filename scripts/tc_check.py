"""Stage check of the tensor-core exact path: raw q·b against float64, then ids/distances against the reference folds."""
import ctypes as C, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import duckdb_vss_b200 as gv
from duckdb_vss_b200 import synth
from oracle import ref
L = gv.lib()
fn = L.gvssdbg_tc_exact
fn.restype = C.c_int
fn.argtypes = [C.c_char, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
for metric, d, n, nq in (("l2sq", 128, 1000, 200), ("ip", 96, 777, 130), ("cosine", 40, 5000, 300), ("l2sq", 128, 100000, 1000)):
    base, q = synth.clustered(n, d, 32, 8, 5, nq)
    b, qq = torch.from_numpy(base).cuda(), torch.from_numpy(q).cuda()
    ids = torch.empty((nq, 10), dtype=torch.int64, device="cuda"); dist = torch.empty((nq, 10), dtype=torch.float32, device="cuda")
    dots = torch.zeros((nq, n), dtype=torch.float32, device="cuda") if n <= 5000 else None
    flagged = C.c_uint64(0)
    t = time.time()
    rc = fn(gv.index.METRIC_CHAR[metric], b.data_ptr(), n, qq.data_ptr(), nq, d, 10, ids.data_ptr(), dist.data_ptr(), dots.data_ptr() if dots is not None else None, C.byref(flagged))
    t = time.time() - t
    if rc: print("ERROR", L.gvss_last_error().decode()); continue
    if dots is not None:
        want = q.astype(np.float64) @ base.astype(np.float64).T
        err = np.abs(dots.cpu().numpy() - want).max() / np.abs(want).max()
        print(f"{metric} d={d} n={n}: max rel dot error {err:.2e}", end="; ")
    wi, wd = ref.duckdb_topk(metric, base, q, 10, threads=8)
    gi, gd = ids.cpu().numpy(), dist.cpu().numpy()
    print(f"{metric} d={d} n={n} nq={nq}: ids equal {np.array_equal(gi, wi)} dist bits equal {np.array_equal(gd.view(np.uint32), wd.view(np.uint32))} flagged {flagged.value} time {t*1e3:.1f} ms")
