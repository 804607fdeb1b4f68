"""GPU: gvss_search_device_noalloc — the search entry point for callers that may not allocate device memory (a host
whose own pool already owns all free HBM, or a latency-critical caller that cannot afford cudaMalloc's device-wide
synchronisation). Its scratch is an array built into the library, so it has to give exactly the results of
gvss_search_device, call after call, although the array still holds the previous call's tables.

Collected last on purpose (file name): a run that stops at the first failure has then already run every other test."""
import numpy as np
import pytest
import duckdb_vss_b200 as gv
from conftest import load_golden

pytestmark = pytest.mark.gpu


def _search(ix, fn, q, k, ef):
    import torch
    nq = q.shape[0]
    keys = torch.full((nq, k), -7, dtype=torch.int64, device="cuda")
    dist = torch.full((nq, k), -7.0, dtype=torch.float32, device="cuda")
    counts = torch.full((nq,), -7, dtype=torch.int32, device="cuda")
    stats = torch.zeros((nq, 4), dtype=torch.int32, device="cuda")
    fn(q.data_ptr(), nq, k, ef, keys.data_ptr(), dist.data_ptr(), counts.data_ptr(), stats.data_ptr(),
       torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return keys.cpu().numpy(), dist.cpu().numpy(), counts.cpu().numpy(), stats.cpu().numpy()


def test_noalloc_search_equals_allocating_search_on_two_calls_in_a_row():
    import torch
    z, blob = load_golden("clustered16_l2sq")
    ix = gv.Index(z["base"].shape[1], "l2sq", M=int(z["M"]), M0=int(z["M0"]), ef_construction=int(z["efc"]))
    ix.load(blob)
    q1 = torch.from_numpy(z["queries"]).cuda()
    q2 = torch.from_numpy(z["base"][100:180].copy()).cuda()       # a second, different batch reusing the same scratch
    for q, ef in ((q1, 64), (q2, 10)):
        want = _search(ix, ix.search_device, q, 10, ef)
        got = _search(ix, ix.search_device_noalloc, q, 10, ef)
        np.testing.assert_array_equal(got[0], want[0])                                   # row ids
        np.testing.assert_array_equal(got[1].view(np.uint32), want[1].view(np.uint32))   # distances, bit for bit
        np.testing.assert_array_equal(got[2], want[2])                                   # counts
        np.testing.assert_array_equal(got[3], want[3])                                   # traversal counters
    np.testing.assert_array_equal(_search(ix, ix.search_device_noalloc, q1, 10, 64)[0], z["keys_ef64"])


def test_noalloc_rejects_what_does_not_fit_and_staged_rows():
    import torch
    z, blob = load_golden("readme_l2sq")
    ix = gv.Index(3, "l2sq")
    ix.load(blob)
    q = torch.from_numpy(z["queries"]).cuda()
    out = torch.empty((q.shape[0], 16000), dtype=torch.int64, device="cuda")
    with pytest.raises(gv.GvssError, match="scratch"):           # beam 16000 -> 512 Ki table slots x 12 B = 6 MiB per query slot > 1 MiB
        ix.search_device_noalloc(q.data_ptr(), q.shape[0], 16000, 16000, out.data_ptr())
    iy = gv.Index(3, "l2sq")
    iy.reserve(8, 1)
    iy.add(1, np.zeros(3, np.float32))
    with pytest.raises(gv.GvssError, match="gvss_flush"):
        iy.search_device_noalloc(q.data_ptr(), 1, 1, 8, out.data_ptr())
