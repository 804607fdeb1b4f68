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


def test_noalloc_search_leaves_free_device_memory_untouched():
    """The property the entry point exists for: once its one-time set-up has run (first call on the device: the address of the
    built-in scratch array is looked up and the search kernel's module is loaded), further calls take no device memory at all —
    free memory as the driver reports it and the index's own accounting stay exactly where they were — also when a larger beam
    or another index re-uses the array the previous call left dirty."""
    import torch
    z, blob = load_golden("clustered40_l2sq")
    ix = gv.Index(z["base"].shape[1], "l2sq", M=int(z["M"]), M0=int(z["M0"]), ef_construction=int(z["efc"]))
    ix.load(blob)
    z2, blob2 = load_golden("clustered16_l2sq")
    iy = gv.Index(z2["base"].shape[1], "l2sq", M=int(z2["M"]), M0=int(z2["M0"]), ef_construction=int(z2["efc"]))
    iy.load(blob2)
    q, q2 = torch.from_numpy(z["queries"]).cuda(), torch.from_numpy(z2["queries"]).cuda()
    nq = max(q.shape[0], q2.shape[0])
    keys = torch.empty((nq, 10), dtype=torch.int64, device="cuda")
    dist = torch.empty((nq, 10), dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for index, queries in ((ix, q), (iy, q2)):   # warm-up: one-time set-up per device, kernel module loaded
        index.search_device_noalloc(queries.data_ptr(), queries.shape[0], 10, 16, keys.data_ptr(), dist.data_ptr(), 0, 0, st)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    use0 = (ix.info()["memory_usage"], iy.info()["memory_usage"])
    for ef in (10, 64, 128, 64):                  # smaller beam, larger beam (another table layout over the dirty array), back
        for index, queries, zz in ((ix, q, z), (iy, q2, z2)):
            index.search_device_noalloc(queries.data_ptr(), queries.shape[0], 10, ef, keys.data_ptr(), dist.data_ptr(), 0, 0, st)
            torch.cuda.synchronize()
            np.testing.assert_array_equal(keys[:queries.shape[0]].cpu().numpy(), zz[f"keys_ef{ef}"])
    assert torch.cuda.mem_get_info()[0] == free0
    assert (ix.info()["memory_usage"], iy.info()["memory_usage"]) == use0
