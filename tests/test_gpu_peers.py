"""GPU (2 or more devices): the exchange of per-shard top-k lists through peer memory over NVLink, one process per GPU,
against the NCCL all-gather path. Skipped on a single-GPU box (bench.py --gpus N makes the same check on every run)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_peer_memory_exchange_equals_nccl_gather_merge():
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least two GPUs")
    world = 2 if n < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29577", os.path.join(ROOT, "tests", "peer_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert f"PEER_EXCHANGE_OK {world}" in r.stdout


def test_one_process_sharded_index_over_several_gpus_matches_global_exact_topk():
    """SURVEY.md 8(e) through the C ABI alone (no torch.distributed, no symmetric memory): gvss_create_sharded over all GPUs
    of the box; gvss_search returns the merged list. On a small set with a wide beam the merged ids are the global exact
    top-k, and they equal the per-shard results merged on the host."""
    import numpy as np
    import torch
    import duckdb_vss_b200 as gv
    from duckdb_vss_b200 import synth
    G = torch.cuda.device_count()
    if G < 2:
        pytest.skip("needs at least two GPUs")
    G = min(G, 8)
    n, d, k = 24000, 32, 10
    base, q = synth.clustered(n, d, 240, 6, 77, 400)
    ix = gv.Index(d, "ip", M=16, devices=list(range(G)))
    ix.reserve(n, 1)
    ix.add_batch(base)
    ix.flush()
    assert ix.size() == n
    keys, dist, counts = ix.ef_search(q, k, 512)
    truth, tdist = gv.exact_topk("ip", base, q, k)
    assert synth.recall_at_k(keys, truth) >= 0.999
    assert (counts == k).all() and (np.diff(dist, axis=1) >= 0).all()
    # the index metric is 1 - a.b, DuckDB's fold is -a.b: same order, distances differ by one
    hit = keys == truth
    np.testing.assert_allclose(dist[hit] - 1.0, tdist[hit], rtol=1e-5, atol=1e-6)
