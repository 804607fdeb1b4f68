"""CPU, world_size 2 over gloo: the N>1 host logic — shard ranges, the all-gather layout of per-shard top-k lists and
max-over-ranks timing. The merge itself is a device kernel (tested in tests/test_gpu_exact.py); here a numpy statement
of its contract (order by (distance, key), padding last) checks that gathered shards reproduce the global top-k."""
import os
import sys
import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    from duckdb_vss_b200 import sharded, synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n, d, nq, k = 1001, 8, 13, 5
    base, q = synth.clustered(n, d, 8, 3, 77, nq)          # same data on every rank; each rank owns a row range
    b, e = sharded.shard_range(n, rank, world)
    dmat = ((q[:, None, :] - base[None, b:e, :]) ** 2).sum(-1)
    order = np.argsort(dmat, axis=1, kind="stable")[:, :k]
    keys = torch.from_numpy((order + b).astype(np.int64))
    dists = torch.from_numpy(np.take_along_axis(dmat, order, axis=1).astype(np.float32))
    if rank == 1:                                          # a short list on one shard: padding (-1, +inf)
        keys[0, 3:] = -1
        dists[0, 3:] = float("inf")
    gk, gd = sharded.gather_topk(keys, dists)
    t = sharded.max_over_ranks(0.5 + rank, torch.device("cpu"))
    if rank == 0:
        out["gk"], out["gd"], out["t"] = gk.numpy().copy(), gd.numpy().copy(), t
        out["ranges"] = [sharded.shard_range(n, r, world) for r in range(world)]
        full = ((q[:, None, :] - base[None, :, :]) ** 2).sum(-1).astype(np.float32)
        out["full"] = full
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gather_and_merge_contract():
    world = 2
    with mp.Manager() as m:
        out = m.dict()
        mp.spawn(_worker, args=(world, 29731, out), nprocs=world, join=True)
        gk, gd, full, ranges, t = out["gk"], out["gd"], out["full"], out["ranges"], out["t"]
    assert ranges == [(0, 501), (501, 1001)]
    assert t == 1.5                                        # max over ranks
    assert gk.shape == (2, 13, 5) and gd.shape == (2, 13, 5)
    for qi in range(13):
        cand = sorted((float(gd[s, qi, i]), int(gk[s, qi, i])) for s in range(2) for i in range(5) if gk[s, qi, i] >= 0)[:5]
        if qi == 0:
            continue                                       # shard 1 was truncated on purpose for query 0
        want = sorted((float(full[qi, j]), j) for j in range(full.shape[1]))[:5]
        assert [c[1] for c in cand] == [w[1] for w in want]
    # query 0: padding never outranks a real candidate
    cand0 = sorted((float(gd[s, 0, i]), int(gk[s, 0, i])) for s in range(2) for i in range(5) if gk[s, 0, i] >= 0)
    assert len(cand0) == 8 and all(np.isfinite(c[0]) for c in cand0)


def test_shard_ranges_cover_everything():
    from duckdb_vss_b200 import sharded
    for n, w in ((10, 3), (1_000_000, 8), (7, 8)):
        r = [sharded.shard_range(n, i, w) for i in range(w)]
        assert r[0][0] == 0 and r[-1][1] == n and all(r[i][1] == r[i + 1][0] for i in range(w - 1))
        assert max(e - b for b, e in r) - min(e - b for b, e in r) <= 1


def test_weak_scaling_shards_are_samples_of_one_data_set():
    """bench.py's weak-scaling shards: rank 0 holds exactly the N=1 set; the other ranks hold other samples of the SAME data set
    (same cluster centres and projection, other points), so the queries are in-distribution for every shard."""
    import numpy as np
    from duckdb_vss_b200 import synth
    n, d, C = 4000, 16, 40
    base0, q = synth.clustered(n, d, C, 4, 42, 50)
    again, _ = synth.clustered(n, d, C, 4, 42, 0, point_seed=None)
    other, _ = synth.clustered(n, d, C, 4, 42, 0, point_seed=1042)
    alien, _ = synth.clustered(n, d, C, 4, 1042, 0)
    assert np.array_equal(base0, again) and not np.array_equal(base0, other)

    def nearest(x, y):  # mean distance from each row of x to its nearest row of y
        d2 = ((x[:, None, :] - y[None, :, :]) ** 2).sum(-1)
        return float(np.sqrt(d2.min(1)).mean())
    # queries of the seed-42 set find close rows in another sample of the same set, not in an unrelated set
    assert nearest(q, other) < 0.6 * nearest(q, alien)
    assert abs(nearest(q, other) - nearest(q, base0)) < 0.25 * nearest(q, base0)
