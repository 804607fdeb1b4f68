"""Worker of tests/test_gpu_peers.py (launched under torch.distributed.run, one rank per GPU): the peer-memory exchange
(sharded.PeerExchange -> gvss_merge_topk_peers_device) against the NCCL all-gather + gvss_merge_topk_device path."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import duckdb_vss_b200 as gv  # noqa: E402
from duckdb_vss_b200.sharded import PeerExchange, gather_topk, merge_on_device  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    nq, k = 777, 10
    ex = PeerExchange(nq, k, dev)
    for step in range(4):  # more than two steps: both buffer sets are reused
        rng = np.random.default_rng(1000 * step + rank)
        d = np.sort(rng.random((nq, k), dtype=np.float32), axis=1)
        keys = (rng.permutation(nq * k).astype(np.int64) * world + rank).reshape(nq, k)  # disjoint key sets per shard
        if rank == 1:
            d[:, 8:] = np.inf  # a shard with short lists: padding entries
            keys[:, 8:] = -1
        kd, dd = torch.from_numpy(keys).to(dev), torch.from_numpy(d).to(dev)
        kp, dp = ex.outputs()
        b = (ex.step + 1) & 1
        ex.keys[b].copy_(kd)   # what the rank's search would have written
        ex.dists[b].copy_(dd)
        assert ex.keys[b].data_ptr() == kp and ex.dists[b].data_ptr() == dp
        mk = torch.empty((nq, k), dtype=torch.int64, device=dev)
        md = torch.empty((nq, k), dtype=torch.float32, device=dev)
        ex.merge(mk, md)
        gk, gd = gather_topk(kd, dd)
        wk, wd = torch.empty_like(mk), torch.empty_like(md)
        merge_on_device(gk.contiguous(), gd.contiguous(), wk, wd)
        torch.cuda.synchronize()
        assert torch.equal(mk, wk) and torch.equal(md, wd), f"rank {rank} step {step}: peer merge differs from the gathered merge"
    dist.barrier()
    if rank == 0:
        print("PEER_EXCHANGE_OK", world, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
