"""Multi-GPU host logic: the base set is partitioned by contiguous row ranges, one independent
sub-graph per rank (no build-time communication); a query batch is searched on every shard, the per-shard top-k lists
are all-gathered (12·nq·k bytes per rank) and merged by (distance, key). One process per GPU, torch.distributed
(NCCL on GPUs; the same code runs over gloo for the CPU tests of the plumbing)."""
import torch
import torch.distributed as dist


def shard_range(n_total: int, rank: int, world: int):
    """Contiguous rows [begin, end) owned by `rank` (sizes differ by at most one)."""
    base, rem = divmod(n_total, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def gather_topk(keys: torch.Tensor, dists: torch.Tensor, group=None):
    """All-gather per-shard results: keys [nq, k] i64, dists [nq, k] f32 -> ([world, nq, k], [world, nq, k])."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return keys.unsqueeze(0), dists.unsqueeze(0)
    nq, k = keys.shape
    gk = torch.empty((world * nq, k), dtype=keys.dtype, device=keys.device)   # rank-major concatenation
    gd = torch.empty((world * nq, k), dtype=dists.dtype, device=dists.device)
    dist.all_gather_into_tensor(gk, keys.contiguous(), group=group)
    dist.all_gather_into_tensor(gd, dists.contiguous(), group=group)
    return gk.view(world, nq, k), gd.view(world, nq, k)


def merge_on_device(gk: torch.Tensor, gd: torch.Tensor, out_keys: torch.Tensor, out_dists: torch.Tensor):
    """Device merge of the gathered lists (kernel k_merge<1>, C ABI gvss_merge_topk_device). CUDA tensors only."""
    from .index import merge_topk_device
    if not gk.is_cuda:
        raise RuntimeError("merge_on_device needs CUDA tensors: the product has no CPU path")
    world, nq, k = gk.shape
    merge_topk_device(gk.data_ptr(), gd.data_ptr(), world, nq, k, out_keys.data_ptr(), out_dists.data_ptr(),
                      torch.cuda.current_stream().cuda_stream)


class PeerExchange:
    """Result buffers in symmetric memory (torch.distributed._symmetric_memory: CUDA VMM handles exchanged once, every
    rank's buffer mapped into every process) so that the merge kernel reads the peers' lists over NVLink itself and no
    collective runs per step (kernel k_merge_peers, C ABI gvss_merge_topk_peers_device). Two buffer sets alternate from
    step to step: a fast rank may already search step s+1 while a slow peer still merges step s.

        ex = PeerExchange(nq, k, device)                  # collective: call on every rank
        keys_ptr, dist_ptr = ex.outputs()                 # where THIS step's search writes its lists
        ex.merge(out_keys, out_dists)                     # merge of all shards' lists of this step; advances the step
    """

    def __init__(self, nq: int, k: int, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        group = group or dist.group.WORLD
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world > 16 or self.world * k > 256:
            raise ValueError("PeerExchange: at most 16 shards and shards * k <= 256")
        self.nq, self.k, self.step = nq, k, 0
        self.keys = symm_mem.empty((2, nq, k), dtype=torch.int64, device=device)
        self.dists = symm_mem.empty((2, nq, k), dtype=torch.float32, device=device)
        self.flags = symm_mem.empty((16,), dtype=torch.int32, device=device)
        self.flags.zero_()
        torch.cuda.synchronize(device)
        hk, hd, hf = (symm_mem.rendezvous(t, group) for t in (self.keys, self.dists, self.flags))
        self._handles = (hk, hd, hf)  # keep the mappings alive
        self._kp, self._dp, self._fp = list(hk.buffer_ptrs), list(hd.buffer_ptrs), list(hf.buffer_ptrs)
        dist.barrier(group)  # every rank has zeroed its flags and mapped its peers before anyone signals

    def outputs(self):
        b = (self.step + 1) & 1
        return self.keys[b].data_ptr(), self.dists[b].data_ptr()

    def merge(self, out_keys: torch.Tensor, out_dists: torch.Tensor):
        from .index import merge_topk_peers_device
        self.step += 1
        b = self.step & 1
        off_k, off_d = b * self.nq * self.k * 8, b * self.nq * self.k * 4
        merge_topk_peers_device([p + off_k for p in self._kp], [p + off_d for p in self._dp], self._fp, self.rank, self.step,
                                self.nq, self.k, out_keys.data_ptr(), out_dists.data_ptr(),
                                torch.cuda.current_stream().cuda_stream)


def max_over_ranks(seconds: float, device) -> float:
    """Timing rule of the benchmark contract: a multi-GPU duration is the MAX over ranks."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return seconds
    t = torch.tensor([seconds], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
