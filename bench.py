#!/usr/bin/env python
"""bench.py — headline benchmark of the HNSW hot path (BASELINE.json: "HNSW search QPS at recall@10>=0.95 (batched) +
index build vectors/sec").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl gvss|reference] [--workload sift1m|deep10m]

Default workload = configs[1]: 1M x FLOAT[128] l2sq, 10 000-query batch, k=10. A step = one batched search of the whole
query batch at the smallest ef of the sweep whose recall@10 is >= 0.95. `value` is device-timed with queries and index
resident in HBM; `e2e` is the same step from HOST buffers (H2D of the queries, search, merge, D2H of the result list inside
the timed region). The bulk build (gvss_add_batch_device + gvss_flush) is timed once, before the steps, and reported under
"build" with its own roofline (from gvss_get_build_stats); the exact top-k that provides the ground truth is reported
under "exact_topk".

N > 1 (one process per GPU, torchrun):
  * --workload sift1m (default, weak scaling): every rank owns an independent 1M-vector shard (own sub-graph, no build-time
    communication). Every query is searched on every shard; the per-shard top-k lists are exchanged through peer memory over
    NVLink and merged by one kernel on a second stream, so the merge of step s overlaps the search of step s+1 (fallback:
    NCCL all-gather + merge). Every shard is another 1M-row sample of the same data set (rank 0's rows are the N=1 set). The
    operating point is the N=1 one (per-GPU work fixed), checked against the gate on the MERGED list's recall vs the exact
    top-k over the union of all shards. `value` counts per-shard query searches summed over ranks;
    config.global_queries_per_s is the batch rate.
  * --workload deep10m (configs[3], strong scaling): ONE 10M x FLOAT[96] ip set partitioned over the N GPUs by contiguous
    row ranges; `value` = queries answered per second by the whole job (every query answered once, from all shards).

`--impl reference` times the UNMODIFIED reference (vendored usearch via oracle/_ref) on the host cores on the SAME config:
the full 1M-vector index built by all host threads with the extension's pattern, all 10 000 queries per step, at the
smallest ef of the same sweep at which the REFERENCE-built graph reaches recall@10 >= 0.95.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "hnsw_search_qps_at_recall10_ge_0.95"
EF_SWEEP = (16, 32, 64, 96, 128, 160, 192, 224, 256, 288, 304, 320, 384, 512)
# with the usearch defaults (M=16, efc=128) recall@10 >= 0.95 is NOT reachable on the 1M set within
# ef <= 512 (reference-built graph: 0.80 at ef=512, profiles/recall_1m_r1.txt), so both arms use M=32 / ef_construction=200.
M_BENCH, M0_BENCH, EFC_BENCH = 32, 64, 200
WORKLOADS = {
    # name: (label, metric, dims, vectors, generator note)
    "sift1m": ("1M x FLOAT[128] l2sq (SIFT-1M shape), 10k-query batch, k=10", "l2sq", 128, 1_000_000,
               "clustered(C=N/100, r=16) seed 42 (SURVEY.md App. D law, numpy)"),
    "deep10m": ("10M x FLOAT[96] ip (Deep10M shape) partitioned over the GPUs, 10k-query batch, k=10", "ip", 96, 10_000_000,
                "clustered(C=N/100, r=24) seed 46, L2-normalised (SURVEY.md §8d config 4, numpy)"),
    "text768": ("1M x FLOAT[768] cosine (text-embedding shape), 10k-query batch, k=10", "cosine", 768, 1_000_000,
                "clustered(C=N/100, r=32) seed 44, rows L2-normalised then scaled by U(0.5, 2) (SURVEY.md §8d config 3, numpy)"),
    "join10m": ("index join / exact top-k: 100k queries x 10M x FLOAT[128] l2sq, k=10", "l2sq", 128, 10_000_000,
                "clustered(C=N/100, r=16) seed 48 (SURVEY.md §8d config 5, numpy)"),
}
DEFAULT_NQ = {"join10m": 100_000}
# the 10M-row sets need wider beams than the 1M sets to reach the gate
EF_SWEEP_WIDE = EF_SWEEP + (768, 1024, 1536, 2048)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return float(j["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy bandwidth)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def tensor_peak_tf32():
    """Dense TF32 rate = half the dense bf16 rate; the bf16 figure is the measured one when the driver wrote it."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return float(j["bf16_tflops"]) / 2.0, "measured bf16 burst (MEASURED_PEAKS.json) / 2 for TF32"
    return 2250.0 / 2.0, "fallback: nominal 2.25 PFLOP/s bf16 / 2 for TF32"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def algorithmic_bytes(stats, d, M, M0, k):
    """B_q = C_q*4d + H0_q*(4+4*M0) + Hup_q*(4+4*M) + 12k, from usearch-compatible counters."""
    c = stats[:, 0].astype(np.float64).sum()
    h0 = stats[:, 2].astype(np.float64).sum()
    hup = (stats[:, 1].astype(np.float64) - stats[:, 2].astype(np.float64)).sum()
    return c * 4 * d + h0 * (4 + 4 * M0) + hup * (4 + 4 * M) + stats.shape[0] * 12.0 * k


def make_data(workload, n_total, nq):
    """(base[n_total, d], queries[nq, d]) of the named workload, deterministic."""
    from duckdb_vss_b200 import synth
    if workload in ("deep10m", "text768", "join10m"):
        c = synth.config(workload, n=n_total, nq=nq)
        return c["base"], c["queries"]
    return synth.clustered(n_total, 128, max(1024, n_total // 100), 16, 42, nq)


def run_reference(args):
    """CPU arm: the reference's own usearch path on the host cores, on the same config as the GPU arm at N=1."""
    import oracle
    from duckdb_vss_b200 import synth
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    label, metric, d, n_default, gen = WORKLOADS["sift1m"]
    n = args.n or n_default
    nq, k = args.nq or 10_000, 10
    if oracle.have_ref():
        from oracle import ref as impl
        kind, Index, threads = "reference", impl.RefIndex, impl.hardware_threads()
    else:  # reference sources absent and no prebuilt oracle/_ref: the pinned C port on a sample it can finish
        oracle.build()
        from oracle import port as impl
        kind, Index, threads = "port", impl.PortIndex, 1
        n, nq = min(n, 20000), min(nq, 200)
    base, q = make_data("sift1m", n, nq)
    ix = Index(d, metric, M=M_BENCH, M0=M0_BENCH, ef_construction=EFC_BENCH)
    ix.reserve(n, threads)
    t0 = time.perf_counter()
    ix.add_bulk(base, threads=threads)   # T threads, shared 2048-row chunk cursor (hnsw_index_physical_create.cpp:144-205)
    t_build = time.perf_counter() - t0
    t0 = time.perf_counter()
    if kind == "reference":
        truth, _ = impl.fast_topk(metric, base, q, k, threads=threads)
    else:
        truth, _ = impl.exact_topk(metric, base, q, k)
    t_truth = time.perf_counter() - t0
    sweep, ef, recall = {}, args.ef, None
    for e in ([args.ef] if args.ef else EF_SWEEP):
        r = synth.recall_at_k(ix.search_batch(q, k, e, threads=threads)["keys"], truth)
        sweep[e] = round(r, 4)
        ef, recall = e, r
        if r >= 0.95:
            break
    for _ in range(args.warmup):
        ix.search_batch(q, k, ef, threads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ix.search_batch(q, k, ef, threads=threads)
    dt = time.perf_counter() - t0
    qps = nq * args.steps / dt
    sample = (f"the whole workload: index over all {n} vectors built by {threads} host threads in {t_build:.0f} s, all {nq} queries per "
              f"step, ef={ef} = smallest ef of the sweep with recall@10 >= 0.95 on the reference-built graph (ground truth: "
              f"vectorised CPU scan + exact re-score, {t_truth:.0f} s, untimed)")
    line = {"impl": "reference", "metric": METRIC, "value": qps, "unit": "queries/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": label, "generator": gen, "n_per_gpu": n, "queries_per_step": nq, "k": k, "ef_search": ef,
                       "recall_at_10": recall, "ef_sweep": sweep, "M": M_BENCH, "M0": M0_BENCH, "ef_construction": EFC_BENCH,
                       "note": "the CPU arm runs the N=1 workload whatever --gpus says (host cores do not grow with the GPU count)"},
            "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": threads, "kind": kind, "sample": sample,
                             "build_vectors_per_s": n / t_build},
            "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "build": {"value": n / t_build, "unit": "vectors/s", "vectors": n, "seconds": t_build}}
    print(json.dumps(line))


def cpu_baseline(args, d, metric, blob_getter, q, ef, truth, gpu_keys=None):
    """Reference usearch (oracle/_ref) on the host cores: search on the SAME (GPU-built) graph, build on a bounded sample."""
    try:
        import oracle
        if oracle.have_ref():
            from oracle import ref as impl
            kind, threads = "reference", impl.hardware_threads()
            Index = impl.RefIndex
        else:
            if not oracle.have_port():
                oracle.build()
            from oracle import port as impl
            kind, threads, Index = "port", 1, impl.PortIndex
        from duckdb_vss_b200 import synth
        ix = Index(d, metric, M=M_BENCH, M0=M0_BENCH, ef_construction=EFC_BENCH)
        t0 = time.perf_counter()
        ix.load(blob_getter())
        t_load = time.perf_counter() - t0
        nq = q.shape[0] if kind == "reference" else min(q.shape[0], 300)
        ix.search_batch(q[:min(nq, 512)], 10, ef, threads=threads)  # warm
        t0 = time.perf_counter()
        r = ix.search_batch(q[:nq], 10, ef, threads=threads)
        dt = time.perf_counter() - t0
        rec = synth.recall_at_k(r["keys"], truth[:nq])
        ids_same = bool(gpu_keys is not None and np.array_equal(r["keys"], gpu_keys[:nq]))
        # build sample: a prefix of the set with the extension's thread pattern (the full build is what --impl reference times)
        nb = min(args.n or 10 ** 9, max(20000, 1500 * threads)) if kind == "reference" else 3000
        base_s, _ = make_data("sift1m", nb, 0)
        b = Index(128, "l2sq", M=M_BENCH, M0=M0_BENCH, ef_construction=EFC_BENCH)
        b.reserve(nb, threads)
        t0 = time.perf_counter()
        b.add_bulk(base_s, threads=threads)
        tb = time.perf_counter() - t0
        return {"value": nq / dt, "unit": "queries/s", "cores": threads, "kind": kind,
                "sample": f"{nq} queries, ef={ef}, on the GPU-built graph loaded from its usearch stream (load {t_load:.1f}s); "
                          f"build timed on the first {nb} vectors of the 1M x 128 set", "recall_at_10": rec,
                "ids_identical_to_gpu_on_same_graph": ids_same,
                "build_vectors_per_s": nb / tb}
    except Exception as e:  # the baseline must never take the product number down
        return {"value": None, "unit": "queries/s", "cores": 0, "kind": "unavailable", "sample": repr(e)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="gvss", choices=["gvss", "reference"])
    ap.add_argument("--workload", default="sift1m", choices=sorted(WORKLOADS))
    ap.add_argument("--vectors", dest="n", type=int, default=0, help="base vectors (per GPU for sift1m, in total for deep10m); 0 = the config's size")
    ap.add_argument("--nq", type=int, default=0, help="queries per step; 0 = the config's batch (10 000; join10m: 100 000)")
    ap.add_argument("--ef", type=int, default=0, help="0 = smallest ef of the sweep with recall@10 >= 0.95")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--M", type=int, default=0, help="connectivity (0 = the bench default 32; M0 = 2M)")
    ap.add_argument("--efc", type=int, default=0, help="ef_construction (0 = the bench default 200)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "gvss" else args.warmup
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import duckdb_vss_b200 as gv
    from duckdb_vss_b200 import synth, index as gvi
    from duckdb_vss_b200.sharded import shard_range

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    label, metric, d, n_default, gen = WORKLOADS[args.workload]
    global M_BENCH, M0_BENCH, EFC_BENCH
    if args.M:
        M_BENCH, M0_BENCH = args.M, 2 * args.M
    if args.efc:
        EFC_BENCH = args.efc
    strong = args.workload != "sift1m"     # ONE named set, partitioned over the ranks when there are several
    nq, k, M, M0 = args.nq or DEFAULT_NQ.get(args.workload, 10_000), 10, M_BENCH, M0_BENCH
    ef_sweep = EF_SWEEP_WIDE if args.workload == "join10m" else EF_SWEEP
    if strong:
        # ONE set partitioned by contiguous row ranges; keys = global row numbers
        n_total = args.n or n_default
        base_all, q_np = make_data(args.workload, n_total, nq)
        lo, hi = shard_range(n_total, rank, world)
        base_np = np.ascontiguousarray(base_all[lo:hi])
        del base_all
        n = hi - lo
        keys_np = np.arange(lo, hi, dtype=np.int64)
    else:
        # weak scaling: per-rank shard of the same law (shard-specific seed), the query batch identical on every rank
        n = args.n or n_default
        n_total = n * world
        # every shard is another 1M-row sample of the SAME data set (centres and projection of seed 42; rank 0's rows are exactly
        # the N=1 set), i.e. the union is one base set that shards naturally and the queries are in-distribution for every shard
        base_np, _ = synth.clustered(n, d, max(1024, n // 100), 16, 42, 0, point_seed=None if rank == 0 else 42 + 1000 * rank)
        _, q_np = synth.clustered(1, d, max(1024, n // 100), 16, 42, nq)
        keys_np = np.arange(n, dtype=np.int64) + rank * n
    base = torch.from_numpy(base_np).to(dev)
    q = torch.from_numpy(q_np).to(dev)
    s_search = torch.cuda.current_stream()
    stream = s_search.cuda_stream

    # ---- bulk build (timed once): rows already in HBM -> graph ------------------------------------------------
    ix = gv.Index(d, metric, M=M, M0=M0, ef_construction=EFC_BENCH, ef_search=64, device=local)
    ix.reserve(n, 1)
    torch.cuda.synchronize()
    l0 = gv.lib().gvss_kernel_launches()
    t0 = time.perf_counter()
    ix.add_batch_device(base.data_ptr(), n, keys=keys_np)
    ix.flush()
    torch.cuda.synchronize()
    t_build = time.perf_counter() - t0
    build_launches = gv.lib().gvss_kernel_launches() - l0
    bst = ix.build_stats()
    # B_v = C_v*4d*w + H_v*(4+4*M0) + E_v*(4+4*M0), w = 1 for query-to-node, 2 for node-to-node distances
    build_bytes = (bst["query_distances"] * 4.0 * d + bst["pair_distances"] * 8.0 * d + bst["base_expansions"] * (4 + 4.0 * M0)
                   + bst["upper_expansions"] * (4 + 4.0 * M) + bst["rows_written"] * (4 + 4.0 * M0))

    # ---- ground truth: exact kernel per shard (parity-checked against DuckDB's folds), merged over the shards ------------
    truth = torch.empty((nq, k), dtype=torch.int64, device=dev)
    tdist = torch.empty((nq, k), dtype=torch.float32, device=dev)
    kdev = torch.from_numpy(keys_np).to(dev)
    t_exact = None
    for _ in range(2):  # second call: pools and modules warm
        t0 = time.perf_counter()
        gvi.exact_topk_device(metric, base.data_ptr(), n, q.data_ptr(), nq, d, k, truth.data_ptr(), tdist.data_ptr(), kdev.data_ptr(), stream)
        torch.cuda.synchronize()
        t_exact = time.perf_counter() - t0
    truth_np = truth.cpu().numpy()            # this shard's exact top-k
    gtruth_np = truth_np                      # exact top-k over the union of all shards
    if world > 1:
        gk = torch.empty((world * nq, k), dtype=torch.int64, device=dev)   # rank-major [world][nq][k]
        gd = torch.empty((world * nq, k), dtype=torch.float32, device=dev)
        dist.all_gather_into_tensor(gk, truth)
        dist.all_gather_into_tensor(gd, tdist)
        gtruth = torch.empty_like(truth)
        gtd = torch.empty_like(tdist)
        gv.merge_topk_device(gk.data_ptr(), gd.data_ptr(), world, nq, k, gtruth.data_ptr(), gtd.data_ptr(), stream)
        torch.cuda.synchronize()
        gtruth_np = gtruth.cpu().numpy()

    okeys = torch.empty((nq, k), dtype=torch.int64, device=dev)
    odist = torch.empty((nq, k), dtype=torch.float32, device=dev)
    ocnt = torch.empty(nq, dtype=torch.int32, device=dev)
    ostat = torch.zeros((nq, 4), dtype=torch.int32, device=dev)
    mk = torch.empty((nq, k), dtype=torch.int64, device=dev)
    md = torch.empty((nq, k), dtype=torch.float32, device=dev)

    def search(ef, stats=False, qptr=None):
        ix.search_device(qptr or q.data_ptr(), nq, k, ef, okeys.data_ptr(), odist.data_ptr(), ocnt.data_ptr(),
                         ostat.data_ptr() if stats else 0, stream)

    def nccl_merge(out_k, out_d):
        dist.all_gather_into_tensor(gk, okeys)
        dist.all_gather_into_tensor(gd, odist)
        gv.merge_topk_device(gk.data_ptr(), gd.data_ptr(), world, nq, k, out_k.data_ptr(), out_d.data_ptr(), stream)

    # ---- operating point: smallest ef of the sweep whose recall@10 (of the merged list, N > 1) is >= 0.95 ---------------
    sweep, shard_sweep = {}, {}
    ef, ef_merged_min = args.ef, None
    for e in ([args.ef] if args.ef else ef_sweep):
        search(e)
        if world > 1:
            nccl_merge(mk, md)
        torch.cuda.synchronize()
        shard_sweep[e] = round(synth.recall_at_k(okeys.cpu().numpy(), truth_np), 4)
        r = synth.recall_at_k(mk.cpu().numpy(), gtruth_np) if world > 1 else shard_sweep[e]
        sweep[e] = round(r, 4)
        ef = e
        ok = r >= 0.95
        if world > 1 and not strong:
            # weak scaling keeps the per-GPU work of the N=1 run: the operating point is rank 0's single-shard operating point
            # (its rows are the N=1 set), raised further only if the merged list were still below the gate
            t0 = torch.tensor([shard_sweep[e] if rank == 0 else 0.0], dtype=torch.float64, device=dev)
            dist.all_reduce(t0)
            ok = ok and float(t0.item()) >= 0.95
            if r >= 0.95 and ef_merged_min is None:
                ef_merged_min = e
        if ok:
            break
    recall = sweep[ef]                         # N > 1: merged list vs exact top-k over the union (identical on every rank)
    recall_min = recall_max = shard_sweep[ef]
    if world > 1:
        tr = torch.tensor([shard_sweep[ef], -shard_sweep[ef]], dtype=torch.float64, device=dev)
        dist.all_reduce(tr, op=dist.ReduceOp.MIN)
        recall_min, recall_max = float(tr[0].item()), -float(tr[1].item())
    search(ef, stats=True)
    torch.cuda.synchronize()
    st = ostat.cpu().numpy().view(np.uint32)
    algo_bytes = algorithmic_bytes(st, d, M, M0, k)
    ref_keys_np = okeys.cpu().numpy()

    # ---- multi-GPU exchange of the per-shard lists (12*nq*k bytes per rank) + device merge ------------------------------
    # Preferred: the merge kernel reads the peers' lists over NVLink out of symmetric memory and the ranks synchronise through
    # flags in peer memory (no collective per step); fallback (GVSS_EXCHANGE=nccl, or symmetric memory unavailable): NCCL
    # all-gather + merge of the gathered copy. The merge runs on its own stream behind an event: step s+1's search does not
    # wait for step s's exchange (two result-buffer sets alternate).
    exchange, peer = "none", None
    s_merge = torch.cuda.Stream(dev) if world > 1 else None
    if world > 1:
        want_p2p = os.environ.get("GVSS_EXCHANGE", "p2p") != "nccl"
        ok = torch.zeros(1, dtype=torch.int32, device=dev)
        if want_p2p:
            try:
                from duckdb_vss_b200.sharded import PeerExchange
                peer = PeerExchange(nq, k, dev)
                ok += 1
            except Exception as e:  # symmetric memory not available on this box / torch build
                peer = None
                if rank == 0:
                    print(f"bench: peer-memory exchange unavailable ({e!r}); using NCCL all-gather", file=sys.stderr)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)  # all ranks or none
        if int(ok.item()) == 0:
            peer = None
        exchange = ("nvlink peer reads + flags in peer memory (k_merge_peers) on a second stream" if peer
                    else "nccl all_gather + k_merge")
    searched = [torch.cuda.Event() for _ in range(2)]      # search of buffer set b has finished
    merged = [torch.cuda.Event() for _ in range(2)]        # merge that read buffer set b has finished
    step_no = [0]

    def step(qptr=None):
        """one step: search (stream A); exchange + merge (stream B behind an event). Returns nothing; results in mk/md."""
        if not peer:
            search(ef, qptr=qptr)
            if world > 1:
                nccl_merge(mk, md)
            return
        b = step_no[0] & 1
        step_no[0] += 1
        s_search.wait_event(merged[b])                  # the merge two steps ago has released this buffer set
        kp, dp = peer.outputs()
        ix.search_device(qptr or q.data_ptr(), nq, k, ef, kp, dp, ocnt.data_ptr(), 0, stream)
        searched[b].record(s_search)
        with torch.cuda.stream(s_merge):
            s_merge.wait_event(searched[b])
            peer.merge(mk, md)
            merged[b].record(s_merge)

    def drain():
        if s_merge is not None:
            s_merge.synchronize()
        torch.cuda.synchronize()

    clocks = ClockSampler(local)   # started before the warm-up: nvidia-smi needs about a second to deliver its first sample
    clocks.__enter__()
    for _ in range(args.warmup):
        step()
    drain()
    # this rank's search launch on its own (no exchange, no waiting for peers): what bounds a step when the shards differ
    iso = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    search(ef)
    iso[0].record(s_search)
    for _ in range(3):
        search(ef)
    iso[1].record(s_search)
    torch.cuda.synchronize()
    t_kernel_alone = iso[0].elapsed_time(iso[1]) / 3.0
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # keep the GPUs under the same load until every rank's sampler delivers (all ranks take the same number of extra steps: the
    # peer exchange counts them)
    t_wait = time.perf_counter()
    while True:
        need = torch.tensor([int(len(clocks.rows) < 2 and time.perf_counter() - t_wait < 5.0)], dtype=torch.int32, device=dev)
        if world > 1:
            dist.all_reduce(need, op=dist.ReduceOp.MAX)
        if not int(need.item()):
            break
        step()
        drain()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks.rows.clear()
    l0 = gv.lib().gvss_kernel_launches()
    t_wall0 = time.perf_counter()
    t_begin.record(s_search)
    for a, b in ev:
        a.record(s_search)
        step()
        b.record(s_search)                           # end of this step's search launch (the merge is on the other stream)
    (s_merge or s_search).wait_stream(s_search)
    t_end.record(s_merge or s_search)                # after the last merge AND the last search
    drain()
    t_wall = time.perf_counter() - t_wall0
    clocks.__exit__(None, None, None)
    launches = gv.lib().gvss_kernel_launches() - l0
    t_dev = t_begin.elapsed_time(t_end) / 1e3                                # whole timed region, device time
    t_kernel = sum(a.elapsed_time(b) for a, b in ev) / 1e3 / args.steps      # search launch alone (roofline)
    per_rank_kernel_ms = [1e3 * t_kernel]
    per_rank_alone_ms = [t_kernel_alone]
    if world > 1:
        tt = torch.tensor([t_dev], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_dev = float(tt.item())
        tk = torch.zeros(world, dtype=torch.float64, device=dev)
        tk[rank] = 1e3 * t_kernel
        dist.all_reduce(tk)
        per_rank_kernel_ms = [round(float(v), 3) for v in tk.tolist()]
        tk.zero_()
        tk[rank] = t_kernel_alone
        dist.all_reduce(tk)
        per_rank_alone_ms = [round(float(v), 3) for v in tk.tolist()]
        dist.barrier()

    # ---- N > 1 checks: peer-memory exchange == NCCL exchange; merged recall after the timed loop ------------------------
    exchange_check, global_recall_after = None, None
    if world > 1:
        step()
        drain()
        global_recall_after = synth.recall_at_k(mk.cpu().numpy(), gtruth_np)
        if peer:
            search(ef)
            mk2, md2 = torch.empty_like(mk), torch.empty_like(md)
            nccl_merge(mk2, md2)
            torch.cuda.synchronize()
            same = torch.tensor([int(torch.equal(mk, mk2) and torch.equal(md, md2))], dtype=torch.int32, device=dev)
            dist.all_reduce(same, op=dist.ReduceOp.MIN)
            exchange_check = "peer-memory merge == nccl all_gather merge on all ranks" if int(same.item()) else "MISMATCH between peer-memory and nccl merge"
            late = ctypes.c_uint32(0)
            gv.capi.check(gv.lib().gvss_peer_exchange_status(ctypes.byref(late)))
            if late.value:
                exchange_check = f"TIMEOUT waiting for ranks (bit mask {late.value:#x})"

    # ---- e2e from HOST buffers: H2D of the queries + search (+ exchange + merge) + D2H of the result list, every step ----
    qh = torch.from_numpy(q_np).pin_memory()
    kh = torch.empty((nq, k), dtype=torch.int64).pin_memory()
    dh = torch.empty((nq, k), dtype=torch.float32).pin_memory()
    ch = torch.empty(nq, dtype=torch.int32).pin_memory()
    q_in = torch.empty_like(q)

    if world == 1:
        def e2e_step():  # the public host-pointer call
            gv.capi.check(gv.lib().gvss_search(ix._h, ctypes.c_void_p(qh.data_ptr()), nq, k, ef, ctypes.c_void_p(kh.data_ptr()),
                                               ctypes.c_void_p(dh.data_ptr()), ctypes.c_void_p(ch.data_ptr()), None))
        e2e_api = "gvss_search (host pointers)"
        h2d, d2h = nq * d * 4, nq * k * 12 + nq * 4 + 256
    else:
        def e2e_step():  # host queries -> this rank's GPU -> shard search -> exchange + merge -> merged list back on the host
            q_in.copy_(qh, non_blocking=True)
            step(qptr=q_in.data_ptr())
            with torch.cuda.stream(s_merge):
                kh.copy_(mk, non_blocking=True)
                dh.copy_(md, non_blocking=True)
            drain()
        e2e_api = "pinned host queries -> gvss_search_device per shard -> gvss_merge_topk_peers_device -> merged list to pinned host"
        h2d, d2h = nq * d * 4, nq * k * 12
    for _ in range(3):
        e2e_step()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    t_e2e = time.perf_counter() - t0
    if world > 1:
        tt = torch.tensor([t_e2e], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_e2e = float(tt.item())
    e2e_recall = synth.recall_at_k(kh.numpy(), gtruth_np)

    # bulk build end to end: rows in (pageable) host memory -> gvss_add_batch -> gvss_flush (staging copy + H2D + graph build)
    build_e2e = None
    if world == 1 and not strong:
        iy = gv.Index(d, metric, M=M, M0=M0, ef_construction=EFC_BENCH, ef_search=64, device=local)
        iy.reserve(n, 1)
        t0 = time.perf_counter()
        iy.add_batch(base_np, keys=keys_np)
        iy.flush()
        torch.cuda.synchronize()
        t_be = time.perf_counter() - t0
        build_e2e = {"value": n / t_be, "unit": "vectors/s", "seconds": t_be, "h2d_bytes": int(n) * d * 4,
                     "api": "gvss_add_batch (host pointers) + gvss_flush"}
        iy.close()

    build_vps = n / t_build
    if world > 1:
        tt = torch.tensor([t_build], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        build_vps = n_total / float(tt.item())

    failed = world > 1 and exchange_check is not None and not exchange_check.startswith("peer-memory merge ==")
    if rank == 0:
        peak, peak_src = peaks()
        achieved = algo_bytes / t_kernel / 1e9
        traffic = None  # dram read+write bytes of one launch from the committed ncu --set full capture of this exact config
        for name in ("r3_search_traffic.json", "r2_search_traffic.json"):
            try:
                tj = json.load(open(os.path.join(ROOT, "profiles", name)))
                c = tj["config"]
                if (c["n"], c["nq"], c["ef_search"], c["M"], c["dims"]) == (n, nq, ef, M, d) and tj.get("kernel_version", "r2") == gv.lib().gvss_version().decode().split()[1]:
                    traffic = tj["traffic_bytes_per_launch"]
                    break
            except Exception:
                continue
        cpu = None
        if world == 1 and not args.no_cpu:
            cpu = cpu_baseline(args, d, metric, ix.save, q_np, ef, truth_np, gpu_keys=ref_keys_np)
        units = (nq if strong else world * nq) * args.steps
        line = {
            "metric": METRIC, "value": units / t_dev, "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * t_dev / args.steps, "higher_is_better": True,
            "scaling": "strong" if strong else "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": label, "generator": gen, "n_per_gpu": n, "n_total": n_total,
                       "queries_per_step": nq, "k": k, "ef_search": ef, "recall_at_10": recall, "ef_sweep": sweep,
                       "M": M, "M0": M0, "ef_construction": EFC_BENCH, "recall_gate_met": bool(recall >= 0.95),
                       "smallest_ef_with_merged_recall_ge_0.95": ef_merged_min,
                       "units": ("queries answered by the whole job (each query is searched on every shard of the one partitioned set)" if strong
                                 else "per-shard query searches summed over ranks (every query visits every shard)"),
                       "global_queries_per_s": nq * args.steps / t_dev,
                       "global_recall_at_10": recall if world > 1 else None,
                       "global_recall_at_10_after_timed_loop": global_recall_after,
                       "recall_definition": ("merged top-10 vs exact top-10 over the union of all shards" if world > 1
                                             else "top-10 vs exact top-10 (gvss_exact_topk, bit-exact vs DuckDB's folds)"),
                       "shard_recall_at_10_min_max": [recall_min, recall_max], "shard_ef_sweep_rank0": shard_sweep,
                       "l2_policy": f"inputs larger than L2: {n * d * 4 / 1e6:.0f} MB vectors + {n * M0 * 4 / 1e6:.0f} MB adjacency per GPU vs 126 MB L2; no flush",
                       "parallelism": f"shard{world}" if world > 1 else "single", "exchange": exchange,
                       "exchange_check": exchange_check, "per_rank_kernel_ms": per_rank_kernel_ms,
                       "per_rank_kernel_ms_without_exchange": per_rank_alone_ms,
                       "per_rank_note": "kernel_ms = search launch inside the timed steps (a fast rank also waits there for the buffer set the slowest peer still merges); without_exchange = the same launch timed alone: max over ranks of the latter is what a step costs"},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src, "kernel": f"k_search<{metric}>",
                         "algorithmic_bytes_per_launch": algo_bytes, "kernel_ms": 1e3 * t_kernel},
            "cpu_baseline": cpu,
            "e2e": {"value": units / t_e2e, "unit": "queries/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "recall_at_10": e2e_recall, "api": e2e_api},
            "gpu_launches": int(launches),
            "clocks": clocks.summary(),
            "build": {"value": build_vps, "unit": "vectors/s", "seconds": t_build, "kernel_launches": int(build_launches),
                      "note": "gvss_add_batch_device + gvss_flush, rows resident in HBM, timed once before the steps (rank 0's counters below)",
                      "per_vector": {"query_distances": bst["query_distances"] / max(1, bst["inserts"]),
                                     "pair_distances": bst["pair_distances"] / max(1, bst["inserts"]),
                                     "expansions": (bst["base_expansions"] + bst["upper_expansions"]) / max(1, bst["inserts"]),
                                     "rows_written": bst["rows_written"] / max(1, bst["inserts"])},
                      "e2e": build_e2e,
                      "roofline": {"bound": "hbm", "achieved": build_bytes / t_build / 1e9, "peak": peak, "unit": "GB/s",
                                   "frac": build_bytes / t_build / 1e9 / peak, "kernel": f"k_build<{metric}>",
                                   "algorithmic_bytes": build_bytes, "note": "wall clock of the whole bulk build incl. uploads and the ramp launches"}},
            "exact_topk": {"seconds": t_exact, "queries": nq, "base": n, "queries_per_s": nq / t_exact,
                           "roofline": {"bound": "tensor", "achieved": 6.0 * nq * n * d / t_exact / 1e12, "peak": tensor_peak_tf32()[0],
                                        "unit": "TFLOP/s", "frac": 6.0 * nq * n * d / t_exact / 1e12 / tensor_peak_tf32()[0],
                                        "fp32_accurate_tflops": 2.0 * nq * n * d / t_exact / 1e12,
                                        "peak_source": tensor_peak_tf32()[1],
                                        "note": "3xTF32: 3 tensor passes x 2*Q*N*d flops over the wall clock of the whole call (split, tiles, merge, re-rank)"},
                           "note": "ground truth: tcgen05 3xTF32 candidate tiles + exact re-rank (bit-exact vs DuckDB folds), 2nd call, wall clock incl. plane split"},
            "wall_s_timed_region": t_wall,
        }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if failed:
        raise SystemExit(3)


if __name__ == "__main__":
    main()
