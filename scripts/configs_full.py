#!/usr/bin/env python
"""Full-size runs of BASELINE.json configs[2] (1M x FLOAT[768] cosine) and configs[3] (10M x FLOAT[96] ip) on one GPU
(or one shard per rank under torchrun): build, exact ground truth, ef sweep with recall@10 and device-timed QPS, and a
bounded identity check against the compiled reference on the same GPU-built graph.

    python scripts/configs_full.py text768 [--vectors N] [--check-queries 300]
    python scripts/configs_full.py deep10m [--vectors N]

Prints one JSON line per config; nothing here is the headline bench (bench.py is)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
M, M0, EFC = 32, 64, 200
EF_SWEEP = (32, 64, 128, 256, 384, 512)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", choices=["text768", "deep10m", "sift1m", "join10m"])
    ap.add_argument("--vectors", type=int, default=0)
    ap.add_argument("--nq", type=int, default=0, help="0 = the config's own query count")
    ap.add_argument("--check-queries", type=int, default=300, help="queries replayed on the compiled reference (0 = skip)")
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()

    import torch
    import duckdb_vss_b200 as gv
    from duckdb_vss_b200 import synth, index as gvi

    t0 = time.perf_counter()
    cfg = synth.config(args.config, n=args.vectors or None, nq=args.nq or None)
    base_np, q_np, metric, k = cfg["base"], cfg["queries"], cfg["metric"], cfg["k"]
    n, d = base_np.shape
    nq = q_np.shape[0]
    t_gen = time.perf_counter() - t0
    dev = torch.device("cuda", 0)
    base = torch.from_numpy(base_np).to(dev)
    q = torch.from_numpy(q_np).to(dev)
    stream = torch.cuda.current_stream().cuda_stream

    ix = gv.Index(d, metric, M=M, M0=M0, ef_construction=EFC, ef_search=64, device=0)
    ix.reserve(n, 1)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ix.add_batch_device(base.data_ptr(), n)
    ix.flush()
    torch.cuda.synchronize()
    t_build = time.perf_counter() - t0

    truth = torch.empty((nq, k), dtype=torch.int64, device=dev)
    tdist = torch.empty((nq, k), dtype=torch.float32, device=dev)
    t_exact = None
    for _ in range(2):
        t0 = time.perf_counter()
        gvi.exact_topk_device(metric, base.data_ptr(), n, q.data_ptr(), nq, d, k, truth.data_ptr(), tdist.data_ptr(), 0, stream)
        torch.cuda.synchronize()
        t_exact = time.perf_counter() - t0
    truth_np = truth.cpu().numpy()

    okeys = torch.empty((nq, k), dtype=torch.int64, device=dev)
    odist = torch.empty((nq, k), dtype=torch.float32, device=dev)
    ocnt = torch.empty(nq, dtype=torch.int32, device=dev)
    ostat = torch.zeros((nq, 4), dtype=torch.int32, device=dev)
    sweep = {}
    for ef in EF_SWEEP:
        def run(stats=False):
            ix.search_device(q.data_ptr(), nq, k, ef, okeys.data_ptr(), odist.data_ptr(), ocnt.data_ptr(),
                             ostat.data_ptr() if stats else 0, stream)
        run()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(args.steps):
            run()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / args.steps
        run(stats=True)
        torch.cuda.synchronize()
        st = ostat.cpu().numpy().view(np.uint32)
        rec = synth.recall_at_k(okeys.cpu().numpy(), truth_np)
        algo = float(st[:, 0].astype(np.float64).sum() * 4 * d + st[:, 2].astype(np.float64).sum() * (4 + 4 * M0)
                     + (st[:, 1].astype(np.float64) - st[:, 2]).sum() * (4 + 4 * M) + nq * 12.0 * k)
        sweep[ef] = {"recall_at_10": round(rec, 4), "qps": round(nq / ms * 1e3), "ms": round(ms, 3),
                     "algorithmic_GBps": round(algo / ms / 1e6, 1), "tie_fraction": float(np.mean((st[:, 3] & 4) != 0)),
                     "bitmap_fraction": float(np.mean((st[:, 3] & 2) != 0))}
        if rec >= 0.95:
            break
    op_ef = max(sweep)

    check = None
    if args.check_queries:
        import oracle
        if oracle.have_ref():
            from oracle import ref as impl
            ncheck = min(args.check_queries, nq)
            ref = impl.RefIndex(d, metric, M=M, M0=M0, ef_construction=EFC)
            t0 = time.perf_counter()
            ref.load(ix.save())
            t_load = time.perf_counter() - t0
            threads = impl.hardware_threads()
            t0 = time.perf_counter()
            r = ref.search_batch(q_np[:ncheck], k, op_ef, threads=threads)
            dt = time.perf_counter() - t0
            run(stats=False)
            torch.cuda.synchronize()
            gk, gd = okeys.cpu().numpy()[:ncheck], odist.cpu().numpy()[:ncheck]
            check = {"queries": ncheck, "ef": op_ef, "ids_identical": bool(np.array_equal(r["keys"], gk)),
                     "distance_bits_identical": bool(np.array_equal(r["dist"].view(np.uint32), gd.view(np.uint32))),
                     "reference_qps": round(ncheck / dt, 1), "reference_threads": threads, "stream_load_s": round(t_load, 1)}
    print(json.dumps({"config": args.config, "n": n, "d": d, "metric": metric, "queries": nq, "M": M, "M0": M0,
                      "ef_construction": EFC, "generate_s": round(t_gen, 1), "build_s": round(t_build, 2),
                      "build_vectors_per_s": round(n / t_build), "exact_topk_s": round(t_exact, 3), "sweep": sweep,
                      "reference_check_on_gpu_built_graph": check}))


if __name__ == "__main__":
    main()
