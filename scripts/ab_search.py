#!/usr/bin/env python
"""A/B harness for k_search: build the bench graph ONCE, then time the same 10k-query batch under several library
builds / environment settings, each in its own process (the knobs are read once per process).

    python scripts/ab_search.py [--vectors 1000000] [--nq 10000] [--ef 384] name[:lib.so][,VAR=val...] ...

e.g.  python scripts/ab_search.py base g4:duckdb_vss_b200/libgvss_g4.so base2,GVSS_STAGE_ROWS=32
Prints per variant: QPS (CUDA events, 3 warm-up + 5 timed launches), recall@10, ids identical to the first variant.
"""
import argparse
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
TMP = "/tmp/gvss_ab"


def child(args):
    import torch
    import duckdb_vss_b200 as gv
    q_np = np.load(f"{TMP}/q.npy")
    truth = np.load(f"{TMP}/truth.npy")
    d, nq, k = q_np.shape[1], q_np.shape[0], 10
    ix = gv.Index(d, "l2sq", M=32, M0=64, ef_construction=200, ef_search=64, device=0)
    ix.load(open(f"{TMP}/graph.usearch", "rb").read())
    step = int(os.environ.get("AB_REMOVE_EVERY", "0"))  # lazy-delete every step-th member before searching
    if step:
        import ctypes
        removed = ctypes.c_uint64(0)
        for key in range(0, ix.size(), step):
            gv.lib().gvss_remove(ix._h, key, ctypes.byref(removed))
    dev = torch.device("cuda", 0)
    q = torch.from_numpy(q_np).to(dev)
    okeys = torch.empty((nq, k), dtype=torch.int64, device=dev)
    odist = torch.empty((nq, k), dtype=torch.float32, device=dev)
    ocnt = torch.empty(nq, dtype=torch.int32, device=dev)
    ostat = torch.zeros((nq, 4), dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream

    def search(stats=False):
        ix.search_device(q.data_ptr(), nq, k, args.ef, okeys.data_ptr(), odist.data_ptr(), ocnt.data_ptr(),
                         ostat.data_ptr() if stats else 0, stream)
    for _ in range(3):
        search()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.steps):
        search()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / args.steps
    search(stats=True)
    torch.cuda.synchronize()
    keys = okeys.cpu().numpy()
    st = ostat.cpu().numpy().view(np.uint32)
    from duckdb_vss_b200 import synth
    rec = synth.recall_at_k(keys, truth)
    same = "first"
    if os.path.exists(f"{TMP}/keys0.npy"):
        same = str(bool(np.array_equal(keys, np.load(f"{TMP}/keys0.npy"))))
    else:
        np.save(f"{TMP}/keys0.npy", keys)
    print(f"{args.child:28s} {nq / ms:8.1f} kQPS  {ms:8.3f} ms  recall {rec:.4f}  ids==first {same}  "
          f"computed/q {st[:, 0].mean():.0f} hops/q {st[:, 2].mean():.0f} flags bitmap {np.mean((st[:, 3] & 2) != 0):.4f} tie {np.mean((st[:, 3] & 4) != 0):.4f}", flush=True)
    try:
        fn = gv.lib().gvssdbg_opcounts
    except AttributeError:
        return
    import ctypes
    buf = (ctypes.c_ulonglong * 32)()
    n = fn(buf, 32)
    names = ["queries", "hops", "new", "hops_le16", "hops_none", "pushes", "prunes", "prune_kept", "ins_batches", "inserted",
             "shifted", "vis_rounds", "maybe_batches", "push_levels", "pop_levels", "clk_select", "clk_gather", "clk_dist",
             "clk_lb", "clk_stage", "clk_wait", "clk_math", "steps", "clk_total", "clk_accept", "clk_shift", "clk_adj", "clk_vis", "clk_shiftloop", "pred_hit"]
    nqr = max(1, buf[0])
    print("   per query: " + "  ".join(f"{names[i]} {buf[i] / nqr:.1f}" for i in range(1, min(n, len(names)))), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--vectors", type=int, default=1_000_000)
    ap.add_argument("--nq", type=int, default=10_000)
    ap.add_argument("--ef", type=int, default=384)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--child", default=None)
    ap.add_argument("variants", nargs="*")
    args = ap.parse_args()
    if args.child:
        return child(args)

    import torch
    import duckdb_vss_b200 as gv
    from duckdb_vss_b200 import synth, index as gvi
    os.makedirs(TMP, exist_ok=True)
    if os.path.exists(f"{TMP}/keys0.npy"):
        os.remove(f"{TMP}/keys0.npy")
    n, nq, d, k = args.vectors, args.nq, 128, 10
    base_np, _ = synth.clustered(n, d, max(1024, n // 100), 16, 42, 0)
    _, q_np = synth.clustered(1, d, max(1024, n // 100), 16, 42, nq)
    dev = torch.device("cuda", 0)
    base = torch.from_numpy(base_np).to(dev)
    q = torch.from_numpy(q_np).to(dev)
    ix = gv.Index(d, "l2sq", M=32, M0=64, ef_construction=200, ef_search=64, device=0)
    ix.reserve(n, 1)
    t0 = time.perf_counter()
    ix.add_batch_device(base.data_ptr(), n)
    ix.flush()
    torch.cuda.synchronize()
    print(f"build {n} vectors: {time.perf_counter() - t0:.2f} s", flush=True)
    truth = torch.empty((nq, k), dtype=torch.int64, device=dev)
    tdist = torch.empty((nq, k), dtype=torch.float32, device=dev)
    gvi.exact_topk_device("l2sq", base.data_ptr(), n, q.data_ptr(), nq, d, k, truth.data_ptr(), tdist.data_ptr(), 0,
                          torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    np.save(f"{TMP}/q.npy", q_np)
    np.save(f"{TMP}/truth.npy", truth.cpu().numpy())
    open(f"{TMP}/graph.usearch", "wb").write(ix.save())
    ix.close()
    del base
    torch.cuda.empty_cache()
    for v in args.variants or ["base"]:
        parts = v.split(",")
        name_lib = parts[0].split(":")
        env = dict(os.environ)
        if len(name_lib) > 1:
            env["GVSS_LIB"] = os.path.join(ROOT, name_lib[1])
        for kv in parts[1:]:
            key, val = kv.split("=")
            env[key] = val
        subprocess.run([sys.executable, os.path.abspath(__file__), "--child", v, "--ef", str(args.ef), "--steps", str(args.steps)],
                       env=env, timeout=900)


if __name__ == "__main__":
    main()
