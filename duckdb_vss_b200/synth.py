"""Seeded synthetic data sets of SURVEY.md §8(d) / Appendix D (numpy, deterministic for a given seed).

``clustered(n, d, C, r)``: pick one of C Gaussian centres in R^d, add a rank-r Gaussian latent through a fixed random
r×d projection (sigma 0.35) plus 0.05·N(0, I) noise — intrinsic dimension ≈ r, which is what SIFT / Deep / text
embeddings have and iid noise lacks (iid data cannot reach recall@10 ≥ 0.95 at M=16)."""
import numpy as np


def readme_grid():
    """Config 1: all (a, b, c), a,b,c in 1..9, lexicographic (README.md:12-47)."""
    return np.array([(a, b, c) for a in range(1, 10) for b in range(1, 10) for c in range(1, 10)], dtype=np.float32)


def clustered(n, d, C, r, seed, n_queries=0, sigma=0.35, noise=0.05, chunk=262144, point_seed=None):
    """``point_seed``: draw the points from a stream of their own while centres and projection still come from ``seed`` —
    another sample of the SAME data set (the shards of one partitioned base set), not another data set."""
    rng = np.random.default_rng(seed)
    centres = rng.standard_normal((C, d), dtype=np.float32)
    proj = (sigma * rng.standard_normal((r, d), dtype=np.float32)).astype(np.float32)
    if point_seed is not None:
        rng = np.random.default_rng(point_seed)

    def draw(m):
        out = np.empty((m, d), dtype=np.float32)
        for b in range(0, m, chunk):
            e = min(m, b + chunk)
            c = rng.integers(0, C, size=e - b)
            z = rng.standard_normal((e - b, r), dtype=np.float32)
            x = centres[c] + z @ proj
            x += noise * rng.standard_normal((e - b, d), dtype=np.float32)
            out[b:e] = x
        return out

    base = draw(n)
    queries = draw(n_queries) if n_queries else np.empty((0, d), dtype=np.float32)
    return base, queries


def config(name, n=None, nq=None):
    """The named workloads of BASELINE.json:configs (sizes can be scaled down for tests)."""
    if name == "readme":
        return {"metric": "l2sq", "base": readme_grid(), "queries": np.array([[1, 2, 3]], np.float32), "k": 3}
    if name == "sift1m":  # 1M x 128 l2sq, clustered(C = max(1024, N/100), r = 16), seeds 42
        n = n or 1_000_000
        base, q = clustered(n, 128, max(1024, n // 100), 16, 42, nq or 10_000)
        return {"metric": "l2sq", "base": base, "queries": q, "k": 10}
    if name == "text768":  # 1M x 768 cosine, clustered(C = N/100, r = 32), rows normalised then scaled by U(0.5, 2)
        n = n or 1_000_000
        base, q = clustered(n, 768, max(64, n // 100), 32, 44, nq or 10_000)
        rng = np.random.default_rng(45)
        for a in (base, q):
            a /= np.linalg.norm(a, axis=1, keepdims=True)
            a *= rng.uniform(0.5, 2.0, size=(a.shape[0], 1)).astype(np.float32)
        return {"metric": "cosine", "base": base, "queries": q, "k": 10}
    if name == "deep10m":  # 10M x 96 ip, clustered(C = N/100, r = 24), L2-normalised
        n = n or 10_000_000
        base, q = clustered(n, 96, max(64, n // 100), 24, 46, nq or 10_000)
        for a in (base, q):
            a /= np.linalg.norm(a, axis=1, keepdims=True)
        return {"metric": "ip", "base": base, "queries": q, "k": 10}
    if name == "join10m":  # config 5: 100k queries x 10M x 128 l2sq, generator as sift1m at 10M (C = N/100), seed 48
        n = n or 10_000_000
        base, q = clustered(n, 128, max(1024, n // 100), 16, 48, nq or 100_000)
        return {"metric": "l2sq", "base": base, "queries": q, "k": 10}
    raise KeyError(name)


def recall_at_k(found, truth):
    """Mean |found ∩ truth| / k over queries (ids, -1 ignored)."""
    hits = 0
    for f, t in zip(found, truth):
        hits += len(set(int(x) for x in f if x >= 0) & set(int(x) for x in t))
    return hits / float(truth.shape[0] * truth.shape[1])
