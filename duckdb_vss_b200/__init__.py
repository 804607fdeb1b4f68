"""duckdb_vss_b200 — B200-native HNSW hot path of DuckDB's vss extension behind a C ABI (include/gvss.h).

The product is ``libgvss.so`` (hand-written sm_100a CUDA + a C ABI); this package is only the thin Python
binding used by tests and bench.py. There is no CPU fallback: importing works without a GPU (so that the
symbol table can be checked), but every compute call fails loudly when no CUDA device / extension is present.
"""
from .capi import lib, LIB_PATH, GvssError  # noqa: F401
from .index import Index, exact_topk, merge_topk_device  # noqa: F401

__all__ = ["lib", "LIB_PATH", "GvssError", "Index", "exact_topk", "merge_topk_device"]
