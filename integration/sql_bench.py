#!/usr/bin/env python
"""SQL-level timing of the drop-in: the same statements through DuckDB + vss built over libgvss (integration/_out/duckdb) and,
when present, through the stock reference build (oracle/_ref/sql/duckdb_ref, CPU usearch). Not the headline metric (bench.py
is) — it shows what the library-level numbers turn into behind the unchanged SQL surface:

    CREATE INDEX ... USING HNSW            (bulk build; N3: rows streamed into the index during the Sink phase)
    ORDER BY array_distance(...) LIMIT k   (HNSW_INDEX_SCAN, one query per statement)
    LATERAL top-k over an outer table      (HNSW_INDEX_JOIN; N1: one batched search per 2048-row chunk)

    python integration/sql_bench.py [--rows 200000] [--dims 128] [--outer 4096] [--threads 16]
Writes one JSON line per engine."""
import argparse
import json
import os
import re
import subprocess
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def script(rows, dims, outer, threads, k, exact_outer=0, gvss=True):
    vec = f"[((hash(i * {dims} + x) % 100000) / 100000.0)::FLOAT for x in range({dims})]::FLOAT[{dims}]"
    const = "[" + ", ".join(f"{((7 * j) % 100) / 100.0:.2f}" for j in range(dims)) + f"]::FLOAT[{dims}]"
    # N2 for tables WITHOUT an index: the lateral top-k join over the un-indexed table — gvss plans it as HNSW_INDEX_JOIN over a
    # transient rows-only index (n2b_exact_join_unindexed.patch), the reference runs DuckDB's cross product + window on the CPU
    unindexed = ""
    if exact_outer:
        lateral = f"SELECT count(*), sum(id) FROM q_small, LATERAL (SELECT id FROM t ORDER BY array_distance(t.vec, q_small.qv) LIMIT {k});"
        unindexed = (f"DROP INDEX idx;\nCREATE TABLE q_small AS SELECT * FROM q WHERE qid < {exact_outer};\n" + ("SET hnsw_exact_search = true;\n" if gvss else "") + ".print ##unindexed_plan\nEXPLAIN " + lateral +
                     "\n.print ##unindexed_exact_join\n" + lateral +
                     ("\n.print ##unindexed_exact_join_again\n" + lateral if gvss else "") + "\n.print ##end\n")
    return f"""
.timer on
SET threads={threads};
CREATE TABLE t AS SELECT i AS id, {vec} AS vec FROM range({rows}) r(i);
CREATE TABLE q AS SELECT i AS qid, {vec.replace('i *', '(i + 77777) *')} AS qv FROM range({outer}) r(i);
.print ##create_index
CREATE INDEX idx ON t USING HNSW (vec) WITH (metric = 'l2sq');
.print ##first_scan
SELECT id FROM t ORDER BY array_distance(vec, {const}) LIMIT {k};
.print ##scan
SELECT id FROM t ORDER BY array_distance(vec, {const}) LIMIT {k};
SELECT id FROM t ORDER BY array_distance(vec, {const}) LIMIT {k};
SELECT id FROM t ORDER BY array_distance(vec, {const}) LIMIT {k};
.print ##plan
EXPLAIN SELECT count(*), sum(id) FROM q, LATERAL (SELECT id FROM t ORDER BY array_distance(t.vec, q.qv) LIMIT {k});
.print ##index_join
SELECT count(*), sum(id) FROM q, LATERAL (SELECT id FROM t ORDER BY array_distance(t.vec, q.qv) LIMIT {k});
.print ##index_join_again
SELECT count(*), sum(id) FROM q, LATERAL (SELECT id FROM t ORDER BY array_distance(t.vec, q.qv) LIMIT {k});
.print ##info
SELECT count, capacity, approx_memory_usage FROM pragma_hnsw_index_info();
{unindexed}
"""


def run(exe, env, sql):
    t0 = time.perf_counter()
    r = subprocess.run([exe], input=sql, capture_output=True, text=True, env=env, timeout=3600)
    wall = time.perf_counter() - t0
    out = r.stdout + "\n" + r.stderr
    sections, cur = {}, "setup"
    for line in out.splitlines():
        if line.startswith("##"):
            cur = line[2:].strip()
            continue
        sections.setdefault(cur, []).append(line)
    times = {name: [float(x) for x in re.findall(r"Run Time \(s\): real ([0-9.]+)", "\n".join(lines))] for name, lines in sections.items()}
    return {"rc": r.returncode, "wall_s": wall, "times_s": times, "index_join_rows": [l for l in sections.get("index_join", []) if "│" in l][-2:],
            "uses_index_join": "HNSW_INDEX_JOIN" in "\n".join(sections.get("plan", [])),
            "unindexed_uses_index_join": "HNSW_INDEX_JOIN" in "\n".join(sections.get("unindexed_plan", [])),
            "unindexed_rows": [l for l in sections.get("unindexed_exact_join", []) if "│" in l][-2:], "info": [l for l in sections.get("info", []) if "│" in l][-2:],
            "stderr_tail": r.stderr[-500:]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=200000)
    ap.add_argument("--dims", type=int, default=128)
    ap.add_argument("--outer", type=int, default=4096)
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 8)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--exact-outer", type=int, default=256, help="outer rows of the exact lateral join on the UN-indexed table (0 = skip)")
    ap.add_argument("--skip-reference", action="store_true")
    args = ap.parse_args()
    engines = [("gvss", os.path.join(HERE, "_out", "duckdb"), {"LD_LIBRARY_PATH": f"{HERE}/_out:{ROOT}/duckdb_vss_b200"})]
    ref = os.path.join(ROOT, "oracle", "_ref", "sql", "duckdb_ref")
    if os.path.exists(ref) and not args.skip_reference:
        engines.append(("reference", ref, {}))
    for name, exe, extra in engines:
        if not os.path.exists(exe):
            print(json.dumps({"engine": name, "unavailable": exe}))
            continue
        env = dict(os.environ)
        env.update(extra)
        res = run(exe, env, script(args.rows, args.dims, args.outer, args.threads, args.k, min(args.exact_outer, args.outer), name == "gvss"))
        t = res["times_s"]
        line = {"engine": name, "rows": args.rows, "dims": args.dims, "outer_rows": args.outer, "k": args.k, "threads": args.threads,
                "create_index_s": (t.get("create_index") or [None])[0],
                "create_index_rows_per_s": args.rows / t["create_index"][0] if t.get("create_index") and t["create_index"][0] > 0 else None,
                "first_scan_s": (t.get("first_scan") or [None])[0], "scan_s": min(t["scan"]) if t.get("scan") else None,
                "index_join_s": (t.get("index_join") or [None])[-1], "index_join_again_s": (t.get("index_join_again") or [None])[-1],
                "index_join_rows_per_s": args.outer / t["index_join_again"][-1] if t.get("index_join_again") and t["index_join_again"][-1] > 0 else None,
                "unindexed_exact_join_outer_rows": min(args.exact_outer, args.outer),
                "unindexed_exact_join_s": (t.get("unindexed_exact_join") or [None])[-1],
                "unindexed_exact_join_again_s": (t.get("unindexed_exact_join_again") or [None])[-1],
                "unindexed_plan_has_hnsw_index_join": res["unindexed_uses_index_join"], "unindexed_result": res["unindexed_rows"],
                "plan_has_hnsw_index_join": res["uses_index_join"], "join_result": res["index_join_rows"], "index_info": res["info"],
                "rc": res["rc"], "stderr_tail": res["stderr_tail"] if res["rc"] else ""}
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
