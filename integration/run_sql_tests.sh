#!/bin/bash
# Runs the reference's own sqllogictests (test/sql/hnsw/*.test, unchanged) against DuckDB + vss built over libgvss
# (integration/build_sql.sh). Needs a B200: without a GPU every CREATE INDEX fails with libgvss's "no CUDA device" error.
#   integration/run_sql_tests.sh [pattern]      default pattern: "test/sql/hnsw/*"
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(dirname "$HERE")"
export LD_LIBRARY_PATH="$HERE/_out:$ROOT/duckdb_vss_b200:${LD_LIBRARY_PATH:-}"
cd "$HERE/_out" || exit 2
# this repository's own sqllogictests for the optional patches run from the same test directory
mkdir -p "$HERE/_out/test/sql" && rm -rf "$HERE/_out/test/sql/gvss" && cp -r "$HERE/test/sql/gvss" "$HERE/_out/test/sql/"
exec ./unittest --test-dir "$HERE/_out" "${1:-test/sql/hnsw/*}"
