#!/bin/bash
# TEST INFRASTRUCTURE ONLY. Builds the UNMODIFIED reference at SQL level — DuckDB (the reference's submodule) with the stock vss
# extension over the vendored usearch — from where the sources lie, out of tree; the shell goes to oracle/_ref/sql/ (git-ignored,
# travels to the GPU box) as the CPU side of integration/sql_bench.py. ~9 min on 8 cores.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${VSS_REFERENCE_DIR:-/root/reference}"
BUILD="${BUILD:-/tmp/ref_sql_build}"
cmake -G Ninja -DCMAKE_POLICY_VERSION_MINIMUM=3.5 -DOVERRIDE_GIT_DESCRIBE=v1.1.3 -DEXTENSION_STATIC_BUILD=1 \
      -DDUCKDB_EXTENSION_CONFIGS="$REF/extension_config.cmake" -DCMAKE_BUILD_TYPE=Release -DENABLE_SANITIZER=0 -DENABLE_UBSAN=0 \
      -S "$REF/duckdb" -B "$BUILD"
ninja -C "$BUILD" -j "${1:-6}" duckdb
mkdir -p "$HERE/_ref/sql"
cp "$BUILD/duckdb" "$HERE/_ref/sql/duckdb_ref"
strip "$HERE/_ref/sql/duckdb_ref"
ls -la "$HERE/_ref/sql"
