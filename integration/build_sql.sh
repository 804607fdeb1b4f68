#!/bin/bash
# Builds DuckDB (the reference's submodule, in place, read-only) with the vss extension compiled over the gvss shim and
# linked to libgvss.so — the SQL-level artefact (SURVEY.md §7.3 artefact 3). Out of tree: objects under $BUILD
# (default /tmp/gvss_sql_build), the two binaries and the reference's sqllogictest scripts (inputs of the run, not sources)
# are copied to integration/_out/ (git-ignored; it travels to the GPU box). ~9 min on 8 cores.
#   integration/build_sql.sh [-j N]          then on a B200:   integration/run_sql_tests.sh
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(dirname "$HERE")"
REF="${VSS_REFERENCE_DIR:-/root/reference}"
BUILD="${BUILD:-/tmp/gvss_sql_build}"
JOBS="${2:-6}"
[ -f "$ROOT/duckdb_vss_b200/libgvss.so" ] || make -C "$ROOT/duckdb_vss_b200/csrc" -j4
# the two extra flags: this cmake rejects cmake_minimum_required(2.8.12); the submodule has no .git for `git describe`
cmake -G Ninja -DCMAKE_POLICY_VERSION_MINIMUM=3.5 -DOVERRIDE_GIT_DESCRIBE=v1.1.3 -DEXTENSION_STATIC_BUILD=1 \
      -DDUCKDB_EXTENSION_CONFIGS="$HERE/extension_config.cmake" -DVSS_REFERENCE_DIR="$REF" \
      -DCMAKE_BUILD_TYPE=Release -DBUILD_UNITTESTS=1 -DENABLE_SANITIZER=0 -DENABLE_UBSAN=0 \
      -S "$REF/duckdb" -B "$BUILD"
ninja -C "$BUILD" -j "$JOBS" unittest duckdb
mkdir -p "$HERE/_out/test/sql"
cp "$BUILD/test/unittest" "$BUILD/duckdb" "$BUILD/src/libduckdb.so" "$HERE/_out/"
strip "$HERE/_out/unittest" "$HERE/_out/duckdb" "$HERE/_out/libduckdb.so"
rm -rf "$HERE/_out/test/sql/hnsw" "$HERE/_out/test/sql/slow"
cp -r "$REF/test/sql/hnsw" "$REF/test/sql/slow" "$HERE/_out/test/sql/"
ls -la "$HERE/_out"
