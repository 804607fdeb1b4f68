"""CPU: the drop-in claim at compile level. The reference's UNCHANGED extension sources must compile against the gvss shim
(duckdb_vss_b200/shim ahead of the vendored usearch on the include path), and the optional operator patches
(integration/n1_*.patch, n2_*.patch, n2b_*.patch, n3_*.patch) must apply to the reference and compile.
Needs the reference checkout (/root/reference): skipped where it is absent (the GPU box). The linked and SQL-tested
artefact is integration/build_sql.sh + integration/run_sql_tests.sh (profiles/r3_sql_tests_gpu.log)."""
import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("VSS_REFERENCE_DIR", "/root/reference")
HNSW = ["hnsw_index.cpp", "hnsw_index_macros.cpp", "hnsw_index_physical_create.cpp", "hnsw_index_pragmas.cpp", "hnsw_index_scan.cpp",
        "hnsw_index_plan.cpp", "hnsw_topk_operator.cpp", "hnsw_optimize_expr.cpp", "hnsw_optimize_join.cpp", "hnsw_optimize_topk.cpp",
        "hnsw_optimize_scan.cpp"]

pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "src", "hnsw")), reason="reference checkout not present")


def _syntax_check(src, extra_includes=()):
    cmd = ["g++", "-std=c++17", "-fsyntax-only", "-DDUCKDB_USEARCH_USE_SIMSIMD=0", "-w"]
    for inc in list(extra_includes) + [f"{ROOT}/duckdb_vss_b200/shim", f"{ROOT}/include", f"{REF}/src/include",
                                      f"{REF}/duckdb/src/include", f"{REF}/duckdb/third_party/re2", f"{REF}/duckdb/third_party/fmt/include",
                                      f"{REF}/duckdb/extension/core_functions/include"]:
        cmd += ["-I", inc]
    r = subprocess.run(cmd + ["-M", src], capture_output=True, text=True)   # which duckdb_usearch.hpp did it resolve?
    assert r.returncode == 0, f"{src}:\n{r.stderr[-3000:]}"
    deps = r.stdout
    if "duckdb_usearch.hpp" in deps:
        assert f"{ROOT}/duckdb_vss_b200/shim/usearch/duckdb_usearch.hpp" in deps, f"{src} resolved the vendored usearch, not the shim"
        assert f"{REF}/src/include/usearch/index.hpp" not in deps
    r = subprocess.run(cmd + [src], capture_output=True, text=True)
    assert r.returncode == 0, f"{src}:\n{r.stderr[-3000:]}"
    return src


def test_unchanged_extension_sources_compile_against_the_shim():
    srcs = [f"{REF}/src/hnsw/{f}" for f in HNSW] + [f"{REF}/src/vss_extension.cpp"]
    with ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        done = list(ex.map(_syntax_check, srcs))
    assert len(done) == len(srcs)


@pytest.mark.parametrize("patches,files", [
    (["n1_batched_index_join.patch"], ["src/hnsw/hnsw_index.cpp", "src/hnsw/hnsw_optimize_join.cpp", "src/include/hnsw/hnsw_index.hpp"]),
    (["n1_batched_index_join.patch", "n2_exact_mode.patch"], ["src/hnsw/hnsw_index.cpp", "src/hnsw/hnsw_optimize_join.cpp", "src/include/hnsw/hnsw_index.hpp"]),
    (["n1_batched_index_join.patch", "n2_exact_mode.patch", "n2b_exact_join_unindexed.patch"],
     ["src/hnsw/hnsw_index.cpp", "src/hnsw/hnsw_optimize_join.cpp", "src/include/hnsw/hnsw_index.hpp"]),
    (["n3_streaming_sink.patch"], ["src/hnsw/hnsw_index_physical_create.cpp"]),
])
def test_operator_patches_apply_to_the_reference_and_compile(tmp_path, patches, files):
    for f in files:
        os.makedirs(tmp_path / os.path.dirname(f), exist_ok=True)
        shutil.copy(os.path.join(REF, f), tmp_path / f)
    for patch in patches:
        r = subprocess.run(["patch", "-p1", "-i", os.path.join(ROOT, "integration", patch)], cwd=tmp_path, capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
    srcs = [str(tmp_path / f) for f in files if f.endswith(".cpp")]
    with ThreadPoolExecutor(max_workers=4) as ex:
        list(ex.map(lambda s: _syntax_check(s, extra_includes=[str(tmp_path / "src" / "include")]), srcs))


def test_sql_bench_script_is_well_formed():
    """integration/sql_bench.py: the generated shell script names every section its parser reads, for both engines (the stock
    build must not see the hnsw_exact_search setting it does not know)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("sql_bench", os.path.join(ROOT, "integration", "sql_bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    for gvss in (True, False):
        sql = mod.script(1000, 8, 64, 4, 10, 16, gvss)
        for section in ("create_index", "first_scan", "scan", "plan", "index_join", "index_join_again", "info", "unindexed_plan", "unindexed_exact_join"):
            assert f".print ##{section}\n" in sql, section
        assert ("hnsw_exact_search" in sql) == gvss
        assert ("##unindexed_exact_join_again" in sql) == gvss
        assert sql.index("CREATE INDEX idx") < sql.index("DROP INDEX idx") < sql.index("##unindexed_plan")
    assert "##unindexed_plan" not in mod.script(1000, 8, 64, 4, 10, 0, True)
