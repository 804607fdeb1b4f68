"""GPU: the C++ drop-in shim (duckdb_vss_b200/shim/usearch/duckdb_usearch.hpp) driven with the extension's own call
patterns (tests/cpp/shim_driver.cpp cites each call site). Compiles with g++ against include/gvss.h + libgvss.so."""
import os
import re
import subprocess
import pytest
from duckdb_vss_b200 import capi

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shim_driver_behaves_like_the_reference_index(tmp_path):
    exe = str(tmp_path / "shim_driver")
    libdir = os.path.dirname(capi.LIB_PATH)
    subprocess.run(["g++", "-std=c++17", "-O1", "-pthread", f"-I{ROOT}/include", f"-I{ROOT}/duckdb_vss_b200/shim",
                    f"{ROOT}/tests/cpp/shim_driver.cpp", f"-L{libdir}", "-lgvss", f"-Wl,-rpath,{libdir}", "-o", exe], check=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True, timeout=120).stdout
    print(out)
    assert 'add_failures 0 size 729 capacity 729 over_capacity_error "Reserve capacity ahead of insertions!"' in out
    m = re.search(r"scan 3: (\d+)/0 (\d+)/1 (\d+)/1", out)          # hnsw_result.test:20-25 (squared distances 0, 1, 1)
    assert m and m.group(1) == "11"
    assert "removed 1 size 728" in out and "level0_nodes 729" in out and "dims 3 metric e" in out
    assert "removed_key_returned 0" in out and "compact_ok 1" in out
    m = re.search(r"stream_bytes (\d+) serialized_length (\d+) load_ok 1 reloaded_size 728 same_results 1", out)
    assert m and m.group(1) == m.group(2)
    assert "after_reset_size 0" in out
    assert "concurrent_insert_scan failures 0 size 729 nearest 5011" in out   # row (1,2,3) is grid row 11
    assert "thread_out_of_range_refused 1" in out


def test_multi_scan_prefetch_serves_the_index_join_from_one_batched_search(tmp_path):
    """per-row ExecuteMultiScan loop vs the prefetching cache — identical ids for all 2048 outer rows."""
    exe = str(tmp_path / "multiscan_driver")
    libdir = os.path.dirname(capi.LIB_PATH)
    subprocess.run(["g++", "-std=c++17", "-O1", "-pthread", f"-I{ROOT}/include", f"-I{ROOT}/duckdb_vss_b200/shim",
                    f"{ROOT}/tests/cpp/multiscan_driver.cpp", f"-L{libdir}", "-lgvss", f"-Wl,-rpath,{libdir}", "-o", exe], check=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True, timeout=300).stdout
    print(out)
    assert "prefetch_ok 1 served 2048/2048 same_ids 1 mismatched_counts 0 foreign 0 other_limit 0 after_reset 0" in out
    m = re.search(r"speedup ([0-9.]+)", out)
    assert m and float(m.group(1)) > 3.0  # one launch + one PCIe round trip instead of 2048 (measured: several hundred; the bar only guards the mechanism)


def test_background_build_overlaps_the_feeding_threads(tmp_path):
    """T threads feed 2048-row chunks (PhysicalCreateHNSWIndex pattern) at a throttled 'table scan' rate.
    With gvss_set_background_build the graph is built while the rows arrive: the streamed run ends soon after the scan,
    the one-shot run pays scan + build; both graphs answer alike."""
    exe = str(tmp_path / "sink_driver")
    libdir = os.path.dirname(capi.LIB_PATH)
    subprocess.run(["g++", "-std=c++17", "-O2", "-pthread", f"-I{ROOT}/include", f"{ROOT}/tests/cpp/sink_driver.cpp", f"-L{libdir}",
                    "-lgvss", f"-Wl,-rpath,{libdir}", "-o", exe], check=True)
    seen = []
    for attempt in range(3):                              # wall-clock comparison on a shared host: one quiet run out of three is enough
        out = subprocess.run([exe, "400000", "128", "8", "400000"], check=True, capture_output=True, text=True, timeout=600).stdout
        print(out)
        seen.append(out)
        m = re.search(r"one_shot_s ([0-9.]+) streamed_s ([0-9.]+) scan_s ([0-9.]+) recall_one_shot ([0-9.]+) recall_streamed ([0-9.]+)", out)
        assert m, out
        one_shot, streamed, scan, r1, r2 = map(float, m.groups())
        assert abs(r1 - r2) < 0.015 and r2 > 0.7, out     # both graphs answer alike, every time
        if streamed < one_shot:                           # the build overlapped the scan
            break
    else:
        raise AssertionError("streamed build never beat the one-shot build:\n" + "\n".join(seen))


def test_persistence_round_trip_through_callbacks_is_pipelined_and_exact(tmp_path):
    """gvss_save / gvss_load through in-memory callbacks (the shape of HNSWIndex::PersistToDisk / the loading constructor,
    hnsw_index.cpp:193-206, 510-532): the stream has exactly serialized_length bytes and the reloaded index answers alike."""
    exe = str(tmp_path / "persist_driver")
    libdir = os.path.dirname(capi.LIB_PATH)
    subprocess.run(["g++", "-std=c++17", "-O2", f"-I{ROOT}/include", f"{ROOT}/tests/cpp/persist_driver.cpp", f"-L{libdir}", "-lgvss",
                    f"-Wl,-rpath,{libdir}", "-o", exe], check=True)
    out = subprocess.run([exe, "150000", "96"], check=True, capture_output=True, text=True, timeout=600).stdout
    print(out)
    m = re.search(r"stream_bytes (\d+) serialized_length (\d+) save_s ([0-9.]+) .* load_s ([0-9.]+) .* same_results 1", out)
    assert m and m.group(1) == m.group(2), out


def test_concurrent_single_query_scans_keep_up_with_the_reference_threads(tmp_path):
    """T host threads each issue one query per call (concurrent HNSW_INDEX_SCANs); the calls overlap on the GPU (a search slot
    each) and together keep up with the compiled reference searching the same graph with T threads; ids identical."""
    import oracle
    exe = str(tmp_path / "scan_driver")
    libdir = os.path.dirname(capi.LIB_PATH)
    subprocess.run(["g++", "-std=c++17", "-O2", "-pthread", f"-I{ROOT}/include", f"{ROOT}/tests/cpp/scan_driver.cpp", f"-L{libdir}", "-lgvss",
                    "-ldl", f"-Wl,-rpath,{libdir}", "-o", exe], check=True)
    ref = oracle.REF_SO if oracle.have_ref() else "-"
    T = min(16, os.cpu_count() or 4)
    out = subprocess.run([exe, "200000", "128", "32", "128", str(T), ref], check=True, capture_output=True, text=True, timeout=900).stdout
    print(out)
    m = re.search(r"gvss single_query_median_ms ([0-9.]+) p95_ms [0-9.]+ threads_qps ([0-9.]+) failures 0 concurrent_peak (\d+)", out)
    assert m, out
    assert int(m.group(3)) >= 2, out
    r = re.search(r"reference single_query_mean_ms ([0-9.]+) threads_qps ([0-9.]+) same_ids (\d)", out)
    if r:
        assert r.group(3) == "1", out
        # measured on the B200 box (16 cores): 1.3-1.4x the reference's T-thread rate, 0.7x its single-thread latency; the bars below
        # only guard against the calls being serialised again
        assert float(m.group(2)) >= 0.5 * float(r.group(2)), out
        assert float(m.group(1)) <= 1.5 * float(r.group(1)), out
