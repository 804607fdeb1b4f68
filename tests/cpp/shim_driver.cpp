// Drives the drop-in shim exactly the way the reference's HNSWIndex does (call sites cited inline) and prints a small
// report that tests/test_gpu_shim.py checks. Build: g++ -std=c++17 -Iinclude -Iduckdb_vss_b200/shim shim_driver.cpp -L... -lgvss
#include "usearch/duckdb_usearch.hpp"

#include <atomic>
#include <cstdio>
#include <thread>
#include <vector>

using row_t = int64_t;
using USearchIndexType = unum::usearch::index_dense_gt<row_t>; // hnsw_index.hpp:28

int main() {
    // HNSWIndex ctor, hnsw_index.cpp:162-189
    unum::usearch::metric_punned_t metric(3, unum::usearch::metric_kind_t::l2sq_k, unum::usearch::scalar_kind_t::f32_k);
    unum::usearch::index_dense_config_t config = {};
    config.enable_key_lookups = false;
    USearchIndexType index = USearchIndexType::make(metric, config);
    if (!index) { std::printf("create failed: %s\n", gvss_last_error()); return 2; }
    index.reserve(32);                                            // hnsw_index.cpp:208

    // PhysicalCreateHNSWIndex::Finalize + HNSWIndexConstructTask, hnsw_index_physical_create.cpp:144-205,296
    std::vector<float> rows;
    for (int a = 1; a <= 9; a++) for (int b = 1; b <= 9; b++) for (int c = 1; c <= 9; c++) { rows.push_back(a); rows.push_back(b); rows.push_back(c); }
    const size_t n = 729, T = 4;
    index.reserve({n, T});
    std::atomic<size_t> cursor{0};
    std::atomic<int> failures{0};
    std::vector<std::thread> tasks;
    for (size_t t = 0; t < T; t++)
        tasks.emplace_back([&, t] {
            for (;;) {
                size_t i = cursor.fetch_add(1);
                if (i >= n) break;
                const auto result = index.add((row_t)i, rows.data() + 3 * i, t);
                if (!result) failures++;
            }
        });
    for (auto &t : tasks) t.join();
    const auto over = index.add(9999, rows.data(), 0);            // beyond capacity -> error text of index.hpp:2332-2335
    std::printf("add_failures %d size %zu capacity %zu over_capacity_error \"%s\"\n", failures.load(), index.size(), index.capacity(),
                over ? "" : over.error.what());

    // HNSWIndex::InitializeScan, hnsw_index.cpp:285-311
    float q[3] = {1, 2, 3};
    auto search_result = index.ef_search(q, 3, index.expansion_search());
    std::vector<row_t> ids(search_result.size());
    std::vector<float> dist(search_result.size());
    search_result.dump_to(ids.data(), dist.data());
    std::printf("scan %zu:", search_result.size());
    for (size_t i = 0; i < ids.size(); i++) std::printf(" %lld/%g", (long long)ids[i], dist[i]);
    std::printf("\n");

    // HNSWIndex::Delete, hnsw_index.cpp:466-482, then GetStats, :262-276
    auto removed = index.remove(ids.empty() ? 0 : ids[0]);
    std::printf("removed %zu size %zu max_level %zu level0_nodes %zu dims %zu metric %c\n", removed.completed, index.size(), index.max_level(),
                index.stats(0).nodes, index.dimensions(), (char)index.metric().metric_kind());
    auto again = index.ef_search(q, 3, 64);
    std::vector<row_t> ids2(again.size());
    again.dump_to(ids2.data());
    bool still_there = false;
    for (auto id : ids2) still_there |= (!ids.empty() && id == ids[0]);
    std::printf("removed_key_returned %d\n", (int)still_there);
    auto compacted = index.compact();                             // hnsw_index.cpp:451-464
    std::printf("compact_ok %d\n", (int)(bool)compacted);

    // PersistToDisk / load, hnsw_index.cpp:510-532, 193-206
    std::vector<unsigned char> blob;
    index.save_to_stream([&](const void *data, size_t size) {
        blob.insert(blob.end(), (const unsigned char *)data, (const unsigned char *)data + size);
        return true;
    });
    USearchIndexType loaded = USearchIndexType::make(metric, config);
    size_t off = 0;
    auto lr = loaded.load_from_stream([&](void *data, size_t size) {
        if (off + size > blob.size()) return false;
        std::memcpy(data, blob.data() + off, size);
        off += size;
        return true;
    });
    auto r3 = loaded.ef_search(q, 3, 64);
    std::vector<row_t> ids3(r3.size());
    r3.dump_to(ids3.data());
    std::printf("stream_bytes %zu serialized_length %zu load_ok %d reloaded_size %zu same_results %d\n", blob.size(), index.serialized_length(),
                (int)(bool)lr, loaded.size(), (int)(ids3 == ids2));
    index.reset();                                                // hnsw_index.cpp:384
    std::printf("after_reset_size %zu\n", index.size());

    // INSERT while scans run: HNSWIndex::Construct adds with any_thread() under the SHARED side of rwlock
    // (hnsw_index.cpp:433-447), so adds and ef_search calls of other connections overlap. Nothing may be lost.
    index.reserve({2048, T});
    std::atomic<bool> stop{false};
    std::atomic<int> bad{0};
    std::thread scanner([&] {
        while (!stop.load()) {
            auto r = index.ef_search(q, 3, 64);
            if (!r) bad++;
        }
    });
    std::vector<std::thread> writers;
    for (size_t t = 0; t < 2; t++)
        writers.emplace_back([&, t] {
            for (size_t i = t; i < n; i += 2) {
                const auto result = index.add((row_t)(5000 + i), rows.data() + 3 * i, USearchIndexType::any_thread());
                if (!result) bad++;
            }
        });
    for (auto &w : writers) w.join();
    stop = true;
    scanner.join();
    auto r4 = index.ef_search(q, 1, 64);
    std::vector<row_t> ids4(r4.size());
    r4.dump_to(ids4.data());
    std::printf("concurrent_insert_scan failures %d size %zu nearest %lld\n", bad.load(), index.size(), ids4.empty() ? -1ll : (long long)ids4[0]);
    const auto wrong_thread = index.add(1, rows.data(), 99);      // thread id beyond reserve({n, T}): refused, not aliased
    std::printf("thread_out_of_range_refused %d\n", (int)!wrong_thread);
    return 0;
}
