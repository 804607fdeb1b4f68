// the index join's per-row ExecuteMultiScan loop (hnsw_optimize_join.cpp:139-151, hnsw_index.cpp:353-367)
// against the prefetching form (duckdb_vss_b200/shim/hnsw/multi_scan_prefetch.hpp): same ids for every outer row, one
// batched search per outer chunk instead of one search per row.
#include "usearch/duckdb_usearch.hpp"
#include "hnsw/multi_scan_prefetch.hpp"
#include <chrono>
#include <cstdio>
#include <random>
#include <vector>

using row_t = int64_t;
using USearchIndexType = unum::usearch::index_dense_gt<row_t>; // hnsw_index.hpp:28

int main() {
    const size_t d = 64, n = 50000, outer = 2048, limit = 10, ef = 64; // one STANDARD_VECTOR_SIZE outer chunk
    unum::usearch::metric_punned_t metric(d, unum::usearch::metric_kind_t::l2sq_k, unum::usearch::scalar_kind_t::f32_k);
    unum::usearch::index_dense_config_t config = {};
    USearchIndexType index = USearchIndexType::make(metric, config);
    if (!index) { std::printf("create failed: %s\n", gvss_last_error()); return 2; }
    index.reserve(n);
    std::mt19937 rng(5);
    std::normal_distribution<float> g(0.f, 1.f);
    std::vector<float> centres(256 * d), rows(n * d), queries(outer * d);
    for (auto &x : centres) x = g(rng);
    auto draw = [&](std::vector<float> &out, size_t m) {
        for (size_t i = 0; i < m; i++) {
            const size_t c = rng() % 256;
            for (size_t j = 0; j < d; j++) out[i * d + j] = centres[c * d + j] + 0.3f * g(rng);
        }
    };
    draw(rows, n);
    draw(queries, outer);
    for (size_t i = 0; i < n; i++)
        if (!index.add((row_t)i, rows.data() + i * d, 0)) { std::printf("add failed\n"); return 3; }

    // the reference's loop: one ExecuteMultiScan (= one ef_search) per outer row, results appended to state.row_ids
    std::vector<row_t> per_row;
    std::vector<size_t> per_row_count(outer);
    index.ef_search(queries.data(), limit, ef); // first search flushes the staged rows (graph build) — not part of the comparison
    auto t0 = std::chrono::steady_clock::now();
    for (size_t i = 0; i < outer; i++) {
        auto r = index.ef_search(queries.data() + i * d, limit, ef);
        const size_t off = per_row.size();
        per_row.resize(off + r.size());
        r.dump_to(per_row.data() + off);
        per_row_count[i] = r.size();
    }
    const double t_rows = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    // N1: the operator announces the chunk once, ExecuteMultiScan is served from the cache
    gvss_shim::multi_scan_prefetch_t<USearchIndexType> cache;
    cache.prefetch(index, queries.data(), outer, d, limit, ef); // like the per-row path above, the batched path is timed warm (its
                                                                // first call loads the throughput kernel and sizes the scratch)
    std::vector<row_t> batched;
    size_t served = 0, mismatched_counts = 0;
    t0 = std::chrono::steady_clock::now();
    const bool ok = cache.prefetch(index, queries.data(), outer, d, limit, ef);
    for (size_t i = 0; i < outer; i++) {
        const row_t *ids = nullptr;
        size_t cnt = 0;
        if (cache.lookup(queries.data() + i * d, limit, ef, ids, cnt)) {
            served++;
            batched.insert(batched.end(), ids, ids + cnt);
            mismatched_counts += cnt != per_row_count[i];
        }
    }
    const double t_batch = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    const row_t *ids = nullptr;
    size_t cnt = 0;
    const bool foreign = cache.lookup(rows.data(), limit, ef, ids, cnt);          // a pointer outside the announced block
    const bool other_limit = cache.lookup(queries.data(), limit + 1, ef, ids, cnt); // other LIMIT: not served
    cache.reset();
    const bool after_reset = cache.lookup(queries.data(), limit, ef, ids, cnt);
    std::printf("prefetch_ok %d served %zu/%zu same_ids %d mismatched_counts %zu foreign %d other_limit %d after_reset %d\n", (int)ok, served,
                outer, (int)(batched == per_row), mismatched_counts, (int)foreign, (int)other_limit, (int)after_reset);
    std::printf("per_row_s %.4f batched_s %.4f speedup %.1f\n", t_rows, t_batch, t_rows / t_batch);
    return 0;
}
