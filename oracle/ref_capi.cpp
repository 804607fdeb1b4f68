// TEST INFRASTRUCTURE ONLY — never linked into, imported by or called from the product path.
//
// extern "C" driver over the UNMODIFIED reference: the vendored usearch v2.12 headers under
// /root/reference/src/include/usearch (configured exactly like src/include/usearch/duckdb_usearch.hpp:6-8:
// SIMSIMD off, FP16LIB on, OPENMP off) and DuckDB's array_kernels.hpp. Nothing is copied: the
// headers are included from where they lie (-I/root/reference/...) by oracle/Makefile and the
// output goes to oracle/_ref/libref_usearch.so (git-ignored, travels to the GPU box).
//
// What it is used for: (1) pinning oracle/hnsw_oracle.c, (2) producing reference-built graphs
// (usearch streams) and reference search results for the GPU parity tests, (3) the timed CPU
// baseline of bench.py (`--impl reference`, `cpu_baseline.kind = "reference"`).
//
// The index type and config mirror src/hnsw/hnsw_index.cpp:162-189 (index_dense_gt<row_t>,
// enable_key_lookups=false); the bulk-add thread pattern mirrors
// src/hnsw/hnsw_index_physical_create.cpp:144-205 (T tasks, shared cursor, add(key, vec, thread)).
#define DUCKDB_USEARCH_USE_SIMSIMD 0
#include "usearch/duckdb_usearch.hpp"
#include "core_functions/array_kernels.hpp"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

using namespace unum::usearch;
using ref_index_t = index_dense_gt<int64_t>;

struct ref_handle {
    ref_index_t index;
    uint32_t threads = 1;
};

static metric_kind_t to_kind(char m) {
    switch (m) {
    case 'c': return metric_kind_t::cos_k;
    case 'i': return metric_kind_t::ip_k;
    default: return metric_kind_t::l2sq_k;
    }
}

extern "C" {

void *ref_create(uint32_t dims, char metric, uint32_t M, uint32_t M0, uint32_t ef_add, uint32_t ef_search) {
    metric_punned_t m(dims, to_kind(metric), scalar_kind_t::f32_k);
    index_dense_config_t config = {};
    config.enable_key_lookups = false;
    if (ef_add) config.expansion_add = ef_add;
    if (ef_search) config.expansion_search = ef_search;
    if (M) { config.connectivity = M; config.connectivity_base = M * 2; }
    if (M0) config.connectivity_base = M0;
    auto *h = new ref_handle();
    h->index = ref_index_t::make(m, config);
    return h;
}

void ref_destroy(void *p) { delete static_cast<ref_handle *>(p); }

int ref_reserve(void *p, uint64_t members, uint32_t threads) {
    auto *h = static_cast<ref_handle *>(p);
    h->threads = threads ? threads : 1;
    return h->index.reserve(index_limits_t(members, h->threads)) ? 0 : 1;
}

// One add; returns 0 on success. Counters are per-call (add_result_t, index.hpp:2133-2139).
int ref_add(void *p, int64_t key, const float *vec, uint32_t thread, uint64_t *computed, uint64_t *visited) {
    auto *h = static_cast<ref_handle *>(p);
    auto r = h->index.add(key, vec, thread);
    if (!r) { r.error.release(); return 1; }
    if (computed) *computed = r.computed_distances;
    if (visited) *visited = r.visited_members;
    return 0;
}

// Bulk add with the extension's pattern: `threads` workers pull 2048-row chunks from a shared cursor.
// Returns the number of failed adds; seconds_out = wall time of the add phase only.
uint64_t ref_add_bulk(void *p, const int64_t *keys, const float *vecs, uint64_t n, uint32_t threads,
                      uint64_t *computed_sum, uint64_t *visited_sum, double *seconds_out) {
    auto *h = static_cast<ref_handle *>(p);
    uint32_t d = (uint32_t)h->index.dimensions();
    std::atomic<uint64_t> cursor{0}, failed{0}, comp{0}, vis{0};
    auto t0 = std::chrono::steady_clock::now();
    auto work = [&](uint32_t tid) {
        uint64_t c = 0, v = 0;
        for (;;) {
            uint64_t b = cursor.fetch_add(2048);
            if (b >= n) break;
            uint64_t e = b + 2048 < n ? b + 2048 : n;
            for (uint64_t i = b; i < e; i++) {
                auto r = h->index.add(keys ? keys[i] : (int64_t)i, vecs + i * d, tid);
                if (!r) { r.error.release(); failed++; continue; }
                c += r.computed_distances; v += r.visited_members;
            }
        }
        comp += c; vis += v;
    };
    if (threads <= 1) work(0);
    else {
        std::vector<std::thread> ts;
        for (uint32_t t = 0; t < threads; t++) ts.emplace_back(work, t);
        for (auto &t : ts) t.join();
    }
    auto t1 = std::chrono::steady_clock::now();
    if (seconds_out) *seconds_out = std::chrono::duration<double>(t1 - t0).count();
    if (computed_sum) *computed_sum = comp.load();
    if (visited_sum) *visited_sum = vis.load();
    return failed.load();
}

// One query through the only search entry the extension uses (index_dense.hpp:619-623).
uint32_t ref_search(void *p, const float *q, uint32_t k, uint32_t ef, int64_t *keys, float *dists,
                    uint64_t *computed, uint64_t *visited) {
    auto *h = static_cast<ref_handle *>(p);
    auto r = h->index.ef_search(q, k, ef, 0);
    if (!r) { r.error.release(); return 0; }
    uint32_t c = (uint32_t)r.size();
    if (dists) r.dump_to(keys, dists); else r.dump_to(keys);
    if (computed) *computed = r.computed_distances;
    if (visited) *visited = r.visited_members;
    return c;
}

// Batched search, static split over `threads` host threads (thread t owns usearch thread id t).
// keys_out/dist_out are nq×k (unused tail: key -1 / dist +inf). Returns wall seconds.
double ref_search_batch(void *p, const float *qs, uint64_t nq, uint32_t k, uint32_t ef, uint32_t threads,
                        int64_t *keys_out, float *dist_out, uint32_t *counts_out, uint64_t *computed_out,
                        uint64_t *visited_out) {
    auto *h = static_cast<ref_handle *>(p);
    uint32_t d = (uint32_t)h->index.dimensions();
    if (!threads) threads = 1;
    auto t0 = std::chrono::steady_clock::now();
    auto work = [&](uint32_t tid) {
        uint64_t b = nq * tid / threads, e = nq * (tid + 1) / threads;
        std::vector<float> dtmp(k);
        for (uint64_t i = b; i < e; i++) {
            auto r = h->index.ef_search(qs + i * d, k, ef, tid);
            uint32_t c = 0;
            if (!r) r.error.release();
            else { c = (uint32_t)r.size(); r.dump_to(keys_out + i * k, dist_out ? dist_out + i * k : dtmp.data()); }
            for (uint32_t j = c; j < k; j++) { keys_out[i * k + j] = -1; if (dist_out) dist_out[i * k + j] = INFINITY; }
            if (counts_out) counts_out[i] = c;
            if (computed_out) computed_out[i] = r ? r.computed_distances : 0;
            if (visited_out) visited_out[i] = r ? r.visited_members : 0;
        }
    };
    if (threads <= 1) work(0);
    else {
        std::vector<std::thread> ts;
        for (uint32_t t = 0; t < threads; t++) ts.emplace_back(work, t);
        for (auto &t : ts) t.join();
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

uint64_t ref_remove(void *p, int64_t key) {
    auto *h = static_cast<ref_handle *>(p);
    auto r = h->index.remove(key);
    if (!r) { r.error.release(); return 0; }
    return r.completed;
}

uint64_t ref_size(void *p) { return static_cast<ref_handle *>(p)->index.size(); }
uint64_t ref_capacity(void *p) { return static_cast<ref_handle *>(p)->index.capacity(); }
uint64_t ref_max_level(void *p) { return static_cast<ref_handle *>(p)->index.max_level(); }
uint64_t ref_memory_usage(void *p) { return static_cast<ref_handle *>(p)->index.memory_usage(); }
void ref_level_stats(void *p, uint32_t level, uint64_t *out4) {
    auto s = static_cast<ref_handle *>(p)->index.stats(level);
    out4[0] = s.nodes; out4[1] = s.edges; out4[2] = s.max_edges; out4[3] = s.allocated_bytes;
}

uint64_t ref_serialized_length(void *p) { return static_cast<ref_handle *>(p)->index.serialized_length(); }

// save_to_stream (index_dense.hpp:811-878) into a caller buffer; returns bytes written or 0 on failure.
uint64_t ref_save(void *p, uint8_t *buf, uint64_t cap) {
    auto *h = static_cast<ref_handle *>(p);
    uint64_t off = 0; bool ok = true;
    auto r = h->index.save_to_stream([&](const void *d, size_t n) {
        if (off + n > cap) { ok = false; return false; }
        std::memcpy(buf + off, d, n); off += n; return true;
    });
    if (!r) { r.error.release(); return 0; }
    return ok ? off : 0;
}

// load_from_stream (index_dense.hpp:900-973); returns 0 on success.
int ref_load(void *p, const uint8_t *buf, uint64_t n) {
    auto *h = static_cast<ref_handle *>(p);
    uint64_t off = 0;
    auto r = h->index.load_from_stream([&](void *d, size_t m) {
        if (off + m > n) return false;
        std::memcpy(d, buf + off, m); off += m; return true;
    });
    if (!r) { r.error.release(); return 1; }
    return 0;
}

// The three metrics exactly as the index calls them (metric_punned_t::operator()).
float ref_metric(char metric, const float *a, const float *b, uint32_t d) {
    metric_punned_t m(d, to_kind(metric), scalar_kind_t::f32_k);
    return m(reinterpret_cast<const byte_t *>(a), reinterpret_cast<const byte_t *>(b));
}

// DuckDB's exact distance folds (array_kernels.hpp:11-105): 'e' array_distance (sqrt!), 'c' array_cosine_distance,
// 'i' array_negative_inner_product.
float ref_duckdb_distance(char metric, const float *a, const float *b, uint32_t d) {
    switch (metric) {
    case 'c': return duckdb::CosineDistanceOp::Operation<float>(a, b, d);
    case 'i': return duckdb::NegativeInnerProductOp::Operation<float>(a, b, d);
    default: return duckdb::DistanceOp::Operation<float>(a, b, d);
    }
}

// Brute-force top-k with DuckDB's folds, ordered by (distance, rowid) — the SQL
// `ORDER BY array_distance(v, q), rowid LIMIT k` comparator of SURVEY §8(c). Multi-threaded over queries.
void ref_duckdb_topk(char metric, const float *base, uint64_t n, const float *qs, uint64_t nq, uint32_t d, uint32_t k,
                     uint32_t threads, int64_t *ids_out, float *dist_out) {
    if (!threads) threads = 1;
    auto work = [&](uint32_t tid) {
        std::vector<std::pair<float, int64_t>> top;
        for (uint64_t qi = nq * tid / threads; qi < nq * (tid + 1) / threads; qi++) {
            top.clear();
            const float *q = qs + qi * d;
            for (uint64_t i = 0; i < n; i++) {
                float dist = ref_duckdb_distance(metric, base + i * d, q, d);
                std::pair<float, int64_t> e(dist, (int64_t)i);
                if (top.size() < k) { top.push_back(e); std::push_heap(top.begin(), top.end()); }
                else if (e < top.front()) { std::pop_heap(top.begin(), top.end()); top.back() = e; std::push_heap(top.begin(), top.end()); }
            }
            std::sort_heap(top.begin(), top.end());
            for (uint32_t j = 0; j < k; j++) {
                ids_out[qi * k + j] = j < top.size() ? top[j].second : -1;
                dist_out[qi * k + j] = j < top.size() ? top[j].first : INFINITY;
            }
        }
    };
    if (threads <= 1) work(0);
    else {
        std::vector<std::thread> ts;
        for (uint32_t t = 0; t < threads; t++) ts.emplace_back(work, t);
        for (auto &t : ts) t.join();
    }
}

// index.compact() as HNSWIndex::Compact calls it (hnsw_index.cpp:451-464 -> index_dense.hpp:1479-1496 -> index.hpp:3007-3096)
int ref_compact(void *p) {
    auto *h = static_cast<ref_handle *>(p);
    auto r = h->index.compact();
    if (!r) { r.error.release(); return 1; }
    return 0;
}

// Ground truth for recall at full size (10k queries x 1M rows is minutes of work for the sequential folds above): candidates
// are ranked with SSE-width partial sums (4 queries per pass over a row, any summation order), the best k + 16 of each query
// are then re-scored with DuckDB's exact fold and ordered by (distance, rowid) — the same final comparator as
// ref_duckdb_topk. Not a product path and not a timed baseline: it only tells the reference arm of bench.py its recall.
void ref_fast_topk(char metric, const float *base, uint64_t n, const float *qs, uint64_t nq, uint32_t d, uint32_t k,
                   uint32_t threads, int64_t *ids_out, float *dist_out) {
    typedef float v4 __attribute__((vector_size(16)));
    if (!threads) threads = 1;
    const uint32_t kc = k + 16, d4 = d / 4;
    std::atomic<uint64_t> cursor{0};
    auto work = [&]() {
        std::vector<std::pair<float, int64_t>> heap[4], fin;
        for (;;) {
            const uint64_t q0 = cursor.fetch_add(4);
            if (q0 >= nq) break;
            const uint32_t nqq = (uint32_t)std::min<uint64_t>(4, nq - q0);
            for (auto &hp : heap) hp.clear();
            for (uint64_t i = 0; i < n; i++) {
                const float *b = base + i * d;
                float sc[4];
                for (uint32_t j = 0; j < nqq; j++) {
                    const float *q = qs + (q0 + j) * d;
                    v4 ab = {0, 0, 0, 0}, a2 = ab, b2 = ab;
                    for (uint32_t t = 0; t < d4; t++) {
                        v4 x, y;
                        std::memcpy(&x, q + 4 * t, 16); std::memcpy(&y, b + 4 * t, 16);
                        if (metric == 'e') { const v4 df = x - y; ab += df * df; }
                        else { ab += x * y; if (metric == 'c') { a2 += x * x; b2 += y * y; } }
                    }
                    float sab = ab[0] + ab[1] + ab[2] + ab[3], sa2 = a2[0] + a2[1] + a2[2] + a2[3], sb2 = b2[0] + b2[1] + b2[2] + b2[3];
                    for (uint32_t t = d4 * 4; t < d; t++) {
                        if (metric == 'e') sab += (q[t] - b[t]) * (q[t] - b[t]);
                        else { sab += q[t] * b[t]; sa2 += q[t] * q[t]; sb2 += b[t] * b[t]; }
                    }
                    sc[j] = metric == 'e' ? sab : metric == 'i' ? -sab : 1.0f - sab / std::sqrt(sa2 * sb2);
                }
                for (uint32_t j = 0; j < nqq; j++) {
                    auto &hp = heap[j];
                    std::pair<float, int64_t> e(sc[j], (int64_t)i);
                    if (hp.size() < kc) { hp.push_back(e); std::push_heap(hp.begin(), hp.end()); }
                    else if (e < hp.front()) { std::pop_heap(hp.begin(), hp.end()); hp.back() = e; std::push_heap(hp.begin(), hp.end()); }
                }
            }
            for (uint32_t j = 0; j < nqq; j++) {
                fin.clear();
                for (auto &e : heap[j]) fin.emplace_back(ref_duckdb_distance(metric, base + (uint64_t)e.second * d, qs + (q0 + j) * d, d), e.second);
                std::sort(fin.begin(), fin.end());
                for (uint32_t t = 0; t < k; t++) {
                    ids_out[(q0 + j) * k + t] = t < fin.size() ? fin[t].second : -1;
                    if (dist_out) dist_out[(q0 + j) * k + t] = t < fin.size() ? fin[t].first : INFINITY;
                }
            }
        }
    };
    std::vector<std::thread> ts;
    for (uint32_t t = 1; t < threads; t++) ts.emplace_back(work);
    work();
    for (auto &t : ts) t.join();
}

uint32_t ref_hardware_threads(void) { return std::thread::hardware_concurrency(); }

} // extern "C"
