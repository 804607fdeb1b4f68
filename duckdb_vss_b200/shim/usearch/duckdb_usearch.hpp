// Drop-in replacement for the vss extension's `usearch/duckdb_usearch.hpp` (reference
// src/include/usearch/duckdb_usearch.hpp:1-15): the same `unum::usearch` names the extension uses, implemented over
// the C ABI of libgvss (include/gvss.h) instead of the header-only usearch. Put this directory ahead of the vendored
// one on the include path and link libgvss.so — no line of src/hnsw/*.cpp or src/include/hnsw/*.hpp changes:
// `using USearchIndexType = unum::usearch::index_dense_gt<row_t>` (hnsw_index.hpp:28) now names the class below.
//
// Only the subset the extension calls is provided (complete list with call sites: SURVEY.md §8b):
//   metric_punned_t(dim, metric_kind_t, scalar_kind_t), index_dense_config_t, index_limits_t,
//   index_dense_gt<key>::make / reserve / add / ef_search / remove / compact / reset / size / capacity / max_level /
//   memory_usage / stats / dimensions / metric / expansion_search / save_to_stream / load_from_stream,
//   index_dense_t::any_thread(), nested types stats_t and search_result_t, result objects with `operator bool` and
//   `error.what()`.
// Differences from usearch, all deliberate: result objects own their data (usearch's alias a per-thread buffer valid
// only until the next call, index.hpp:2472) and an unhandled error never throws from a destructor (index.hpp:351-363).
#pragma once

#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "gvss.h"

namespace gvss_shim {
// The first CUDA call of a process creates the context — about a second on a B200, 44 % of a 1M-row CREATE INDEX when it
// happens inside it. One helper thread starts it when the extension is loaded (static initialisation), joined at exit.
// GVSS_WARMUP=0 switches it off. Without a GPU gvss_warmup fails at once and nothing else happens.
struct warmup_at_load_t {
    std::thread helper;
    warmup_at_load_t() {
        const char *e = std::getenv("GVSS_WARMUP");
        if (e && e[0] == '0') return;
        helper = std::thread([] { gvss_warmup(-1); });
    }
    ~warmup_at_load_t() {
        if (helper.joinable()) helper.join();
    }
};
// C++11 stand-in for an inline variable: one instance per program, initialised before main once make() below names it
template <class T = void> struct warmup_holder_t { static warmup_at_load_t instance; };
template <class T> warmup_at_load_t warmup_holder_t<T>::instance;
} // namespace gvss_shim

namespace unum {
namespace usearch {

// numeric values are part of the file format (index_plugins.hpp:103-142)
enum class metric_kind_t : std::uint8_t {
    unknown_k = 0, ip_k = 'i', cos_k = 'c', l2sq_k = 'e', pearson_k = 'p', haversine_k = 'h', divergence_k = 'd',
    jaccard_k = 'j', hamming_k = 'b', tanimoto_k = 't', sorensen_k = 's',
};
enum class scalar_kind_t : std::uint8_t {
    unknown_k = 0, b1x8_k = 1, u40_k = 2, uuid_k = 3, f64_k = 10, f32_k = 11, f16_k = 12, f8_k = 13,
    u64_k = 14, u32_k = 15, u16_k = 16, u8_k = 17, i64_k = 20, i32_k = 21, i16_k = 22, i8_k = 23,
};

// error_t stand-in: a message that is empty on success
struct error_t {
    std::string message;
    const char *what() const noexcept { return message.c_str(); }
    explicit operator bool() const noexcept { return !message.empty(); }
};

class metric_punned_t {
    std::size_t dimensions_ = 0;
    metric_kind_t metric_kind_ = metric_kind_t::l2sq_k;
    scalar_kind_t scalar_kind_ = scalar_kind_t::f32_k;

  public:
    metric_punned_t() = default;
    metric_punned_t(std::size_t dimensions, metric_kind_t metric_kind = metric_kind_t::l2sq_k,
                    scalar_kind_t scalar_kind = scalar_kind_t::f32_k) noexcept
        : dimensions_(dimensions), metric_kind_(metric_kind), scalar_kind_(scalar_kind) {}
    std::size_t dimensions() const noexcept { return dimensions_; }
    metric_kind_t metric_kind() const noexcept { return metric_kind_; }
    scalar_kind_t scalar_kind() const noexcept { return scalar_kind_; }
};

// index_config_t + index_dense_config_t (index.hpp:1097-1131, index_dense.hpp:98-110); 0 = usearch default
struct index_dense_config_t {
    std::size_t connectivity = 16;
    std::size_t connectivity_base = 32;
    std::size_t expansion_add = 128;
    std::size_t expansion_search = 64;
    bool exclude_vectors = false;
    bool multi = false;
    bool enable_key_lookups = true;
};

struct index_limits_t { // index.hpp:1134-1140
    std::size_t members = 0, threads_add = 0, threads_search = 0;
    index_limits_t(std::size_t n, std::size_t t) noexcept : members(n), threads_add(t), threads_search(t) {}
    index_limits_t(std::size_t n = 0) noexcept : index_limits_t(n, std::thread::hardware_concurrency()) {}
    std::size_t threads() const noexcept { return threads_add > threads_search ? threads_add : threads_search; }
};

template <typename key_at = std::int64_t> class index_dense_gt {
    static_assert(sizeof(key_at) == 8, "the vss extension instantiates index_dense_gt<row_t> with a 64-bit key");
    gvss_index *handle_ = nullptr;
    metric_punned_t metric_{};
    index_dense_config_t config_{};
    std::string create_error_; // why make() failed (the thread that calls add() later has no message of its own)

    static std::string last_error() { return std::string(gvss_last_error()); }

  public:
    using vector_key_t = key_at;

    struct stats_t { // index.hpp:2612-2635
        std::size_t nodes = 0, edges = 0, max_edges = 0, allocated_bytes = 0;
    };
    struct add_result_t {
        error_t error{};
        std::size_t new_size = 0, visited_members = 0, computed_distances = 0;
        explicit operator bool() const noexcept { return !error; }
    };
    struct labeling_result_t {
        error_t error{};
        std::size_t completed = 0;
        explicit operator bool() const noexcept { return !error; }
    };
    struct compaction_result_t {
        error_t error{};
        explicit operator bool() const noexcept { return !error; }
    };
    struct serialization_result_t {
        error_t error{};
        explicit operator bool() const noexcept { return !error; }
    };
    class search_result_t {
        friend class index_dense_gt;
        std::vector<vector_key_t> keys_;
        std::vector<float> distances_;

      public:
        error_t error{};
        std::size_t visited_members = 0, computed_distances = 0;
        explicit operator bool() const noexcept { return !error; }
        std::size_t size() const noexcept { return keys_.size(); }
        std::size_t dump_to(vector_key_t *keys) const noexcept {
            if (!keys_.empty()) std::memcpy(keys, keys_.data(), keys_.size() * sizeof(vector_key_t));
            return keys_.size();
        }
        std::size_t dump_to(vector_key_t *keys, float *distances) const noexcept {
            if (!keys_.empty()) std::memcpy(distances, distances_.data(), distances_.size() * sizeof(float));
            return dump_to(keys);
        }
    };

    index_dense_gt() = default;
    index_dense_gt(index_dense_gt const &) = delete;
    index_dense_gt &operator=(index_dense_gt const &) = delete;
    index_dense_gt(index_dense_gt &&o) noexcept { swap(o); }
    index_dense_gt &operator=(index_dense_gt &&o) noexcept {
        swap(o);
        return *this;
    }
    ~index_dense_gt() { gvss_destroy(handle_); }
    void swap(index_dense_gt &o) noexcept {
        std::swap(handle_, o.handle_);
        std::swap(metric_, o.metric_);
        std::swap(config_, o.config_);
        std::swap(create_error_, o.create_error_);
    }

    // index_dense_gt::make(metric, config) (hnsw_index.cpp:189). A failed creation leaves an index whose calls fail.
    static index_dense_gt make(metric_punned_t metric, index_dense_config_t config = {}) {
        (void)&gvss_shim::warmup_holder_t<>::instance; // names the start-up helper so that it exists
        index_dense_gt r;
        r.metric_ = metric;
        r.config_ = config;
        gvss_config c{};
        c.dimensions = (uint32_t)metric.dimensions();
        c.metric = (char)metric.metric_kind();
        c.connectivity = (uint32_t)config.connectivity;
        c.connectivity_base = (uint32_t)config.connectivity_base;
        c.expansion_add = (uint32_t)config.expansion_add;
        c.expansion_search = (uint32_t)config.expansion_search;
        c.device = -1;
        if (metric.scalar_kind() != scalar_kind_t::f32_k) { r.create_error_ = "only f32 vectors are supported"; return r; }
        // GVSS_DEVICES=0,1,2,3: partition every index over those GPUs (gvss_create_sharded). In-memory databases only: a
        // sharded index is not one usearch stream, so PersistToDisk fails with the library's message.
        std::vector<int32_t> devices;
        if (const char *e = std::getenv("GVSS_DEVICES"))
            for (const char *p = e; *p;) {
                char *end = nullptr;
                const long v = std::strtol(p, &end, 10);
                if (end == p) break;
                devices.push_back((int32_t)v);
                p = *end == ',' ? end + 1 : end;
            }
        const int rc = devices.size() > 1 ? gvss_create_sharded(&c, devices.data(), (uint32_t)devices.size(), &r.handle_) : gvss_create(&c, &r.handle_);
        if (rc) { r.create_error_ = last_error(); r.handle_ = nullptr; return r; }
        // pipelined bulk load: rows are uploaded and inserted in the background as soon as this many are staged,
        // while the construct tasks (or, with n3_streaming_sink.patch, the Sink phase) keep adding. GVSS_BACKGROUND_BUILD=0: off
        unsigned long long bg = 32768;
        if (const char *e = std::getenv("GVSS_BACKGROUND_BUILD")) bg = std::strtoull(e, nullptr, 10);
        if (bg) gvss_set_background_build(r.handle_, bg);
        return r;
    }
    explicit operator bool() const noexcept { return handle_ != nullptr; }
    static std::size_t any_thread() noexcept { return std::numeric_limits<std::size_t>::max(); }
    static vector_key_t free_key() noexcept { return (vector_key_t)GVSS_FREE_KEY; }

    bool reserve(index_limits_t limits) { return handle_ && !gvss_reserve(handle_, limits.members, (uint32_t)limits.threads()); }

    add_result_t add(vector_key_t key, const float *vector, std::size_t thread = any_thread(), bool = true) {
        add_result_t r;
        if (!handle_) { r.error.message = "index was not created: " + create_error_; return r; }
        if (gvss_add(handle_, (int64_t)key, vector, thread == any_thread() ? GVSS_ANY_THREAD : (uint32_t)thread)) r.error.message = last_error();
        return r;
    }

    // the only search entry the extension uses (index_dense.hpp:619-623)
    search_result_t ef_search(const float *query, std::size_t wanted, std::size_t ef, std::size_t = any_thread(), bool = false) const {
        search_result_t r;
        if (!handle_) { r.error.message = "index was not created: " + create_error_; return r; }
        if (!wanted) return r;
        r.keys_.resize(wanted);
        r.distances_.resize(wanted);
        uint32_t count = 0;
        gvss_search_stats st{};
        if (gvss_search(handle_, query, 1, (uint32_t)wanted, (uint32_t)ef, (int64_t *)r.keys_.data(), r.distances_.data(), &count, &st)) {
            r.error.message = last_error();
            count = 0;
        }
        r.keys_.resize(count);
        r.distances_.resize(count);
        r.visited_members = st.visited_members;
        r.computed_distances = st.computed_distances;
        return r;
    }
    // batched form used by the index join once HNSWIndex::ExecuteMultiScan is re-pointed at it
    bool ef_search_batch(const float *queries, std::size_t nq, std::size_t wanted, std::size_t ef, vector_key_t *keys, float *distances,
                         uint32_t *counts) const {
        return handle_ && !gvss_search(handle_, queries, nq, (uint32_t)wanted, (uint32_t)ef, (int64_t *)keys, distances, counts, nullptr);
    }

    // rows only, no graph: the transient row store behind the exact join of a table WITHOUT an HNSW index
    // (integration/n2b_exact_join_unindexed.patch). Only exact_search / exact_search_batch answer afterwards.
    bool rows_only(bool on = true) { return handle_ && !gvss_set_rows_only(handle_, on ? 1 : 0); }

    // exact (brute-force) top-k over the rows the index holds in HBM (integration/n2_exact_mode.patch)
    bool exact_search_batch(const float *queries, std::size_t nq, std::size_t wanted, vector_key_t *keys, float *distances, uint32_t *counts) const {
        return handle_ && !gvss_exact_search(handle_, queries, nq, (uint32_t)wanted, (int64_t *)keys, distances, counts);
    }
    search_result_t exact_search(const float *query, std::size_t wanted) const {
        search_result_t r;
        if (!handle_) { r.error.message = "index was not created: " + create_error_; return r; }
        if (!wanted) return r;
        r.keys_.resize(wanted);
        r.distances_.resize(wanted);
        uint32_t count = 0;
        if (gvss_exact_search(handle_, query, 1, (uint32_t)wanted, (int64_t *)r.keys_.data(), r.distances_.data(), &count)) {
            r.error.message = last_error();
            count = 0;
        }
        r.keys_.resize(count);
        r.distances_.resize(count);
        return r;
    }

    labeling_result_t remove(vector_key_t key) {
        labeling_result_t r;
        uint64_t n = 0;
        if (!handle_ || gvss_remove(handle_, (int64_t)key, &n)) r.error.message = last_error();
        r.completed = n;
        return r;
    }
    // PRAGMA hnsw_compact_index (hnsw_index.cpp:451-464). Default: usearch's compact() — nodes re-ordered, removed members kept.
    // GVSS_COMPACT=rebuild: the deleting compaction the extension's README.md:69 describes (gvss_compact_rebuild).
    compaction_result_t compact() {
        compaction_result_t r;
        const char *mode = std::getenv("GVSS_COMPACT");
        const bool rebuild = mode && std::strcmp(mode, "rebuild") == 0;
        if (!handle_ || (rebuild ? gvss_compact_rebuild(handle_, nullptr) : gvss_compact(handle_))) r.error.message = last_error();
        return r;
    }
    // deleting variant (not in usearch): rebuilds the graph over the live members and drops the removed ones
    compaction_result_t compact_rebuild() {
        compaction_result_t r;
        if (!handle_ || gvss_compact_rebuild(handle_, nullptr)) r.error.message = last_error();
        return r;
    }
    void reset() { if (handle_) gvss_reset(handle_); }

    gvss_info info_() const {
        gvss_info i{};
        if (handle_) gvss_get_info(handle_, &i);
        return i;
    }
    std::size_t size() const { return info_().size; }
    std::size_t capacity() const { return info_().capacity; }
    std::size_t max_level() const { return info_().max_level; }
    std::size_t memory_usage() const { return info_().memory_usage; }
    std::size_t dimensions() const { return metric_.dimensions(); }
    std::size_t connectivity() const { return info_().connectivity; }
    std::size_t expansion_add() const { return config_.expansion_add; }
    std::size_t expansion_search() const { return config_.expansion_search; }
    metric_punned_t const &metric() const { return metric_; }
    stats_t stats(std::size_t level) const {
        stats_t s;
        gvss_level_stats g{};
        if (handle_ && !gvss_get_level_stats(handle_, (uint32_t)level, &g)) {
            s.nodes = g.nodes; s.edges = g.edges; s.max_edges = g.max_edges; s.allocated_bytes = g.allocated_bytes;
        }
        return s;
    }
    std::size_t serialized_length() const {
        uint64_t n = 0;
        if (handle_) gvss_serialized_length(handle_, &n);
        return n;
    }

    // callbacks are `bool(const void*, size_t)` / `bool(void*, size_t)` like usearch's (hnsw_index.cpp:204-205, 526-529)
    template <typename output_callback_at> serialization_result_t save_to_stream(output_callback_at &&output) const {
        serialization_result_t r;
        auto tramp = [](void *user, const void *data, size_t size) -> int { return (*static_cast<typename std::remove_reference<output_callback_at>::type *>(user))(data, size) ? 1 : 0; };
        if (!handle_ || gvss_save(handle_, tramp, (void *)&output)) r.error.message = last_error();
        return r;
    }
    template <typename input_callback_at> serialization_result_t load_from_stream(input_callback_at &&input) {
        serialization_result_t r;
        auto tramp = [](void *user, void *data, size_t size) -> int { return (*static_cast<typename std::remove_reference<input_callback_at>::type *>(user))(data, size) ? 1 : 0; };
        if (!handle_ || gvss_load(handle_, tramp, (void *)&input)) r.error.message = last_error();
        else {
            gvss_info i = info_();
            metric_ = metric_punned_t(metric_.dimensions(), (metric_kind_t)i.metric, scalar_kind_t::f32_k);
        }
        return r;
    }
};

using index_dense_t = index_dense_gt<std::int64_t>;

} // namespace usearch
} // namespace unum
