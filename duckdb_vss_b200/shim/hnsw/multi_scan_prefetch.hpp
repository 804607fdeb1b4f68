// SURVEY.md §8(f) N1 — batched multi-scan for HNSW_INDEX_JOIN.
//
// The reference's index join asks the index one outer row at a time: PhysicalHNSWIndexJoin::Execute
// (src/hnsw/hnsw_optimize_join.cpp:113-170) loops over the rows of the flattened outer chunk and calls
// HNSWIndex::ExecuteMultiScan (src/hnsw/hnsw_index.cpp:353-367) = one index.ef_search per row. On a GPU that is one
// kernel launch plus two PCIe round trips per row. The outer chunk's ARRAY child vector is one contiguous
// [rows x d] fp32 block (`rhs_vector_ptr`, hnsw_optimize_join.cpp:126-129), i.e. exactly gvss_search's query layout,
// so the whole chunk (<= STANDARD_VECTOR_SIZE = 2048 rows) can be searched in ONE call when the chunk arrives and
// every ExecuteMultiScan served from the result.
//
// multi_scan_prefetch_t is that cache. It lives in MultiScanState; the operator announces a chunk once
// (prefetch), ExecuteMultiScan asks it first (lookup). A query pointer outside the announced block — or a different
// limit / ef — is not served (lookup returns false) and the caller falls back to the single-query path, so behaviour
// is unchanged wherever the prefetch does not apply. See INTEGRATION.md for the 3 edits in the extension.
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace gvss_shim {

template <typename index_at> class multi_scan_prefetch_t {
  public:
    using key_t = typename index_at::vector_key_t;

    // Search rows [0, count) of the contiguous block `queries` (row stride = dims floats) with one batched call.
    // Returns false (and caches nothing) if the index refuses, e.g. count == 0.
    // exact = true: brute-force top-k over the index's rows instead of the graph search
    bool prefetch(const index_at &index, const float *queries, std::size_t count, std::size_t dims, std::size_t limit,
                  std::size_t ef_search, bool exact = false) {
        reset();
        if (!queries || !count || !limit) return false;
        keys_.resize(count * limit);
        counts_.resize(count);
        const bool ok = exact ? index.exact_search_batch(queries, count, limit, keys_.data(), nullptr, counts_.data())
                              : index.ef_search_batch(queries, count, limit, ef_search, keys_.data(), nullptr, counts_.data());
        if (!ok) {
            reset();
            return false;
        }
        base_ = queries; count_ = count; dims_ = dims; limit_ = limit; ef_ = ef_search;
        return true;
    }

    // Result of the row that starts at `query_vector`, if that row belongs to the announced block and was searched with
    // the same limit and ef: ids[0..n) ascending by distance, exactly what ef_search(query_vector, limit, ef) returns.
    bool lookup(const float *query_vector, std::size_t limit, std::size_t ef_search, const key_t *&ids, std::size_t &n) const {
        if (!base_ || limit != limit_ || ef_search != ef_ || query_vector < base_) return false;
        const std::size_t off = (std::size_t)(query_vector - base_);
        if (off % dims_ != 0) return false;
        const std::size_t row = off / dims_;
        if (row >= count_) return false;
        ids = keys_.data() + row * limit_;
        n = counts_[row];
        return true;
    }

    void reset() { base_ = nullptr; count_ = dims_ = limit_ = ef_ = 0; }
    std::size_t rows() const { return count_; }

  private:
    const float *base_ = nullptr;
    std::size_t count_ = 0, dims_ = 0, limit_ = 0, ef_ = 0;
    std::vector<key_t> keys_;
    std::vector<uint32_t> counts_;
};

} // namespace gvss_shim
