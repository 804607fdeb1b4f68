// Multi-GPU index behind the C ABI (no counterpart in the reference, which is single-process CPU code):
// one process, one single-GPU index ("shard") per device. Rows are dealt out to the shards in chunks of 2048 in arrival
// order, keys stay global row ids, every shard builds its own sub-graph with no build-time communication. A search uploads
// the batch once to every GPU, all GPUs search at the same time, and the per-shard top-k lists are merged by ONE kernel on
// the first device that reads the other GPUs' lists in place over NVLink (peer access; ordering by CUDA events, no flags
// and no collective library). Without peer access the lists are copied to the first device and merged there.
// A sharded index cannot be persisted as one usearch stream (the reference cannot merge indexes either,
// hnsw_index.cpp:561-563): gvss_save / gvss_load refuse.
#include "router.h"

#include <algorithm>
#include <atomic>
#include <cstring>
#include <string>
#include <mutex>
#include <thread>
#include <vector>

namespace gvss {

constexpr uint64_t kDealChunk = 2048; // rows per deal = STANDARD_VECTOR_SIZE, the unit the extension hands over (hnsw_index_physical_create.cpp:144-205)

struct Router {
    gvss_config cfg{};
    std::vector<gvss_index *> shards;
    std::vector<int> devices;
    std::atomic<uint64_t> arrivals{0}; // rows handed to the router so far (decides the shard of the next row)
    std::atomic<uint64_t> capacity{0}, removed{0};
    std::mutex mu;                     // one sharded search at a time (pinned query block, merge buffers)
    cudaStream_t stream = nullptr;     // merge stream on devices[0]
    std::vector<cudaEvent_t> done;     // shard s's search has finished
    std::vector<char> direct;          // devices[0] can read shard s's memory in place
    float *h_q = nullptr; size_t h_q_bytes = 0;
    uint8_t *d_merge = nullptr; size_t d_merge_bytes = 0;   // merged keys | distances on devices[0]
    uint8_t *h_merge = nullptr; size_t h_merge_bytes = 0;
    uint8_t *d_gather = nullptr; size_t d_gather_bytes = 0; // [shards][nq][k] keys | distances, only without peer access
};

#define RCU(x)                                                                                                        \
    do {                                                                                                              \
        cudaError_t e_ = (x);                                                                                         \
        if (e_ != cudaSuccess) return set_error("CUDA error %s at %s:%d (%s)", cudaGetErrorString(e_), __FILE__, __LINE__, #x); \
    } while (0)

static uint32_t shard_of(const Router *r, uint64_t arrival) { return (uint32_t)((arrival / kDealChunk) % r->shards.size()); }

// runs fn(shard index) on every shard at the same time (each call blocks on its own GPU); first failure wins
template <typename F> static int on_all_shards(Router *r, F fn) {
    const size_t n = r->shards.size();
    std::vector<int> rc(n, 0);
    std::vector<std::string> msg(n);
    std::vector<std::thread> ts;
    for (size_t s = 0; s < n; s++)
        ts.emplace_back([&, s] {
            rc[s] = fn(s);
            if (rc[s]) msg[s] = gvss_last_error(); // the message is thread-local: carry it over
        });
    for (auto &t : ts) t.join();
    for (size_t s = 0; s < n; s++)
        if (rc[s]) return set_error("shard %zu (device %d): %s", s, r->devices[s], msg[s].c_str());
    return 0;
}

int router_create(const gvss_config *cfg, const int32_t *devices, uint32_t n_devices, Router **out) {
    if (!cfg || !devices || !out) return set_error("gvss_create_sharded: null argument");
    if (!n_devices || n_devices > 16) return set_error("gvss_create_sharded: 1..16 devices");
    auto *r = new Router();
    r->cfg = *cfg;
    for (uint32_t s = 0; s < n_devices; s++) {
        gvss_config c = *cfg;
        c.device = devices[s];
        gvss_index *ix = nullptr;
        if (gvss_create(&c, &ix)) { router_destroy(r); return 1; }
        r->shards.push_back(ix);
        r->devices.push_back(devices[s]);
    }
    // the merge runs on the first device and reads the other shards' lists where they are, if the hardware allows it
    r->direct.assign(n_devices, 0);
    r->done.assign(n_devices, nullptr);
    cudaSetDevice(r->devices[0]);
    if (cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking) != cudaSuccess) { router_destroy(r); return set_error("gvss_create_sharded: cannot create a stream"); }
    for (uint32_t s = 0; s < n_devices; s++) {
        if (r->devices[s] == r->devices[0]) r->direct[s] = 1;
        else {
            int ok = 0;
            if (cudaDeviceCanAccessPeer(&ok, r->devices[0], r->devices[s]) == cudaSuccess && ok) {
                const cudaError_t e = cudaDeviceEnablePeerAccess(r->devices[s], 0);
                if (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled) r->direct[s] = 1;
                cudaGetLastError();
            }
        }
    }
    for (uint32_t s = 0; s < n_devices; s++) {
        cudaSetDevice(r->devices[s]);
        if (cudaEventCreateWithFlags(&r->done[s], cudaEventDisableTiming) != cudaSuccess) { router_destroy(r); return set_error("gvss_create_sharded: cannot create an event"); }
    }
    *out = r;
    return 0;
}

void router_destroy(Router *r) {
    if (!r) return;
    for (size_t s = 0; s < r->shards.size(); s++) {
        if (s < r->done.size() && r->done[s]) { cudaSetDevice(r->devices[s]); cudaEventDestroy(r->done[s]); }
        gvss_destroy(r->shards[s]);
    }
    if (!r->devices.empty()) cudaSetDevice(r->devices[0]);
    cudaFree(r->d_merge); cudaFree(r->d_gather);
    if (r->h_q) cudaFreeHost(r->h_q);
    if (r->h_merge) cudaFreeHost(r->h_merge);
    if (r->stream) cudaStreamDestroy(r->stream);
    delete r;
}

int router_reset(Router *r) {
    for (auto *s : r->shards) if (gvss_reset(s)) return 1;
    r->arrivals = 0; r->capacity = 0; r->removed = 0;
    return 0;
}

int router_reserve(Router *r, uint64_t members, uint32_t host_threads) {
    // shard s receives the chunks s, s+G, s+2G, ...: at most ceil(chunks / G) of them, plus one chunk of slack for rows that
    // re-use capacity freed by gvss_remove on another shard
    const uint64_t G = r->shards.size(), chunks = (members + kDealChunk - 1) / kDealChunk;
    const uint64_t per = std::min<uint64_t>(members, ((chunks + G - 1) / G) * kDealChunk) + (G > 1 ? kDealChunk : 0);
    for (auto *s : r->shards) if (gvss_reserve(s, per, host_threads)) return 1;
    uint64_t cur = r->capacity.load();
    while (members > cur && !r->capacity.compare_exchange_weak(cur, members)) {}
    return 0;
}

int router_add(Router *r, int64_t key, const float *vec, uint32_t host_thread) {
    const uint64_t a = r->arrivals.fetch_add(1);
    if (a - r->removed.load() >= r->capacity.load()) { // index.hpp:2329-2335
        r->arrivals.fetch_sub(1);
        return set_error("Reserve capacity ahead of insertions!");
    }
    return gvss_add(r->shards[shard_of(r, a)], key, vec, host_thread);
}

int router_add_batch(Router *r, const int64_t *keys, const float *vecs, uint64_t n) {
    const uint32_t d = r->cfg.dimensions;
    std::lock_guard<std::mutex> lk(r->mu); // a batch is dealt out as one run of arrivals
    const uint64_t a0 = r->arrivals.load();
    if (a0 + n - r->removed.load() > r->capacity.load()) return set_error("Reserve capacity ahead of insertions!");
    std::vector<int64_t> own;
    for (uint64_t i = 0; i < n;) {
        const uint64_t a = a0 + i;
        const uint64_t len = std::min<uint64_t>(n - i, kDealChunk - a % kDealChunk);
        const int64_t *kp = keys ? keys + i : nullptr;
        if (!keys) { own.resize(len); for (uint64_t j = 0; j < len; j++) own[j] = (int64_t)(a + j); kp = own.data(); }
        if (gvss_add_batch(r->shards[shard_of(r, a)], kp, vecs + i * d, len)) return 1;
        i += len;
    }
    r->arrivals += n;
    return 0;
}

int router_flush(Router *r) { return on_all_shards(r, [&](size_t s) { return gvss_flush(r->shards[s]); }); }

int router_set_build_concurrency(Router *r, uint32_t n) {
    for (auto *s : r->shards) if (gvss_set_build_concurrency(s, n)) return 1;
    return 0;
}

int router_set_background_build(Router *r, uint64_t min_rows) {
    for (auto *s : r->shards) if (gvss_set_background_build(s, min_rows)) return 1;
    return 0;
}

int router_remove(Router *r, int64_t key, uint64_t *removed) {
    uint64_t total = 0;
    for (auto *s : r->shards) { // keys are global: whichever shard holds the row removes it
        uint64_t n = 0;
        if (gvss_remove(s, key, &n)) return 1;
        total += n;
    }
    r->removed += total;
    if (removed) *removed = total;
    return 0;
}

int router_compact(Router *r, bool rebuild, uint64_t *dropped) {
    std::vector<uint64_t> d(r->shards.size(), 0);
    const int rc = on_all_shards(r, [&](size_t s) { return rebuild ? gvss_compact_rebuild(r->shards[s], &d[s]) : gvss_compact(r->shards[s]); });
    if (dropped) { *dropped = 0; for (auto v : d) *dropped += v; }
    return rc;
}

static int grow_pinned(void **p, size_t &have, size_t want) {
    if (want <= have) return 0;
    if (*p) cudaFreeHost(*p);
    *p = nullptr; have = 0;
    RCU(cudaMallocHost(p, want));
    have = want;
    return 0;
}
static int grow_device(uint8_t *&p, size_t &have, size_t want) {
    if (want <= have) return 0;
    cudaFree(p); p = nullptr; have = 0;
    RCU(cudaMalloc(&p, want));
    have = want;
    return 0;
}

int router_search(Router *r, const float *queries, uint64_t nq, uint32_t k, uint32_t ef, int64_t *keys_out, float *dist_out,
                  uint32_t *counts_out, gvss_search_stats *stats_out) {
    if (!nq) return 0;
    if (!k) { if (counts_out) memset(counts_out, 0, nq * 4); return 0; }
    const uint32_t G = (uint32_t)r->shards.size(), d = r->cfg.dimensions;
    if ((uint64_t)G * k > 256 && G > 1) return set_error("gvss_search on a sharded index: shards * k must not exceed 256 (%u x %u)", G, k);
    std::lock_guard<std::mutex> lk(r->mu);
    const size_t kb = nq * k * 8ull, db = nq * k * 4ull;
    cudaSetDevice(r->devices[0]);
    if (grow_pinned((void **)&r->h_q, r->h_q_bytes, nq * d * 4ull) || grow_pinned((void **)&r->h_merge, r->h_merge_bytes, kb + db) ||
        grow_device(r->d_merge, r->d_merge_bytes, kb + db)) return 1;
    memcpy(r->h_q, queries, nq * d * 4ull); // one pinned copy of the batch feeds every GPU's upload
    // 1. every shard: upload + launch (asynchronous; the GPUs now search at the same time)
    std::vector<HostSearch> hs(G);
    struct Rel { std::vector<HostSearch> &h; ~Rel() { for (auto &x : h) host_search_release(x); } } rel{hs};
    for (uint32_t s = 0; s < G; s++) {
        if (host_search_launch(r->shards[s], r->h_q, nq, k, ef, stats_out != nullptr, hs[s])) return 1;
        RCU(cudaEventRecord(r->done[s], hs[s].stream));
    }
    // 2. merge on the first device, behind the shards' events
    cudaSetDevice(r->devices[0]);
    for (uint32_t s = 0; s < G; s++) RCU(cudaStreamWaitEvent(r->stream, r->done[s], 0));
    int64_t *mk = reinterpret_cast<int64_t *>(r->d_merge);
    float *md = reinterpret_cast<float *>(r->d_merge + kb);
    const bool all_direct = std::all_of(r->direct.begin(), r->direct.end(), [](char c) { return c != 0; });
    if (all_direct) {
        const int64_t *pk[16];
        const float *pd[16];
        for (uint32_t s = 0; s < G; s++) { pk[s] = hs[s].d_keys; pd[s] = hs[s].d_dist; }
        RCU(launch_merge_topk_peers(pk, pd, nullptr, 0, G, 0, nq, k, mk, md, r->stream)); // step 0: ordered by the events above
    } else {
        if (grow_device(r->d_gather, r->d_gather_bytes, (size_t)G * (kb + db))) return 1;
        int64_t *gk = reinterpret_cast<int64_t *>(r->d_gather);
        float *gd = reinterpret_cast<float *>(r->d_gather + (size_t)G * kb);
        for (uint32_t s = 0; s < G; s++) {
            RCU(cudaMemcpyPeerAsync(gk + (size_t)s * nq * k, r->devices[0], hs[s].d_keys, r->devices[s], kb, r->stream));
            RCU(cudaMemcpyPeerAsync(gd + (size_t)s * nq * k, r->devices[0], hs[s].d_dist, r->devices[s], db, r->stream));
        }
        RCU(launch_merge_topk(gk, gd, G, nq, k, mk, md, r->stream));
    }
    RCU(cudaMemcpyAsync(r->h_merge, r->d_merge, kb + db, cudaMemcpyDeviceToHost, r->stream));
    // 3. per-shard work words (overflow report) and, when asked for, the traversal counters
    unsigned long long failed = 0;
    for (uint32_t s = 0; s < G; s++) {
        cudaSetDevice(r->devices[s]);
        RCU(cudaMemcpyAsync(hs[s].h_res, hs[s].res, 16, cudaMemcpyDeviceToHost, hs[s].stream));
        if (stats_out) RCU(cudaMemcpyAsync(hs[s].h_res + hs[s].o_stats, hs[s].res + hs[s].o_stats, nq * sizeof(gvss_search_stats), cudaMemcpyDeviceToHost, hs[s].stream));
    }
    for (uint32_t s = 0; s < G; s++) {
        cudaSetDevice(r->devices[s]);
        RCU(cudaStreamSynchronize(hs[s].stream));
        failed = std::max(failed, reinterpret_cast<const unsigned long long *>(hs[s].h_res)[1]);
    }
    cudaSetDevice(r->devices[0]);
    RCU(cudaStreamSynchronize(r->stream));
    memcpy(keys_out, r->h_merge, kb);
    if (dist_out) memcpy(dist_out, r->h_merge + kb, db);
    if (counts_out)
        for (uint64_t i = 0; i < nq; i++) {
            uint32_t c = 0;
            while (c < k && keys_out[i * k + c] >= 0) c++;
            counts_out[i] = c;
        }
    if (stats_out) { // counters add up over the shards, flags are OR-ed
        memset(stats_out, 0, nq * sizeof(gvss_search_stats));
        for (uint32_t s = 0; s < G; s++) {
            const auto *st = reinterpret_cast<const gvss_search_stats *>(hs[s].h_res + hs[s].o_stats);
            for (uint64_t i = 0; i < nq; i++) {
                stats_out[i].computed_distances += st[i].computed_distances; stats_out[i].visited_members += st[i].visited_members;
                stats_out[i].base_expansions += st[i].base_expansions; stats_out[i].flags |= st[i].flags;
            }
        }
    }
    if (failed) return set_error("gvss_search: candidate queue overflow on query %llu", failed - 1);
    return 0;
}

int router_info(Router *r, gvss_info *out) {
    memset(out, 0, sizeof *out);
    for (auto *s : r->shards) {
        gvss_info i{};
        if (gvss_get_info(s, &i)) return 1;
        out->size += i.size; out->node_count += i.node_count; out->memory_usage += i.memory_usage;
        out->max_level = std::max(out->max_level, i.max_level);
        out->concurrent_searches_peak = std::max(out->concurrent_searches_peak, i.concurrent_searches_peak);
        out->dimensions = i.dimensions; out->connectivity = i.connectivity; out->connectivity_base = i.connectivity_base;
        out->expansion_add = i.expansion_add; out->expansion_search = i.expansion_search; out->metric = i.metric;
    }
    out->capacity = r->capacity.load();
    return 0;
}

int router_build_stats(Router *r, gvss_build_stats *out) {
    memset(out, 0, sizeof *out);
    for (auto *s : r->shards) {
        gvss_build_stats b{};
        if (gvss_get_build_stats(s, &b)) return 1;
        out->inserts += b.inserts; out->query_distances += b.query_distances; out->pair_distances += b.pair_distances;
        out->base_expansions += b.base_expansions; out->upper_expansions += b.upper_expansions; out->rows_written += b.rows_written;
    }
    return 0;
}

int router_level_stats(Router *r, uint32_t level, gvss_level_stats *out) {
    memset(out, 0, sizeof *out);
    for (auto *s : r->shards) {
        gvss_level_stats l{};
        if (gvss_get_level_stats(s, level, &l)) return 1;
        out->nodes += l.nodes; out->edges += l.edges; out->max_edges += l.max_edges; out->allocated_bytes += l.allocated_bytes;
    }
    return 0;
}

} // namespace gvss
