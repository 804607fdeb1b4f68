// Internal interface between the single-GPU index (gvss_api.cu) and the multi-GPU router (router.cu). Not part of the ABI.
#pragma once
#include "kernels.h"
#include <shared_mutex>

struct gvss_index;
struct SearchSlot;

namespace gvss {

// One host-pointer search in flight on one single-GPU index: the slot it runs on and where its packed results lie
// (device: res + o_*, pinned host mirror: h_res + o_*). The index's reader lock is held from launch to release.
struct HostSearch {
    gvss_index *x = nullptr;
    SearchSlot *slot = nullptr;
    cudaStream_t stream = nullptr;
    std::shared_lock<std::shared_mutex> lock;
    size_t q_bytes = 0, o_counts = 0, o_keys = 0, o_dist = 0, o_stats = 0, r_bytes = 0;
    uint8_t *res = nullptr, *h_res = nullptr;
    int64_t *d_keys = nullptr;
    float *d_dist = nullptr;
};
// upload + launch on a slot of `x` (inserts staged rows first, makes x's device current, takes the reader lock);
// fetch = ONE packed D2H copy + stream sync; release gives slot and lock back (safe to call twice)
int host_search_launch(gvss_index *x, const float *queries, uint64_t nq, uint32_t k, uint32_t ef, bool want_stats, HostSearch &hs);
int host_search_fetch(HostSearch &hs, unsigned long long *failed);
void host_search_release(HostSearch &hs);
int set_error(const char *fmt, ...); // thread-local message of gvss_last_error; returns 1

// gvss_create_sharded: one logical index = one single-GPU index per device, rows dealt out in chunks, every query searched
// on every shard, per-shard lists merged on the first device
struct Router;
int router_create(const gvss_config *cfg, const int32_t *devices, uint32_t n_devices, Router **out);
void router_destroy(Router *r);
int router_reset(Router *r);
int router_reserve(Router *r, uint64_t members, uint32_t host_threads);
int router_add(Router *r, int64_t key, const float *vec, uint32_t host_thread);
int router_add_batch(Router *r, const int64_t *keys, const float *vecs, uint64_t n);
int router_flush(Router *r);
int router_set_build_concurrency(Router *r, uint32_t n);
int router_set_background_build(Router *r, uint64_t min_rows);
int router_remove(Router *r, int64_t key, uint64_t *removed);
int router_compact(Router *r, bool rebuild, uint64_t *dropped);
int router_search(Router *r, const float *queries, uint64_t nq, uint32_t k, uint32_t ef, int64_t *keys_out, float *dist_out,
                  uint32_t *counts_out, gvss_search_stats *stats_out);
int router_info(Router *r, gvss_info *out);
int router_build_stats(Router *r, gvss_build_stats *out);
int router_level_stats(Router *r, uint32_t level, gvss_level_stats *out);

} // namespace gvss
