// Host-visible launch interface of the CUDA kernels (internal to libgvss; the public boundary is include/gvss.h).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/gvss.h"
#include "hnsw_device.cuh"

namespace gvss {

// shared-memory carve-out of one warp (one CTA = one warp, so a CTA never needs __syncthreads)
struct SmemLayout {
    uint32_t off_mbar, off_stage0, off_stage1, off_c, off_q, off_heap, off_top_d, off_top_s, off_nb, off_nd, off_misc;
    uint32_t rstride, stage_rows, chunk, nbuf, heap_smem, top_cap, bytes;
};
// stage_rows: 32 or 16; with_left: room for a streamed left-hand row chunk (node-to-node distances, build only);
// list_cap: longest neighbour list (max(M, M0)); g4_alias (build): the whole-row stage is 1024-byte aligned so that its first
// 4 KB can serve as the gather4 stage of the query-to-node distances
// chunk: floats of a row staged per step (0 = default min(dpad, 128)); double_buffer: two stage buffers when a row spans
// several chunks (build), else the chunks are staged one after the other into one buffer (search: half the footprint)
SmemLayout make_smem_layout(uint32_t dpad, uint32_t top_cap, uint32_t heap_smem, uint32_t misc_bytes, uint32_t stage_rows,
                            bool with_left, uint32_t list_cap, uint32_t chunk = 0, bool double_buffer = true, bool gather4 = false,
                            bool g4_alias = false, uint32_t g4_stages = 1);
// tensor map for the gather4 staging path (exact_tc.cu holds the driver entry point); false when the driver refuses
bool encode_row_gather_map(CUtensorMap *m, const float *vectors, uint64_t rows, uint32_t dpad);
uint32_t default_heap_smem(uint32_t beam);

// global scratch of one warp: visited table (vis_slots u32) followed by the queue overflow (heap_ovf Cand)
struct ScratchLayout {
    uint32_t vis_slots, vis_limit, heap_ovf, bitmap_words;
    size_t bytes_per_warp;
};
ScratchLayout make_scratch_layout(uint32_t beam, uint64_t nodes, uint32_t list_cap);

struct SearchParams {
    const float *queries;
    uint64_t nq;
    uint32_t k, limit;
    int64_t *keys_out;
    float *dist_out;
    uint32_t *counts_out;
    gvss_search_stats *stats_out;
    uint32_t entry;
    int32_t max_level;
    unsigned long long *work_counter;
    unsigned long long *fail_word; // optional: receives 1 + the largest index of a query whose candidate queue overflowed
    uint32_t query_stride;         // floats between queries (0 = dimensions)
    uint32_t *cluster_out;         // optional: descent-only mode for compact — cluster_out[q] = result of the greedy descent, no beam search
    uint8_t *scratch;
    ScratchLayout sl;
    SmemLayout sm;
    alignas(64) CUtensorMap vmap; // 2-D map over `vectors` ([capacity][dpad], box 32 floats x 1 row, 128-byte swizzle) for row gathers
    long long trace_query; // -1 = off
    int direct;            // bit0: stream rows into registers instead of staging them; bit1: speculative row prefetch;
                           // bit2: stage with cp.async instead of TMA; bit3: always use usearch's heap as the queue;
                           // bit4: rows staged by TMA gather4 (4 rows per instruction) through `vmap`, 32 rows x 32 floats per step;
                           // bit5: no whole-row L2 prefetch at the start of a hop; bit6: no visited-window prefetch;
                           // bit7: latency kernel — 16 KB gather4 stage, batches that fit the GPU at once (k_search<METRIC, 2>)
};
// grid = number of resident warps; returns cudaError
cudaError_t launch_search(int metric, const GraphView &g, const SearchParams &p, uint32_t grid, cudaStream_t stream);
// size of the built-in scratch array behind gvss_search_device_noalloc (held by every GPU that loads the library)
#define GVSS_STATIC_SCRATCH_BYTES (64u << 20)
// largest scratch one query slot may need there (so that at least 63 queries are always in flight)
#define GVSS_STATIC_SLOT_MAX_BYTES (1u << 20)
cudaError_t launch_search_static_scratch(int metric, int device, const GraphView &g, SearchParams p, uint32_t resident,
                                         cudaStream_t stream, char *err, size_t err_len);
// how many single-warp CTAs of the search / build kernel fit the device at this shared-memory size
uint32_t resident_warps(int device, uint32_t smem_bytes, bool build);

struct BuildParams {
    const uint32_t *slots;   // nodes to insert, in order
    uint32_t begin, end;     // this launch inserts slots[begin..end)
    const int16_t *levels;   // [capacity]
    unsigned long long *entry_state; // (max_level << 32) | entry_slot ; ~0ull while the graph is empty
    unsigned long long *work_counter;
    uint32_t *locks;         // [capacity] per-node spin locks (index.hpp:3214-3224 nodes_mutexes_)
    uint32_t ef_add;
    int direct;
    uint32_t *inflight;       // [grid] slot each warp is inserting right now (kInvalid = idle); nullptr disables the peer pass
    unsigned long long *stats; // [6] inserts, query distances, pair distances, level-0 expansions, upper expansions, rows written
    unsigned long long *prof; // optional phase clocks: [0] greedy [1] beam search [2] refine [3] connect [4] reconnect [5] inserts
    uint8_t *scratch;
    ScratchLayout sl;
    SmemLayout sm;
};
cudaError_t launch_build(int metric, const GraphView &g, const BuildParams &p, uint32_t grid, cudaStream_t stream);

// multi-GPU merge that reads the peers' lists over NVLink and synchronises through flags in peer memory (exact.cu)
cudaError_t launch_merge_topk_peers(const int64_t *const *peer_keys, const float *const *peer_dist, uint32_t *const *peer_flags,
                                    uint32_t rank, uint32_t shards, uint32_t step, uint64_t nq, uint32_t k, int64_t *keys_out,
                                    float *dist_out, cudaStream_t stream);
// bit r set = some merge on this device gave up waiting for rank r (read and cleared; synchronises the device)
cudaError_t read_peer_timeouts(uint32_t *out);
cudaError_t launch_exact_topk(char metric, const float *base, uint64_t n, const int64_t *keys, const float *queries,
                              uint64_t nq, uint32_t d, uint32_t k, int64_t *ids_out, float *dist_out,
                              cudaStream_t stream);
// tensor-core candidate generation + exact re-rank (exact_tc.cu); cudaErrorNotSupported = shape outside that path
cudaError_t launch_exact_topk_tc(char metric, const float *base, uint64_t n, const int64_t *keys, const float *queries, uint64_t nq,
                                 uint32_t d, uint32_t k, int64_t *ids_out, float *dist_out, float *dbg_dots, uint64_t *flagged_out,
                                 cudaStream_t stream);
cudaError_t launch_merge_partials(const unsigned long long *pk, const float *ps, uint32_t splits, uint64_t nq, uint32_t kc, int64_t *rows_out,
                                  float *scores_out, cudaStream_t stream);
cudaError_t launch_exact_topk_flagged(char metric, const float *base, uint64_t n, const int64_t *keys, const float *queries, uint64_t nq,
                                      uint32_t d, uint32_t k, int64_t *ids_out, float *dist_out, const uint32_t *flags, uint64_t *flagged_out,
                                      cudaStream_t stream);
cudaError_t launch_merge_topk(const int64_t *gk, const float *gd, uint32_t shards, uint64_t nq, uint32_t k,
                              int64_t *keys_out, float *dist_out, cudaStream_t stream);

void count_launch(uint64_t n = 1);

} // namespace gvss
