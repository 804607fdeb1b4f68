/* gvss — C ABI of the B200-native HNSW hot path that replaces usearch's index_dense_gt<row_t>
 * inside DuckDB's vss extension (reference: duckdb/duckdb_vss).
 *
 * The drop-in boundary is the value member `HNSWIndex::index`
 * (reference src/include/hnsw/hnsw_index.hpp:28,39). Every entry point below cites the reference
 * call site / usearch method it stands in for. Rules of the ABI:
 *   - plain C: opaque handle, plain pointers and sizes, no C++/torch types, no exceptions;
 *   - every function returns 0 on success, non-zero on failure; gvss_last_error() gives the
 *     message of the calling thread's last failure (static lifetime until that thread's next call);
 *   - input pointers are borrowed for the duration of the call, outputs are caller-allocated;
 *   - pointers are HOST memory unless the function name ends in `_device`;
 *   - there is no CPU fallback: with no CUDA device every compute entry point fails with an error.
 *
 * Threading (mirrors how the extension calls usearch):
 *   - gvss_add may be called from `host_threads` threads at once, each with its own thread id 0..host_threads-1
 *     (PhysicalCreateHNSWIndex, hnsw_index_physical_create.cpp:144-205), and from any number of threads with
 *     GVSS_ANY_THREAD (HNSWIndex::Construct under the SHARED side of HNSWIndex::rwlock, hnsw_index.cpp:433-447) — also
 *     while searches run: a row is only staged, behind a per-buffer mutex, and enters the graph in the next flush;
 *   - gvss_search / gvss_search_device / gvss_save / gvss_get_* hold the index's reader lock: calls from different host
 *     threads run concurrently, each gvss_search on a stream and scratch buffer of its own;
 *   - everything that changes the graph (flush — explicit or triggered by a search that finds staged rows —, remove,
 *     compact, reserve, reset, load) takes the writer lock and waits for running searches. The library does its own
 *     locking: no call sequence from any number of threads corrupts the index.
 */
#ifndef GVSS_H
#define GVSS_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GVSS_METRIC_L2SQ 'e' /* metric_kind_t::l2sq_k, index_plugins.hpp:103-109 (values are part of the file format) */
#define GVSS_METRIC_COS 'c'  /* metric_kind_t::cos_k */
#define GVSS_METRIC_IP 'i'   /* metric_kind_t::ip_k  */

#define GVSS_FREE_KEY INT64_MAX /* key written into removed nodes, index_dense.hpp:1246-1250 */
#define GVSS_ANY_THREAD 0xFFFFFFFFu /* index_dense_t::any_thread(), hnsw_index.cpp:485,497: "whichever thread I am" */

typedef struct gvss_index gvss_index;

/* index_dense_config_t + metric_punned_t as filled by the HNSWIndex ctor (hnsw_index.cpp:140-189).
 * Zero means "usearch default": M=16, M0=2*M, ef_construction=128, ef_search=64 (index.hpp:1097-1107,1125). */
typedef struct gvss_config {
    uint32_t dimensions;
    char metric;              /* GVSS_METRIC_* */
    uint32_t connectivity;      /* M  */
    uint32_t connectivity_base; /* M0 */
    uint32_t expansion_add;     /* ef_construction */
    uint32_t expansion_search;  /* ef_search */
    int32_t device;             /* CUDA ordinal; -1 = current device */
} gvss_config;

/* index_dense_gt::stats_t per level (index.hpp:2612-2635; hnsw_index_pragmas.cpp:121-135) */
typedef struct gvss_level_stats {
    uint64_t nodes, edges, max_edges, allocated_bytes;
} gvss_level_stats;

typedef struct gvss_info {
    uint64_t size;         /* present (non-removed) members: index.size(), hnsw_index.cpp:210,463,481 */
    uint64_t capacity;     /* index.capacity(), hnsw_index.cpp:416,427 */
    uint64_t node_count;   /* graph nodes incl. removed slots */
    uint64_t max_level;    /* index.max_level(), hnsw_index.cpp:266 */
    uint64_t memory_usage; /* index.memory_usage(), hnsw_index.cpp:272,558 */
    uint64_t dimensions;   /* index.dimensions(), hnsw_index.cpp:216 */
    uint32_t connectivity, connectivity_base, expansion_add, expansion_search;
    char metric;           /* index.metric().metric_kind(), hnsw_index.cpp:220 */
    uint32_t concurrent_searches_peak; /* most gvss_search calls that were in flight on the GPU at the same time so far */
} gvss_info;

/* per-query traversal counters, same meaning as usearch's search_result_t (index.hpp:2187-2192) */
typedef struct gvss_search_stats {
    uint32_t computed_distances; /* context_t::computed_distances_count delta */
    uint32_t visited_members;    /* context_t::iteration_cycles delta = level-0 expansions + upper-level hops */
    uint32_t base_expansions;    /* level-0 expansions only */
    uint32_t flags;              /* bit1: hashed visited table overfilled, level 0 redone with the exact bitmap (results exact);
                                    bit2: two candidates tied exactly for the head of the queue, the rest of the query ran with
                                          usearch's binary heap as the queue to keep its pop order (results exact);
                                    bit0: candidate queue overflow (results of this query are not reliable) */
} gvss_search_stats;

/* --- lifetime ----------------------------------------------------------------------------------------------- */
/* index_dense_gt<row_t>::make(metric, config), hnsw_index.cpp:162-189 */
/* Optional: create the CUDA context of `device` (-1 = the current one) now instead of inside the first gvss_create. On a B200
 * the first CUDA call of a process costs about a second; a host that knows it will use the index (DuckDB loading the vss
 * extension) calls this from a helper thread at start-up so that the first CREATE INDEX does not wait for it. Blocking,
 * idempotent, thread-safe. */
int gvss_warmup(int32_t device);
int gvss_create(const gvss_config *config, gvss_index **out);
/* One logical index partitioned over several GPUs of this process (no counterpart in the reference, which is
 * single-process CPU code): one sub-graph per device, rows dealt out in chunks of 2048 in arrival order, keys stay the
 * caller's global row ids, no build-time communication. Every other entry point takes the handle unchanged: gvss_add* route
 * rows, gvss_flush builds all shards at the same time, gvss_search uploads the batch to every GPU, searches all shards
 * concurrently and merges the per-shard lists by (distance, key) in one kernel on devices[0] that reads the other GPUs' lists
 * over NVLink (peer access; a staged copy without it). devices: n_devices CUDA ordinals (1..16, repeats allowed);
 * shards * k <= 256 per search; searches on one sharded handle are serialised. NOT available on a sharded handle:
 * gvss_save / gvss_load / gvss_serialized_length (a sharded index is not one usearch stream — the reference cannot merge
 * indexes either, hnsw_index.cpp:561-563 — so persisted indexes stay single-GPU), gvss_search_device*, gvss_add_batch_device. */
int gvss_create_sharded(const gvss_config *config, const int32_t *devices, uint32_t n_devices, gvss_index **out);
void gvss_destroy(gvss_index *index);
/* index.reset(), hnsw_index.cpp:384 */
int gvss_reset(gvss_index *index);
/* index.reserve(n) / index.reserve({members, threads}), hnsw_index.cpp:208,429; hnsw_index_physical_create.cpp:296.
 * Grow only; thread ids 0..host_threads-1 become valid for gvss_add. Not concurrent with gvss_add. */
int gvss_reserve(gvss_index *index, uint64_t members, uint32_t host_threads);

/* --- construction ------------------------------------------------------------------------------------------- */
/* index.add(key, vec, thread), hnsw_index_physical_create.cpp:184 and hnsw_index.cpp:443. Copies `vec` before
 * returning; only stages the row (graph work happens in gvss_flush or the next call that needs a consistent
 * graph). host_thread: an id below the thread count given to gvss_reserve, or GVSS_ANY_THREAD; any other value is an
 * error. Fails with "Reserve capacity ahead of insertions!" (index.hpp:2332-2335) when full. Duplicate keys are NOT
 * rejected ("Duplicate keys not allowed in high-level wrappers", index_dense.hpp:1752-1753, never fires in the
 * extension either: enable_key_lookups=false makes contains() always false). */
int gvss_add(gvss_index *index, int64_t key, const float *vec, uint32_t host_thread);
/* n rows, row-major n×dimensions — the layout HNSWIndex::Construct already has after Flatten (hnsw_index.cpp:399-406) */
int gvss_add_batch(gvss_index *index, const int64_t *keys, const float *vecs, uint64_t n);
/* Same with rows already resident on the index's device (bulk-build benchmark path: H2D excluded). */
int gvss_add_batch_device(gvss_index *index, const int64_t *keys_host, const float *vecs_device, uint64_t n);
/* Insert everything staged: level draw (index.hpp:3315-3319), per-level beam search (search_to_insert_,
 * index.hpp:3430-3496), neighbour-selection heuristic (refine_, :3602-3638), reverse links with re-pruning
 * (reconnect_neighbor_nodes_, :3270-3313) — as sm_100a kernels. */
int gvss_flush(gvss_index *index);
/* Upper bound on vectors inserted concurrently by gvss_flush (0 = as many as the GPU holds). The role the thread count
 * plays in index.reserve({members, threads}) (hnsw_index_physical_create.cpp:296; index_limits_t, index.hpp:1134-1140):
 * with 1 the build is sequential and reproduces the reference's single-thread graph byte for byte. */
int gvss_set_build_concurrency(gvss_index *index, uint32_t max_concurrent_inserts);
/* Rows only (SURVEY.md 8(f) N2, tables without an HNSW index): with on != 0 the index keeps rows, keys and removals in HBM
 * and never builds adjacency lists — gvss_flush stops after the uploads. gvss_exact_search is the one search it answers;
 * the graph calls (gvss_search*, gvss_save / gvss_load, gvss_compact*) fail with a message. Choose before the first add.
 * It is the transient row store behind the index join of an un-indexed table (integration/n2b_exact_join_unindexed.patch:
 * the brute-force lateral join of hnsw_index_macros.cpp:10-78 without the CPU cross product). */
int gvss_set_rows_only(gvss_index *index, int on);
/* Pipelined bulk load (the reference collects every row first and builds after Finalize,
 * hnsw_index_physical_create.cpp:98-128,283-306): with min_rows > 0 a library-owned thread uploads and inserts the staged
 * rows whenever at least min_rows of them wait, while the caller goes on adding — table scan, H2D and graph construction
 * overlap, and gvss_flush only drains the rest. 0 switches it off (default). Row order within one staging buffer is kept,
 * so a single adder with gvss_set_build_concurrency(1) still gets the reference's sequential graph. A failure of a
 * background insert is reported by the next gvss_flush. */
int gvss_set_background_build(gvss_index *index, uint64_t min_rows);
/* index.remove(key), hnsw_index.cpp:478: lazy delete (index_dense.hpp:1228-1255); unknown key: *removed = 0. */
int gvss_remove(gvss_index *index, int64_t key, uint64_t *removed);
/* index.compact(), hnsw_index.cpp:458 = index_gt::compact (index.hpp:3007-3096): re-orders the nodes by (level descending,
 * nearest upper-level node ascending) and renames every link; removed members are kept (this usearch version does not
 * drop them). The stream saved afterwards equals the reference's after its compact(), byte for byte. */
int gvss_compact(gvss_index *index);
/* Deleting compaction (the extension's README.md:69 says PRAGMA hnsw_compact_index prunes deleted
 * items; index_gt::compact in this usearch version, index.hpp:3007-3096, only re-orders and keeps them): builds a fresh graph
 * over the live members on the GPU and drops the removed ones, so searches stop traversing tombstones. Capacity and
 * configuration stay; slot numbers and the graph change, keys do not. *dropped (optional) = members removed. Not called
 * by gvss_compact: a maintainer opts in by calling it from HNSWIndex::Compact instead. */
int gvss_compact_rebuild(gvss_index *index, uint64_t *dropped);

/* --- search -------------------------------------------------------------------------------------------------- */
/* index.ef_search(query, k, ef) for nq queries at once (hnsw_index.cpp:303-309 single query, :359-364 index join).
 * keys_out: nq×k row ids ascending by distance (unused tail = -1); dist_out (nq×k, may be NULL): the index metric
 * (l2sq / 1-cos / 1-ip), unused tail = +inf; counts_out (nq, may be NULL): results per query (<= k);
 * stats_out (nq, may be NULL). ef == 0 means 64 (index.hpp:2493-2494); beam = max(ef, k) (index.hpp:2512). */
int gvss_search(gvss_index *index, const float *queries, uint64_t nq, uint32_t k, uint32_t ef, int64_t *keys_out,
                float *dist_out, uint32_t *counts_out, gvss_search_stats *stats_out);
/* Same with every buffer resident on the index's device; runs on `cuda_stream` (a cudaStream_t, 0 = default stream)
 * and returns without synchronising the stream. queries: nq×dimensions, 16-byte aligned. */
int gvss_search_device(gvss_index *index, const float *queries, uint64_t nq, uint32_t k, uint32_t ef,
                       int64_t *keys_out, float *dist_out, uint32_t *counts_out, gvss_search_stats *stats_out,
                       void *cuda_stream);

/* Same search as gvss_search_device, for callers that must not allocate device memory while they search. No counterpart
 * in the reference (usearch allocates freely on the host); it exists for hosts that have given all free HBM to their
 * own memory pool, where a cudaMalloc from this library would fail or make that pool shrink, and for latency-critical
 * callers, because cudaMalloc / cudaFree synchronise the whole device. This call runs no cudaMalloc, cudaFree or other
 * allocating routine: the per-query scratch (visited tables, candidate-queue overflow) lives in a 64 MiB array built
 * into the library, which every GPU that loads the library holds until the process exits.
 *  - Limit: one query slot may use at most 1 MiB of the array: 12 bytes per visited-table slot (max(ef,k) * max(M,M0)
 *    + 2048 slots rounded up to a power of two, at least 4096) plus one bit per graph node — e.g. with M0 = 32: ef <= 512 on
 *    a 4M-node index, ef <= 1024 on 1M nodes. Beyond that the call fails with an error that states the sizes. When fewer slots fit
 *    than the GPU could run at once, the call is correct but slower than gvss_search_device.
 *  - Rows staged by gvss_add* must have been inserted with gvss_flush beforehand (flushing allocates; this call
 *    refuses instead).
 *  - The first call on each device resolves the array's device address and, like any first launch, loads the search kernel's
 *    module (the driver may allocate for that): make one warm-up call before any stream capture and before counting on
 *    "no allocation".
 *  - All calls on one GPU share the array and it keeps the previous call's contents: the library clears what must
 *    start clean, but two calls that overlap on the GPU (different streams or threads) race and corrupt each other.
 *    Serialise them. */
int gvss_search_device_noalloc(gvss_index *index, const float *queries, uint64_t nq, uint32_t k, uint32_t ef,
                               int64_t *keys_out, float *dist_out, uint32_t *counts_out, gvss_search_stats *stats_out,
                               void *cuda_stream);

/* --- persistence: usearch v2.12 stream, byte-compatible in both directions ------------ */
/* index.save_to_stream(cb), hnsw_index.cpp:526-529 (index_dense.hpp:811-878, index.hpp:2709-2749).
 * `write` returns non-zero on success (like the usearch callback returning true). */
int gvss_save(gvss_index *index, int (*write)(void *user, const void *data, size_t size), void *user);
/* index.load_from_stream(cb), hnsw_index.cpp:204-205 (index_dense.hpp:900-973, index.hpp:2755-2807) */
int gvss_load(gvss_index *index, int (*read)(void *user, void *data, size_t size), void *user);
/* index.serialized_length() (index_dense.hpp:883-891) */
int gvss_serialized_length(gvss_index *index, uint64_t *bytes);

/* --- introspection ------------------------------------------------------------------------------------------- */
int gvss_get_info(gvss_index *index, gvss_info *out);
/* Traversal counters of the graph construction, summed over every insert since gvss_create / gvss_reset: the bulk form of
 * add_result_t::computed_distances / visited_members (index.hpp:2133-2139), split the way SURVEY.md 8(d)'s build formula
 * needs them: B_v = C_v*4d*w + H_v*(4+4*M0) + E_v*(4+4*M0). Flushes first. */
typedef struct gvss_build_stats {
    uint64_t inserts;             /* nodes linked into the graph */
    uint64_t query_distances;     /* distances from the new vector to a node (one row read each, w = 1) */
    uint64_t pair_distances;      /* node-to-node distances of the neighbour-selection heuristic and re-pruning (two rows, w = 2);
                                     evaluated 32 candidates at a time: a few per cent above the reference's count on the same graph */
    uint64_t base_expansions;     /* level-0 expansions of the insert-time beam searches (one adjacency row read each) */
    uint64_t upper_expansions;    /* greedy hops and beam-search expansions on levels > 0 */
    uint64_t rows_written;        /* adjacency rows written (own lists, appended and re-pruned reverse links) */
} gvss_build_stats;
int gvss_get_build_stats(gvss_index *index, gvss_build_stats *out);
/* index.stats(level), hnsw_index.cpp:271-273 */
int gvss_get_level_stats(gvss_index *index, uint32_t level, gvss_level_stats *out);

/* --- exact (brute-force) top-k: the comparator behind `ORDER BY array_distance(v, q), rowid LIMIT k`, the vss_join /
 * vss_match macros (hnsw_index_macros.cpp:10-78) and the ground truth for recall. Distances follow DuckDB's folds
 * bit for bit (array_kernels.hpp:11-105: 'e' = sqrt(sum (x-y)^2), 'c' = 1 - clamp(xy / sqrt(xx*yy)), 'i' = -sum xy);
 * ties are broken by the smaller row index. ids_out: nq×k indices into `base` (or keys[idx] when keys != NULL). */
int gvss_exact_topk(char metric, const float *base, uint64_t n, const int64_t *keys, const float *queries,
                    uint64_t nq, uint32_t dimensions, uint32_t k, int64_t *ids_out, float *dist_out);
int gvss_exact_topk_device(char metric, const float *base, uint64_t n, const int64_t *keys, const float *queries,
                           uint64_t nq, uint32_t dimensions, uint32_t k, int64_t *ids_out, float *dist_out,
                           void *cuda_stream);

/* The same comparator over the rows an index already holds in HBM (no copy of the base set): exact top-k of nq host queries
 * among the index's live members, distances = DuckDB's fold of the index metric (array_distance / array_cosine_distance /
 * array_negative_inner_product, array_kernels.hpp:11-105), keys ascending by (distance, slot); keys_out nq×k (unused tail -1),
 * dist_out nq×k or NULL, counts_out nq or NULL. It is what `SET hnsw_exact_search = true` (integration/n2_exact_mode.patch)
 * puts behind HNSW_INDEX_SCAN / HNSW_INDEX_JOIN, i.e. behind the vss_join / vss_match / min_by(.., k) shapes
 * (hnsw_index_macros.cpp:10-78, hnsw_optimize_topk.cpp:64-219) once an index exists. k + removed members <= 1024. */
int gvss_exact_search(gvss_index *index, const float *queries, uint64_t nq, uint32_t k, int64_t *keys_out, float *dist_out,
                      uint32_t *counts_out);

/* --- multi-GPU: merge of per-shard top-k lists after the all-gather (no reference counterpart:
 * the reference is single-process). gathered_*: shards×nq×k on the device, each list ascending; output nq×k
 * ordered by (distance, key). */
int gvss_merge_topk_device(const int64_t *gathered_keys, const float *gathered_dist, uint32_t shards, uint64_t nq,
                           uint32_t k, int64_t *keys_out, float *dist_out, void *cuda_stream);

/* The same merge WITHOUT a preceding all-gather: the kernel reads every shard's nq×k list out of the owning GPU's memory
 * over NVLink and the ranks synchronise through flags in peer memory (no reference counterpart).
 * peer_keys[s] / peer_dist[s]: shard s's result buffers as mapped into THIS process (CUDA IPC / symmetric memory; entry
 * [rank] is the local buffer); peer_flags[s]: shard s's flag array (16 zero-initialised u32). All three are HOST arrays
 * of `shards` device pointers. Call it on every rank, after that rank's search in stream order, with the same `step`
 * = 1, 2, 3, ... ; the caller alternates between two sets of result buffers from step to step. shards <= 16,
 * shards * k <= 256. A rank that does not arrive within GVSS_PEER_TIMEOUT_MS (environment, default 20000) makes the
 * others skip the merge of that step (outputs untouched) and record it: see gvss_peer_exchange_status. */
int gvss_merge_topk_peers_device(const int64_t *const *peer_keys, const float *const *peer_dist, uint32_t *const *peer_flags,
                                 uint32_t rank, uint32_t shards, uint32_t step, uint64_t nq, uint32_t k, int64_t *keys_out,
                                 float *dist_out, void *cuda_stream);

/* Bit r of *timed_out is set when a gvss_merge_topk_peers_device launch on the current device gave up waiting for rank r
 * since the last call (the word is cleared by the call; it synchronises the device). 0 = every merge so far was complete. */
int gvss_peer_exchange_status(uint32_t *timed_out);

/* --- misc ---------------------------------------------------------------------------------------------------- */
const char *gvss_last_error(void);
const char *gvss_version(void);
/* number of kernels this library has launched in the calling process (bench.py's `gpu_launches`) */
uint64_t gvss_kernel_launches(void);

#ifdef __cplusplus
}
#endif
#endif /* GVSS_H */
