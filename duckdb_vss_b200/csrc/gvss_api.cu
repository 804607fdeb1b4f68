// libgvss host side: device-resident index container, usearch v2.12 stream (de)serialisation and the C ABI of
// include/gvss.h. Stands in for usearch's index_dense_gt<row_t> (index_dense.hpp:305-1996) behind
// HNSWIndex::index (reference src/include/hnsw/hnsw_index.hpp:28,39). No CPU fallback: every compute entry
// point needs a CUDA device and fails loudly without one.
#include "kernels.h"
#include "router.h"

#include <algorithm>
#include <chrono>
#include <atomic>
#include <condition_variable>
#include <exception>
#include <memory>
#include <shared_mutex>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

namespace gvss {

static std::atomic<uint64_t> g_launches{0};
void count_launch(uint64_t n) { g_launches += n; }

static thread_local std::string t_error;
static int fail(const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    t_error = buf;
    return 1;
}
int set_error(const char *fmt, ...) { // for router.cu
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    t_error = buf;
    return 1;
}
#define CU(x)                                                                                                         \
    do {                                                                                                              \
        cudaError_t e_ = (x);                                                                                         \
        if (e_ != cudaSuccess) return fail("CUDA error %s at %s:%d (%s)", cudaGetErrorString(e_), __FILE__, __LINE__, #x); \
    } while (0)

static int metric_id(char m) { return m == GVSS_METRIC_COS ? kCos : m == GVSS_METRIC_IP ? kIp : kL2sq; }

template <typename T> static cudaError_t dev_resize(T *&p, size_t old_n, size_t new_n, int fill_byte, cudaStream_t st) {
    T *np = nullptr;
    cudaError_t e = cudaMalloc(&np, std::max<size_t>(new_n, 1) * sizeof(T));
    if (e != cudaSuccess) return e;
    if (old_n && p) e = cudaMemcpyAsync(np, p, std::min(old_n, new_n) * sizeof(T), cudaMemcpyDeviceToDevice, st);
    if (e == cudaSuccess && new_n > old_n) e = cudaMemsetAsync(np + old_n, fill_byte, (new_n - old_n) * sizeof(T), st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (p) cudaFree(p);
    p = np;
    return e;
}

} // namespace gvss

using namespace gvss;

// One concurrent host-pointer search in flight: its own stream, work counter, scratch and I/O buffers, so that gvss_search
// calls from several host threads (the extension holds only the shared side of HNSWIndex::rwlock around a scan,
// hnsw_index.cpp:285-311) overlap on the GPU instead of queueing behind one mutex.
struct SearchSlot {
    cudaStream_t stream = nullptr;
    unsigned long long *d_work = nullptr; // [0] work counter, [1] 1 + index of a query that overflowed (0 = none)
    uint8_t *d_scratch = nullptr;
    size_t scratch_bytes = 0;
    uint8_t *d_io = nullptr;              // queries | packed results
    size_t io_bytes = 0;
    uint8_t *h_io = nullptr;              // pinned mirror of the packed results
    size_t h_io_bytes = 0;
    bool busy = false;
};

struct gvss_index {
    gvss_config cfg{};
    int device = 0;
    uint32_t dpad = 0, stride0 = 0, strideU = 0;
    cudaStream_t stream = nullptr;

    // device state
    float *d_vectors = nullptr;
    int64_t *d_keys = nullptr;
    int16_t *d_levels = nullptr;
    uint32_t *d_nbr0 = nullptr, *d_upper = nullptr, *d_upper_off = nullptr, *d_locks = nullptr;
    unsigned long long *d_entry_state = nullptr, *d_work = nullptr;
    unsigned long long *d_build_stats = nullptr; // 6 counters summed by k_build (gvss_build_stats order)
    uint64_t capacity = 0, count = 0;           // count = graph nodes (incl. removed slots)
    uint64_t upper_cap = 0, upper_used = 0;      // in blocks of strideU u32
    uint8_t *d_scratch = nullptr;                // build + gvss_search_device (caller's stream)
    size_t scratch_bytes = 0;
    cudaEvent_t scratch_free = nullptr; // recorded after every launch that uses d_scratch / d_work

    // host mirrors
    std::vector<int64_t> h_keys;
    std::vector<int16_t> h_levels;
    std::vector<uint32_t> h_upper_off;
    uint32_t entry = 0;
    int32_t max_level = -1;
    uint64_t n_deleted = 0;
    bool lookup_valid = false;
    std::unordered_multimap<int64_t, uint32_t> key_to_slot;
    std::deque<uint32_t> free_slots; // ring of removed slots, FIFO like usearch's free_keys_ (index_dense.hpp:1766-1771)
    uint32_t build_conc = 0;         // gvss_set_build_concurrency
    bool flat = false;               // gvss_set_rows_only: rows, keys and removals only, no graph is ever built
    uint64_t level_rng = 1;          // minstd_rand0 state (std::default_random_engine, index.hpp:1892)
    std::string poisoned;            // non-empty after a CUDA failure in the middle of a graph update: every later call fails with it

    // staging (gvss_add from T threads): one buffer per host thread id plus one for GVSS_ANY_THREAD callers, each behind its
    // own mutex (uncontended for the bulk build; serialises any_thread adders and the flush that takes the rows away)
    struct Stage {
        std::mutex mu;
        std::vector<float> vecs;
        std::vector<int64_t> keys;
    };
    uint32_t host_threads = 1;
    std::vector<std::unique_ptr<Stage>> stages; // [host_threads + kAnyStages]; the last kAnyStages belong to GVSS_ANY_THREAD callers
    static constexpr size_t kAnyStages = 8;     // (picked by a hash of the calling thread: concurrent INSERT / Sink threads do not queue on one mutex)
    std::vector<std::pair<const float *, std::vector<int64_t>>> st_device; // rows already on the device (guarded by rw, exclusive)
    std::atomic<uint64_t> staged{0};
    std::atomic<int64_t> room{0}; // rows that may still be added: capacity + free slots - nodes - staged (index.hpp:2329-2335)
    // what gvss_get_info reports, republished by every writer before it lets go of `rw`: size() / capacity() are asked for by
    // HNSWIndex::Construct before every chunk (hnsw_index.cpp:413-431) and must not wait for a flush that holds the writer lock
    std::atomic<uint64_t> pub_capacity{0}, pub_count{0}, pub_deleted{0}, pub_upper_cap{0}, pub_scratch{0};
    std::atomic<uint64_t> pub_inflight{0}; // rows a flush has taken from staging and not yet published in pub_count
    std::atomic<int32_t> pub_max_level{-1};

    // Locking. `rw` plays the role usearch's per-node locks and per-thread contexts play in the reference: searches hold it
    // shared and run concurrently, everything that changes the graph holds it exclusively. `dev_mu` orders the enqueue
    // sections of gvss_search_device calls (they share d_scratch / d_work through the scratch_free event).
    std::shared_mutex rw;
    std::mutex dev_mu;
    std::mutex slot_mu;
    std::condition_variable slot_cv;
    std::vector<std::unique_ptr<SearchSlot>> slots;
    uint32_t slots_peak = 0; // most slots busy at the same time

    // background build (gvss_set_background_build): a library-owned thread inserts staged rows while the caller adds more
    std::thread bg;
    std::mutex bg_mu;
    std::condition_variable bg_cv;
    bool bg_stop = false;
    std::atomic<uint64_t> bg_min_rows{0};
    std::string bg_error; // first failure of a background flush (guarded by bg_mu), reported by the next gvss_flush

    Router *router = nullptr; // gvss_create_sharded: this handle only routes to one single-GPU index per device (router.cu)
};

namespace gvss {

static GraphView view_of(const gvss_index *x) {
    GraphView g{};
    g.vectors = x->d_vectors; g.keys = x->d_keys; g.nbr0 = x->d_nbr0; g.upper = x->d_upper; g.upper_off = x->d_upper_off;
    g.dpad = x->dpad; g.dims = x->cfg.dimensions; g.M = x->cfg.connectivity; g.M0 = x->cfg.connectivity_base;
    g.stride0 = x->stride0; g.strideU = x->strideU; g.count = (uint32_t)x->count; g.has_deleted = x->n_deleted ? 1u : 0u;
    return g;
}

static void free_device(gvss_index *x) {
    cudaFree(x->d_vectors); cudaFree(x->d_keys); cudaFree(x->d_levels); cudaFree(x->d_nbr0); cudaFree(x->d_upper);
    cudaFree(x->d_upper_off); cudaFree(x->d_locks); cudaFree(x->d_scratch);
    x->d_vectors = nullptr; x->d_keys = nullptr; x->d_levels = nullptr; x->d_nbr0 = nullptr; x->d_upper = nullptr;
    x->d_upper_off = nullptr; x->d_locks = nullptr; x->d_scratch = nullptr; x->scratch_bytes = 0;
    x->capacity = 0; x->count = 0; x->upper_cap = 0; x->upper_used = 0;
}

// rows that may still be added (gvss_add's lock-free capacity test)
static void recompute_room(gvss_index *x) {
    x->room = (int64_t)x->capacity + (int64_t)x->free_slots.size() - (int64_t)x->count - (int64_t)x->staged.load();
}

static void publish_info(gvss_index *x) {
    // order matters to gvss_get_info's lock-free size(): deleted, then count, then the in-flight rows are cleared
    x->pub_capacity = x->capacity; x->pub_deleted = x->n_deleted; x->pub_count = x->count; x->pub_inflight = 0;
    x->pub_upper_cap = x->upper_cap; x->pub_scratch = x->scratch_bytes; x->pub_max_level = x->max_level;
}
struct Publish { // declared right after a writer lock: runs before the lock is released
    gvss_index *x;
    ~Publish() { publish_info(x); }
};

static int ensure_capacity(gvss_index *x, uint64_t members) {
    if (members <= x->capacity) return 0;
    if (members >= 0x7FFFFFF0ull) return fail("gvss: more than 2^31-16 members are not addressable (u32 slots, top bit reserved)");
    const uint64_t oldc = x->capacity;
    CU(dev_resize(x->d_vectors, oldc * x->dpad, members * x->dpad, 0, x->stream));
    CU(dev_resize(x->d_keys, oldc, members, 0, x->stream));
    CU(dev_resize(x->d_levels, oldc, members, 0, x->stream));
    CU(dev_resize(x->d_nbr0, oldc * x->stride0, members * x->stride0, 0xFF, x->stream));
    CU(dev_resize(x->d_upper_off, oldc, members, 0xFF, x->stream));
    CU(dev_resize(x->d_locks, oldc, members, 0, x->stream));
    x->capacity = members;
    x->room += (int64_t)(members - oldc);
    x->h_keys.reserve(members); x->h_levels.reserve(members); x->h_upper_off.reserve(members);
    return 0;
}
static int ensure_upper(gvss_index *x, uint64_t blocks) {
    if (blocks <= x->upper_cap) return 0;
    uint64_t want = std::max<uint64_t>(blocks, x->upper_cap * 2);
    want = std::max<uint64_t>(want, 1024);
    CU(dev_resize(x->d_upper, x->upper_cap * x->strideU, want * x->strideU, 0xFF, x->stream));
    x->upper_cap = want;
    return 0;
}
static int ensure_bytes(uint8_t *&p, size_t &have, size_t bytes) {
    if (bytes <= have) return 0;
    cudaFree(p); p = nullptr; have = 0;
    CU(cudaMalloc(&p, bytes));
    have = bytes;
    return 0;
}
static int ensure_scratch(gvss_index *x, size_t bytes) { return ensure_bytes(x->d_scratch, x->scratch_bytes, bytes); }

// choose_random_level_ (index.hpp:3315-3319) over libstdc++'s default_random_engine + uniform_real_distribution<double>
static int16_t draw_level(uint64_t &state, uint32_t M) {
    const double R = 2147483646.0;
    double sum = 0, mul = 1;
    for (int i = 0; i < 2; i++) {
        state = state * 16807ull % 2147483647ull;
        sum += double((uint32_t)state - 1u) * mul;
        mul *= R;
    }
    double u = sum / mul;
    if (u >= 1.0) u = std::nextafter(1.0, 0.0);
    return (int16_t)(-std::log(u) * (1.0 / std::log((double)M)));
}

static int flush_locked(gvss_index *x);
static int compact_locked(gvss_index *x);

// GVSS_BUILD_PROF=1: milliseconds since the library was first asked, for the host-side timeline printed on stderr
static const bool g_build_prof = [] { const char *e = getenv("GVSS_BUILD_PROF"); return e && atoi(e); }();
static double prof_ms() {
    static const auto t_origin = std::chrono::steady_clock::now();
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_origin).count();
}

// frees a temporary device allocation on every exit path (the CU() macro returns early)
struct TempDev {
    void *p = nullptr;
    ~TempDev() { if (p) cudaFree(p); }
    template <typename T> cudaError_t alloc(T *&out, size_t n) {
        const cudaError_t e = cudaMalloc(&p, n ? n * sizeof(T) : sizeof(T));
        out = static_cast<T *>(p);
        return e;
    }
};

// two pinned bounce buffers with an event each (persistence: copies overlap the stream callback)
struct PinnedPair {
    void *buf[2] = {nullptr, nullptr};
    cudaEvent_t ev[2] = {nullptr, nullptr};
    ~PinnedPair() {
        for (int i = 0; i < 2; i++) { if (buf[i]) cudaFreeHost(buf[i]); if (ev[i]) cudaEventDestroy(ev[i]); }
    }
    cudaError_t alloc(size_t bytes) {
        for (int i = 0; i < 2; i++) {
            cudaError_t e = cudaMallocHost(&buf[i], bytes ? bytes : 1);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming);
            if (e != cudaSuccess) return e;
        }
        return cudaSuccess;
    }
};

// heap entries held in shared memory on the search path (GVSS_HEAP_SMEM overrides default_heap_smem for experiments)
// With the list-as-queue search (no removed members) the heap only serves the rare exact-tie replay and lives in global
// scratch entirely: the 2 KB it would take per query buy more resident queries.
static bool list_queue();
static uint32_t search_heap_smem(const gvss_index *x, uint32_t limit) {
    static int v = [] { const char *e = getenv("GVSS_HEAP_SMEM"); return e ? atoi(e) : -1; }();
    if (v >= 0) return (uint32_t)v / 32u * 32u;
    (void)x;
    return list_queue() ? 32u : default_heap_smem(limit); // 32 entries = the 5 top levels, fits the slack
}
// rows of the staging buffer on the search path: 16 halves the shared memory per query (more resident warps); 32 serves
// a whole 32-neighbour batch in one go. GVSS_STAGE_ROWS overrides for experiments.
static uint32_t search_stage_rows() {
    static uint32_t rows = [] {
        const char *e = getenv("GVSS_STAGE_ROWS");
        const uint32_t v = e ? (uint32_t)atoi(e) : 16u;
        return (v == 32u || v == 16u || v == 8u) ? v : 16u;
    }();
    return rows;
}
// floats of a row staged per step on the search path (GVSS_STAGE_CHUNK: 32/64/128)
static uint32_t search_stage_chunk() {
    static uint32_t v = [] { const char *e = getenv("GVSS_STAGE_CHUNK"); const uint32_t c = e ? (uint32_t)atoi(e) : 64u; return (c == 32u || c == 64u || c == 128u) ? c : 64u; }();
    return v;
}
static bool list_queue() { // GVSS_LIST_QUEUE=0: k_search always replays usearch's heap (A/B, debugging)
    static bool on = [] { const char *e = getenv("GVSS_LIST_QUEUE"); return e ? atoi(e) != 0 : true; }();
    return on;
}
// query-to-node distances: 0 = rows staged in shared memory by TMA (default: faster once the base set exceeds L2,
// profiles/r1_search_ncu_summary.txt), 1 = rows streamed straight into registers (256-bit loads, no staging; A/B builds only)
static bool direct_rows() {
    static bool on = [] { const char *e = getenv("GVSS_DIRECT_ROWS"); return GVSS_DIRECT_PATH && e ? atoi(e) != 0 : false; }();
    return on;
}
// row staging engine: 1 = TMA bulk copies (one UBLKCP per row chunk), 0 = coalesced 16-byte cp.async (GVSS_USE_TMA)
static bool stage_with_tma() {
    static bool on = [] { const char *e = getenv("GVSS_USE_TMA"); return e ? atoi(e) != 0 : true; }();
    return on;
}
// prefetch every listed neighbour's row to L2 before the visited test resolves (GVSS_SPEC_PREFETCH=0/1)
static bool spec_prefetch() {
    static bool on = [] { const char *e = getenv("GVSS_SPEC_PREFETCH"); return e ? atoi(e) != 0 : false; }();
    return on;
}

// ------------------------------------------------------------------------------------------------------------
// search
// ------------------------------------------------------------------------------------------------------------
static bool gather4_rows() { // GVSS_GATHER4=0: stage rows with one 1-D bulk copy per row chunk instead of TMA gather4
    static bool on = [] { const char *e = getenv("GVSS_GATHER4"); return e ? atoi(e) != 0 : true; }();
    return on;
}
static int search_mode_bits() {
    static const bool pf = [] { const char *e = getenv("GVSS_PREFETCH_ROWS"); return e ? atoi(e) != 0 : true; }();
    static const bool pv = [] { const char *e = getenv("GVSS_PREFETCH_VIS"); return e ? atoi(e) != 0 : true; }();
    return (direct_rows() ? 1 : 0) | (spec_prefetch() ? 2 : 0) | (stage_with_tma() ? 0 : 4) | (list_queue() ? 0 : 8) | (pf ? 0 : 32) | (pv ? 0 : 64);
}
// gather4 staging needs rows of at least one 128-byte line and a tensor map over the vector array
static bool setup_gather4(const gvss_index *x, SearchParams &p) {
    if (!gather4_rows() || direct_rows() || !stage_with_tma() || x->dpad < 32 || !x->capacity) return false;
    if (!encode_row_gather_map(&p.vmap, x->d_vectors, x->capacity, x->dpad)) return false;
    p.direct |= 16;
    return true;
}

// Everything of a search launch that does not depend on where its scratch lives. `grid` = resident query warps wanted.
// for the calls that read or write adjacency lists
static int check_graph(const gvss_index *x, const char *what) {
    if (x->flat) return fail("%s: this index holds rows only (gvss_set_rows_only), it has no graph; gvss_exact_search answers it", what);
    return 0;
}
static int plan_search(const gvss_index *x, const float *d_queries, uint64_t nq, uint32_t k, uint32_t ef, int64_t *d_keys_out,
                       float *d_dist_out, uint32_t *d_counts_out, gvss_search_stats *d_stats_out, SearchParams &p, uint32_t &grid) {
    if (check_graph(x, "gvss_search")) return 1;
    if (!ef) ef = 64;                        // default_expansion_search, index.hpp:1125,2493-2494
    const uint32_t limit = std::max(ef, k);  // index.hpp:2512
    if (limit > 16384) return fail("gvss: ef/k above 16384 not supported");
    p = SearchParams{};
    p.queries = d_queries; p.nq = nq; p.k = k; p.limit = limit;
    p.keys_out = d_keys_out; p.dist_out = d_dist_out; p.counts_out = d_counts_out; p.stats_out = d_stats_out;
    p.entry = x->entry; p.max_level = x->max_level < 0 ? 0 : x->max_level;
    p.trace_query = -1;
    static const long long trace = [] { const char *e = getenv("GVSS_TRACE_QUERY"); return e ? atoll(e) : -1ll; }();
    p.trace_query = trace;
    const uint32_t list_cap = std::max(x->cfg.connectivity_base, x->cfg.connectivity);
    p.sl = make_scratch_layout(limit, std::max<uint64_t>(x->count, x->capacity), list_cap); // by capacity: inserts between searches do not regrow it
    p.direct = search_mode_bits();
    const bool g4 = setup_gather4(x, p);
    // Latency kernel (GVSS_SEARCH_LATENCY=0 switches it off): when the whole batch fits the GPU at once even with a 16 KB row
    // stage per query, shared memory buys fewer fetch-wait rounds instead of more resident warps (index scans, small joins)
    static const bool latency_on = [] { const char *e = getenv("GVSS_SEARCH_LATENCY"); return e ? atoi(e) != 0 : true; }();
    bool latency = false;
    if (g4 && latency_on && p.direct == 16 && x->dpad > 32) {
        const SmemLayout big = make_smem_layout(x->dpad, std::max<uint32_t>(limit, 1), search_heap_smem(x, limit), 0, 32u, false, list_cap, 0, false, true, false, 4);
        if (nq <= resident_warps(x->device, big.bytes, false)) { p.sm = big; p.direct |= 128; latency = true; }
    }
    if (!latency)
        p.sm = make_smem_layout(x->dpad, std::max<uint32_t>(limit, 1), search_heap_smem(x, limit), 0, direct_rows() ? 0u : search_stage_rows(), false, list_cap, search_stage_chunk(), false, g4);
    grid = resident_warps(x->device, p.sm.bytes, false);
    if (!grid) return fail("gvss: no CUDA device");
    static const uint32_t max_warps = [] { const char *e = getenv("GVSS_SEARCH_MAX_WARPS"); return e ? (uint32_t)atoi(e) : 0u; }();
    if (max_warps && grid > max_warps) grid = max_warps; // experiment knob: fewer resident queries = smaller visited-table footprint
    if (grid > nq) grid = (uint32_t)nq;
    return 0;
}

// gvss_search_device: the launch runs on the CALLER's stream and the call returns before it finishes, so these launches share
// the index-owned scratch and work counter and are ordered on the device through the scratch_free event. Caller holds
// rw (shared) and dev_mu.
static int search_device_enqueue(gvss_index *x, const float *d_queries, uint64_t nq, uint32_t k, uint32_t ef,
                                 int64_t *d_keys_out, float *d_dist_out, uint32_t *d_counts_out,
                                 gvss_search_stats *d_stats_out, cudaStream_t st) {
    if (!nq) return 0;
    SearchParams p;
    uint32_t grid = 0;
    if (plan_search(x, d_queries, nq, k, ef, d_keys_out, d_dist_out, d_counts_out, d_stats_out, p, grid)) return 1;
    if (p.sl.bytes_per_warp * grid > x->scratch_bytes) {
        if (x->scratch_free) CU(cudaEventSynchronize(x->scratch_free)); // an earlier launch may still use the buffer about to be replaced
        if (ensure_scratch(x, p.sl.bytes_per_warp * grid)) return 1;
    }
    p.scratch = x->d_scratch;
    p.work_counter = x->d_work;
    if (!x->scratch_free) CU(cudaEventCreateWithFlags(&x->scratch_free, cudaEventDisableTiming));
    else CU(cudaStreamWaitEvent(st, x->scratch_free, 0));
    CU(cudaMemsetAsync(x->d_work, 0, sizeof(unsigned long long), st));
    GraphView g = view_of(x);
    CU(launch_search(metric_id(x->cfg.metric), g, p, grid, st));
    CU(cudaEventRecord(x->scratch_free, st));
    return 0;
}

// exclusive operations wait for searches that gvss_search_device left running on callers' streams
static void wait_device_searches(gvss_index *x) {
    if (x->scratch_free) cudaEventSynchronize(x->scratch_free);
}

constexpr size_t kMaxSearchSlots = 32;
constexpr size_t kSlotScratchKeep = 64u << 20; // slots other than the first give larger scratch buffers back when they finish

static SearchSlot *acquire_slot(gvss_index *x) {
    std::unique_lock<std::mutex> lk(x->slot_mu);
    auto took = [&](SearchSlot *s) {
        s->busy = true;
        uint32_t busy = 0;
        for (auto &o : x->slots) busy += o->busy ? 1u : 0u;
        x->slots_peak = std::max(x->slots_peak, busy);
        return s;
    };
    for (;;) {
        for (auto &s : x->slots)
            if (!s->busy) return took(s.get());
        if (x->slots.size() < kMaxSearchSlots) {
            auto s = std::make_unique<SearchSlot>();
            if (cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) != cudaSuccess) {
                fail("gvss_search: cannot create a stream: %s", cudaGetErrorString(cudaGetLastError()));
                return nullptr;
            }
            x->slots.push_back(std::move(s));
            return took(x->slots.back().get());
        }
        x->slot_cv.wait(lk);
    }
}
static void release_slot(gvss_index *x, SearchSlot *s) {
    std::lock_guard<std::mutex> lk(x->slot_mu);
    if (s != x->slots.front().get() && s->scratch_bytes > kSlotScratchKeep) {
        cudaFree(s->d_scratch); s->d_scratch = nullptr; s->scratch_bytes = 0;
    }
    s->busy = false;
    x->slot_cv.notify_one();
}
static void free_slots_all(gvss_index *x) {
    for (auto &s : x->slots) {
        cudaFree(s->d_scratch); cudaFree(s->d_io);
        if (s->h_io) cudaFreeHost(s->h_io);
        if (s->stream) cudaStreamDestroy(s->stream);
    }
    x->slots.clear();
}
struct SlotGuard {
    gvss_index *x; SearchSlot *s;
    ~SlotGuard() { if (s) release_slot(x, s); }
};

static int check_usable(const gvss_index *x) {
    if (!x) return fail("null index");
    if (!x->poisoned.empty()) return fail("gvss: index unusable after an earlier failure: %s", x->poisoned.c_str());
    return 0;
}

// Inserts staged rows if there are any. Takes rw exclusively for that; returns with no lock held.
static int flush_if_staged(gvss_index *x) {
    if (!x->staged.load()) return 0;
    if (g_build_prof) fprintf(stderr, "[gvss flush_if_staged] at %.1f ms: %llu rows staged\n", prof_ms(), (unsigned long long)x->staged.load());
    std::unique_lock<std::shared_mutex> lk(x->rw);
    Publish pub{x};
    return flush_locked(x);
}

// graph state of two indexes changes places (build-then-swap of gvss_load and gvss_compact_rebuild); streams, locks, slots,
// staging and scratch stay with their owner
static void swap_graph_state(gvss_index *a, gvss_index *b) {
    std::swap(a->cfg.metric, b->cfg.metric);
    std::swap(a->cfg.connectivity, b->cfg.connectivity);
    std::swap(a->cfg.connectivity_base, b->cfg.connectivity_base);
    std::swap(a->stride0, b->stride0); std::swap(a->strideU, b->strideU);
    std::swap(a->d_vectors, b->d_vectors); std::swap(a->d_keys, b->d_keys); std::swap(a->d_levels, b->d_levels);
    std::swap(a->d_nbr0, b->d_nbr0); std::swap(a->d_upper, b->d_upper); std::swap(a->d_upper_off, b->d_upper_off);
    std::swap(a->d_locks, b->d_locks); std::swap(a->d_build_stats, b->d_build_stats);
    std::swap(a->capacity, b->capacity); std::swap(a->count, b->count);
    std::swap(a->upper_cap, b->upper_cap); std::swap(a->upper_used, b->upper_used);
    a->h_keys.swap(b->h_keys); a->h_levels.swap(b->h_levels); a->h_upper_off.swap(b->h_upper_off);
    std::swap(a->entry, b->entry); std::swap(a->max_level, b->max_level); std::swap(a->n_deleted, b->n_deleted);
    std::swap(a->lookup_valid, b->lookup_valid); a->key_to_slot.swap(b->key_to_slot); a->free_slots.swap(b->free_slots);
    std::swap(a->level_rng, b->level_rng);
}

static void clear_staging(gvss_index *x) {
    for (auto &st : x->stages) {
        std::lock_guard<std::mutex> lk(st->mu);
        st->vecs.clear(); st->keys.clear();
    }
    x->st_device.clear();
    x->staged = 0;
}

// the body of the background builder. It sleeps until `bg_min_rows` rows are staged, then takes the writer lock
// and inserts whatever is staged at that moment — upload and graph construction overlap the caller's table scan. Adding
// never waits for it (gvss_add only touches its staging buffer); searches wait for the chunk in flight like for any flush.
static void background_builder(gvss_index *x) {
    cudaSetDevice(x->device);
    for (;;) {
        {
            std::unique_lock<std::mutex> lk(x->bg_mu);
            x->bg_cv.wait(lk, [&] { return x->bg_stop || (x->bg_min_rows.load() && x->staged.load() >= x->bg_min_rows.load()); });
            if (x->bg_stop) return;
        }
        std::unique_lock<std::shared_mutex> lk(x->rw);
        Publish pub{x};
        if (!x->poisoned.empty()) { x->bg_min_rows = 0; continue; } // nothing more to do for a broken index: go back to sleep
        if (!x->bg_min_rows.load() || x->staged.load() < x->bg_min_rows.load()) continue;
        if (flush_locked(x)) {
            std::lock_guard<std::mutex> el(x->bg_mu);
            if (x->bg_error.empty()) x->bg_error = t_error;
            x->bg_min_rows = 0; // stop trying; the foreground flush reports the error
        }
    }
}
static void stop_background_builder(gvss_index *x) {
    if (!x->bg.joinable()) return;
    {
        std::lock_guard<std::mutex> lk(x->bg_mu);
        x->bg_stop = true;
    }
    x->bg_cv.notify_all();
    x->bg.join();
    x->bg_stop = false;
}
static inline void poke_background_builder(gvss_index *x) {
    const uint64_t m = x->bg_min_rows.load(std::memory_order_relaxed);
    if (m && x->staged.load(std::memory_order_relaxed) >= m) x->bg_cv.notify_one();
}

} // namespace gvss

// every ABI body runs inside this pair: nothing may throw across the C boundary (std::bad_alloc, std::length_error ...)
#define API_BEGIN try {
#define API_END                                                                                                       \
    }                                                                                                                 \
    catch (const std::exception &e_) { return fail("gvss: %s", e_.what()); }                                         \
    catch (...) { return fail("gvss: unknown C++ exception"); }

// ------------------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------------------
extern "C" {

const char *gvss_last_error(void) { return t_error.c_str(); }
const char *gvss_version(void) { return "gvss 0.2 (sm_100a; usearch v2.12 stream)"; }
uint64_t gvss_kernel_launches(void) { return g_launches.load(); }

int gvss_warmup(int32_t device) {
    API_BEGIN
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail("gvss: no CUDA device available (this library has no CPU fallback)");
    if (device >= ndev) return fail("gvss_warmup: device %d out of range", device);
    const double t0 = prof_ms();
    if (device >= 0) CU(cudaSetDevice(device));
    CU(cudaFree(nullptr)); // creates the primary context: the second or so a first CUDA call costs on a B200
    if (g_build_prof) fprintf(stderr, "[gvss warmup] device %d: context ready after %.1f ms (started at %.1f ms)\n", device, prof_ms() - t0, t0);
    return 0;
    API_END
}

int gvss_create(const gvss_config *config, gvss_index **out) {
    API_BEGIN
    if (!config || !out) return fail("gvss_create: null argument");
    if (!config->dimensions) return fail("gvss_create: dimensions must be > 0");
    if (g_build_prof) fprintf(stderr, "[gvss create] at %.1f ms\n", prof_ms());
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail("gvss: no CUDA device available (this library has no CPU fallback)");
    auto x = std::make_unique<gvss_index>();
    x->cfg = *config;
    if (!x->cfg.metric) x->cfg.metric = GVSS_METRIC_L2SQ;
    if (x->cfg.metric != GVSS_METRIC_L2SQ && x->cfg.metric != GVSS_METRIC_COS && x->cfg.metric != GVSS_METRIC_IP)
        return fail("gvss_create: unknown metric '%c'", config->metric);
    if (!x->cfg.connectivity) x->cfg.connectivity = 16;                              // index.hpp:1097-1107
    if (!x->cfg.connectivity_base) x->cfg.connectivity_base = 2 * x->cfg.connectivity; // hnsw_index.cpp:178-187
    if (!x->cfg.expansion_add) x->cfg.expansion_add = 128;
    if (!x->cfg.expansion_search) x->cfg.expansion_search = 64;
    if (x->cfg.connectivity < 2 || x->cfg.connectivity > 4096 || x->cfg.connectivity_base < 2 || x->cfg.connectivity_base > 4096)
        return fail("gvss_create: connectivity %u / %u outside 2..4096", x->cfg.connectivity, x->cfg.connectivity_base);
    if (x->cfg.device < 0) cudaGetDevice(&x->device); else x->device = x->cfg.device;
    if (x->device >= ndev) return fail("gvss_create: device %d out of range", x->device);
    cudaSetDevice(x->device);
    x->dpad = (x->cfg.dimensions + 7) / 8 * 8; // 32-byte rows: TMA bulk copies need 16 B, the 256-bit row loads 32 B
    x->stride0 = (x->cfg.connectivity_base + 7) / 8 * 8;
    x->strideU = (x->cfg.connectivity + 7) / 8 * 8;
    if (cudaStreamCreateWithFlags(&x->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaMalloc(&x->d_entry_state, 8) != cudaSuccess || cudaMalloc(&x->d_work, 8) != cudaSuccess ||
        cudaMalloc(&x->d_build_stats, 48) != cudaSuccess || cudaMemset(x->d_build_stats, 0, 48) != cudaSuccess) {
        const char *why = cudaGetErrorString(cudaGetLastError());
        cudaFree(x->d_entry_state); cudaFree(x->d_work); cudaFree(x->d_build_stats);
        if (x->stream) cudaStreamDestroy(x->stream);
        return fail("gvss_create: CUDA initialisation failed: %s", why);
    }
    for (size_t i = 0; i < 1 + gvss_index::kAnyStages; i++) x->stages.push_back(std::make_unique<gvss_index::Stage>()); // thread 0 + GVSS_ANY_THREAD
    *out = x.release();
    return 0;
    API_END
}

int gvss_create_sharded(const gvss_config *config, const int32_t *devices, uint32_t n_devices, gvss_index **out) {
    API_BEGIN
    if (!config || !devices || !out) return fail("gvss_create_sharded: null argument");
    Router *r = nullptr;
    if (router_create(config, devices, n_devices, &r)) return 1;
    auto *x = new gvss_index();
    x->cfg = *config;
    x->router = r;
    *out = x;
    return 0;
    API_END
}

void gvss_destroy(gvss_index *x) {
    if (!x) return;
    if (x->router) { router_destroy(x->router); delete x; return; }
    stop_background_builder(x);
    cudaSetDevice(x->device);
    wait_device_searches(x);
    free_device(x);
    free_slots_all(x);
    cudaFree(x->d_entry_state); cudaFree(x->d_work); cudaFree(x->d_build_stats);
    if (x->scratch_free) cudaEventDestroy(x->scratch_free);
    if (x->stream) cudaStreamDestroy(x->stream);
    delete x;
}

static void reset_locked(gvss_index *x) {
    wait_device_searches(x);
    free_device(x);
    x->h_keys.clear(); x->h_levels.clear(); x->h_upper_off.clear();
    x->entry = 0; x->max_level = -1; x->n_deleted = 0; x->lookup_valid = false; x->key_to_slot.clear(); x->free_slots.clear();
    cudaMemset(x->d_build_stats, 0, 48);
    clear_staging(x);
    x->level_rng = 1;
    x->poisoned.clear();
    recompute_room(x);
}
int gvss_reset(gvss_index *x) {
    API_BEGIN
    if (!x) return fail("null index");
    if (x->router) return router_reset(x->router);
    std::unique_lock<std::shared_mutex> lk(x->rw);
    Publish pub{x};
    cudaSetDevice(x->device);
    reset_locked(x);
    return 0;
    API_END
}

// rows of the live members, densely packed (row stride = dims), for gvss_compact_rebuild
__global__ void k_pack_rows(const float *__restrict__ vectors, uint32_t dpad, uint32_t dims, const uint32_t *__restrict__ slots, uint64_t m,
                            float *__restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m * dims) return;
    const uint64_t r = i / dims;
    out[i] = vectors[(size_t)slots[r] * dpad + (uint32_t)(i - r * dims)];
}

int gvss_compact_rebuild(gvss_index *x, uint64_t *dropped) {
    API_BEGIN
    if (check_usable(x)) return 1;
    if (check_graph(x, "gvss_compact_rebuild")) return 1;
    if (x->router) return router_compact(x->router, true, dropped);
    std::unique_lock<std::shared_mutex> lk(x->rw);
    Publish pub{x};
    cudaSetDevice(x->device);
    if (flush_locked(x)) return 1;
    wait_device_searches(x);
    if (dropped) *dropped = x->n_deleted;
    if (!x->n_deleted) return 0; // nothing to drop: the graph stays as it is
    std::vector<uint32_t> live;
    std::vector<int64_t> keys;
    live.reserve(x->count); keys.reserve(x->count);
    for (uint64_t s = 0; s < x->count; s++)
        if (x->h_keys[s] != INT64_MAX) { live.push_back((uint32_t)s); keys.push_back(x->h_keys[s]); }
    const uint64_t m = live.size(), cap = x->capacity;
    const uint32_t d = x->cfg.dimensions;
    TempDev rows_mem, slots_mem;
    float *d_rows = nullptr;
    uint32_t *d_live = nullptr;
    if (m) {
        CU(rows_mem.alloc(d_rows, m * d));
        CU(slots_mem.alloc(d_live, m));
        CU(cudaMemcpyAsync(d_live, live.data(), m * 4, cudaMemcpyHostToDevice, x->stream));
        k_pack_rows<<<(unsigned)((m * d + 255) / 256), 256, 0, x->stream>>>(x->d_vectors, x->dpad, d, d_live, m, d_rows);
        count_launch();
        CU(cudaGetLastError());
        CU(cudaStreamSynchronize(x->stream));
    }
    // The new graph is built in an index of its own and changes places with the old one only when it is complete: a failure
    // on the way (out of memory, most likely — both graphs exist for a moment) leaves the index as it was.
    gvss_config cfg = x->cfg;
    cfg.device = x->device;
    gvss_index *y = nullptr;
    if (gvss_create(&cfg, &y)) return 1;
    struct Drop { gvss_index *p; ~Drop() { gvss_destroy(p); } } drop{y};
    y->build_conc = x->build_conc;
    if (ensure_capacity(y, cap)) return 1; // same capacity as before
    if (m) {
        y->st_device.emplace_back(d_rows, std::move(keys));
        y->staged += m;
        y->room -= (int64_t)m;
        if (flush_locked(y)) return 1;     // bulk build of the live members only
    }
    swap_graph_state(x, y);
    recompute_room(x);
    return 0;
    API_END
}

int gvss_reserve(gvss_index *x, uint64_t members, uint32_t host_threads) {
    API_BEGIN
    if (check_usable(x)) return 1;
    if (x->router) return router_reserve(x->router, members, host_threads);
    std::unique_lock<std::shared_mutex> lk(x->rw);
    Publish pub{x};
    cudaSetDevice(x->device);
    if (host_threads == 0) host_threads = 1;
    if (host_threads > 65536) return fail("gvss_reserve: %u host threads", host_threads);
    if (host_threads > x->host_threads) {
        // thread ids 0..host_threads-1 become valid; the GVSS_ANY_THREAD buffer stays last
        std::vector<std::unique_ptr<gvss_index::Stage>> any; // the GVSS_ANY_THREAD buffers stay last
        for (size_t i = 0; i < gvss_index::kAnyStages; i++) { any.push_back(std::move(x->stages.back())); x->stages.pop_back(); }
        while (x->stages.size() < host_threads) x->stages.push_back(std::make_unique<gvss_index::Stage>());
        while (!any.empty()) { x->stages.push_back(std::move(any.back())); any.pop_back(); }
        x->host_threads = host_threads;
    }
    x->level_rng = 1; // reserve() allocates fresh thread contexts => the level sequence restarts (index.hpp:2113,2124)
    wait_device_searches(x);
    if (g_build_prof) fprintf(stderr, "[gvss reserve] at %.1f ms: %llu members (capacity %llu), %u threads\n", prof_ms(), (unsigned long long)members, (unsigned long long)x->capacity, host_threads);
    return ensure_capacity(x, members);
    API_END
}

int gvss_add(gvss_index *x, int64_t key, const float *vec, uint32_t host_thread) {
    API_BEGIN
    if (!x || !vec) return fail("gvss_add: null argument");
    if (x->router) return router_add(x->router, key, vec, host_thread);
    if (!x->poisoned.empty()) return check_usable(x);
    size_t t;
    if (host_thread == GVSS_ANY_THREAD) t = x->stages.size() - gvss_index::kAnyStages + std::hash<std::thread::id>()(std::this_thread::get_id()) % gvss_index::kAnyStages;
    else if (host_thread < x->host_threads) t = host_thread;
    else return fail("gvss_add: thread id %u is not below the %u reserved by gvss_reserve (or GVSS_ANY_THREAD)", host_thread, x->host_threads);
    // capacity check as usearch: nodes_count_.fetch_add(1) >= capacity fails (index.hpp:2329-2335); removed slots are reusable
    if (x->room.fetch_sub(1) < 1) {
        x->room.fetch_add(1);
        return fail("Reserve capacity ahead of insertions!");
    }
    gvss_index::Stage &st = *x->stages[t];
    std::lock_guard<std::mutex> lk(st.mu);
    st.vecs.insert(st.vecs.end(), vec, vec + x->cfg.dimensions);
    st.keys.push_back(key);
    x->staged.fetch_add(1); // inside the lock: whoever takes the row away has seen it counted
    poke_background_builder(x);
    return 0;
    API_END
}

int gvss_add_batch(gvss_index *x, const int64_t *keys, const float *vecs, uint64_t n) {
    API_BEGIN
    if (!x || (!vecs && n)) return fail("gvss_add_batch: null argument");
    if (check_usable(x)) return 1;
    if (x->router) return router_add_batch(x->router, keys, vecs, n);
    // default keys number the rows in arrival order: no flush may run between reading the node count and the append
    std::shared_lock<std::shared_mutex> lk(x->rw, std::defer_lock);
    if (!keys) lk.lock();
    if (x->room.fetch_sub((int64_t)n) < (int64_t)n) {
        x->room.fetch_add((int64_t)n);
        return fail("Reserve capacity ahead of insertions!");
    }
    gvss_index::Stage &st = *x->stages.front(); // same buffer as thread 0: call order is insertion order
    {
        std::lock_guard<std::mutex> sl(st.mu);
        st.vecs.insert(st.vecs.end(), vecs, vecs + n * x->cfg.dimensions);
        const int64_t first = keys ? 0 : (int64_t)(x->count + x->staged.load());
        for (uint64_t i = 0; i < n; i++) st.keys.push_back(keys ? keys[i] : first + (int64_t)i);
        x->staged += n;
    }
    poke_background_builder(x);
    return 0;
    API_END
}

int gvss_add_batch_device(gvss_index *x, const int64_t *keys_host, const float *vecs_device, uint64_t n) {
    API_BEGIN
    if (!x || (!vecs_device && n)) return fail("gvss_add_batch_device: null argument");
    if (check_usable(x)) return 1;
    if (x->router) return fail("gvss_add_batch_device: not available on a sharded index (the rows live on one GPU): pass host rows to gvss_add_batch");
    std::unique_lock<std::shared_mutex> lk(x->rw);
    Publish pub{x};
    if (x->room.fetch_sub((int64_t)n) < (int64_t)n) {
        x->room.fetch_add((int64_t)n);
        return fail("Reserve capacity ahead of insertions!");
    }
    std::vector<int64_t> k(n);
    const int64_t first = (int64_t)(x->count + x->staged.load());
    for (uint64_t i = 0; i < n; i++) k[i] = keys_host ? keys_host[i] : first + (int64_t)i;
    x->st_device.emplace_back(vecs_device, std::move(k));
    x->staged += n;
    return 0;
    API_END
}

int gvss_flush(gvss_index *x) {
    API_BEGIN
    if (check_usable(x)) return 1;
    if (x->router) return router_flush(x->router);
    {
        std::lock_guard<std::mutex> el(x->bg_mu);
        if (!x->bg_error.empty()) {
            const std::string why = x->bg_error;
            x->bg_error.clear();
            return fail("background build failed: %s", why.c_str());
        }
    }
    if (g_build_prof) fprintf(stderr, "[gvss_flush] called at %.1f ms\n", prof_ms());
    std::unique_lock<std::shared_mutex> lk(x->rw);
    Publish pub{x};
    cudaSetDevice(x->device);
    const int rc = flush_locked(x);
    if (g_build_prof) fprintf(stderr, "[gvss_flush] returns at %.1f ms\n", prof_ms());
    return rc;
    API_END
}

int gvss_set_background_build(gvss_index *x, uint64_t min_rows) {
    API_BEGIN
    if (check_usable(x)) return 1;
    if (x->router) return router_set_background_build(x->router, min_rows);
    x->bg_min_rows = min_rows;
    if (min_rows && !x->bg.joinable()) x->bg = std::thread(background_builder, x);
    if (!min_rows) stop_background_builder(x);
    else poke_background_builder(x);
    return 0;
    API_END
}

int gvss_set_rows_only(gvss_index *x, int on) {
    API_BEGIN
    if (check_usable(x)) return 1;
    if (x->router) return fail("gvss_set_rows_only: not available on a sharded index");
    std::unique_lock<std::shared_mutex> lk(x->rw);
    Publish pub{x};
    if (x->count || x->staged.load()) return fail("gvss_set_rows_only: the index already has members; choose before the first add");
    x->flat = on != 0;
    return 0;
    API_END
}

int gvss_set_build_concurrency(gvss_index *x, uint32_t n) {
    API_BEGIN
    if (!x) return fail("null index");
    if (x->router) return router_set_build_concurrency(x->router, n);
    std::unique_lock<std::shared_mutex> lk(x->rw);
    Publish pub{x};
    x->build_conc = n;
    return 0;
    API_END
}

int gvss_remove(gvss_index *x, int64_t key, uint64_t *removed) {
    API_BEGIN
    if (check_usable(x)) return 1;
    if (x->router) return router_remove(x->router, key, removed);
    std::unique_lock<std::shared_mutex> lk(x->rw);
    Publish pub{x};
    cudaSetDevice(x->device);
    if (flush_locked(x)) return 1;
    wait_device_searches(x);
    if (removed) *removed = 0;
    if (key == GVSS_FREE_KEY) return 0;
    if (!x->lookup_valid) { // rebuilt on demand (the reference never rebuilds it after a load, SURVEY App. A — strictly better here)
        x->key_to_slot.clear();
        for (uint32_t s = 0; s < x->count; s++) if (x->h_keys[s] != GVSS_FREE_KEY) x->key_to_slot.emplace(x->h_keys[s], s);
        x->lookup_valid = true;
    }
    auto range = x->key_to_slot.equal_range(key);
    uint64_t n = 0;
    const int64_t freek = GVSS_FREE_KEY;
    for (auto it = range.first; it != range.second; ++it) {
        const uint32_t s = it->second;
        x->h_keys[s] = freek;
        CU(cudaMemcpyAsync(x->d_keys + s, &freek, 8, cudaMemcpyHostToDevice, x->stream));
        x->free_slots.push_back(s);
        n++;
    }
    CU(cudaStreamSynchronize(x->stream));
    x->key_to_slot.erase(key);
    x->n_deleted += n;
    x->room += (int64_t)n;
    if (removed) *removed = n;
    return 0;
    API_END
}

int gvss_compact(gvss_index *x) {
    API_BEGIN
    if (check_usable(x)) return 1;
    if (check_graph(x, "gvss_compact")) return 1;
    if (x->router) return router_compact(x->router, false, nullptr);
    std::unique_lock<std::shared_mutex> lk(x->rw);
    Publish pub{x};
    cudaSetDevice(x->device);
    if (flush_locked(x)) return 1;
    wait_device_searches(x);
    return compact_locked(x);
    API_END
}

int gvss_search_device(gvss_index *x, const float *queries, uint64_t nq, uint32_t k, uint32_t ef, int64_t *keys_out,
                       float *dist_out, uint32_t *counts_out, gvss_search_stats *stats_out, void *cuda_stream) {
    API_BEGIN
    if (check_usable(x)) return 1;
    if (x->router) return fail("gvss_search_device: not available on a sharded index (one device pointer cannot name every GPU): use gvss_search");
    if (nq && (!queries || !keys_out)) return fail("gvss_search_device: null buffer");
    cudaSetDevice(x->device);
    if (flush_if_staged(x)) return 1;
    std::shared_lock<std::shared_mutex> lk(x->rw);
    if (!k) { if (counts_out) CU(cudaMemsetAsync(counts_out, 0, nq * 4, (cudaStream_t)cuda_stream)); return 0; } // index.hpp:2489-2490
    std::lock_guard<std::mutex> dl(x->dev_mu);
    return search_device_enqueue(x, queries, nq, k, ef, keys_out, dist_out, counts_out, stats_out, (cudaStream_t)cuda_stream);
    API_END
}

int gvss_search_device_noalloc(gvss_index *x, const float *queries, uint64_t nq, uint32_t k, uint32_t ef, int64_t *keys_out,
                               float *dist_out, uint32_t *counts_out, gvss_search_stats *stats_out, void *cuda_stream) {
    API_BEGIN
    if (check_usable(x)) return 1;
    if (x->router) return fail("gvss_search_device_noalloc: not available on a sharded index");
    if (nq && (!queries || !keys_out)) return fail("gvss_search_device_noalloc: null buffer");
    std::shared_lock<std::shared_mutex> lk(x->rw);
    cudaSetDevice(x->device);
    // inserting staged rows allocates; this entry point never does
    if (x->staged.load()) return fail("gvss_search_device_noalloc: rows are staged; call gvss_flush first (it allocates, this call must not)");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    if (!nq) return 0;
    if (!k) { if (counts_out) CU(cudaMemsetAsync(counts_out, 0, nq * 4, st)); return 0; }
    SearchParams p;
    uint32_t resident = 0;
    if (plan_search(x, queries, nq, k, ef, keys_out, dist_out, counts_out, stats_out, p, resident)) return 1;
    // the built-in scratch is shared out in fixed slots: size them by the members there are, not by the reserved capacity
    p.sl = make_scratch_layout(p.limit, x->count, std::max(x->cfg.connectivity_base, x->cfg.connectivity));
    char err[384];
    err[0] = 0;
    if (launch_search_static_scratch(metric_id(x->cfg.metric), x->device, view_of(x), p, resident, st, err, sizeof err) != cudaSuccess)
        return fail("%s", err);
    return 0;
    API_END
}

} // extern "C"
namespace gvss {
// Host-pointer search, first half: takes a slot of `x` (stream + scratch + I/O buffers), uploads the queries and launches
// k_search on the slot's stream. Results stay in the slot's device buffer: [work words | counts | keys | distances | stats].
// The caller holds x->rw shared until host_search_release.
int host_search_launch(gvss_index *x, const float *queries, uint64_t nq, uint32_t k, uint32_t ef, bool want_stats, HostSearch &hs) {
    hs = HostSearch{};
    hs.x = x;
    cudaSetDevice(x->device);
    if (flush_if_staged(x)) return 1;
    hs.lock = std::shared_lock<std::shared_mutex>(x->rw);
    SearchSlot *s = acquire_slot(x);
    if (!s) return 1;
    hs.slot = s;
    hs.stream = s->stream;
    const uint32_t d = x->cfg.dimensions;
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };
    hs.q_bytes = up(nq * d * 4ull);
    hs.o_counts = 256; hs.o_keys = hs.o_counts + up(nq * 4); hs.o_dist = hs.o_keys + up(nq * k * 8ull);
    hs.o_stats = hs.o_dist + up(nq * k * 4ull);
    hs.r_bytes = hs.o_stats + (want_stats ? up(nq * sizeof(gvss_search_stats)) : 0);
    if (ensure_bytes(s->d_io, s->io_bytes, hs.q_bytes + hs.r_bytes)) return 1;
    if (hs.r_bytes > s->h_io_bytes) {
        if (s->h_io) cudaFreeHost(s->h_io);
        s->h_io = nullptr; s->h_io_bytes = 0;
        CU(cudaMallocHost(&s->h_io, hs.r_bytes));
        s->h_io_bytes = hs.r_bytes;
    }
    hs.res = s->d_io + hs.q_bytes;
    hs.h_res = s->h_io;
    hs.d_keys = reinterpret_cast<int64_t *>(hs.res + hs.o_keys);
    hs.d_dist = reinterpret_cast<float *>(hs.res + hs.o_dist);
    SearchParams p;
    uint32_t grid = 0;
    if (plan_search(x, reinterpret_cast<const float *>(s->d_io), nq, k, ef, hs.d_keys, hs.d_dist,
                    reinterpret_cast<uint32_t *>(hs.res + hs.o_counts),
                    want_stats ? reinterpret_cast<gvss_search_stats *>(hs.res + hs.o_stats) : nullptr, p, grid)) return 1;
    if (ensure_bytes(s->d_scratch, s->scratch_bytes, p.sl.bytes_per_warp * grid)) return 1;
    p.scratch = s->d_scratch;
    p.work_counter = reinterpret_cast<unsigned long long *>(hs.res);
    p.fail_word = p.work_counter + 1;
    CU(cudaMemcpyAsync(s->d_io, queries, nq * d * 4ull, cudaMemcpyHostToDevice, s->stream));
    CU(cudaMemsetAsync(hs.res, 0, 16, s->stream));
    CU(launch_search(metric_id(x->cfg.metric), view_of(x), p, grid, s->stream));
    return 0;
}
// second half: everything packed comes back in ONE copy; returns 1 + the index of a query that overflowed in *failed
int host_search_fetch(HostSearch &hs, unsigned long long *failed) {
    cudaSetDevice(hs.x->device);
    CU(cudaMemcpyAsync(hs.h_res, hs.res, hs.r_bytes, cudaMemcpyDeviceToHost, hs.stream));
    CU(cudaStreamSynchronize(hs.stream));
    *failed = reinterpret_cast<const unsigned long long *>(hs.h_res)[1];
    return 0;
}
void host_search_release(HostSearch &hs) {
    if (hs.slot) release_slot(hs.x, hs.slot);
    hs.slot = nullptr;
    if (hs.lock.owns_lock()) hs.lock.unlock();
}

} // namespace gvss
extern "C" {
// Calls from different host threads run concurrently: each takes a slot and holds rw only shared. One H2D copy (queries),
// one launch, one D2H copy per call.
int gvss_search(gvss_index *x, const float *queries, uint64_t nq, uint32_t k, uint32_t ef, int64_t *keys_out,
                float *dist_out, uint32_t *counts_out, gvss_search_stats *stats_out) {
    API_BEGIN
    if (check_usable(x)) return 1;
    if (nq && (!queries || (!keys_out && k))) return fail("gvss_search: null buffer");
    if (x->router) return router_search(x->router, queries, nq, k, ef, keys_out, dist_out, counts_out, stats_out);
    if (!nq) return flush_if_staged(x);
    if (!k) { if (counts_out) memset(counts_out, 0, nq * 4); return flush_if_staged(x); }
    HostSearch hs;
    struct Rel { HostSearch &h; ~Rel() { host_search_release(h); } } rel{hs};
    if (host_search_launch(x, queries, nq, k, ef, stats_out != nullptr, hs)) return 1;
    unsigned long long failed = 0;
    if (host_search_fetch(hs, &failed)) return 1;
    memcpy(keys_out, hs.h_res + hs.o_keys, nq * k * 8ull);
    if (dist_out) memcpy(dist_out, hs.h_res + hs.o_dist, nq * k * 4ull);
    if (counts_out) memcpy(counts_out, hs.h_res + hs.o_counts, nq * 4);
    if (stats_out) memcpy(stats_out, hs.h_res + hs.o_stats, nq * sizeof(gvss_search_stats));
    // a query whose candidate queue ran out of room was cut short: the synchronous path cannot return it
    if (failed) return fail("gvss_search: candidate queue overflow on query %llu (beam %u)", failed - 1, std::max(ef ? ef : 64u, k));
    return 0;
    API_END
}

// Exact (brute-force) top-k over the rows the index already holds in HBM — no copy of the base set: the comparator of the
// index join / `ORDER BY array_distance LIMIT k` at SQL level. Rows are zero-padded to `dpad` floats in
// the index and the queries are padded the same way: a trailing +0 term changes no bit of any of the three folds.
int gvss_exact_search(gvss_index *x, const float *queries, uint64_t nq, uint32_t k, int64_t *keys_out, float *dist_out, uint32_t *counts_out) {
    API_BEGIN
    if (check_usable(x)) return 1;
    if (nq && (!queries || !keys_out)) return fail("gvss_exact_search: null buffer");
    if (x->router) return fail("gvss_exact_search: not available on a sharded index");
    if (!k || k > 1024) return fail("gvss_exact_search: k must be in 1..1024");
    cudaSetDevice(x->device);
    if (flush_if_staged(x)) return 1;
    if (!nq) return 0;
    std::shared_lock<std::shared_mutex> lk(x->rw);
    const uint64_t n = x->count, live = n - x->n_deleted;
    // removed members cannot be masked inside the kernels: ask for enough extra candidates to drop them afterwards
    const uint64_t kk = std::min<uint64_t>(n, (uint64_t)k + x->n_deleted);
    if (kk > 1024) return fail("gvss_exact_search: %llu removed members are in the way of an exact top-%u; compact the index first", (unsigned long long)x->n_deleted, k);
    SearchSlot *s = acquire_slot(x);
    if (!s) return 1;
    SlotGuard guard{x, s};
    const uint32_t d = x->cfg.dimensions, dp = x->dpad;
    auto up = [](size_t v) { return (v + 255) / 256 * 256; };
    const size_t q_bytes = up(nq * dp * 4ull), o_keys = q_bytes, o_dist = o_keys + up(nq * std::max<uint64_t>(kk, 1) * 8ull),
                 total = o_dist + up(nq * std::max<uint64_t>(kk, 1) * 4ull);
    if (ensure_bytes(s->d_io, s->io_bytes, total)) return 1;
    if (total > s->h_io_bytes) {
        if (s->h_io) cudaFreeHost(s->h_io);
        s->h_io = nullptr; s->h_io_bytes = 0;
        CU(cudaMallocHost(&s->h_io, total));
        s->h_io_bytes = total;
    }
    for (uint64_t i = 0; i < nq; i++) {
        if (counts_out) counts_out[i] = 0;
        for (uint32_t j = 0; j < k; j++) { keys_out[i * k + j] = -1; if (dist_out) dist_out[i * k + j] = INFINITY; }
    }
    if (!kk) return 0; // empty index
    float *hq = reinterpret_cast<float *>(s->h_io);
    if (dp != d) memset(hq, 0, nq * dp * 4ull);
    for (uint64_t i = 0; i < nq; i++) memcpy(hq + i * dp, queries + i * d, d * 4ull);
    CU(cudaMemcpyAsync(s->d_io, hq, nq * dp * 4ull, cudaMemcpyHostToDevice, s->stream));
    if (gvss_exact_topk_device(x->cfg.metric, x->d_vectors, n, x->d_keys, reinterpret_cast<const float *>(s->d_io), nq, dp, (uint32_t)kk,
                               reinterpret_cast<int64_t *>(s->d_io + o_keys), reinterpret_cast<float *>(s->d_io + o_dist), s->stream)) return 1;
    CU(cudaMemcpyAsync(s->h_io + o_keys, s->d_io + o_keys, total - o_keys, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    const int64_t *hk = reinterpret_cast<const int64_t *>(s->h_io + o_keys);
    const float *hd = reinterpret_cast<const float *>(s->h_io + o_dist);
    for (uint64_t i = 0; i < nq; i++) {
        uint32_t c = 0;
        for (uint64_t j = 0; j < kk && c < k && c < live; j++) {
            if (hk[i * kk + j] == GVSS_FREE_KEY || hk[i * kk + j] < 0) continue;
            keys_out[i * k + c] = hk[i * kk + j];
            if (dist_out) dist_out[i * k + c] = hd[i * kk + j];
            c++;
        }
        if (counts_out) counts_out[i] = c;
    }
    return 0;
    API_END
}

// ---- persistence -------------------------------------------------------------------------------------------------
int gvss_serialized_length(gvss_index *x, uint64_t *bytes) {
    API_BEGIN
    if (check_usable(x) || !bytes) return x && bytes ? 1 : fail("null argument");
    if (check_graph(x, "gvss_serialized_length")) return 1;
    if (x->router) return fail("a sharded index is not one usearch stream: persisted indexes are single-GPU (the reference cannot merge indexes either, hnsw_index.cpp:561-563)");
    cudaSetDevice(x->device);
    if (flush_if_staged(x)) return 1;
    std::shared_lock<std::shared_mutex> lk(x->rw);
    const uint64_t M = x->cfg.connectivity, M0 = x->cfg.connectivity_base;
    uint64_t n = 8 + x->count * x->cfg.dimensions * 4ull + 64 + 40 + 2 * x->count;
    for (uint64_t i = 0; i < x->count; i++) n += 10 + (4 + 4 * M0) + (uint64_t)x->h_levels[i] * (4 + 4 * M);
    *bytes = n;
    return 0;
    API_END
}

int gvss_save(gvss_index *x, int (*write)(void *, const void *, size_t), void *user) {
    API_BEGIN
    if (!x || !write) return fail("gvss_save: null argument");
    if (check_usable(x)) return 1;
    if (check_graph(x, "gvss_save")) return 1;
    if (x->router) return fail("a sharded index is not one usearch stream: persisted indexes are single-GPU (the reference cannot merge indexes either, hnsw_index.cpp:561-563)");
    cudaSetDevice(x->device);
    if (flush_if_staged(x)) return 1;
    std::shared_lock<std::shared_mutex> lk(x->rw);
    SearchSlot *s = acquire_slot(x); // a stream of its own: saving may overlap searches
    if (!s) return 1;
    SlotGuard guard{x, s};
    cudaStream_t st = s->stream;
    const uint32_t d = x->cfg.dimensions, M = x->cfg.connectivity, M0 = x->cfg.connectivity_base;
    const uint64_t n = x->count;
    // (1) matrix header + vectors in slot order (index_dense.hpp:822-845)
    uint32_t hdr[2] = {(uint32_t)n, d * 4};
    if (!write(user, hdr, 8)) return fail("gvss_save: output callback failed");
    {
        // two pinned buffers: while the callback consumes one chunk the next one is already on its way from the GPU
        const uint64_t rows_per = std::max<uint64_t>(1, (16ull << 20) / (d * 4ull));
        const uint64_t chunk_rows = std::min(rows_per, std::max<uint64_t>(n, 1));
        PinnedPair pin;
        CU(pin.alloc(chunk_rows * d * 4ull));
        auto fetch = [&](uint64_t r, int b) -> cudaError_t {
            const uint64_t m = std::min(chunk_rows, n - r);
            cudaError_t e = cudaMemcpy2DAsync(pin.buf[b], d * 4ull, x->d_vectors + r * x->dpad, x->dpad * 4ull, d * 4ull, m, cudaMemcpyDeviceToHost, st);
            return e == cudaSuccess ? cudaEventRecord(pin.ev[b], st) : e;
        };
        if (n) CU(fetch(0, 0));
        int b = 0;
        for (uint64_t r = 0; r < n; r += chunk_rows, b ^= 1) {
            const uint64_t m = std::min(chunk_rows, n - r);
            if (r + chunk_rows < n) CU(fetch(r + chunk_rows, b ^ 1));
            CU(cudaEventSynchronize(pin.ev[b]));
            if (!write(user, pin.buf[b], m * d * 4ull)) return fail("gvss_save: output callback failed");
        }
    }
    // (2) 64-byte dense header (index_dense.hpp:46-83,850-874)
    uint8_t head[64];
    memset(head, 0, 64);
    memcpy(head, "usearch", 7);
    const uint16_t ver[3] = {2, 12, 0};
    memcpy(head + 7, ver, 6);
    head[13] = (uint8_t)x->cfg.metric; head[14] = 11; head[15] = 20; head[16] = 15;
    const uint64_t present = n - x->n_deleted, deleted = x->n_deleted, dims = d;
    memcpy(head + 17, &present, 8); memcpy(head + 25, &deleted, 8); memcpy(head + 33, &dims, 8);
    if (!write(user, head, 64)) return fail("gvss_save: output callback failed");
    // (3) graph header + levels (index.hpp:2714-2737)
    const uint64_t gh[5] = {n, M, M0, (uint64_t)(x->max_level < 0 ? 0 : x->max_level), n ? x->entry : 0};
    if (!write(user, gh, 40)) return fail("gvss_save: output callback failed");
    if (n && !write(user, x->h_levels.data(), n * 2)) return fail("gvss_save: output callback failed");
    // (4) node tapes (index.hpp:2740-2746): key, level, [count + M0 slots], level × [count + M slots]
    std::vector<uint32_t> nbr0(n * x->stride0), upper(x->upper_used * x->strideU);
    if (n) CU(cudaMemcpyAsync(nbr0.data(), x->d_nbr0, nbr0.size() * 4, cudaMemcpyDeviceToHost, st));
    if (!upper.empty()) CU(cudaMemcpyAsync(upper.data(), x->d_upper, upper.size() * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    std::vector<uint8_t> out;
    out.reserve(8u << 20);
    auto emit_list = [&](const uint32_t *row, uint32_t cap) {
        uint32_t cnt = 0;
        while (cnt < cap && row[cnt] != kInvalid) cnt++;
        const size_t o = out.size();
        out.resize(o + 4 + 4ull * cap, 0);
        memcpy(out.data() + o, &cnt, 4);
        memcpy(out.data() + o + 4, row, 4ull * cnt);
    };
    for (uint64_t i = 0; i < n; i++) {
        const size_t o = out.size();
        out.resize(o + 10);
        memcpy(out.data() + o, &x->h_keys[i], 8);
        memcpy(out.data() + o + 8, &x->h_levels[i], 2);
        emit_list(nbr0.data() + i * x->stride0, M0);
        for (int l = 1; l <= x->h_levels[i]; l++) emit_list(upper.data() + ((uint64_t)x->h_upper_off[i] + l - 1) * x->strideU, M);
        if (out.size() >= (8u << 20) - 4096) {
            if (!write(user, out.data(), out.size())) return fail("gvss_save: output callback failed");
            out.clear();
        }
    }
    if (!out.empty() && !write(user, out.data(), out.size())) return fail("gvss_save: output callback failed");
    return 0;
    API_END
}

// Reads a usearch v2.12 stream into `y` (an empty index of x's configuration). Everything the stream says is checked before
// it is used: a corrupt file gives an error, never an out-of-bounds slot in a kernel.
static int load_into(gvss_index *y, int (*read)(void *, void *, size_t), void *user) {
    uint32_t hdr[2];
    if (!read(user, hdr, 8)) return fail("gvss_load: failed to read the matrix header");
    const uint64_t n = hdr[0];
    if (hdr[1] != y->cfg.dimensions * 4) return fail("gvss_load: stream has %u bytes per vector, index expects %u", hdr[1], y->cfg.dimensions * 4);
    const uint32_t d = y->cfg.dimensions;
    if (ensure_capacity(y, std::max<uint64_t>(n, 1))) return 1;
    {
        // the stream is read straight into pinned memory, two buffers: the upload of one chunk overlaps the read of the next
        const uint64_t rows_per = std::max<uint64_t>(1, (16ull << 20) / (d * 4ull));
        const uint64_t chunk_rows = std::min(rows_per, std::max<uint64_t>(n, 1));
        PinnedPair pin;
        CU(pin.alloc(chunk_rows * d * 4ull));
        int b = 0;
        for (uint64_t r = 0; r < n; r += chunk_rows, b ^= 1) {
            const uint64_t m = std::min(chunk_rows, n - r);
            CU(cudaEventSynchronize(pin.ev[b])); // the upload that read this buffer two chunks ago
            if (!read(user, pin.buf[b], m * d * 4ull)) return fail("gvss_load: truncated vectors");
            CU(cudaMemcpy2DAsync(y->d_vectors + r * y->dpad, y->dpad * 4ull, pin.buf[b], d * 4ull, d * 4ull, m, cudaMemcpyHostToDevice, y->stream));
            CU(cudaEventRecord(pin.ev[b], y->stream));
        }
        CU(cudaStreamSynchronize(y->stream));
    }
    uint8_t head[64];
    if (!read(user, head, 64)) return fail("gvss_load: truncated header");
    if (memcmp(head, "usearch", 7)) return fail("Magic header mismatch - the file isn't an index"); // index_dense.hpp:945-947
    uint16_t major;
    memcpy(&major, head + 7, 2);
    if (major != 2) return fail("File format may be different, please rebuild");                         // :949
    if (head[15] != 20) return fail("Key type doesn't match, consider rebuilding");                       // :951-953
    if (head[16] != 15) return fail("Slot type doesn't match, consider rebuilding");                      // :955-957
    if (head[14] != 11) return fail("gvss_load: only f32 scalars are supported (scalar kind %u)", head[14]);
    const char metric = (char)head[13];
    if (metric != GVSS_METRIC_L2SQ && metric != GVSS_METRIC_COS && metric != GVSS_METRIC_IP)
        return fail("gvss_load: metric kind '%c' (%u) is not one of l2sq / cos / ip", metric >= 32 && metric < 127 ? metric : '?', head[13]);
    y->cfg.metric = metric;
    uint64_t gh[5];
    if (!read(user, gh, 40)) return fail("gvss_load: truncated graph header");
    if (gh[0] != n) return fail("Index size and the number of vectors doesn't match");                     // :965-969
    if (gh[1] < 2 || gh[1] > 4096 || gh[2] < 2 || gh[2] > 4096) return fail("gvss_load: connectivity %llu / %llu outside 2..4096", (unsigned long long)gh[1], (unsigned long long)gh[2]);
    if (n && (gh[3] > 64 || gh[4] >= n)) return fail("gvss_load: corrupt graph header (max level %llu, entry %llu of %llu nodes)", (unsigned long long)gh[3], (unsigned long long)gh[4], (unsigned long long)n);
    if (gh[1] != y->cfg.connectivity || gh[2] != y->cfg.connectivity_base) {
        // the stream's connectivity wins (index.hpp:2779-2783); adjacency strides change => reallocate
        y->cfg.connectivity = (uint32_t)gh[1]; y->cfg.connectivity_base = (uint32_t)gh[2];
        const uint32_t s0 = (y->cfg.connectivity_base + 7) / 8 * 8, sU = (y->cfg.connectivity + 7) / 8 * 8;
        if (s0 != y->stride0) {
            cudaFree(y->d_nbr0); y->d_nbr0 = nullptr;
            y->stride0 = s0;
            CU(dev_resize(y->d_nbr0, 0, y->capacity * y->stride0, 0xFF, y->stream));
        }
        y->strideU = sU;
    }
    const uint32_t M = y->cfg.connectivity, M0 = y->cfg.connectivity_base;
    y->max_level = n ? (int32_t)gh[3] : -1;
    y->entry = (uint32_t)gh[4];
    y->h_levels.resize(n);
    if (n && !read(user, y->h_levels.data(), n * 2)) return fail("gvss_load: truncated levels");
    y->h_keys.resize(n); y->h_upper_off.assign(n, kInvalid);
    uint64_t blocks = 0;
    for (uint64_t i = 0; i < n; i++) {
        const int16_t lvl = y->h_levels[i];
        if (lvl < 0 || lvl > y->max_level) return fail("gvss_load: node %llu has level %d outside 0..%d", (unsigned long long)i, (int)lvl, y->max_level);
        if (lvl > 0) { y->h_upper_off[i] = (uint32_t)blocks; blocks += (uint64_t)lvl; }
    }
    if (n && y->h_levels[y->entry] != y->max_level) return fail("gvss_load: entry node %u has level %d, graph header says %d", y->entry, (int)y->h_levels[y->entry], y->max_level);
    if (blocks >= 0xFFFFFFFFull) return fail("gvss_load: too many upper-level lists");
    if (ensure_upper(y, blocks)) return 1;
    y->upper_used = blocks;
    std::vector<uint32_t> nbr0(n * y->stride0, kInvalid), upper(blocks * y->strideU, kInvalid);
    std::vector<uint8_t> buf;
    uint64_t i = 0;
    while (i < n) { // read node tapes in ~8 MB pieces that end on node boundaries
        uint64_t j = i, bytes = 0;
        while (j < n && bytes < (8u << 20)) { bytes += 10 + (4 + 4ull * M0) + (uint64_t)y->h_levels[j] * (4 + 4ull * M); j++; }
        buf.resize(bytes);
        if (!read(user, buf.data(), bytes)) return fail("gvss_load: truncated node tapes");
        const uint8_t *p = buf.data();
        for (; i < j; i++) {
            int16_t lvl;
            memcpy(&y->h_keys[i], p, 8); memcpy(&lvl, p + 8, 2); p += 10;
            if (lvl != y->h_levels[i]) return fail("gvss_load: node %llu level mismatch", (unsigned long long)i);
            bool bad = false;
            auto take = [&](uint32_t *row, uint32_t cap, int level) {
                uint32_t cnt; memcpy(&cnt, p, 4);
                if (cnt > cap) cnt = cap;
                memcpy(row, p + 4, 4ull * cnt);
                for (uint32_t e = 0; e < cnt; e++) bad |= row[e] >= n || y->h_levels[row[e]] < level; // a neighbour must exist on that level
                p += 4 + 4ull * cap;
            };
            take(nbr0.data() + i * y->stride0, M0, 0);
            for (int l = 1; l <= lvl; l++) take(upper.data() + ((uint64_t)y->h_upper_off[i] + l - 1) * y->strideU, M, l);
            if (bad) return fail("gvss_load: node %llu links to a slot that does not exist on that level (corrupt stream)", (unsigned long long)i);
            if (y->h_keys[i] == GVSS_FREE_KEY) { y->n_deleted++; y->free_slots.push_back((uint32_t)i); } // reindex_keys_, index_dense.hpp:1901-1929
        }
    }
    if (n) {
        CU(cudaMemcpyAsync(y->d_nbr0, nbr0.data(), nbr0.size() * 4, cudaMemcpyHostToDevice, y->stream));
        if (!upper.empty()) CU(cudaMemcpyAsync(y->d_upper, upper.data(), upper.size() * 4, cudaMemcpyHostToDevice, y->stream));
        CU(cudaMemcpyAsync(y->d_keys, y->h_keys.data(), n * 8, cudaMemcpyHostToDevice, y->stream));
        CU(cudaMemcpyAsync(y->d_levels, y->h_levels.data(), n * 2, cudaMemcpyHostToDevice, y->stream));
        CU(cudaMemcpyAsync(y->d_upper_off, y->h_upper_off.data(), n * 4, cudaMemcpyHostToDevice, y->stream));
        CU(cudaStreamSynchronize(y->stream));
    }
    y->count = n;
    return 0;
}

int gvss_load(gvss_index *x, int (*read)(void *, void *, size_t), void *user) {
    API_BEGIN
    if (!x || !read) return fail("gvss_load: null argument");
    if (check_graph(x, "gvss_load")) return 1;
    if (x->router) return fail("a sharded index is not one usearch stream: persisted indexes are single-GPU (the reference cannot merge indexes either, hnsw_index.cpp:561-563)");
    std::unique_lock<std::shared_mutex> lk(x->rw);
    Publish pub{x};
    cudaSetDevice(x->device);
    // parsed and uploaded into an index of its own, which changes places with this one only when the whole stream was good
    gvss_config cfg = x->cfg;
    cfg.device = x->device;
    gvss_index *y = nullptr;
    if (gvss_create(&cfg, &y)) return 1;
    struct Drop { gvss_index *p; ~Drop() { gvss_destroy(p); } } drop{y};
    if (load_into(y, read, user)) return 1;
    wait_device_searches(x);
    clear_staging(x); // load_from_stream replaces the index: rows that were only staged go with the old contents
    swap_graph_state(x, y);
    cudaMemset(x->d_build_stats, 0, 48);
    x->level_rng = 1;
    x->poisoned.clear();
    recompute_room(x);
    return 0;
    API_END
}

// ---- introspection -----------------------------------------------------------------------------------------------
int gvss_get_info(gvss_index *x, gvss_info *out) {
    API_BEGIN
    if (!x || !out) return fail("null argument");
    if (x->router) return router_info(x->router, out);
    // no lock: the figures are the ones the last writer published (a flush in progress does not make size() / capacity() wait)
    memset(out, 0, sizeof *out);
    // read order staged -> inflight -> count mirrors the writer's order (inflight set, staged dropped ... count raised, inflight
    // cleared): during a flush the sum can run ahead by the rows in flight, it never falls behind
    const uint64_t staged = x->staged.load() + x->pub_inflight.load(), count = x->pub_count.load(), cap = x->pub_capacity.load();
    out->size = count - x->pub_deleted.load() + staged; // size() counts rows whose add() has returned (SyncSize after the bulk build)
    out->capacity = cap;
    out->node_count = count + staged;
    const int32_t ml = x->pub_max_level.load();
    out->max_level = ml < 0 ? 0 : (uint64_t)ml;
    out->dimensions = x->cfg.dimensions;
    out->connectivity = x->cfg.connectivity; out->connectivity_base = x->cfg.connectivity_base;
    out->expansion_add = x->cfg.expansion_add; out->expansion_search = x->cfg.expansion_search;
    out->metric = x->cfg.metric;
    size_t slot_bytes = 0;
    {
        std::lock_guard<std::mutex> sl(x->slot_mu);
        for (auto &s : x->slots) slot_bytes += s->scratch_bytes + s->io_bytes;
        out->concurrent_searches_peak = x->slots_peak;
    }
    out->memory_usage = cap * (x->dpad * 4ull + 8 + 2 + x->stride0 * 4ull + 4 + 4) + x->pub_upper_cap.load() * x->strideU * 4ull + x->pub_scratch.load() + slot_bytes;
    return 0;
    API_END
}

int gvss_get_build_stats(gvss_index *x, gvss_build_stats *out) {
    API_BEGIN
    if (!x || !out) return fail("null argument");
    if (check_usable(x)) return 1;
    if (x->router) return router_build_stats(x->router, out);
    cudaSetDevice(x->device);
    if (flush_if_staged(x)) return 1;
    std::shared_lock<std::shared_mutex> lk(x->rw);
    unsigned long long h[6] = {};
    CU(cudaMemcpy(h, x->d_build_stats, sizeof h, cudaMemcpyDeviceToHost));
    out->inserts = h[0]; out->query_distances = h[1]; out->pair_distances = h[2];
    out->base_expansions = h[3]; out->upper_expansions = h[4]; out->rows_written = h[5];
    return 0;
    API_END
}

int gvss_get_level_stats(gvss_index *x, uint32_t level, gvss_level_stats *out) {
    API_BEGIN
    if (!x || !out) return fail("null argument");
    if (check_usable(x)) return 1;
    if (x->router) return router_level_stats(x->router, level, out);
    cudaSetDevice(x->device);
    if (flush_if_staged(x)) return 1;
    std::shared_lock<std::shared_mutex> lk(x->rw);
    memset(out, 0, sizeof *out);
    const uint64_t n = x->count;
    const uint32_t M = x->cfg.connectivity, M0 = x->cfg.connectivity_base;
    std::vector<uint32_t> adj;
    if (level == 0) { adj.resize(n * x->stride0); if (n) CU(cudaMemcpy(adj.data(), x->d_nbr0, adj.size() * 4, cudaMemcpyDeviceToHost)); }
    else { adj.resize(x->upper_used * x->strideU); if (!adj.empty()) CU(cudaMemcpy(adj.data(), x->d_upper, adj.size() * 4, cudaMemcpyDeviceToHost)); }
    for (uint64_t i = 0; i < n; i++) {
        if ((uint32_t)x->h_levels[i] < level) continue;
        const uint32_t *row = level == 0 ? adj.data() + i * x->stride0 : adj.data() + ((uint64_t)x->h_upper_off[i] + level - 1) * x->strideU;
        const uint32_t cap = level == 0 ? M0 : M;
        uint32_t cnt = 0;
        while (cnt < cap && row[cnt] != kInvalid) cnt++;
        out->nodes++; out->edges += cnt;
        out->allocated_bytes += 4 + 4ull * cap;
    }
    out->max_edges = out->nodes * (level ? M0 : M); // swapped on purpose: index.hpp:2628-2629 (stats quirks)
    return 0;
    API_END
}

int gvss_merge_topk_device(const int64_t *gk, const float *gd, uint32_t shards, uint64_t nq, uint32_t k, int64_t *keys_out,
                           float *dist_out, void *cuda_stream) {
    API_BEGIN
    if (!gk || !gd || !keys_out || !dist_out) return fail("gvss_merge_topk_device: null buffer");
    CU(launch_merge_topk(gk, gd, shards, nq, k, keys_out, dist_out, (cudaStream_t)cuda_stream));
    return 0;
    API_END
}

int gvss_merge_topk_peers_device(const int64_t *const *peer_keys, const float *const *peer_dist, uint32_t *const *peer_flags,
                                 uint32_t rank, uint32_t shards, uint32_t step, uint64_t nq, uint32_t k, int64_t *keys_out,
                                 float *dist_out, void *cuda_stream) {
    API_BEGIN
    if (!peer_keys || !peer_dist || !peer_flags || !keys_out) return fail("gvss_merge_topk_peers_device: null buffer");
    if (!shards || shards > 16 || rank >= shards) return fail("gvss_merge_topk_peers_device: 1..16 shards, rank < shards");
    if ((uint64_t)shards * k > 256) return fail("gvss_merge_topk_peers_device: shards * k must not exceed 256");
    for (uint32_t s = 0; s < shards; s++)
        if (!peer_keys[s] || !peer_dist[s] || !peer_flags[s]) return fail("gvss_merge_topk_peers_device: null peer pointer (shard %u)", s);
    CU(launch_merge_topk_peers(peer_keys, peer_dist, peer_flags, rank, shards, step, nq, k, keys_out, dist_out, (cudaStream_t)cuda_stream));
    return 0;
    API_END
}

int gvss_peer_exchange_status(uint32_t *timed_out) {
    API_BEGIN
    if (!timed_out) return fail("gvss_peer_exchange_status: null argument");
    CU(read_peer_timeouts(timed_out));
    return 0;
    API_END
}

int gvss_exact_topk_device(char metric, const float *base, uint64_t n, const int64_t *keys, const float *queries, uint64_t nq,
                           uint32_t d, uint32_t k, int64_t *ids_out, float *dist_out, void *cuda_stream) {
    API_BEGIN
    if (!base || !queries || !ids_out) return fail("gvss_exact_topk_device: null buffer");
    if (!k || k > 1024) return fail("gvss_exact_topk: k must be in 1..1024");
    // tensor-core candidates (3xTF32) + exact re-rank when the shape fits (k <= 16, enough work to fill the tiles);
    // GVSS_EXACT_TC=0 forces the all-pairs FP32 kernel, =2 forces the tensor-core path even for tiny inputs
    static const int tc_mode = [] { const char *e = getenv("GVSS_EXACT_TC"); return e ? atoi(e) : 1; }();
    if (tc_mode && (tc_mode == 2 || (uint64_t)n * nq >= (1ull << 22))) {
        const cudaError_t e = launch_exact_topk_tc(metric, base, n, keys, queries, nq, d, k, ids_out, dist_out, nullptr, nullptr, (cudaStream_t)cuda_stream);
        if (e == cudaSuccess) return 0;
        if (e != cudaErrorNotSupported) return fail("CUDA error %s in the tensor-core exact top-k", cudaGetErrorString(e));
        cudaGetLastError();
    }
    CU(launch_exact_topk(metric, base, n, keys, queries, nq, d, k, ids_out, dist_out, (cudaStream_t)cuda_stream));
    return 0;
    API_END
}

// test hook (not part of include/gvss.h): raw q·b of the tensor-core path plus the number of queries that needed the fallback
int gvssdbg_tc_exact(char metric, const float *base, uint64_t n, const float *queries, uint64_t nq, uint32_t d, uint32_t k, int64_t *ids_out,
                     float *dist_out, float *dots_out, uint64_t *flagged) {
    const cudaError_t e = launch_exact_topk_tc(metric, base, n, nullptr, queries, nq, d, k, ids_out, dist_out, dots_out, flagged, nullptr);
    if (e != cudaSuccess) return fail("tensor-core exact top-k: %s", cudaGetErrorString(e));
    CU(cudaDeviceSynchronize());
    return 0;
}

int gvss_exact_topk(char metric, const float *base, uint64_t n, const int64_t *keys, const float *queries, uint64_t nq, uint32_t d,
                    uint32_t k, int64_t *ids_out, float *dist_out) {
    API_BEGIN
    if (!base || !queries || !ids_out) return fail("gvss_exact_topk: null buffer");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail("gvss: no CUDA device available (this library has no CPU fallback)");
    float *db = nullptr, *dq = nullptr, *dd = nullptr; int64_t *dk = nullptr, *di = nullptr;
    int rc = 0;
    auto cleanup = [&]() { cudaFree(db); cudaFree(dq); cudaFree(dd); cudaFree(dk); cudaFree(di); };
#define CUX(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { cleanup(); return fail("CUDA error %s (%s)", cudaGetErrorString(e_), #x); } } while (0)
    CUX(cudaMalloc(&db, std::max<uint64_t>(n, 1) * d * 4)); CUX(cudaMalloc(&dq, std::max<uint64_t>(nq, 1) * d * 4));
    CUX(cudaMalloc(&dd, std::max<uint64_t>(nq, 1) * k * 4)); CUX(cudaMalloc(&di, std::max<uint64_t>(nq, 1) * k * 8));
    if (keys) { CUX(cudaMalloc(&dk, std::max<uint64_t>(n, 1) * 8)); CUX(cudaMemcpy(dk, keys, n * 8, cudaMemcpyHostToDevice)); }
    CUX(cudaMemcpy(db, base, n * d * 4, cudaMemcpyHostToDevice));
    CUX(cudaMemcpy(dq, queries, nq * d * 4, cudaMemcpyHostToDevice));
    rc = gvss_exact_topk_device(metric, db, n, dk, dq, nq, d, k, di, dd, nullptr);
    if (!rc) {
        CUX(cudaDeviceSynchronize());
        CUX(cudaMemcpy(ids_out, di, nq * k * 8, cudaMemcpyDeviceToHost));
        if (dist_out) CUX(cudaMemcpy(dist_out, dd, nq * k * 4, cudaMemcpyDeviceToHost));
    }
#undef CUX
    cleanup();
    return rc;
    API_END
}

} // extern "C"

// ------------------------------------------------------------------------------------------------------------
// bulk insert driver
// ------------------------------------------------------------------------------------------------------------
namespace gvss {

// what one flush is going to do, decided on local copies before any state of the index changes
struct FlushPlan {
    std::vector<uint32_t> slots;                              // insertion order
    std::vector<std::pair<uint32_t, const float *>> host_rows; // (slot, source row in a taken staging buffer)
    struct DevRun { uint32_t first; const float *src; uint64_t rows; };
    std::vector<DevRun> dev_runs;                             // maximal runs of consecutive slots of device-resident batches
    std::vector<std::pair<uint32_t, int64_t>> recycled;       // (slot, new key) of reused removed slots
    std::vector<int64_t> fresh_keys;
    std::vector<int16_t> fresh_levels;
    std::vector<uint32_t> fresh_off;
    uint64_t rng = 1, next_fresh = 0, blocks = 0;
    size_t free_taken = 0;
};

// uploads + kernel launches of a flush whose host mirrors are already committed; any failure here poisons the index
static int flush_device_work(gvss_index *x, const FlushPlan &pl, uint64_t old_count, uint32_t *d_slots, uint32_t *d_inflight,
                             unsigned long long *d_prof, uint32_t max_warps, BuildParams p) {
    const uint32_t d = x->cfg.dimensions;
    const size_t row_bytes = (size_t)x->dpad * 4;
    if (!pl.host_rows.empty()) {
        // staged host rows go up through two pinned bounce buffers: while one chunk is on its way to the GPU the next one is
        // padded into the other buffer (no pinned allocation of the size of the whole batch, copy and padding overlap)
        SearchSlot *s = x->slots.front().get(); // slot 0 lends its pinned buffer (rw is held exclusively: no search runs)
        const size_t chunk_rows = s->h_io_bytes / (2 * row_bytes);
        float *pin = reinterpret_cast<float *>(s->h_io);
        cudaEvent_t done[2] = {nullptr, nullptr};
        struct EvGuard { cudaEvent_t *e; ~EvGuard() { for (int i = 0; i < 2; i++) if (e[i]) cudaEventDestroy(e[i]); } } guard{done};
        CU(cudaEventCreateWithFlags(&done[0], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&done[1], cudaEventDisableTiming));
        const bool padded = x->dpad != d;
        int b = 0;
        for (size_t c0 = 0; c0 < pl.host_rows.size(); c0 += chunk_rows, b ^= 1) {
            const size_t c1 = std::min(pl.host_rows.size(), c0 + chunk_rows);
            float *buf = pin + (size_t)b * chunk_rows * x->dpad;
            CU(cudaEventSynchronize(done[b])); // the copies that read this buffer two chunks ago (no-op for a fresh event)
            if (padded) memset(buf, 0, (c1 - c0) * row_bytes);
            for (size_t r = c0; r < c1; r++) memcpy(buf + (r - c0) * x->dpad, pl.host_rows[r].second, d * 4ull);
            size_t i = c0;
            while (i < c1) { // runs of consecutive slots go up as one copy
                size_t j = i + 1;
                while (j < c1 && pl.host_rows[j].first == pl.host_rows[j - 1].first + 1) j++;
                CU(cudaMemcpyAsync(x->d_vectors + (size_t)pl.host_rows[i].first * x->dpad, buf + (i - c0) * x->dpad, (j - i) * row_bytes, cudaMemcpyHostToDevice, x->stream));
                i = j;
            }
            CU(cudaEventRecord(done[b], x->stream));
        }
        CU(cudaEventSynchronize(done[0]));
        CU(cudaEventSynchronize(done[1]));
    }
    for (auto &run : pl.dev_runs)
        CU(cudaMemcpy2DAsync(x->d_vectors + (size_t)run.first * x->dpad, row_bytes, run.src, d * 4ull, d * 4ull, run.rows, cudaMemcpyDeviceToDevice, x->stream));
    for (auto &rc : pl.recycled) { // memset of the tape keeping the level (index.hpp:2441-2444)
        const uint32_t s = rc.first;
        CU(cudaMemsetAsync(x->d_nbr0 + (size_t)s * x->stride0, 0xFF, x->stride0 * 4ull, x->stream));
        if (x->h_levels[s] > 0) CU(cudaMemsetAsync(x->d_upper + (size_t)x->h_upper_off[s] * x->strideU, 0xFF, (size_t)x->h_levels[s] * x->strideU * 4ull, x->stream));
        CU(cudaMemcpyAsync(x->d_keys + s, &x->h_keys[s], 8, cudaMemcpyHostToDevice, x->stream));
    }
    if (pl.next_fresh > old_count) {
        const uint64_t b = old_count, m = pl.next_fresh - b;
        CU(cudaMemcpyAsync(x->d_keys + b, x->h_keys.data() + b, m * 8, cudaMemcpyHostToDevice, x->stream));
        CU(cudaMemcpyAsync(x->d_levels + b, x->h_levels.data() + b, m * 2, cudaMemcpyHostToDevice, x->stream));
        CU(cudaMemcpyAsync(x->d_upper_off + b, x->h_upper_off.data() + b, m * 4, cudaMemcpyHostToDevice, x->stream));
    }
    CU(cudaMemcpyAsync(d_slots, pl.slots.data(), pl.slots.size() * 4, cudaMemcpyHostToDevice, x->stream));
    unsigned long long es = x->max_level < 0 ? ~0ull : (((unsigned long long)(uint32_t)x->max_level << 32) | x->entry);
    CU(cudaMemcpyAsync(x->d_entry_state, &es, 8, cudaMemcpyHostToDevice, x->stream));
    if (d_prof) CU(cudaMemsetAsync(d_prof, 0, 64, x->stream));
    if (d_inflight) CU(cudaMemsetAsync(d_inflight, 0xFF, (size_t)max_warps * 4, x->stream));
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;
    if (d_prof) { cudaEventCreate(&ev0); cudaEventCreate(&ev1); cudaEventCreate(&ev2); cudaEventRecord(ev0, x->stream); }
    // graph construction: concurrency ramps with the graph size (a launch never runs more concurrent inserts than
    // linked nodes / ramp), then one persistent launch for the rest. Vectors inserted at the same time cannot see each other
    // through the graph; k_build's peer pass links them explicitly, which makes the cap almost irrelevant for quality (recall
    // equal to the reference-built graph from ramp 2 to 256 at 30k and 1M vectors, profiles/build_ramp_r1.txt). 4 keeps the
    // very first launches small. GVSS_BUILD_RAMP overrides.
    uint64_t ramp = 4;
    if (const char *e = getenv("GVSS_BUILD_RAMP")) ramp = std::max<uint64_t>(1, strtoull(e, nullptr, 10));
    uint64_t done = 0;
    const uint64_t total = pl.slots.size(), linked = old_count;
    const int metric = metric_id(x->cfg.metric);
    while (done < total && !x->flat) { // a rows-only index stops at the uploads: there is no graph to link the rows into
        const uint64_t have = linked + done;
        uint64_t conc = std::max<uint64_t>(1, have / ramp);
        conc = std::min<uint64_t>(conc, max_warps);
        uint64_t len = conc >= max_warps ? total - done : std::max<uint64_t>(conc * 2, 16);
        len = std::min(len, total - done);
        p.begin = (uint32_t)done; p.end = (uint32_t)(done + len);
        const unsigned long long start = done;
        CU(cudaMemcpyAsync(x->d_work, &start, 8, cudaMemcpyHostToDevice, x->stream));
        GraphView g = view_of(x);
        g.count = (uint32_t)pl.next_fresh;
        g.has_deleted = 0;
        if (d_prof && conc >= max_warps) cudaEventRecord(ev1, x->stream);
        CU(launch_build(metric, g, p, (uint32_t)std::min<uint64_t>(conc, len), x->stream));
        done += len;
    }
    if (d_prof) {
        cudaEventRecord(ev2, x->stream);
        cudaEventSynchronize(ev2);
        unsigned long long h[8] = {0};
        cudaMemcpy(h, d_prof, 64, cudaMemcpyDeviceToHost);
        float ms_all = 0, ms_last = 0;
        cudaEventElapsedTime(&ms_all, ev0, ev2);
        if (cudaEventQuery(ev1) == cudaSuccess) cudaEventElapsedTime(&ms_last, ev1, ev2);
        const double n = h[5] ? (double)h[5] : 1.0;
        fprintf(stderr, "[gvss build] %llu inserts, %u warps max: total %.1f ms (final persistent launch %.1f ms); clocks/insert: greedy %.0f beam %.0f refine %.0f connect %.0f reconnect %.0f\n",
                h[5], max_warps, ms_all, ms_last, h[0] / n, h[1] / n, h[2] / n, h[3] / n, h[4] / n);
        cudaEventDestroy(ev0); cudaEventDestroy(ev1); cudaEventDestroy(ev2);
    }
    CU(cudaMemcpyAsync(&es, x->d_entry_state, 8, cudaMemcpyDeviceToHost, x->stream));
    CU(cudaStreamSynchronize(x->stream));
    x->entry = (uint32_t)(es & 0xFFFFFFFFull);
    x->max_level = (int32_t)(es >> 32);
    x->count = pl.next_fresh;
    return 0;
}

// Inserts everything staged. Caller holds rw exclusively. Order of work: (1) take the staged rows, (2) plan slots / levels on
// local copies, (3) make every allocation that can fail — up to here a failure puts the rows back and leaves the index
// untouched, so the call can be retried — (4) commit the host mirrors, (5) upload and build; a CUDA failure in (5) leaves the
// graph half-written, so it marks the index unusable instead of pretending otherwise.
static int flush_locked(gvss_index *x) {
    if (check_usable(x)) return 1;
    if (!x->staged.load()) return 0;
    auto now_ms = [] { return prof_ms(); }; // GVSS_BUILD_PROF: host-side phases of every flush
    const double t_enter = now_ms();
    wait_device_searches(x);
    const uint32_t d = x->cfg.dimensions;
    struct Taken { std::vector<float> vecs; std::vector<int64_t> keys; };
    std::vector<Taken> host(x->stages.size());
    uint64_t n_new = 0;
    for (size_t t = 0; t < x->stages.size(); t++) {
        std::lock_guard<std::mutex> lk(x->stages[t]->mu);
        host[t].vecs.swap(x->stages[t]->vecs);
        host[t].keys.swap(x->stages[t]->keys);
        n_new += host[t].keys.size();
    }
    auto dev = std::move(x->st_device);
    x->st_device.clear();
    for (auto &b : dev) n_new += b.second.size();
    auto give_back = [&]() { // rows return to the front of their buffers, ahead of anything added meanwhile
        for (size_t t = 0; t < x->stages.size(); t++) {
            std::lock_guard<std::mutex> lk(x->stages[t]->mu);
            auto &sv = x->stages[t]->vecs; auto &sk = x->stages[t]->keys;
            host[t].vecs.insert(host[t].vecs.end(), sv.begin(), sv.end());
            host[t].keys.insert(host[t].keys.end(), sk.begin(), sk.end());
            sv.swap(host[t].vecs); sk.swap(host[t].keys);
        }
        for (auto &b : x->st_device) dev.push_back(std::move(b));
        x->st_device = std::move(dev);
        return 1;
    };
    if (!n_new) return 0;

    // (2) plan: recycled slots first (FIFO, index_dense.hpp:1766-1793: a removed node is reused through index_gt::update), then
    // fresh ones; thread-major staging order, device-resident batches last
    FlushPlan pl;
    pl.rng = x->level_rng; pl.next_fresh = x->count; pl.blocks = x->upper_used;
    pl.slots.reserve(n_new);
    auto plan_slot = [&](int64_t key) -> uint32_t {
        if (pl.free_taken < x->free_slots.size()) {
            const uint32_t s = x->free_slots[pl.free_taken++];
            pl.recycled.emplace_back(s, key);
            return s;
        }
        const uint32_t s = (uint32_t)pl.next_fresh++;
        const int16_t lvl = x->flat ? (int16_t)0 : draw_level(pl.rng, x->cfg.connectivity); // rows only: no upper lists to size
        pl.fresh_keys.push_back(key);
        pl.fresh_levels.push_back(lvl);
        pl.fresh_off.push_back(lvl > 0 ? (uint32_t)pl.blocks : kInvalid);
        if (lvl > 0) pl.blocks += (uint64_t)lvl;
        return s;
    };
    for (size_t t = 0; t < host.size(); t++)
        for (size_t i = 0; i < host[t].keys.size(); i++) {
            const uint32_t s = plan_slot(host[t].keys[i]);
            pl.slots.push_back(s);
            pl.host_rows.emplace_back(s, host[t].vecs.data() + i * d);
        }
    for (auto &b : dev)
        for (size_t i = 0; i < b.second.size(); i++) {
            const uint32_t s = plan_slot(b.second[i]);
            pl.slots.push_back(s);
            if (!pl.dev_runs.empty() && i && pl.dev_runs.back().first + pl.dev_runs.back().rows == s) pl.dev_runs.back().rows++;
            else pl.dev_runs.push_back({s, b.first + i * d, 1});
        }
    if (pl.next_fresh > x->capacity) { give_back(); return fail("Reserve capacity ahead of insertions!"); }
    if (pl.blocks >= 0xFFFFFFFFull) { give_back(); return fail("gvss: too many upper-level lists"); }

    // (3) allocations
    if (ensure_upper(x, pl.blocks)) return give_back();
    uint32_t *d_slots = nullptr, *d_inflight = nullptr;
    unsigned long long *d_prof = nullptr;
    TempDev slots_mem, inflight_mem, prof_mem;
    if (slots_mem.alloc(d_slots, pl.slots.size()) != cudaSuccess) { give_back(); return fail("gvss_flush: out of device memory (slot list)"); }
    BuildParams p{};
    p.slots = d_slots; p.levels = x->d_levels; p.entry_state = x->d_entry_state; p.work_counter = x->d_work; p.locks = x->d_locks;
    p.ef_add = x->cfg.expansion_add;
    const uint32_t cmax = std::max(x->cfg.connectivity_base, x->cfg.connectivity);
    const uint32_t top_cap = std::max(cmax + 1, p.ef_add) + 1;
    // sized by the capacity, not by the count: a run of flushes into one reservation (the background builder) keeps its scratch
    p.sl = make_scratch_layout(std::max(p.ef_add, cmax), std::max<uint64_t>(pl.next_fresh, x->capacity), cmax);
    static const uint32_t build_rows = [] { const char *e = getenv("GVSS_BUILD_STAGE_ROWS"); const uint32_t v = e ? (uint32_t)atoi(e) : 32u; return v == 16u ? 16u : 32u; }();
    p.direct = search_mode_bits();
    // with the list-as-queue beam search the heap only serves exact ties: 32 entries (its 5 top levels) stay in shared memory
    const uint32_t build_heap = (p.direct & 8) ? default_heap_smem(p.ef_add) : 32u;
    p.direct &= ~16; // gather4 staging is a search path
    p.sm = make_smem_layout(x->dpad, top_cap, build_heap, (2 * cmax + 2) * 8, build_rows, true, cmax);
    uint32_t max_warps = resident_warps(x->device, p.sm.bytes, true);
    if (build_heap == 32u) { // half of those entries go to the overflow area when that lets one more warp per SM in (M=16: 10 -> 11)
        const SmemLayout tight = make_smem_layout(x->dpad, top_cap, 16u, (2 * cmax + 2) * 8, build_rows, true, cmax);
        const uint32_t more = resident_warps(x->device, tight.bytes, true);
        if (more > max_warps) { p.sm = tight; max_warps = more; }
    }
    if (!max_warps) { give_back(); return fail("gvss: no CUDA device"); }
    if (x->build_conc && x->build_conc < max_warps) max_warps = x->build_conc;
    if (!x->flat && ensure_scratch(x, p.sl.bytes_per_warp * max_warps)) return give_back();
    p.scratch = x->d_scratch;
    // GVSS_BUILD_PROF=1: per-phase SM clocks summed over all inserts, and wall time of the ramp vs the final launch (stderr)
    static const bool prof_on = [] { const char *e = getenv("GVSS_BUILD_PROF"); return e && atoi(e); }();
    if (prof_on && prof_mem.alloc(d_prof, 8) != cudaSuccess) { give_back(); return fail("gvss_flush: out of device memory"); }
    p.prof = d_prof;
    p.stats = x->d_build_stats;
    // in-flight registry for the peer pass of k_build (GVSS_BUILD_PEERS=0 disables it)
    static const bool peers_on = [] { const char *e = getenv("GVSS_BUILD_PEERS"); return e ? atoi(e) != 0 : true; }();
    if (peers_on && !x->flat && inflight_mem.alloc(d_inflight, max_warps) != cudaSuccess) { give_back(); return fail("gvss_flush: out of device memory"); }
    p.inflight = d_inflight;
    if (!pl.host_rows.empty()) { // pinned bounce buffers (slot 0's pinned block, grown to two 16 MB chunks)
        if (x->slots.empty()) { SearchSlot *s = acquire_slot(x); if (!s) return give_back(); release_slot(x, s); }
        SearchSlot *s = x->slots.front().get();
        const size_t row_bytes = (size_t)x->dpad * 4;
        size_t chunk_rows = std::max<size_t>(1, (16u << 20) / row_bytes);
        chunk_rows = std::min(chunk_rows, pl.host_rows.size());
        const size_t need = 2 * chunk_rows * row_bytes;
        if (need > s->h_io_bytes) {
            if (s->h_io) cudaFreeHost(s->h_io);
            s->h_io = nullptr; s->h_io_bytes = 0;
            if (cudaMallocHost(&s->h_io, need) != cudaSuccess) { give_back(); return fail("gvss_flush: out of pinned host memory"); }
            s->h_io_bytes = need;
        }
    }

    // (4) commit the host mirrors
    const double t_planned = now_ms();
    const uint64_t old_count = x->count;
    for (auto &rc : pl.recycled) {
        x->h_keys[rc.first] = rc.second;
        if (x->lookup_valid) x->key_to_slot.emplace(rc.second, rc.first);
    }
    x->free_slots.erase(x->free_slots.begin(), x->free_slots.begin() + (std::ptrdiff_t)pl.free_taken);
    x->n_deleted -= pl.free_taken;
    for (size_t i = 0; i < pl.fresh_keys.size(); i++) {
        if (x->lookup_valid) x->key_to_slot.emplace(pl.fresh_keys[i], (uint32_t)(old_count + i));
        x->h_keys.push_back(pl.fresh_keys[i]);
        x->h_levels.push_back(pl.fresh_levels[i]);
        x->h_upper_off.push_back(pl.fresh_off[i]);
    }
    x->level_rng = pl.rng;
    x->upper_used = pl.blocks;
    x->pub_inflight = n_new; // before staged drops: a lock-free reader may count these rows twice for a moment, never miss them
    x->staged -= n_new;

    // (5) device work
    const double t_committed = now_ms();
    if (flush_device_work(x, pl, old_count, d_slots, d_inflight, d_prof, max_warps, p)) {
        x->poisoned = t_error;
        cudaGetLastError();
        return 1;
    }
    if (prof_on)
        fprintf(stderr, "[gvss flush] entered at %.1f ms: %llu rows; take+plan+alloc %.1f ms, host mirrors %.1f ms, upload+build %.1f ms, left at %.1f ms\n",
                t_enter, (unsigned long long)n_new, t_planned - t_enter, t_committed - t_planned, now_ms() - t_committed, now_ms());
    return 0;
}

// ------------------------------------------------------------------------------------------------------------
// compact: index_gt::compact (index.hpp:3007-3096) through index_dense_gt::compact (index_dense.hpp:1479-1496)
// ------------------------------------------------------------------------------------------------------------
__global__ void k_permute_rows(const float *__restrict__ src, const uint32_t *__restrict__ old_of_new, uint64_t n, uint32_t dpad4,
                               float4 *__restrict__ dst) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * dpad4) return;
    const uint64_t r = i / dpad4;
    dst[i] = reinterpret_cast<const float4 *>(src)[(uint64_t)old_of_new[r] * dpad4 + (i - r * dpad4)];
}
// list `r` of the new layout = list src_row[r] of the old one with every slot renamed (sentinels stay)
__global__ void k_permute_lists(const uint32_t *__restrict__ src, const uint32_t *__restrict__ src_row, const uint32_t *__restrict__ new_of_old,
                                uint64_t rows, uint32_t stride, uint32_t *__restrict__ dst) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * stride) return;
    const uint64_t r = i / stride;
    const uint32_t s = src[(uint64_t)src_row[r] * stride + (i - r * stride)];
    dst[i] = s == kInvalid ? kInvalid : new_of_old[s];
}

// Re-orders the nodes exactly like the reference: every node's "cluster" is the result of the greedy descent for its own
// vector through levels max_level..1 (search_for_one_, index.hpp:3043-3046), nodes are sorted by (level descending, cluster
// ascending) with std::sort on the same array in the same initial order (so the permutation among equal elements is the
// reference's too), links are renamed, removed members are kept (this usearch version does not drop them). The stream
// written afterwards equals the reference's byte for byte (tests/test_gpu_build.py). Deliberate difference: the reference
// leaves its free-slot ring pointing at pre-compaction slots (latent bug); here the ring is renamed too.
static int compact_locked(gvss_index *x) {
    const uint64_t n = x->count;
    if (n < 2) return 0;
    // 1. clusters: k_search in descent-only mode, queries = the rows themselves
    uint32_t *d_cluster = nullptr;
    TempDev cluster_mem;
    CU(cluster_mem.alloc(d_cluster, n));
    {
        SearchParams p;
        uint32_t grid = 0;
        if (plan_search(x, x->d_vectors, n, 1, 1, nullptr, nullptr, nullptr, nullptr, p, grid)) return 1;
        p.query_stride = x->dpad;
        p.cluster_out = d_cluster;
        if (ensure_scratch(x, p.sl.bytes_per_warp * grid)) return 1;
        p.scratch = x->d_scratch;
        p.work_counter = x->d_work;
        CU(cudaMemsetAsync(x->d_work, 0, sizeof(unsigned long long), x->stream));
        CU(launch_search(metric_id(x->cfg.metric), view_of(x), p, grid, x->stream));
    }
    std::vector<uint32_t> cluster(n);
    CU(cudaMemcpyAsync(cluster.data(), d_cluster, n * 4, cudaMemcpyDeviceToHost, x->stream));
    CU(cudaStreamSynchronize(x->stream));
    // 2. the permutation (index.hpp:3059-3066)
    struct slot_level_t { uint32_t old_slot, cluster; int16_t level; };
    std::vector<slot_level_t> sl(n);
    for (uint64_t s = 0; s < n; s++) sl[s] = {(uint32_t)s, cluster[s], x->h_levels[s]};
    std::sort(sl.begin(), sl.end(), [](slot_level_t const &a, slot_level_t const &b) { return a.level == b.level ? a.cluster < b.cluster : a.level > b.level; });
    std::vector<uint32_t> old_of_new(n), new_of_old(n);
    for (uint64_t s = 0; s < n; s++) { old_of_new[s] = sl[s].old_slot; new_of_old[sl[s].old_slot] = (uint32_t)s; }
    // 3. new host mirrors and the row maps of the two list arrays
    std::vector<int64_t> keys(n);
    std::vector<int16_t> levels(n);
    std::vector<uint32_t> upper_off(n, kInvalid), upper_src;
    upper_src.reserve(x->upper_used);
    for (uint64_t s = 0; s < n; s++) {
        const uint32_t o = old_of_new[s];
        keys[s] = x->h_keys[o]; levels[s] = x->h_levels[o];
        if (levels[s] > 0) {
            upper_off[s] = (uint32_t)upper_src.size();
            for (int l = 0; l < levels[s]; l++) upper_src.push_back(x->h_upper_off[o] + (uint32_t)l);
        }
    }
    // 4. permuted copies on the device; the old arrays are released only when all of them exist
    const uint64_t cap = x->capacity, ublocks = upper_src.size();
    float *nv = nullptr; int64_t *nk = nullptr; int16_t *nl = nullptr; uint32_t *nn = nullptr, *nu = nullptr, *no = nullptr;
    uint32_t *d_oon = nullptr, *d_noo = nullptr, *d_usrc = nullptr;
    TempDev m1, m2, m3;
    auto drop_new = [&]() { cudaFree(nv); cudaFree(nk); cudaFree(nl); cudaFree(nn); cudaFree(nu); cudaFree(no); };
#define CUN(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { drop_new(); return fail("gvss_compact: CUDA error %s (%s)", cudaGetErrorString(e_), #call); } } while (0)
    CUN(m1.alloc(d_oon, n)); CUN(m2.alloc(d_noo, n)); CUN(m3.alloc(d_usrc, std::max<uint64_t>(ublocks, 1)));
    CUN(cudaMalloc(&nv, cap * x->dpad * 4ull)); CUN(cudaMalloc(&nk, cap * 8)); CUN(cudaMalloc(&nl, cap * 2));
    CUN(cudaMalloc(&nn, cap * x->stride0 * 4ull)); CUN(cudaMalloc(&no, cap * 4));
    CUN(cudaMalloc(&nu, std::max<uint64_t>(x->upper_cap, 1) * x->strideU * 4ull));
    cudaStream_t st = x->stream;
    CUN(cudaMemcpyAsync(d_oon, old_of_new.data(), n * 4, cudaMemcpyHostToDevice, st));
    CUN(cudaMemcpyAsync(d_noo, new_of_old.data(), n * 4, cudaMemcpyHostToDevice, st));
    if (ublocks) CUN(cudaMemcpyAsync(d_usrc, upper_src.data(), ublocks * 4, cudaMemcpyHostToDevice, st));
    CUN(cudaMemsetAsync(nv + n * x->dpad, 0, (cap - n) * x->dpad * 4ull, st));
    CUN(cudaMemsetAsync(nk + n, 0, (cap - n) * 8, st)); CUN(cudaMemsetAsync(nl + n, 0, (cap - n) * 2, st));
    CUN(cudaMemsetAsync(nn + n * x->stride0, 0xFF, (cap - n) * x->stride0 * 4ull, st));
    CUN(cudaMemsetAsync(no + n, 0xFF, (cap - n) * 4, st));
    CUN(cudaMemsetAsync(nu, 0xFF, std::max<uint64_t>(x->upper_cap, 1) * x->strideU * 4ull, st));
    const uint32_t dpad4 = x->dpad / 4;
    k_permute_rows<<<(unsigned)((n * dpad4 + 255) / 256), 256, 0, st>>>(x->d_vectors, d_oon, n, dpad4, reinterpret_cast<float4 *>(nv));
    k_permute_lists<<<(unsigned)((n * x->stride0 + 255) / 256), 256, 0, st>>>(x->d_nbr0, d_oon, d_noo, n, x->stride0, nn);
    if (ublocks) k_permute_lists<<<(unsigned)((ublocks * x->strideU + 255) / 256), 256, 0, st>>>(x->d_upper, d_usrc, d_noo, ublocks, x->strideU, nu);
    count_launch(ublocks ? 3 : 2);
    CUN(cudaGetLastError());
    CUN(cudaMemcpyAsync(nk, keys.data(), n * 8, cudaMemcpyHostToDevice, st));
    CUN(cudaMemcpyAsync(nl, levels.data(), n * 2, cudaMemcpyHostToDevice, st));
    CUN(cudaMemcpyAsync(no, upper_off.data(), n * 4, cudaMemcpyHostToDevice, st));
    CUN(cudaStreamSynchronize(st));
#undef CUN
    // 5. swap in
    cudaFree(x->d_vectors); cudaFree(x->d_keys); cudaFree(x->d_levels); cudaFree(x->d_nbr0); cudaFree(x->d_upper); cudaFree(x->d_upper_off);
    x->d_vectors = nv; x->d_keys = nk; x->d_levels = nl; x->d_nbr0 = nn; x->d_upper = nu; x->d_upper_off = no;
    x->h_keys.swap(keys); x->h_levels.swap(levels); x->h_upper_off.swap(upper_off);
    x->entry = new_of_old[x->entry];
    for (auto &s : x->free_slots) s = new_of_old[s];
    x->lookup_valid = false; x->key_to_slot.clear();
    return 0;
}

} // namespace gvss
