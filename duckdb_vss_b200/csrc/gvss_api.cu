// libgvss host side: device-resident index container, usearch v2.12 stream (de)serialisation and the C ABI of
// include/gvss.h. Stands in for usearch's index_dense_gt<row_t> (index_dense.hpp:305-1996) behind
// HNSWIndex::index (reference src/include/hnsw/hnsw_index.hpp:28,39). No CPU fallback: every compute entry
// point needs a CUDA device and fails loudly without one.
#include "kernels.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <deque>
#include <mutex>
#include <string>
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
    uint8_t *d_scratch = nullptr;
    size_t scratch_bytes = 0;
    cudaEvent_t scratch_free = nullptr; // recorded after every launch that uses the scratch / work counter

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
    uint64_t level_rng = 1;          // minstd_rand0 state (std::default_random_engine, index.hpp:1892)

    // staging (gvss_add from T threads)
    uint32_t host_threads = 1;
    std::vector<std::vector<float>> st_vecs;
    std::vector<std::vector<int64_t>> st_keys;
    std::vector<std::pair<const float *, std::vector<int64_t>>> st_device; // rows already on the device
    std::vector<uint64_t> st_device_rows;
    std::atomic<uint64_t> staged{0};
    std::mutex mu; // serialises everything that touches the graph

    // reusable device buffers for the host-pointer search path
    float *d_q = nullptr; int64_t *d_ok = nullptr; float *d_od = nullptr; uint32_t *d_oc = nullptr; gvss_search_stats *d_os = nullptr;
    uint64_t io_nq = 0; uint32_t io_k = 0;
    float *h_pin = nullptr; size_t h_pin_bytes = 0;
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
static int ensure_scratch(gvss_index *x, size_t bytes) {
    if (bytes <= x->scratch_bytes) return 0;
    cudaFree(x->d_scratch); x->d_scratch = nullptr; x->scratch_bytes = 0;
    CU(cudaMalloc(&x->d_scratch, bytes));
    x->scratch_bytes = bytes;
    return 0;
}

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
// query-to-node distances: 0 = rows staged in shared memory by TMA bulk copies (default: faster once the base set
// exceeds L2, profiles/r1_search_ncu_summary.txt), 1 = rows streamed straight into registers (256-bit loads, no staging). GVSS_DIRECT_ROWS overrides for A/B measurements.
static bool list_queue() { // GVSS_LIST_QUEUE=0: k_search always replays usearch's heap (A/B, debugging)
    static bool on = [] { const char *e = getenv("GVSS_LIST_QUEUE"); return e ? atoi(e) != 0 : true; }();
    return on;
}
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

static int search_device_locked(gvss_index *x, const float *d_queries, uint64_t nq, uint32_t k, uint32_t ef,
                                int64_t *d_keys_out, float *d_dist_out, uint32_t *d_counts_out,
                                gvss_search_stats *d_stats_out, cudaStream_t st) {
    if (!nq) return 0;
    if (!ef) ef = 64;                        // default_expansion_search, index.hpp:1125,2493-2494
    const uint32_t limit = std::max(ef, k);  // index.hpp:2512
    if (limit > 16384) return fail("gvss: ef/k above 16384 not supported");
    SearchParams p{};
    p.queries = d_queries; p.nq = nq; p.k = k; p.limit = limit;
    p.keys_out = d_keys_out; p.dist_out = d_dist_out; p.counts_out = d_counts_out; p.stats_out = d_stats_out;
    p.entry = x->entry; p.max_level = x->max_level < 0 ? 0 : x->max_level;
    p.trace_query = -1;
    if (const char *e = getenv("GVSS_TRACE_QUERY")) p.trace_query = atoll(e);
    const uint32_t list_cap = std::max(x->cfg.connectivity_base, x->cfg.connectivity);
    p.sl = make_scratch_layout(limit, x->count, list_cap);
    p.direct = search_mode_bits();
    const bool g4 = setup_gather4(x, p);
    p.sm = make_smem_layout(x->dpad, std::max<uint32_t>(limit, 1), search_heap_smem(x, limit), 0, direct_rows() ? 0u : search_stage_rows(), false, list_cap, search_stage_chunk(), false, g4);
    uint32_t grid = resident_warps(x->device, p.sm.bytes, false);
    if (!grid) return fail("gvss: no CUDA device");
    static const uint32_t max_warps = [] { const char *e = getenv("GVSS_SEARCH_MAX_WARPS"); return e ? (uint32_t)atoi(e) : 0u; }();
    if (max_warps && grid > max_warps) grid = max_warps; // experiment knob: fewer resident queries = smaller visited-table footprint
    if (grid > nq) grid = (uint32_t)nq;
    if (ensure_scratch(x, p.sl.bytes_per_warp * grid)) return 1;
    p.scratch = x->d_scratch;
    p.work_counter = x->d_work;
    // searches issued on different streams share the scratch and the work counter: order them on the device
    if (!x->scratch_free) CU(cudaEventCreateWithFlags(&x->scratch_free, cudaEventDisableTiming));
    else CU(cudaStreamWaitEvent(st, x->scratch_free, 0));
    CU(cudaMemsetAsync(x->d_work, 0, sizeof(unsigned long long), st));
    GraphView g = view_of(x);
    CU(launch_search(metric_id(x->cfg.metric), g, p, grid, st));
    CU(cudaEventRecord(x->scratch_free, st));
    return 0;
}

} // namespace gvss

// ------------------------------------------------------------------------------------------------------------
// C ABI
// ------------------------------------------------------------------------------------------------------------
extern "C" {

const char *gvss_last_error(void) { return t_error.c_str(); }
const char *gvss_version(void) { return "gvss 0.1 (sm_100a; usearch v2.12 stream)"; }
uint64_t gvss_kernel_launches(void) { return g_launches.load(); }

int gvss_create(const gvss_config *config, gvss_index **out) {
    if (!config || !out) return fail("gvss_create: null argument");
    if (!config->dimensions) return fail("gvss_create: dimensions must be > 0");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail("gvss: no CUDA device available (this library has no CPU fallback)");
    auto *x = new gvss_index();
    x->cfg = *config;
    if (!x->cfg.metric) x->cfg.metric = GVSS_METRIC_L2SQ;
    if (x->cfg.metric != GVSS_METRIC_L2SQ && x->cfg.metric != GVSS_METRIC_COS && x->cfg.metric != GVSS_METRIC_IP) {
        delete x;
        return fail("gvss_create: unknown metric '%c'", config->metric);
    }
    if (!x->cfg.connectivity) x->cfg.connectivity = 16;                              // index.hpp:1097-1107
    if (!x->cfg.connectivity_base) x->cfg.connectivity_base = 2 * x->cfg.connectivity; // hnsw_index.cpp:178-187
    if (!x->cfg.expansion_add) x->cfg.expansion_add = 128;
    if (!x->cfg.expansion_search) x->cfg.expansion_search = 64;
    if (x->cfg.device < 0) cudaGetDevice(&x->device); else x->device = x->cfg.device;
    if (x->device >= ndev) { delete x; return fail("gvss_create: device %d out of range", x->device); }
    cudaSetDevice(x->device);
    x->dpad = (x->cfg.dimensions + 7) / 8 * 8; // 32-byte rows: TMA bulk copies need 16 B, the 256-bit row loads 32 B
    x->stride0 = (x->cfg.connectivity_base + 7) / 8 * 8;
    x->strideU = (x->cfg.connectivity + 7) / 8 * 8;
    if (cudaStreamCreateWithFlags(&x->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaMalloc(&x->d_entry_state, 8) != cudaSuccess || cudaMalloc(&x->d_work, 8) != cudaSuccess ||
        cudaMalloc(&x->d_build_stats, 48) != cudaSuccess || cudaMemset(x->d_build_stats, 0, 48) != cudaSuccess) {
        delete x;
        return fail("gvss_create: CUDA initialisation failed: %s", cudaGetErrorString(cudaGetLastError()));
    }
    x->st_vecs.resize(1); x->st_keys.resize(1);
    *out = x;
    return 0;
}

void gvss_destroy(gvss_index *x) {
    if (!x) return;
    cudaSetDevice(x->device);
    free_device(x);
    cudaFree(x->d_entry_state); cudaFree(x->d_work); cudaFree(x->d_build_stats);
    cudaFree(x->d_q); cudaFree(x->d_ok); cudaFree(x->d_od); cudaFree(x->d_oc); cudaFree(x->d_os);
    if (x->h_pin) cudaFreeHost(x->h_pin);
    if (x->scratch_free) cudaEventDestroy(x->scratch_free);
    if (x->stream) cudaStreamDestroy(x->stream);
    delete x;
}

static void reset_locked(gvss_index *x);
int gvss_reset(gvss_index *x) {
    if (!x) return fail("null index");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaSetDevice(x->device);
    reset_locked(x);
    return 0;
}
static void reset_locked(gvss_index *x) {
    free_device(x);
    x->h_keys.clear(); x->h_levels.clear(); x->h_upper_off.clear();
    x->entry = 0; x->max_level = -1; x->n_deleted = 0; x->lookup_valid = false; x->key_to_slot.clear(); x->free_slots.clear();
    cudaMemset(x->d_build_stats, 0, 48);
    for (auto &v : x->st_vecs) v.clear();
    for (auto &v : x->st_keys) v.clear();
    x->st_device.clear(); x->st_device_rows.clear();
    x->staged = 0; x->level_rng = 1;
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
    if (!x) return fail("null index");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaSetDevice(x->device);
    if (flush_locked(x)) return 1;
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
    const uint32_t threads = x->host_threads;
    reset_locked(x);                       // frees the old graph (removed members included)
    if (ensure_capacity(x, cap)) return 1; // same capacity as before
    x->host_threads = threads;
    if (m) {
        x->st_device.emplace_back(d_rows, std::move(keys));
        x->st_device_rows.push_back(m);
        x->staged += m;
        if (flush_locked(x)) return 1;     // bulk build of the live members only
    }
    return 0;
}

int gvss_reserve(gvss_index *x, uint64_t members, uint32_t host_threads) {
    if (!x) return fail("null index");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaSetDevice(x->device);
    if (host_threads == 0) host_threads = 1;
    if (host_threads > x->st_vecs.size()) { x->st_vecs.resize(host_threads); x->st_keys.resize(host_threads); }
    x->host_threads = std::max(x->host_threads, host_threads);
    x->level_rng = 1; // reserve() allocates fresh thread contexts => the level sequence restarts (index.hpp:2113,2124)
    return ensure_capacity(x, members);
}

int gvss_add(gvss_index *x, int64_t key, const float *vec, uint32_t host_thread) {
    if (!x || !vec) return fail("gvss_add: null argument");
    if (host_thread >= x->st_vecs.size()) host_thread = 0;
    // capacity check as usearch: nodes_count_.fetch_add(1) >= capacity fails (index.hpp:2329-2335); removed slots are reusable
    const uint64_t n = x->staged.fetch_add(1);
    if (x->count + n + 1 > x->capacity + x->free_slots.size()) {
        x->staged.fetch_sub(1);
        return fail("Reserve capacity ahead of insertions!");
    }
    auto &v = x->st_vecs[host_thread];
    v.insert(v.end(), vec, vec + x->cfg.dimensions);
    x->st_keys[host_thread].push_back(key);
    return 0;
}

int gvss_add_batch(gvss_index *x, const int64_t *keys, const float *vecs, uint64_t n) {
    if (!x || (!vecs && n)) return fail("gvss_add_batch: null argument");
    std::lock_guard<std::mutex> lk(x->mu);
    if (x->count + x->staged.load() + n > x->capacity + x->free_slots.size()) return fail("Reserve capacity ahead of insertions!");
    auto &v = x->st_vecs[0];
    v.insert(v.end(), vecs, vecs + n * x->cfg.dimensions);
    auto &k = x->st_keys[0];
    for (uint64_t i = 0; i < n; i++) k.push_back(keys ? keys[i] : (int64_t)(x->count + x->staged.load() + i));
    x->staged += n;
    return 0;
}

int gvss_add_batch_device(gvss_index *x, const int64_t *keys_host, const float *vecs_device, uint64_t n) {
    if (!x || (!vecs_device && n)) return fail("gvss_add_batch_device: null argument");
    std::lock_guard<std::mutex> lk(x->mu);
    if (x->count + x->staged.load() + n > x->capacity + x->free_slots.size()) return fail("Reserve capacity ahead of insertions!");
    std::vector<int64_t> k(n);
    for (uint64_t i = 0; i < n; i++) k[i] = keys_host ? keys_host[i] : (int64_t)(x->count + x->staged.load() + i);
    x->st_device.emplace_back(vecs_device, std::move(k));
    x->st_device_rows.push_back(n);
    x->staged += n;
    return 0;
}

int gvss_flush(gvss_index *x) {
    if (!x) return fail("null index");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaSetDevice(x->device);
    return flush_locked(x);
}

int gvss_set_build_concurrency(gvss_index *x, uint32_t n) {
    if (!x) return fail("null index");
    std::lock_guard<std::mutex> lk(x->mu);
    x->build_conc = n;
    return 0;
}

int gvss_remove(gvss_index *x, int64_t key, uint64_t *removed) {
    if (!x) return fail("null index");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaSetDevice(x->device);
    if (flush_locked(x)) return 1;
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
    if (removed) *removed = n;
    return 0;
}

int gvss_compact(gvss_index *x) {
    if (!x) return fail("null index");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaSetDevice(x->device);
    // index_gt::compact (index.hpp:3007-3096) only re-orders nodes for locality and keeps removed slots; slot order is
    // not observable through the extension, so the device layout is left as is. Pending rows are inserted.
    return flush_locked(x);
}

int gvss_search_device(gvss_index *x, const float *queries, uint64_t nq, uint32_t k, uint32_t ef, int64_t *keys_out,
                       float *dist_out, uint32_t *counts_out, gvss_search_stats *stats_out, void *cuda_stream) {
    if (!x) return fail("null index");
    if (nq && (!queries || !keys_out)) return fail("gvss_search_device: null buffer");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaSetDevice(x->device);
    if (flush_locked(x)) return 1;
    if (!k) { if (counts_out) CU(cudaMemsetAsync(counts_out, 0, nq * 4, (cudaStream_t)cuda_stream)); return 0; } // index.hpp:2489-2490
    return search_device_locked(x, queries, nq, k, ef, keys_out, dist_out, counts_out, stats_out, (cudaStream_t)cuda_stream);
}

int gvss_search_device_noalloc(gvss_index *x, const float *queries, uint64_t nq, uint32_t k, uint32_t ef, int64_t *keys_out,
                               float *dist_out, uint32_t *counts_out, gvss_search_stats *stats_out, void *cuda_stream) {
    if (!x) return fail("null index");
    if (nq && (!queries || !keys_out)) return fail("gvss_search_device_noalloc: null buffer");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaSetDevice(x->device);
    // inserting staged rows allocates; this entry point never does
    if (x->staged.load()) return fail("gvss_search_device_noalloc: rows are staged; call gvss_flush first (it allocates, this call must not)");
    cudaStream_t st = (cudaStream_t)cuda_stream;
    if (!nq) return 0;
    if (!k) { if (counts_out) CU(cudaMemsetAsync(counts_out, 0, nq * 4, st)); return 0; }
    if (!ef) ef = 64;
    const uint32_t limit = std::max(ef, k);
    if (limit > 16384) return fail("gvss: ef/k above 16384 not supported");
    SearchParams p{};
    p.queries = queries; p.nq = nq; p.k = k; p.limit = limit;
    p.keys_out = keys_out; p.dist_out = dist_out; p.counts_out = counts_out; p.stats_out = stats_out;
    p.entry = x->entry; p.max_level = x->max_level < 0 ? 0 : x->max_level;
    p.trace_query = -1;
    const uint32_t list_cap = std::max(x->cfg.connectivity_base, x->cfg.connectivity);
    p.sl = make_scratch_layout(limit, x->count, list_cap);
    p.direct = search_mode_bits();
    const bool g4 = setup_gather4(x, p);
    p.sm = make_smem_layout(x->dpad, std::max<uint32_t>(limit, 1), search_heap_smem(x, limit), 0, direct_rows() ? 0u : search_stage_rows(), false, list_cap, search_stage_chunk(), false, g4);
    const uint32_t resident = resident_warps(x->device, p.sm.bytes, false);
    if (!resident) return fail("gvss: no CUDA device");
    char err[384];
    err[0] = 0;
    if (launch_search_static_scratch(metric_id(x->cfg.metric), x->device, view_of(x), p, resident, st, err, sizeof err) != cudaSuccess)
        return fail("%s", err);
    return 0;
}

int gvss_search(gvss_index *x, const float *queries, uint64_t nq, uint32_t k, uint32_t ef, int64_t *keys_out,
                float *dist_out, uint32_t *counts_out, gvss_search_stats *stats_out) {
    if (!x) return fail("null index");
    if (nq && (!queries || (!keys_out && k))) return fail("gvss_search: null buffer");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaSetDevice(x->device);
    if (flush_locked(x)) return 1;
    if (!nq) return 0;
    if (!k) { if (counts_out) memset(counts_out, 0, nq * 4); return 0; }
    const uint32_t d = x->cfg.dimensions;
    if (nq > x->io_nq || k > x->io_k || !x->d_q) {
        cudaFree(x->d_q); cudaFree(x->d_ok); cudaFree(x->d_od); cudaFree(x->d_oc); cudaFree(x->d_os);
        x->d_q = nullptr; x->d_ok = nullptr; x->d_od = nullptr; x->d_oc = nullptr; x->d_os = nullptr;
        const uint64_t cq = std::max<uint64_t>(nq, x->io_nq);
        const uint32_t ck = std::max<uint32_t>(k, x->io_k);
        CU(cudaMalloc(&x->d_q, cq * d * 4)); CU(cudaMalloc(&x->d_ok, cq * ck * 8)); CU(cudaMalloc(&x->d_od, cq * ck * 4));
        CU(cudaMalloc(&x->d_oc, cq * 4)); CU(cudaMalloc(&x->d_os, cq * sizeof(gvss_search_stats)));
        x->io_nq = cq; x->io_k = ck;
    }
    CU(cudaMemcpyAsync(x->d_q, queries, nq * d * 4, cudaMemcpyHostToDevice, x->stream));
    if (search_device_locked(x, x->d_q, nq, k, ef, x->d_ok, x->d_od, x->d_oc, x->d_os, x->stream)) return 1;
    CU(cudaMemcpyAsync(keys_out, x->d_ok, nq * k * 8, cudaMemcpyDeviceToHost, x->stream));
    if (dist_out) CU(cudaMemcpyAsync(dist_out, x->d_od, nq * k * 4, cudaMemcpyDeviceToHost, x->stream));
    if (counts_out) CU(cudaMemcpyAsync(counts_out, x->d_oc, nq * 4, cudaMemcpyDeviceToHost, x->stream));
    std::vector<gvss_search_stats> tmp;
    gvss_search_stats *st = stats_out;
    if (!st) { tmp.resize(nq); st = tmp.data(); }
    CU(cudaMemcpyAsync(st, x->d_os, nq * sizeof(gvss_search_stats), cudaMemcpyDeviceToHost, x->stream));
    CU(cudaStreamSynchronize(x->stream));
    // a query whose visited table would have passed 75 % load was cut short: the synchronous path cannot return it
    for (uint64_t i = 0; i < nq; i++)
        if (st[i].flags & 1u) return fail("gvss_search: candidate queue overflow on query %llu (beam %u)", (unsigned long long)i, std::max(ef ? ef : 64u, k));
    return 0;
}

// ---- persistence -------------------------------------------------------------------------------------------------
int gvss_serialized_length(gvss_index *x, uint64_t *bytes) {
    if (!x || !bytes) return fail("null argument");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaSetDevice(x->device);
    if (flush_locked(x)) return 1;
    const uint64_t M = x->cfg.connectivity, M0 = x->cfg.connectivity_base;
    uint64_t n = 8 + x->count * x->cfg.dimensions * 4ull + 64 + 40 + 2 * x->count;
    for (uint64_t i = 0; i < x->count; i++) n += 10 + (4 + 4 * M0) + (uint64_t)x->h_levels[i] * (4 + 4 * M);
    *bytes = n;
    return 0;
}

int gvss_save(gvss_index *x, int (*write)(void *, const void *, size_t), void *user) {
    if (!x || !write) return fail("gvss_save: null argument");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaSetDevice(x->device);
    if (flush_locked(x)) return 1;
    const uint32_t d = x->cfg.dimensions, M = x->cfg.connectivity, M0 = x->cfg.connectivity_base;
    const uint64_t n = x->count;
    // (1) matrix header + vectors in slot order (index_dense.hpp:822-845)
    uint32_t hdr[2] = {(uint32_t)n, d * 4};
    if (!write(user, hdr, 8)) return fail("gvss_save: output callback failed");
    {
        const uint64_t rows_per = std::max<uint64_t>(1, (32ull << 20) / (d * 4ull));
        std::vector<float> buf(std::min(rows_per, std::max<uint64_t>(n, 1)) * d);
        for (uint64_t r = 0; r < n; r += rows_per) {
            const uint64_t m = std::min(rows_per, n - r);
            CU(cudaMemcpy2DAsync(buf.data(), d * 4ull, x->d_vectors + r * x->dpad, x->dpad * 4ull, d * 4ull, m, cudaMemcpyDeviceToHost, x->stream));
            CU(cudaStreamSynchronize(x->stream));
            if (!write(user, buf.data(), m * d * 4ull)) return fail("gvss_save: output callback failed");
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
    if (n) CU(cudaMemcpyAsync(nbr0.data(), x->d_nbr0, nbr0.size() * 4, cudaMemcpyDeviceToHost, x->stream));
    if (!upper.empty()) CU(cudaMemcpyAsync(upper.data(), x->d_upper, upper.size() * 4, cudaMemcpyDeviceToHost, x->stream));
    CU(cudaStreamSynchronize(x->stream));
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
}

int gvss_load(gvss_index *x, int (*read)(void *, void *, size_t), void *user) {
    if (!x || !read) return fail("gvss_load: null argument");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaSetDevice(x->device);
    uint32_t hdr[2];
    if (!read(user, hdr, 8)) return fail("gvss_load: failed to read the matrix header");
    const uint64_t n = hdr[0];
    if (hdr[1] != x->cfg.dimensions * 4) return fail("gvss_load: stream has %u bytes per vector, index expects %u", hdr[1], x->cfg.dimensions * 4);
    const uint32_t d = x->cfg.dimensions;
    // drop current contents
    free_device(x);
    x->h_keys.clear(); x->h_levels.clear(); x->h_upper_off.clear(); x->free_slots.clear(); x->key_to_slot.clear();
    x->lookup_valid = false; x->n_deleted = 0; x->entry = 0; x->max_level = -1; x->staged = 0;
    for (auto &v : x->st_vecs) v.clear();
    for (auto &v : x->st_keys) v.clear();
    if (ensure_capacity(x, std::max<uint64_t>(n, 1))) return 1;
    {
        const uint64_t rows_per = std::max<uint64_t>(1, (32ull << 20) / (d * 4ull));
        std::vector<float> buf(std::min(rows_per, std::max<uint64_t>(n, 1)) * d);
        for (uint64_t r = 0; r < n; r += rows_per) {
            const uint64_t m = std::min(rows_per, n - r);
            if (!read(user, buf.data(), m * d * 4ull)) return fail("gvss_load: truncated vectors");
            CU(cudaMemcpy2DAsync(x->d_vectors + r * x->dpad, x->dpad * 4ull, buf.data(), d * 4ull, d * 4ull, m, cudaMemcpyHostToDevice, x->stream));
            CU(cudaStreamSynchronize(x->stream));
        }
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
    x->cfg.metric = (char)head[13];
    uint64_t gh[5];
    if (!read(user, gh, 40)) return fail("gvss_load: truncated graph header");
    if (gh[0] != n) return fail("Index size and the number of vectors doesn't match");                     // :965-969
    if (gh[1] != x->cfg.connectivity || gh[2] != x->cfg.connectivity_base) {
        // the stream's connectivity wins (index.hpp:2779-2783); adjacency strides change => reallocate
        x->cfg.connectivity = (uint32_t)gh[1]; x->cfg.connectivity_base = (uint32_t)gh[2];
        const uint32_t s0 = (x->cfg.connectivity_base + 7) / 8 * 8, sU = (x->cfg.connectivity + 7) / 8 * 8;
        if (s0 != x->stride0) {
            cudaFree(x->d_nbr0); x->d_nbr0 = nullptr;
            x->stride0 = s0;
            CU(dev_resize(x->d_nbr0, 0, x->capacity * x->stride0, 0xFF, x->stream));
        }
        x->strideU = sU;
    }
    const uint32_t M = x->cfg.connectivity, M0 = x->cfg.connectivity_base;
    x->max_level = n ? (int32_t)gh[3] : -1;
    x->entry = (uint32_t)gh[4];
    x->h_levels.resize(n);
    if (n && !read(user, x->h_levels.data(), n * 2)) return fail("gvss_load: truncated levels");
    x->h_keys.resize(n); x->h_upper_off.assign(n, kInvalid);
    uint64_t blocks = 0;
    for (uint64_t i = 0; i < n; i++) if (x->h_levels[i] > 0) { x->h_upper_off[i] = (uint32_t)blocks; blocks += x->h_levels[i]; }
    if (ensure_upper(x, blocks)) return 1;
    x->upper_used = blocks;
    std::vector<uint32_t> nbr0(n * x->stride0, kInvalid), upper(blocks * x->strideU, kInvalid);
    std::vector<uint8_t> buf;
    uint64_t i = 0;
    while (i < n) { // read node tapes in ~8 MB pieces that end on node boundaries
        uint64_t j = i, bytes = 0;
        while (j < n && bytes < (8u << 20)) { bytes += 10 + (4 + 4ull * M0) + (uint64_t)x->h_levels[j] * (4 + 4ull * M); j++; }
        buf.resize(bytes);
        if (!read(user, buf.data(), bytes)) return fail("gvss_load: truncated node tapes");
        const uint8_t *p = buf.data();
        for (; i < j; i++) {
            int16_t lvl;
            memcpy(&x->h_keys[i], p, 8); memcpy(&lvl, p + 8, 2); p += 10;
            if (lvl != x->h_levels[i]) return fail("gvss_load: node %llu level mismatch", (unsigned long long)i);
            auto take = [&](uint32_t *row, uint32_t cap) {
                uint32_t cnt; memcpy(&cnt, p, 4);
                if (cnt > cap) cnt = cap;
                memcpy(row, p + 4, 4ull * cnt);
                p += 4 + 4ull * cap;
            };
            take(nbr0.data() + i * x->stride0, M0);
            for (int l = 1; l <= lvl; l++) take(upper.data() + ((uint64_t)x->h_upper_off[i] + l - 1) * x->strideU, M);
            if (x->h_keys[i] == GVSS_FREE_KEY) { x->n_deleted++; x->free_slots.push_back((uint32_t)i); } // reindex_keys_, index_dense.hpp:1901-1929
        }
    }
    if (n) {
        CU(cudaMemcpyAsync(x->d_nbr0, nbr0.data(), nbr0.size() * 4, cudaMemcpyHostToDevice, x->stream));
        if (!upper.empty()) CU(cudaMemcpyAsync(x->d_upper, upper.data(), upper.size() * 4, cudaMemcpyHostToDevice, x->stream));
        CU(cudaMemcpyAsync(x->d_keys, x->h_keys.data(), n * 8, cudaMemcpyHostToDevice, x->stream));
        CU(cudaMemcpyAsync(x->d_levels, x->h_levels.data(), n * 2, cudaMemcpyHostToDevice, x->stream));
        CU(cudaMemcpyAsync(x->d_upper_off, x->h_upper_off.data(), n * 4, cudaMemcpyHostToDevice, x->stream));
        CU(cudaStreamSynchronize(x->stream));
    }
    x->count = n;
    return 0;
}

// ---- introspection -----------------------------------------------------------------------------------------------
int gvss_get_info(gvss_index *x, gvss_info *out) {
    if (!x || !out) return fail("null argument");
    std::lock_guard<std::mutex> lk(x->mu);
    memset(out, 0, sizeof *out);
    const uint64_t staged = x->staged.load();
    out->size = x->count - x->n_deleted + staged; // size() counts rows whose add() has returned (SyncSize after the bulk build)
    out->capacity = x->capacity;
    out->node_count = x->count + staged;
    out->max_level = x->max_level < 0 ? 0 : (uint64_t)x->max_level;
    out->dimensions = x->cfg.dimensions;
    out->connectivity = x->cfg.connectivity; out->connectivity_base = x->cfg.connectivity_base;
    out->expansion_add = x->cfg.expansion_add; out->expansion_search = x->cfg.expansion_search;
    out->metric = x->cfg.metric;
    out->memory_usage = x->capacity * (x->dpad * 4ull + 8 + 2 + x->stride0 * 4ull + 4 + 4) + x->upper_cap * x->strideU * 4ull + x->scratch_bytes;
    return 0;
}

int gvss_get_build_stats(gvss_index *x, gvss_build_stats *out) {
    if (!x || !out) return fail("null argument");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaSetDevice(x->device);
    if (flush_locked(x)) return 1;
    unsigned long long h[6] = {};
    CU(cudaMemcpy(h, x->d_build_stats, sizeof h, cudaMemcpyDeviceToHost));
    out->inserts = h[0]; out->query_distances = h[1]; out->pair_distances = h[2];
    out->base_expansions = h[3]; out->upper_expansions = h[4]; out->rows_written = h[5];
    return 0;
}

int gvss_get_level_stats(gvss_index *x, uint32_t level, gvss_level_stats *out) {
    if (!x || !out) return fail("null argument");
    std::lock_guard<std::mutex> lk(x->mu);
    cudaSetDevice(x->device);
    if (flush_locked(x)) return 1;
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
}

int gvss_merge_topk_device(const int64_t *gk, const float *gd, uint32_t shards, uint64_t nq, uint32_t k, int64_t *keys_out,
                           float *dist_out, void *cuda_stream) {
    if (!gk || !gd || !keys_out || !dist_out) return fail("gvss_merge_topk_device: null buffer");
    CU(launch_merge_topk(gk, gd, shards, nq, k, keys_out, dist_out, (cudaStream_t)cuda_stream));
    return 0;
}

int gvss_merge_topk_peers_device(const int64_t *const *peer_keys, const float *const *peer_dist, uint32_t *const *peer_flags,
                                 uint32_t rank, uint32_t shards, uint32_t step, uint64_t nq, uint32_t k, int64_t *keys_out,
                                 float *dist_out, void *cuda_stream) {
    if (!peer_keys || !peer_dist || !peer_flags || !keys_out) return fail("gvss_merge_topk_peers_device: null buffer");
    if (!shards || shards > 16 || rank >= shards) return fail("gvss_merge_topk_peers_device: 1..16 shards, rank < shards");
    if ((uint64_t)shards * k > 256) return fail("gvss_merge_topk_peers_device: shards * k must not exceed 256");
    for (uint32_t s = 0; s < shards; s++)
        if (!peer_keys[s] || !peer_dist[s] || !peer_flags[s]) return fail("gvss_merge_topk_peers_device: null peer pointer (shard %u)", s);
    CU(launch_merge_topk_peers(peer_keys, peer_dist, peer_flags, rank, shards, step, nq, k, keys_out, dist_out, (cudaStream_t)cuda_stream));
    return 0;
}

int gvss_exact_topk_device(char metric, const float *base, uint64_t n, const int64_t *keys, const float *queries, uint64_t nq,
                           uint32_t d, uint32_t k, int64_t *ids_out, float *dist_out, void *cuda_stream) {
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
}

} // extern "C"

// ------------------------------------------------------------------------------------------------------------
// bulk insert driver
// ------------------------------------------------------------------------------------------------------------
namespace gvss {

static int flush_locked(gvss_index *x) {
    const uint64_t n_new = x->staged.load();
    if (!n_new) return 0;
    const uint32_t d = x->cfg.dimensions;
    // 1. assign slots: recycled slots first (FIFO), then fresh ones; thread-major staging order
    std::vector<uint32_t> slots;
    slots.reserve(n_new);
    uint64_t next_fresh = x->count;
    auto take_slot = [&](int64_t key) -> uint32_t {
        uint32_t s;
        if (!x->free_slots.empty()) { // index_dense.hpp:1766-1793: reuse a removed node through index_gt::update
            s = x->free_slots.front(); x->free_slots.pop_front();
            x->n_deleted--;
            x->h_keys[s] = key;
        } else {
            s = (uint32_t)next_fresh++;
            x->h_keys.push_back(key);
            x->h_levels.push_back(draw_level(x->level_rng, x->cfg.connectivity));
            x->h_upper_off.push_back(kInvalid);
        }
        if (x->lookup_valid) x->key_to_slot.emplace(key, s);
        return s;
    };
    if (next_fresh + n_new > x->capacity + x->free_slots.size()) return fail("Reserve capacity ahead of insertions!");
    // host-staged rows -> pinned buffer -> device rows
    std::vector<std::pair<uint32_t, const float *>> host_rows; // (slot, src)
    for (size_t t = 0; t < x->st_vecs.size(); t++) {
        const auto &keys = x->st_keys[t];
        for (size_t i = 0; i < keys.size(); i++) {
            const uint32_t s = take_slot(keys[i]);
            slots.push_back(s);
            host_rows.emplace_back(s, x->st_vecs[t].data() + i * d);
        }
    }
    std::vector<std::pair<uint32_t, std::pair<const float *, uint64_t>>> dev_runs; // first slot of a contiguous run, (src, rows)
    for (size_t b = 0; b < x->st_device.size(); b++) {
        const auto &keys = x->st_device[b].second;
        uint32_t first = kInvalid;
        bool contiguous = x->free_slots.empty();
        for (size_t i = 0; i < keys.size(); i++) {
            const uint32_t s = take_slot(keys[i]);
            if (i == 0) first = s;
            else if (s != first + i) contiguous = false;
            slots.push_back(s);
        }
        if (!contiguous) return fail("gvss_add_batch_device: cannot be mixed with recycled slots");
        dev_runs.push_back({first, {x->st_device[b].first, keys.size()}});
    }
    // fresh nodes: upper-level blocks
    uint64_t blocks = x->upper_used;
    for (uint64_t s = x->count; s < next_fresh; s++) if (x->h_levels[s] > 0) { x->h_upper_off[s] = (uint32_t)blocks; blocks += x->h_levels[s]; }
    if (ensure_upper(x, blocks)) return 1;
    x->upper_used = blocks;

    // 2. upload rows (16-byte padded), keys, levels, upper offsets; reset lists of recycled nodes
    if (!host_rows.empty()) {
        // staged host rows go up through two pinned bounce buffers: while one chunk is on its way to the GPU the next one is
        // padded into the other buffer (no pinned allocation of the size of the whole batch, copy and padding overlap)
        const size_t row_bytes = (size_t)x->dpad * 4;
        size_t chunk_rows = (16u << 20) / row_bytes;
        if (chunk_rows < 1) chunk_rows = 1;
        if (chunk_rows > host_rows.size()) chunk_rows = host_rows.size();
        const size_t need = 2 * chunk_rows * row_bytes;
        if (need > x->h_pin_bytes) {
            if (x->h_pin) cudaFreeHost(x->h_pin);
            x->h_pin = nullptr; x->h_pin_bytes = 0;
            CU(cudaMallocHost(&x->h_pin, need));
            x->h_pin_bytes = need;
        }
        cudaEvent_t done[2] = {nullptr, nullptr};
        struct EvGuard { cudaEvent_t *e; ~EvGuard() { for (int i = 0; i < 2; i++) if (e[i]) cudaEventDestroy(e[i]); } } guard{done};
        CU(cudaEventCreateWithFlags(&done[0], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&done[1], cudaEventDisableTiming));
        const bool padded = x->dpad != d;
        int b = 0;
        for (size_t c0 = 0; c0 < host_rows.size(); c0 += chunk_rows, b ^= 1) {
            const size_t c1 = std::min(host_rows.size(), c0 + chunk_rows);
            float *buf = x->h_pin + (size_t)b * chunk_rows * x->dpad;
            CU(cudaEventSynchronize(done[b])); // the copies that read this buffer two chunks ago (no-op for a fresh event)
            if (padded) memset(buf, 0, (c1 - c0) * row_bytes);
            for (size_t r = c0; r < c1; r++) memcpy(buf + (r - c0) * x->dpad, host_rows[r].second, d * 4ull);
            size_t i = c0;
            while (i < c1) { // runs of consecutive slots go up as one copy
                size_t j = i + 1;
                while (j < c1 && host_rows[j].first == host_rows[j - 1].first + 1) j++;
                CU(cudaMemcpyAsync(x->d_vectors + (size_t)host_rows[i].first * x->dpad, buf + (i - c0) * x->dpad, (j - i) * row_bytes, cudaMemcpyHostToDevice, x->stream));
                i = j;
            }
            CU(cudaEventRecord(done[b], x->stream));
        }
        CU(cudaEventSynchronize(done[0]));
        CU(cudaEventSynchronize(done[1]));
    }
    for (auto &run : dev_runs)
        CU(cudaMemcpy2DAsync(x->d_vectors + (size_t)run.first * x->dpad, x->dpad * 4ull, run.second.first, d * 4ull, d * 4ull, run.second.second, cudaMemcpyDeviceToDevice, x->stream));
    for (uint32_t s : slots) {
        if (s < x->count) { // recycled: memset of the tape keeping the level (index.hpp:2441-2444)
            CU(cudaMemsetAsync(x->d_nbr0 + (size_t)s * x->stride0, 0xFF, x->stride0 * 4ull, x->stream));
            if (x->h_levels[s] > 0) CU(cudaMemsetAsync(x->d_upper + (size_t)x->h_upper_off[s] * x->strideU, 0xFF, (size_t)x->h_levels[s] * x->strideU * 4ull, x->stream));
            CU(cudaMemcpyAsync(x->d_keys + s, &x->h_keys[s], 8, cudaMemcpyHostToDevice, x->stream));
        }
    }
    if (next_fresh > x->count) {
        const uint64_t b = x->count, m = next_fresh - b;
        CU(cudaMemcpyAsync(x->d_keys + b, x->h_keys.data() + b, m * 8, cudaMemcpyHostToDevice, x->stream));
        CU(cudaMemcpyAsync(x->d_levels + b, x->h_levels.data() + b, m * 2, cudaMemcpyHostToDevice, x->stream));
        CU(cudaMemcpyAsync(x->d_upper_off + b, x->h_upper_off.data() + b, m * 4, cudaMemcpyHostToDevice, x->stream));
    }

    // 3. graph construction: concurrency ramps with the graph size (a launch never runs more concurrent inserts than
    //    1/16 of the nodes already linked), then one persistent launch for the rest
    uint32_t *d_slots = nullptr;
    TempDev slots_mem, inflight_mem, prof_mem;
    CU(slots_mem.alloc(d_slots, slots.size()));
    CU(cudaMemcpyAsync(d_slots, slots.data(), slots.size() * 4, cudaMemcpyHostToDevice, x->stream));
    BuildParams p{};
    p.slots = d_slots; p.levels = x->d_levels; p.entry_state = x->d_entry_state; p.work_counter = x->d_work; p.locks = x->d_locks;
    p.ef_add = x->cfg.expansion_add;
    const uint32_t cmax = std::max(x->cfg.connectivity_base, x->cfg.connectivity);
    const uint32_t top_cap = std::max(cmax + 1, p.ef_add) + 1;
    p.sl = make_scratch_layout(std::max(p.ef_add, cmax), next_fresh, cmax);
    static const uint32_t build_rows = [] { const char *e = getenv("GVSS_BUILD_STAGE_ROWS"); const uint32_t v = e ? (uint32_t)atoi(e) : 32u; return v == 16u ? 16u : 32u; }();
    p.direct = search_mode_bits();
    // with the list-as-queue beam search the heap only serves exact ties: 32 entries (its 5 top levels) stay in shared memory
    const uint32_t build_heap = (p.direct & 8) ? default_heap_smem(p.ef_add) : 32u;
    p.sm = make_smem_layout(x->dpad, top_cap, build_heap, (2 * cmax + 2) * 8, build_rows, true, cmax);
    uint32_t max_warps = resident_warps(x->device, p.sm.bytes, true);
    if (x->build_conc && x->build_conc < max_warps) max_warps = x->build_conc;
    if (ensure_scratch(x, p.sl.bytes_per_warp * max_warps)) return 1;
    p.scratch = x->d_scratch;
    unsigned long long es = x->max_level < 0 ? ~0ull : (((unsigned long long)(uint32_t)x->max_level << 32) | x->entry);
    CU(cudaMemcpyAsync(x->d_entry_state, &es, 8, cudaMemcpyHostToDevice, x->stream));
    // GVSS_BUILD_PROF=1: per-phase SM clocks summed over all inserts, and wall time of the ramp vs the final launch (stderr)
    static const bool prof_on = [] { const char *e = getenv("GVSS_BUILD_PROF"); return e && atoi(e); }();
    unsigned long long *d_prof = nullptr;
    if (prof_on) { CU(prof_mem.alloc(d_prof, 8)); CU(cudaMemsetAsync(d_prof, 0, 64, x->stream)); }
    p.prof = d_prof;
    p.stats = x->d_build_stats;
    // in-flight registry for the peer pass of k_build (GVSS_BUILD_PEERS=0 disables it)
    static const bool peers_on = [] { const char *e = getenv("GVSS_BUILD_PEERS"); return e ? atoi(e) != 0 : true; }();
    uint32_t *d_inflight = nullptr;
    if (peers_on) {
        CU(inflight_mem.alloc(d_inflight, max_warps));
        CU(cudaMemsetAsync(d_inflight, 0xFF, (size_t)max_warps * 4, x->stream));
    }
    p.inflight = d_inflight;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;
    if (prof_on) { cudaEventCreate(&ev0); cudaEventCreate(&ev1); cudaEventCreate(&ev2); cudaEventRecord(ev0, x->stream); }
    uint64_t linked = x->count - 0; // nodes already in the graph
    // In-flight inserts are capped at (linked nodes / ramp). Vectors inserted at the same time cannot see each other through
    // the graph; k_build's peer pass links them explicitly, which makes the cap almost irrelevant for quality (recall equal to
    // the reference-built graph from ramp 2 to 256 at 30k and 1M vectors, profiles/build_ramp_r1.txt). 4 keeps the very first
    // launches small. GVSS_BUILD_RAMP overrides.
    uint64_t ramp = 4;
    if (const char *e = getenv("GVSS_BUILD_RAMP")) ramp = std::max<uint64_t>(1, strtoull(e, nullptr, 10));
    uint64_t done = 0;
    const uint64_t total = slots.size();
    const int metric = metric_id(x->cfg.metric);
    while (done < total) {
        const uint64_t have = linked + done;
        uint64_t conc = std::max<uint64_t>(1, have / ramp);
        conc = std::min<uint64_t>(conc, max_warps);
        uint64_t len = conc >= max_warps ? total - done : std::max<uint64_t>(conc * 2, 16);
        len = std::min(len, total - done);
        p.begin = (uint32_t)done; p.end = (uint32_t)(done + len);
        const unsigned long long start = done;
        CU(cudaMemcpyAsync(x->d_work, &start, 8, cudaMemcpyHostToDevice, x->stream));
        GraphView g = view_of(x);
        g.count = (uint32_t)next_fresh;
        g.has_deleted = 0;
        if (prof_on && conc >= max_warps) cudaEventRecord(ev1, x->stream);
        CU(launch_build(metric, g, p, (uint32_t)std::min<uint64_t>(conc, len), x->stream));
        done += len;
    }
    if (prof_on) {
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
    x->count = next_fresh;
    for (auto &v : x->st_vecs) v.clear();
    for (auto &v : x->st_keys) v.clear();
    x->st_device.clear(); x->st_device_rows.clear();
    x->staged = 0;
    return 0;
}

} // namespace gvss
