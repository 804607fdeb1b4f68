// K1 — batched HNSW search: one warp per query, persistent warps pulling queries from a work counter.
// Replaces index_gt::search (usearch index.hpp:2480-2534) = search_for_one_ (:3384-3422) + search_to_find_in_base_
// (:3504-3573) as called by index_dense_gt::search_ (index_dense.hpp:1797-1827) from HNSWIndex::InitializeScan /
// ExecuteMultiScan (hnsw_index.cpp:285-311, 353-367).
#include "kernels.h"
#include <cstdlib>

namespace gvss {

__device__ __forceinline__ void carve(WarpWs &w, unsigned char *smem, const SmemLayout &sm, uint8_t *scratch,
                                      const ScratchLayout &sl, int direct, const CUtensorMap *vmap) {
    w.g4 = (direct & 16) != 0;
    w.prefetch_rows = (direct & 32) == 0;
    w.prefetch_vis = (direct & 64) == 0;
    w.vmap = vmap;
    w.direct = (direct & 1) != 0;
    w.spec_prefetch = (direct & 2) != 0;
    w.use_tma = (direct & 4) == 0;
    w.mbar = reinterpret_cast<uint64_t *>(smem + sm.off_mbar);
    w.stage[0] = reinterpret_cast<float *>(smem + sm.off_stage0);
    w.stage[1] = reinterpret_cast<float *>(smem + sm.off_stage1);
    w.c = reinterpret_cast<float *>(smem + sm.off_c);
    w.q = reinterpret_cast<float *>(smem + sm.off_q);
    w.heap = reinterpret_cast<Cand *>(smem + sm.off_heap);
    w.top_d = reinterpret_cast<float *>(smem + sm.off_top_d);
    w.top_s = reinterpret_cast<uint32_t *>(smem + sm.off_top_s);
    w.rstride = sm.rstride;
    w.stage_rows = sm.stage_rows ? sm.stage_rows : 32u;
    w.chunk = sm.chunk;
    w.nbuf = sm.nbuf;
    w.nb = reinterpret_cast<uint32_t *>(smem + sm.off_nb);
    w.nd = reinterpret_cast<float *>(smem + sm.off_nd);
    w.heap_smem = sm.heap_smem;
    uint8_t *base = scratch + (size_t)blockIdx.x * sl.bytes_per_warp;
    w.vis = reinterpret_cast<uint32_t *>(base);
    w.vis_mask = sl.vis_slots - 1;
    w.vis_limit = sl.vis_limit;
    w.heap_ovf = reinterpret_cast<Cand *>(base + (size_t)sl.vis_slots * 4);
    w.heap_total = sm.heap_smem + sl.heap_ovf;
    w.bitmap = reinterpret_cast<uint32_t *>(base + (size_t)sl.vis_slots * 4 + (size_t)sl.heap_ovf * 8);
    w.bitmap_words = sl.bitmap_words;
    w.use_bitmap = false;
    w.phase[0] = w.phase[1] = 0;
    if (lane_id() == 0) {
        mbar_init(&w.mbar[0], 1);
        mbar_init(&w.mbar[1], 1);
        mbar_fence_init();
    }
    __syncwarp();
}

#ifndef GVSS_SEARCH_MIN_BLOCKS
#define GVSS_SEARCH_MIN_BLOCKS 24
#endif
// MODE: see warp_distances (hnsw_device.cuh). launch_search picks MODE 1 whenever the run-time configuration is the default one.
template <int METRIC, int MODE> __global__ void __launch_bounds__(32, MODE == 2 ? 8 : GVSS_SEARCH_MIN_BLOCKS) k_search(GraphView g, const __grid_constant__ SearchParams p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    WarpWs w;
    carve(w, smem, p.sm, p.scratch, p.sl, p.direct, &p.vmap);
    const uint32_t lane = lane_id();
    uint32_t staged = kInvalid;

    for (;;) {
        unsigned long long qi = 0;
        if (lane == 0) qi = atomicAdd(p.work_counter, 1ull);
        qi = __shfl_sync(kFull, qi, 0);
        if (qi >= p.nq) break;

        const float *qsrc = p.queries + qi * (p.query_stride ? p.query_stride : g.dims);
        for (uint32_t i = lane; i < g.dpad; i += 32) w.q[i] = i < g.dims ? __ldg(qsrc + i) : 0.0f;
        __syncwarp();
        if (METRIC == kCos && (MODE != 0 || w.g4)) w.q_a2 = fold_a2(w.q, g.dpad);

        Counters cnt = {0, 0, 0, (long long)qi == p.trace_query ? 0x80000000u : 0u};
        uint32_t top_size = 0;
        OPC(kOpQueries, 1);
#if GVSS_OPCOUNT
        for (int i = 0; i < kOpCount; i++) w.opc[i] = i == kOpQueries ? 1ull : 0ull;
#endif
        if (p.cluster_out) { // compact (index.hpp:3043-3046): only the node the descent through levels max_level..1 ends at
            float closest_d;
            const uint32_t closest = greedy_descent<METRIC, false, MODE>(g, w, p.entry, p.max_level, 0, cnt, staged, closest_d);
            if (lane == 0) p.cluster_out[qi] = closest;
            __syncwarp();
            continue;
        }
        if (g.count && p.k) {
            float closest_d;
            const uint32_t closest = greedy_descent<METRIC, false, MODE>(g, w, p.entry, p.max_level, 0, cnt, staged, closest_d);
            const Counters before = cnt;
            // list-as-queue first (hashed visited table); a query that would overfill the table is redone with the exact
            // bitmap (flag bit1); exact ties switch to usearch's heap mid-flight (flag bit2), or — if the operation log ran
            // out of room — repeat the query with the heap from the start
            const bool lq_ok = MODE != 0 || (p.direct & 8) == 0;
            bool lq = lq_ok, bitmap = false;
            OPCLK(t_q0);
            for (;;) {
                cnt = before;
                w.use_bitmap = bitmap;
                beam_search<METRIC, false, false, MODE>(g, w, closest, closest_d, p.limit, 0, kInvalid, cnt, top_size, staged, lq);
                if ((cnt.flags & 1u) && !bitmap) { bitmap = true; continue; }
                if (cnt.flags & 8u) { lq = false; continue; }
                break;
            }
            OPCLK(t_q1);
            OPC(kOpClkTotal, t_q1 - t_q0);
            w.use_bitmap = false;
            cnt.flags = (cnt.flags & ~0xAu) | (bitmap ? 2u : 0u) | (!lq && lq_ok ? 4u : 0u);
        }
        const uint32_t n = min(top_size, p.k); // top.shrink(wanted), index.hpp:2526-2527
        for (uint32_t i = lane; i < p.k; i += 32) {
            const bool have = i < n;
            p.keys_out[qi * p.k + i] = have ? __ldg(&g.keys[w.top_s[i] & kSlotMask]) : -1; // dump_to: slot -> key, index.hpp:2252-2258
            if (p.dist_out) p.dist_out[qi * p.k + i] = have ? w.top_d[i] : CUDART_INF_F;
        }
#if GVSS_OPCOUNT
        if (lane == 0) for (int i = 0; i < kOpCount; i++) atomicAdd(&g_opcount[i], w.opc[i]);
#endif
        if (lane == 0) {
            if ((cnt.flags & 1u) && p.fail_word) atomicMax(p.fail_word, qi + 1ull);
            if (p.counts_out) p.counts_out[qi] = n;
            if (p.stats_out) p.stats_out[qi] = gvss_search_stats{cnt.computed, cnt.cycles, cnt.base, cnt.flags & 0x7FFFFFFFu};
        }
        __syncwarp();
    }
}

SmemLayout make_smem_layout(uint32_t dpad, uint32_t top_cap, uint32_t heap_smem, uint32_t misc_bytes, uint32_t stage_rows,
                            bool with_left, uint32_t list_cap, uint32_t chunk, bool double_buffer, bool gather4, bool g4_alias, uint32_t g4_stages) {
    SmemLayout s{};
    if (g4_stages != 4) g4_stages = 1;
    if (chunk == 0 || chunk > (uint32_t)kChunk) chunk = (uint32_t)kChunk;
    if (gather4) { chunk = kG4Chunk; stage_rows = 32; double_buffer = false; }
    const uint32_t chunkf = dpad < chunk ? dpad : chunk;
    s.chunk = chunkf;
    s.rstride = chunkf + (((chunkf / 4) % 2 == 0) ? 4u : 0u); // odd number of 16-byte granules per row
    if (gather4) s.rstride = kG4Chunk;                         // dense 128-byte rows, conflicts removed by the TMA swizzle
    s.stage_rows = stage_rows;
    const uint32_t stage_bytes = stage_rows * s.rstride * 4u * (gather4 ? g4_stages : 1u);
    const bool two = double_buffer && dpad > chunkf;
    s.nbuf = two ? 2u : 1u;
    uint32_t off = 0;
    auto take = [&](uint32_t bytes, uint32_t align) { off = (off + align - 1) / align * align; uint32_t o = off; off += bytes; return o; };
    const bool stage_first = gather4 || g4_alias;
    if (stage_first) s.off_stage0 = take(stage_bytes < 4096u && g4_alias ? 4096u : stage_bytes, 1024); // swizzle atoms are 1024-byte aligned: first in the carve-out
    s.off_mbar = take(16, 16);
    if (!stage_first) s.off_stage0 = take(stage_bytes, 128);
    s.off_stage1 = two ? take(stage_bytes, 128) : s.off_stage0;
    s.off_c = take(with_left ? (two ? 2u : 1u) * kChunk * 4 : 0, 128); // streamed left-hand chunk, double-buffered with the stage
    s.off_q = take(dpad * 4, 16);
    s.heap_smem = heap_smem;
    s.off_heap = take(heap_smem * 8, 16);
    s.top_cap = top_cap;
    s.off_top_d = take(top_cap * 4, 16);
    s.off_top_s = take(top_cap * 4, 16);
    const uint32_t lc = (list_cap + 31) / 32 * 32;
    s.off_nb = take(lc * 4, 16);
    s.off_nd = take(lc * 4, 16);
    s.off_misc = take(misc_bytes, 16);
    s.bytes = (off + 127) / 128 * 128;
    return s;
}

// Queue entries kept in shared memory. Live entries (distance <= radius) are the not-yet-expanded part of the result
// list plus ties, in practice well under half the beam, so beam/2 + 64 (at least 256) is enough once dead entries are
// pruned when it fills up (heap_prune); measured best among 128/256/512 at beam 384 (profiles/r1_search_ncu_summary.txt).
uint32_t default_heap_smem(uint32_t beam) {
    uint32_t h = beam / 2 + 64;
    if (h < 256) h = 256;
    return (h + 31) / 32 * 32;
}

ScratchLayout make_scratch_layout(uint32_t beam, uint64_t nodes, uint32_t list_cap) {
    // A search expands about `beam` nodes and meets ~0.6 new members per listed neighbour on the measured data sets
    // (1M x 128 at M0=64: 14.2k visited at beam 384; 1M x 768: 2.6k at beam 64), so beam * list_cap + 2048 slots keep the
    // table around 45 % load; the kernel flags a query that would exceed 75 % and redoes it with the exact bitmap.
    ScratchLayout s{};
    static const uint32_t factor = [] { const char *e = getenv("GVSS_VIS_FACTOR"); return e ? (uint32_t)atoi(e) : 0u; }();
    const uint32_t per_beam = factor ? factor : (list_cap < 16u ? 16u : list_cap);
    uint64_t want64 = (uint64_t)beam * per_beam + 2048u;
    if (want64 > (1ull << 30)) want64 = 1ull << 30;
    uint32_t want = (uint32_t)want64;
    if (want < 4096u) want = 4096u;
    uint32_t v = 4096u;
    while (v < want) v <<= 1;
    s.vis_slots = v;
    s.vis_limit = v / 4 * 3 - 32;
    s.heap_ovf = v; // every queue entry was a newly visited node, so the queue cannot outgrow the hashed table
    s.bitmap_words = (uint32_t)((nodes + 127) / 128 * 4);
    s.bytes_per_warp = (size_t)v * 4 + (size_t)s.heap_ovf * 8 + (size_t)s.bitmap_words * 4;
    return s;
}

uint32_t resident_warps(int device, uint32_t smem_bytes, bool build) {
    // cudaGetDeviceProperties costs milliseconds: ask once per device
    static size_t smem_per_sm[64];
    static int sm_count[64];
    if (device < 0 || device >= 64) return 0;
    if (!sm_count[device]) {
        int sms = 0, smem = 0;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return 0;
        if (cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerMultiprocessor, device) != cudaSuccess) return 0;
        smem_per_sm[device] = (size_t)smem;
        sm_count[device] = sms;
    }
    uint32_t by_smem = (uint32_t)(smem_per_sm[device] / (smem_bytes + 1024)); // 1 KB is reserved per CTA
    // 32 = CTA-per-SM limit; build warps also hold a node lock at times; k_search is compiled for 80 registers
    // (__launch_bounds__(32, GVSS_SEARCH_MIN_BLOCKS)): 65536 / (80 * 32) = 25 warps fit the register file
#ifndef GVSS_SEARCH_WARPS_CAP
#define GVSS_SEARCH_WARPS_CAP 25
#endif
    const uint32_t cap = build ? 16u : (uint32_t)GVSS_SEARCH_WARPS_CAP;
    if (by_smem > cap) by_smem = cap;
    if (by_smem < 1) by_smem = 1;
    return by_smem * (uint32_t)sm_count[device];
}

cudaError_t launch_search(int metric, const GraphView &g, const SearchParams &p, uint32_t grid, cudaStream_t st);

template <int METRIC, int MODE> static cudaError_t launch_search_tm(const GraphView &g, const SearchParams &p, uint32_t grid, cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(k_search<METRIC, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.sm.bytes);
    if (e != cudaSuccess) return e;
    k_search<METRIC, MODE><<<grid, 32, p.sm.bytes, st>>>(g, p);
    count_launch();
    return cudaGetLastError();
}
template <int METRIC> static cudaError_t launch_search_t(const GraphView &g, const SearchParams &p, uint32_t grid, cudaStream_t st) {
    // p.direct == 16: gather4 staging and nothing else switched away from its default (GVSS_SEARCH_GENERIC=1 forces MODE 0)
    static const bool generic = [] { const char *e = getenv("GVSS_SEARCH_GENERIC"); return e && atoi(e); }();
    if ((p.direct & ~128) == 16 && !generic) // bit7: latency kernel (the host laid out a 16 KB stage)
        return (p.direct & 128) ? launch_search_tm<METRIC, 2>(g, p, grid, st) : launch_search_tm<METRIC, 1>(g, p, grid, st);
    return launch_search_tm<METRIC, 0>(g, p, grid, st);
}

// ---------------------------------------------------------------------------------------------------------------
// Search variant whose scratch memory is a fixed array inside this module instead of a cudaMalloc'ed buffer.
//
// Who needs it: callers that must not allocate device memory while they search — a host process that has already
// handed every free byte of HBM to its own pool (a cudaMalloc from here would fail, or force that pool to give
// memory back), and callers on a latency-critical path, because cudaMalloc / cudaFree synchronise the whole device
// and would stall work the caller has queued on other streams.
//
// The array is shared by every call on that GPU and keeps whatever the previous call left in it: the kernel clears
// each visited table before it uses it and the work counter is zeroed on the stream before every launch. Two calls
// that overlap on the GPU (different streams, or different host threads) would use the same tables at the same
// time and corrupt each other's results — callers must serialise them.
// ---------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[GVSS_STATIC_SCRATCH_BYTES];

// The name above is only a host-side handle; the kernel needs the device address, resolved once per device.
static unsigned char *g_static_scratch_ptr[64];

cudaError_t launch_search_static_scratch(int metric, int device, const GraphView &g, SearchParams p, uint32_t resident,
                                         cudaStream_t st, char *err, size_t err_len) {
    if (device < 0 || device >= 64) { snprintf(err, err_len, "device ordinal %d out of range", device); return cudaErrorInvalidDevice; }
    if (!g_static_scratch_ptr[device]) { // first call on this device (that device is current): do it before any stream capture
        void *ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
        if (e != cudaSuccess) { snprintf(err, err_len, "cudaGetSymbolAddress(static scratch) failed: %s", cudaGetErrorString(e)); return e; }
        g_static_scratch_ptr[device] = static_cast<unsigned char *>(ptr);
    }
    const size_t header = 256; // work counter
    const size_t avail = (size_t)GVSS_STATIC_SCRATCH_BYTES - header;
    if (p.sl.bytes_per_warp > GVSS_STATIC_SLOT_MAX_BYTES) {
        snprintf(err, err_len,
                 "gvss_search_device_noalloc: one query slot needs %zu bytes of scratch (beam %u, %u nodes), the built-in "
                 "scratch array allows %u per slot; use a smaller ef/k or gvss_search_device",
                 p.sl.bytes_per_warp, p.limit, g.count, (unsigned)GVSS_STATIC_SLOT_MAX_BYTES);
        return cudaErrorInvalidValue;
    }
    uint32_t grid = (uint32_t)(avail / p.sl.bytes_per_warp);
    if (grid > resident) grid = resident;
    if (grid > p.nq) grid = (uint32_t)p.nq;
    p.work_counter = reinterpret_cast<unsigned long long *>(g_static_scratch_ptr[device]);
    p.scratch = g_static_scratch_ptr[device] + header;
    cudaError_t e = cudaMemsetAsync(p.work_counter, 0, sizeof(unsigned long long), st);
    if (e == cudaSuccess) e = launch_search(metric, g, p, grid, st);
    if (e != cudaSuccess) snprintf(err, err_len, "gvss_search_device_noalloc: launch failed: %s", cudaGetErrorString(e));
    return e;
}

#if GVSS_OPCOUNT
// debug hook (not part of the ABI): read and reset the operation counters of an OPCOUNT build
extern "C" int gvssdbg_opcounts(unsigned long long *out, int n) {
    unsigned long long h[kOpCount] = {};
    if (cudaMemcpyFromSymbol(h, g_opcount, sizeof(h)) != cudaSuccess) return -1;
    for (int i = 0; i < n && i < kOpCount; i++) out[i] = h[i];
    unsigned long long z[kOpCount] = {};
    cudaMemcpyToSymbol(g_opcount, z, sizeof(z));
    return kOpCount;
}
#endif

cudaError_t launch_search(int metric, const GraphView &g, const SearchParams &p, uint32_t grid, cudaStream_t st) {
    switch (metric) {
    case kCos: return launch_search_t<kCos>(g, p, grid, st);
    case kIp: return launch_search_t<kIp>(g, p, grid, st);
    default: return launch_search_t<kL2sq>(g, p, grid, st);
    }
}

} // namespace gvss
