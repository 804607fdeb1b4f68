// K2/K3/K4 — bulk graph construction: one warp per inserted vector, many inserts in flight on the live graph.
// Restates index_gt::add / update (usearch index.hpp:2297-2378, 2405-2463) = connect_node_across_levels_ (:3227-3244):
// greedy descent (search_for_one_), per-level beam search (search_to_insert_, :3430-3496), the neighbour-selection
// heuristic (refine_, :3602-3638; connect_new_node_, :3247-3267) and reverse links with re-pruning
// (reconnect_neighbor_nodes_, :3270-3313), as driven by PhysicalCreateHNSWIndex (hnsw_index_physical_create.cpp:144-205).
//
// Concurrency model: the reference runs T host threads that take per-node spin locks while reading and writing
// neighbour lists. Here thousands of warps play the threads. Lists are read WITHOUT locks (every 4-byte entry is
// always a valid slot or the sentinel, so a torn read only yields a slightly stale list), and a node's lock is held
// only while its list is rewritten; a warp never holds two locks, so there is no lock-order deadlock.
// With one warp (concurrency 1) the kernel reproduces the reference's single-thread build byte for byte.
#include "kernels.h"
#include <cstdlib>

namespace gvss {

__device__ __forceinline__ void carve_build(WarpWs &w, unsigned char *smem, const SmemLayout &sm, uint8_t *scratch,
                                            const ScratchLayout &sl, int direct) {
    w.g4 = false; // the build stages whole rows (neighbour selection re-reads them); gather4 staging for its query-to-node
                  // distances was measured (+1.6 % at best, profiles/r3_build_ab.txt) and taken out again
    w.vmap = nullptr;
    w.q_a2 = 0.0f;
    w.prefetch_rows = false;
    w.prefetch_vis = (direct & 64) == 0;
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
    w.stage_rows = sm.stage_rows;
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

__device__ __forceinline__ void lock_node(uint32_t *locks, uint32_t s) {
    if (lane_id() == 0) {
        while (atomicCAS(&locks[s], 0u, 1u) != 0u) __nanosleep(64);
    }
    __syncwarp();
    __threadfence();
}
__device__ __forceinline__ void unlock_node(uint32_t *locks, uint32_t s) {
    __threadfence();
    __syncwarp();
    if (lane_id() == 0) atomicExch(&locks[s], 0u);
    __syncwarp();
}

// refine_ (index.hpp:3602-3638). Candidates = w.top_* [0, ncand) ascending. Keeps c iff for every already kept s:
// !(d(c, s) < d(c, q)). Lanes hold 32 consecutive candidates; each kept node costs one batch of distances.
template <int METRIC, int MODE>
__device__ __forceinline__ uint32_t refine(const GraphView &g, WarpWs &w, uint32_t ncand, uint32_t needed,
                                           uint32_t *kept_s, float *kept_d, uint32_t &staged) {
    const uint32_t lane = lane_id();
    if (ncand < needed) {
        for (uint32_t i = lane; i < ncand; i += 32) { kept_s[i] = w.top_s[i]; kept_d[i] = w.top_d[i]; }
        __syncwarp();
        return ncand;
    }
    uint32_t nkept = 0;
    for (uint32_t base = 0; base < ncand && nkept < needed; base += 32) {
        const uint32_t my = base + lane;
        bool alive = my < ncand;
        const uint32_t cs = alive ? w.top_s[my] : kInvalid;
        const float cd = alive ? w.top_d[my] : 0.0f;
        const uint32_t nprev = nkept;
        for (uint32_t k = 0; k < nprev; k++) {
            const uint32_t am0 = __ballot_sync(kFull, alive);
            if (!am0) break;
            w.n_pair += __popc(am0);
            const float d = warp_distances<METRIC, MODE>(g, w, LeftSrc{nullptr, kept_s[k]}, cs, alive, staged);
            if (alive && d < cd) alive = false;
        }
        uint32_t amask = __ballot_sync(kFull, alive);
        while (amask && nkept < needed) {
            const uint32_t j = __ffs(amask) - 1;
            const uint32_t sj = __shfl_sync(kFull, cs, j);
            const float dj = __shfl_sync(kFull, cd, j);
            if (lane == 0) { kept_s[nkept] = sj; kept_d[nkept] = dj; }
            __syncwarp();
            nkept++;
            if (nkept == needed) break;
            const bool act = alive && lane > j;
            const uint32_t am1 = __ballot_sync(kFull, act);
            w.n_pair += __popc(am1);
            if (am1) {
                // the kept candidate's own row usually sits in the stage already (row j, single-chunk rows): then the
                // step needs no memory access at all
                const bool resident = g.dpad <= (uint32_t)kChunk && w.stage_rows == 32 && __shfl_sync(kFull, staged, j) == sj;
                const LeftSrc left = resident ? LeftSrc{w.stage[0] + j * w.rstride, 0} : LeftSrc{nullptr, sj};
                const float d = warp_distances<METRIC, MODE>(g, w, left, cs, act, staged);
                if (act && d < cd) alive = false;
            }
            amask = __ballot_sync(kFull, alive && lane > j);
        }
    }
    return nkept;
}

// MODE 3 = the production configuration fixed at compile time (32-row stage, TMA bulk copies, prefetches as shipped): the A/B
// paths stay out of the kernel, which is otherwise large enough to stall on instruction fetch; MODE 0 = run-time knobs.
template <int METRIC, int MODE> __global__ void __launch_bounds__(32) k_build(GraphView g, BuildParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    WarpWs w;
    carve_build(w, smem, p.sm, p.scratch, p.sl, p.direct);
    const uint32_t lane = lane_id();
    const uint32_t cmax = max(g.M, g.M0);
    uint32_t *mine_s = reinterpret_cast<uint32_t *>(smem + p.sm.off_misc);
    uint32_t *kept_s = mine_s + (cmax + 1);
    float *kept_d = reinterpret_cast<float *>(kept_s + (cmax + 1));
    float *mine_d = kept_d + (cmax + 1);
    const uint32_t top_limit_all = p.sm.top_cap; // never reached: plain insert_reserved
    uint32_t staged = kInvalid;

    for (;;) {
        unsigned long long idx = 0;
        if (lane == 0) idx = atomicAdd(p.work_counter, 1ull);
        idx = __shfl_sync(kFull, idx, 0);
        if (idx >= p.end) break;
        const uint32_t slot = p.slots[idx];
        const int level = p.levels[slot];

        for (uint32_t i = lane; i < g.dpad; i += 32) w.q[i] = g.vectors[(size_t)slot * g.dpad + i];
        if (p.inflight && lane == 0) __stcg(p.inflight + blockIdx.x, slot);
        __syncwarp();

        unsigned long long es = 0;
        if (lane == 0) {
            es = __ldcg(p.entry_state);
            if (es == ~0ull) { // first node of the graph: index.hpp:2352-2357
                const unsigned long long mineEs = ((unsigned long long)(uint32_t)level << 32) | slot;
                const unsigned long long old = atomicCAS(p.entry_state, ~0ull, mineEs);
                es = (old == ~0ull) ? ~0ull : old;
            }
        }
        es = __shfl_sync(kFull, es, 0);
        if (es == ~0ull) { // this node became the first node of the graph
            if (p.inflight && lane == 0) __stcg(p.inflight + blockIdx.x, kInvalid);
            continue;
        }
        const int max_level = (int)(es >> 32);
        const uint32_t entry = (uint32_t)(es & 0xFFFFFFFFull);

        Counters cnt = {0, 0, 0, 0};
        w.n_pair = 0;
        w.n_qdist = 0;
        w.n_rows_written = 0;
        uint32_t n_qdist = 0, n_base = 0;
        long long t0 = clock64(), t_beam = 0, t_refine = 0, t_conn = 0, t_recon = 0;
        float closest_d;
        uint32_t closest = greedy_descent<METRIC, true, MODE>(g, w, entry, max_level, level, cnt, staged, closest_d);
        const long long t_greedy = clock64() - t0;

        for (int l = min(level, max_level); l >= 0; --l) {
            uint32_t top_size = 0;
            t0 = clock64();
            // list-as-queue first; a level that would overfill the visited table is redone with the exact bitmap; exact ties
            // switch to usearch's heap mid-flight (same scheme as k_search)
            bool lq = MODE == 3 || (p.direct & 8) == 0, bitmap = false;
            const uint32_t base_before = cnt.base;
            for (;;) {
                cnt.flags = 0;
                w.use_bitmap = bitmap;
                beam_search<METRIC, true, true, MODE>(g, w, closest, closest_d, p.ef_add, l, slot, cnt, top_size, staged, lq);
                if ((cnt.flags & 1u) && !bitmap) { bitmap = true; continue; }
                if (cnt.flags & 8u) { lq = false; continue; }
                break;
            }
            w.use_bitmap = false;
            if (l == 0) n_base += cnt.base - base_before;
            for (uint32_t i = lane; i < top_size; i += 32) w.top_s[i] &= kSlotMask; // drop the expanded flags
            __syncwarp();

            // Everything that follows on this level — the node's own list (connect_new_node_), the reverse links of its links
            // (reconnect_neighbor_nodes_) and, on level 0, the peer pass — is ONE loop with ONE refine site and one list-rewrite
            // site: the routines are large and every further inlined copy made the kernel stall on instruction fetch
            // (profiles/r3_build_ncu_summary.txt). Order of the link updates: own list; reverse links in list order; then the
            // peers. Peer pass: inserts that are in flight at the same time cannot find each other through the graph, so before
            // this node leaves it measures its distance to every vector other warps are inserting right now and links both ways
            // with those that are closer than its farthest own link and survive the heuristic against its kept links (same
            // append / re-prune rule as any reverse link). With one insert in flight the registry is empty and the build stays
            // byte-identical to the reference's single-thread build.
            const uint32_t cap_l = l ? g.M : g.M0; // own links: up to M on every level, clamped to the list size when M0 < M
            t_beam += clock64() - t0; t0 = clock64();
            const bool peers = l == 0 && p.inflight && gridDim.x > 1;
            bool own_done = false, second = false; // second: the slot <- xs half of a peer link is still to do
            uint32_t nk = 0, ri = 0, pbase = 0, cm = 0, px = kInvalid, xs = kInvalid;
            float a_worst = CUDART_INF_F, pd = 0.0f, dx = 0.0f;
            for (;;) {
                uint32_t target = slot, ncand, needed, capc = 0;
                uint32_t *clist = nullptr;
                if (!own_done) {
                    ncand = top_size; needed = g.M; // candidates = the beam search's result list
                } else {
                    uint32_t nbr;
                    float dtn;
                    int lvl = l;
                    bool count_dtn = false;
                    if (ri < nk) {
                        const uint32_t close = mine_s[ri++];
                        if (close == slot) continue;
                        float dn = warp_distances<METRIC, MODE>(g, w, LeftSrc{w.q, 0}, close, lane == 0, staged);
                        dn = __shfl_sync(kFull, dn, 0);
                        target = close; nbr = slot; dtn = dn; count_dtn = true;
                    } else if (second) {
                        target = slot; nbr = xs; dtn = dx; lvl = 0; second = false;
                    } else {
                        if (!peers) break;
                        bool found = false;
                        while (!found) {
                            if (!cm) { // next 32 registry entries
                                if (pbase >= gridDim.x) break;
                                const uint32_t wi = pbase + lane;
                                pbase += 32;
                                px = wi < gridDim.x ? __ldcg(p.inflight + wi) : kInvalid;
                                const bool act = px != kInvalid && px != slot && px < g.count;
                                const uint32_t pm = __ballot_sync(kFull, act);
                                if (!pm) continue;
                                n_qdist += __popc(pm);
                                pd = warp_distances<METRIC, MODE>(g, w, LeftSrc{w.q, 0}, px, act, staged);
                                cm = __ballot_sync(kFull, act && pd < a_worst);
                                continue;
                            }
                            const uint32_t j = __ffs(cm) - 1;
                            cm &= cm - 1;
                            xs = __shfl_sync(kFull, px, j);
                            dx = __shfl_sync(kFull, pd, j);
                            // the heuristic's own test for a candidate at distance dx: dropped if one of this node's links that
                            // is at least as close dominates it (refine_, index.hpp:3617-3627)
                            bool dominated = false;
                            for (uint32_t b2 = 0; b2 < nk; b2 += 32) {
                                const bool a2 = b2 + lane < nk && mine_d[b2 + lane] <= dx;
                                const uint32_t dm = __ballot_sync(kFull, a2);
                                if (!dm) continue;
                                w.n_pair += __popc(dm);
                                const float dxs = warp_distances<METRIC, MODE>(g, w, LeftSrc{nullptr, xs}, a2 ? mine_s[b2 + lane] : kInvalid, a2, staged);
                                dominated |= __ballot_sync(kFull, a2 && dxs < dx) != 0u;
                            }
                            found = !dominated;
                        }
                        if (!found) break;
                        target = xs; nbr = slot; dtn = dx; lvl = 0; second = true;
                    }
                    // one directed link update under the target's lock: append `nbr` to `target`'s list, or, when the list is
                    // full, re-run the heuristic over {(dtn, nbr)} U {(d(target, x), x)} with needed = capacity
                    // (reconnect_neighbor_nodes_, index.hpp:3279-3311)
                    clist = list_ptr(g, target, lvl, capc);
                    lock_node(p.locks, target);
                    uint32_t cntc = 0;
                    bool present = false;
                    for (uint32_t base = 0; base < capc; base += 32) {
                        const uint32_t e = base + lane < capc ? __ldcg(clist + base + lane) : kInvalid;
                        cntc += __popc(__ballot_sync(kFull, e != kInvalid));
                        present |= __ballot_sync(kFull, e == nbr) != 0;
                    }
                    if (present) { unlock_node(p.locks, target); continue; }
                    if (cntc < capc) {
                        if (lane == 0) clist[cntc] = nbr;
                        w.n_rows_written++;
                        unlock_node(p.locks, target);
                        continue;
                    }
                    uint32_t tsz = 0;
                    if (count_dtn) w.n_qdist++; // the reference measures d(target, new node) only here, when the list has to be re-pruned
                    top_insert(w, tsz, top_limit_all, dtn, nbr);
                    for (uint32_t base = 0; base < capc; base += 32) {
                        const uint32_t e = base + lane < capc ? __ldcg(clist + base + lane) : kInvalid;
                        const bool act = e != kInvalid;
                        const uint32_t am = __ballot_sync(kFull, act);
                        if (!am) break;
                        w.n_pair += __popc(am);
                        const float d = warp_distances<METRIC, MODE>(g, w, LeftSrc{nullptr, target}, e, act, staged);
                        for (uint32_t m = am; m; m &= m - 1) {
                            const uint32_t j = __ffs(m) - 1;
                            top_insert(w, tsz, top_limit_all, __shfl_sync(kFull, d, j), __shfl_sync(kFull, e, j));
                        }
                    }
                    ncand = tsz; needed = capc;
                }
                const uint32_t nk2 = refine<METRIC, MODE>(g, w, ncand, needed, kept_s, kept_d, staged); // the single site
                if (own_done) {
                    for (uint32_t j = lane; j < capc; j += 32) clist[j] = j < nk2 ? kept_s[j] : kInvalid;
                    w.n_rows_written++;
                    unlock_node(p.locks, target);
                    continue;
                }
                // connect_new_node_ (index.hpp:3247-3267)
                nk = min(nk2, cap_l);
                t_refine += clock64() - t0; t0 = clock64();
                for (uint32_t i = lane; i < nk; i += 32) { mine_s[i] = kept_s[i]; mine_d[i] = kept_d[i]; }
                __syncwarp();
                uint32_t capx;
                uint32_t *mylist = list_ptr(g, slot, l, capx);
                lock_node(p.locks, slot);
                {
                    // the list is normally blank; links appended meanwhile by concurrent inserts are kept behind ours
                    uint32_t extra = kInvalid;
                    if (lane < capx) extra = __ldcg(mylist + lane);
                    bool keep_extra = extra != kInvalid;
                    for (uint32_t i = 0; i < nk; i++) if (extra == mine_s[i]) keep_extra = false;
                    const uint32_t emask = __ballot_sync(kFull, keep_extra);
                    const uint32_t epos = nk + __popc(emask & ((1u << lane) - 1u));
                    __syncwarp();
                    for (uint32_t i = lane; i < capx; i += 32) mylist[i] = i < nk ? mine_s[i] : kInvalid;
                    __syncwarp();
                    if (keep_extra && epos < capx) mylist[epos] = extra;
                }
                unlock_node(p.locks, slot);
                w.n_rows_written++;
                if (nk) { closest = mine_s[0]; closest_d = mine_d[0]; } // the next level starts at this level's best link
                a_worst = nk ? kept_d[nk - 1] : CUDART_INF_F; // distance of this node's farthest own link
                own_done = true;
                t_conn += clock64() - t0; t0 = clock64();
            }
            t_recon += clock64() - t0;
        }
        if (p.inflight && lane == 0) __stcg(p.inflight + blockIdx.x, kInvalid);
        if (p.stats && lane == 0) {
            atomicAdd(p.stats + 0, 1ull);
            atomicAdd(p.stats + 1, (unsigned long long)(cnt.computed + n_qdist + w.n_qdist));
            atomicAdd(p.stats + 2, (unsigned long long)w.n_pair);
            atomicAdd(p.stats + 3, (unsigned long long)n_base);
            atomicAdd(p.stats + 4, (unsigned long long)(cnt.cycles - n_base));
            atomicAdd(p.stats + 5, (unsigned long long)w.n_rows_written);
        }
        if (p.prof && lane == 0) {
            atomicAdd(p.prof + 0, (unsigned long long)t_greedy); atomicAdd(p.prof + 1, (unsigned long long)t_beam);
            atomicAdd(p.prof + 2, (unsigned long long)t_refine); atomicAdd(p.prof + 3, (unsigned long long)t_conn);
            atomicAdd(p.prof + 4, (unsigned long long)t_recon); atomicAdd(p.prof + 5, 1ull);
        }
        // a node above the current top level becomes the entry point once it is connected: index.hpp:2372-2376
        if (level > max_level && lane == 0) {
            const unsigned long long mineEs = ((unsigned long long)(uint32_t)level << 32) | slot;
            unsigned long long cur = __ldcg(p.entry_state);
            while ((int)(cur >> 32) < level) {
                const unsigned long long old = atomicCAS(p.entry_state, cur, mineEs);
                if (old == cur) break;
                cur = old;
            }
        }
        __syncwarp();
    }
}

template <int METRIC, int MODE> static cudaError_t launch_build_tm(const GraphView &g, const BuildParams &p, uint32_t grid, cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(k_build<METRIC, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.sm.bytes);
    if (e != cudaSuccess) return e;
    k_build<METRIC, MODE><<<grid, 32, p.sm.bytes, st>>>(g, p);
    count_launch();
    return cudaGetLastError();
}
template <int METRIC> static cudaError_t launch_build_t(const GraphView &g, const BuildParams &p, uint32_t grid, cudaStream_t st) {
    // every knob at its default (p.direct == 0, 32-row stage): the specialised kernel; GVSS_BUILD_GENERIC=1 forces MODE 0
    static const bool generic = [] { const char *e = getenv("GVSS_BUILD_GENERIC"); return e && atoi(e); }();
    return (p.direct == 0 && p.sm.stage_rows == 32 && !generic) ? launch_build_tm<METRIC, 3>(g, p, grid, st) : launch_build_tm<METRIC, 0>(g, p, grid, st);
}

cudaError_t launch_build(int metric, const GraphView &g, const BuildParams &p, uint32_t grid, cudaStream_t st) {
    switch (metric) {
    case kCos: return launch_build_t<kCos>(g, p, grid, st);
    case kIp: return launch_build_t<kIp>(g, p, grid, st);
    default: return launch_build_t<kL2sq>(g, p, grid, st);
    }
}

} // namespace gvss
