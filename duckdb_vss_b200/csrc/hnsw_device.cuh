// Warp-level device primitives of the HNSW hot path (sm_100a). One warp owns one query / one insert.
//
// Numerics contract: every distance is a strict left-to-right fp32 sum with separate multiply and add
// (__fmul_rn/__fadd_rn or the packed sub.rn.f32x2 / mul.rn.f32x2, never contracted into FMA) computed by ONE lane over a
// whole row, so its bits equal the reference's scalar loops on x86-64 (usearch index_plugins.hpp:981-993 ip, :1006-1026
// cos, :1037-1052 l2sq). Rows are brought from HBM into shared memory by TMA, completion on an mbarrier:
//   * search: warp_distances_g4 — cp.async.bulk.tensor ... tile::gather4 (one instruction = the same 128-byte line of four
//     arbitrary rows) into a 128-byte-swizzled 4 KB stage; lane i <-> row i for all 32 lanes;
//   * build (and GVSS_GATHER4=0): warp_distances — one 1-D bulk copy (cp.async.bulk) per row chunk into rows with an odd
//     number of 16-byte granules; coalesced 16-byte cp.async with GVSS_USE_TMA=0 (A/B only).
// Both layouts make the per-lane LDS.128 reads bank-conflict free.
//
// Control flow contract (beam_search): the result list (`top`) is usearch's sorted buffer (index.hpp:655-766) and the
// acceptance test replays search_to_find_in_base_ / search_to_insert_ (:3504-3573, :3430-3496) for a whole batch of
// neighbours at once. The candidate queue (`next`, max_heap_gt :516-645) is not materialised on the fast path: the result
// list doubles as the queue, and usearch's binary heap is rebuilt from a log of the queue operations only when two live
// candidates tie exactly for the head of the queue. Results, their order among ties and the traversal counters are
// identical to the reference on the same graph.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>
#include <stdio.h>

// Alternative row path kept for A/B measurements only (build with -DGVSS_DIRECT_PATH=1, run with GVSS_DIRECT_ROWS=1):
// every lane streams its own row into registers with 256-bit loads, no shared-memory staging. It costs 128 registers per
// thread and lost to TMA staging once the base set exceeds L2 (profiles/r1_search_ncu_summary.txt).
#ifndef GVSS_DIRECT_PATH
#define GVSS_DIRECT_PATH 0
#endif

namespace gvss {

constexpr uint32_t kInvalid = 0xFFFFFFFFu; // adjacency sentinel / empty visited slot
constexpr uint32_t kFull = 0xFFFFFFFFu;
constexpr int kChunk = 128;                // floats of a row staged per step (512 B)

enum Metric : int { kL2sq = 0, kCos = 1, kIp = 2 };

// Device view of the index. Adjacency rows are sentinel-terminated (kInvalid), so a level-0 list of M0=32
// neighbours is exactly one 128-byte line and needs no separate count.
struct GraphView {
    const float *vectors;      // [capacity][dpad] fp32, rows padded with zeros to a multiple of 4 floats (16 B)
    const int64_t *keys;       // [capacity]
    uint32_t *nbr0;            // [capacity][stride0]
    uint32_t *upper;           // [upper blocks][strideU]
    const uint32_t *upper_off; // [capacity] first upper block of the node (levels 1..L are consecutive blocks)
    uint32_t dpad, dims, M, M0, stride0, strideU;
    uint32_t count;
    uint32_t has_deleted;      // any key == FREE_KEY present (skips the predicate load otherwise)
};

struct Cand {
    float d;
    uint32_t s;
};

// Operation counts for kernel tuning (build with -DGVSS_OPCOUNT=1; read with the debug hook gvssdbg_opcounts). Off in the
// shipped library: OPC() compiles to nothing.
#ifndef GVSS_OPCOUNT
#define GVSS_OPCOUNT 0
#endif
#ifndef GVSS_VSHIFT
#define GVSS_VSHIFT 1 // result-list insertion moves 128-element blocks (4 per lane) instead of 32-element windows (+5 % QPS)
#endif
#if GVSS_OPCOUNT
enum { kOpQueries, kOpHops, kOpNew, kOpHopsLe16, kOpHopsNone, kOpPushes, kOpPrunes, kOpPruneKept, kOpInsBatches, kOpInserted,
       kOpShifted, kOpVisRounds, kOpMaybeBatches, kOpPushLevels, kOpPopLevels, kOpClkSelect, kOpClkGather, kOpClkDist, kOpClkMerge,
       kOpClkStage, kOpClkWait, kOpClkMath, kOpSteps, kOpClkTotal, kOpClkAccept, kOpClkShift, kOpClkAdj, kOpClkVis, kOpClkShiftLoop, kOpPredHit, kOpCount };
static __device__ unsigned long long g_opcount[kOpCount];
// per-warp, flushed once per query by k_search; GVSS_OPCOUNT=2 keeps only the phase clocks (less distortion)
#define OPC(i, v) do { if (GVSS_OPCOUNT == 1 || ((i) >= kOpClkSelect && (i) != kOpPredHit) || (i) == kOpQueries) w.opc[i] += (unsigned long long)(v); } while (0)
#define OPCLK(var) const long long var = clock64()
#else
#define OPC(i, v) do { } while (0)
#define OPCLK(var) do { } while (0)
#endif

// Per-warp workspace: shared-memory carve-out + global scratch (visited table, queue overflow).
struct WarpWs {
    float *q;            // dpad floats: the query / the vector being inserted
    float *c;            // 2 × kChunk floats: streamed "left-hand" row chunk for node-to-node distances
    float *stage[2];     // 2 × stage_rows × rstride floats (second buffer only when dpad > kChunk)
    uint32_t rstride;    // floats between staged rows (odd multiple of 4 floats -> conflict-free LDS.128)
    uint32_t stage_rows; // rows of one stage buffer: 32 (lane <-> row) or 16
    uint32_t chunk;      // floats of a row staged per step (<= kChunk): 64 halves the footprint of a 32-row stage
    uint32_t nbuf;       // 2 = chunks double-buffered (rows longer than one chunk), 1 = chunks staged one after the other
    bool use_tma;        // rows staged by TMA bulk copies (cp.async.bulk) or by coalesced 16-byte cp.async
    bool spec_prefetch;  // prefetch every neighbour's row to L2 before the visited test resolves (trades HBM traffic for latency)
    bool direct;         // query-to-node distances stream each lane's row straight into registers (no staging)
    bool g4;             // query-to-node distances stage 32 rows x 32 floats per step by TMA gather4 (stage[0]: 4 KB, 1024-aligned)
    const CUtensorMap *vmap; // ... through this 2-D map over `vectors`
    float q_a2;          // cosine, g4: sum of squares of the query, folded once per query in the reference's order (identical bits)
    bool prefetch_rows;  // g4: L2 prefetch of every line of every new row at the start of a hop
    bool prefetch_vis;   // L2 prefetch of the visited windows of neighbours 32.. while neighbours 0..31 are tested
    uint32_t *nb;        // new (unvisited) neighbours of the node being expanded, in list order
    float *nd;           // ... and their distances
    float *top_d;        // result list, ascending
    uint32_t *top_s;
    Cand *heap;          // candidate queue: first heap_smem entries in shared memory ...
    Cand *heap_ovf;      // ... the rest in global scratch
    uint32_t heap_smem, heap_total;
    uint32_t *vis;       // visited table (open addressing) in global scratch
    uint32_t vis_mask, vis_limit;
    uint32_t *bitmap;    // exact fallback: one bit per node, used when a search would overfill the table
    uint32_t bitmap_words;
    bool use_bitmap;
#if GVSS_OPCOUNT
    unsigned long long opc[kOpCount];
#endif
    uint32_t n_pair, n_qdist, n_rows_written; // build statistics of this insert: node-to-node distances, extra query-to-node distances, adjacency rows written
    uint64_t *mbar;      // 2 mbarriers (one per stage buffer)
    uint32_t phase[2];
};

struct Counters {
    uint32_t computed, cycles, base;
    uint32_t flags; // bit0: visited table / queue overflow; bit31: trace this query (debug aid, GVSS_TRACE_QUERY)
};


__device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---------------------------------------------------------------------------------------------------------
// async row staging
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// TMA 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_row(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// TMA gather: the same 128-byte column window [col, col+32) of FOUR arbitrary rows of a 2-D tensor in one instruction
// (SASS UTMALDG...GATHER4); the rows land 128 B apart, 16-byte granules XOR-swizzled with the row index (SWIZZLE_128B).
// A row index outside the tensor reads as zeros without touching memory.
__device__ __forceinline__ void tma_gather4(void *dst, const CUtensorMap *map, int col, int r0, int r1, int r2, int r3,
                                            uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(smem_u32(bar)), "r"(col), "r"(r0), "r"(r1), "r"(r2), "r"(r3)
        : "memory");
}
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ void cp_async16(void *dst, const void *src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// ---------------------------------------------------------------------------------------------------------
// distances
// ---------------------------------------------------------------------------------------------------------
struct Acc {
    float ab, a2, b2;
};

template <int METRIC> __device__ __forceinline__ void acc_step(Acc &r, float a, float b) {
    if (METRIC == kL2sq) {
        float t = __fsub_rn(a, b);
        r.ab = __fadd_rn(r.ab, __fmul_rn(t, t));
    } else if (METRIC == kIp) {
        r.ab = __fadd_rn(r.ab, __fmul_rn(a, b));
    } else {
        r.ab = __fadd_rn(r.ab, __fmul_rn(a, b));
        r.a2 = __fadd_rn(r.a2, __fmul_rn(a, a));
        r.b2 = __fadd_rn(r.b2, __fmul_rn(b, b));
    }
}
template <int METRIC> __device__ __forceinline__ void acc_pair(Acc &r, float2 a, float2 b);
template <int METRIC> __device__ __forceinline__ void acc_chunk(Acc &r, const float4 *a, const float4 *b, int n4) {
#pragma unroll 4
    for (int i = 0; i < n4; i++) {
        const float4 x = a[i], y = b[i];
        acc_pair<METRIC>(r, make_float2(x.x, x.y), make_float2(y.x, y.y));
        acc_pair<METRIC>(r, make_float2(x.z, x.w), make_float2(y.z, y.w));
    }
}
template <int METRIC> __device__ __forceinline__ float acc_finish(const Acc &r) {
    if (METRIC == kL2sq) return r.ab;
    if (METRIC == kIp) return __fsub_rn(1.0f, r.ab);
    // metric_cos_gt: result_if_zero[a2 == 0][b2 == 0]
    if (r.a2 == 0.0f && r.b2 == 0.0f) return 0.0f;
    if (r.a2 == 0.0f || r.b2 == 0.0f) return 1.0f;
    return __fsub_rn(1.0f, __fdiv_rn(r.ab, __fmul_rn(__fsqrt_rn(r.a2), __fsqrt_rn(r.b2))));
}

// ---- register-streamed rows: each lane reads ITS row with 256-bit loads (one 32-byte sector per load, SASS LDG.E.ENL2.256)
// and folds it against the query held in shared memory. Differences and products of two neighbouring dimensions are
// formed with packed FP32x2 instructions (FADD2 / FMUL2, round-to-nearest per element), the sum itself stays one strict
// left-to-right chain — so the bits are the same as in acc_step.
__device__ __forceinline__ void ldg256(const float *p, float (&v)[8]) {
    asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "l"(p));
}
__device__ __forceinline__ float2 sub2(float2 a, float2 b) {
    unsigned long long r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(*reinterpret_cast<unsigned long long *>(&a)), "l"(*reinterpret_cast<unsigned long long *>(&b)));
    return *reinterpret_cast<float2 *>(&r);
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
    unsigned long long r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(*reinterpret_cast<unsigned long long *>(&a)), "l"(*reinterpret_cast<unsigned long long *>(&b)));
    return *reinterpret_cast<float2 *>(&r);
}
template <int METRIC> __device__ __forceinline__ void acc_pair(Acc &r, float2 a, float2 b) {
    if (METRIC == kL2sq) {
        const float2 t = sub2(a, b);
        const float2 p = mul2(t, t);
        r.ab = __fadd_rn(__fadd_rn(r.ab, p.x), p.y);
    } else {
        const float2 p = mul2(a, b);
        r.ab = __fadd_rn(__fadd_rn(r.ab, p.x), p.y);
        if (METRIC == kCos) {
            const float2 pa = mul2(a, a), pb = mul2(b, b);
            r.a2 = __fadd_rn(__fadd_rn(r.a2, pa.x), pa.y);
            r.b2 = __fadd_rn(__fadd_rn(r.b2, pb.x), pb.y);
        }
    }
}
// cosine with a known left-hand norm: only ab and b2 are folded per row (the reference recomputes a2 for every row, always
// from the same query in the same order, so the value cached in WarpWs::q_a2 has the same bits)
template <int METRIC, bool SKIP_A2> __device__ __forceinline__ void acc_pair_q(Acc &r, float2 a, float2 b) {
    if (METRIC == kCos && SKIP_A2) {
        const float2 p = mul2(a, b), pb = mul2(b, b);
        r.ab = __fadd_rn(__fadd_rn(r.ab, p.x), p.y);
        r.b2 = __fadd_rn(__fadd_rn(r.b2, pb.x), pb.y);
    } else acc_pair<METRIC>(r, a, b);
}
// left-hand norm of a vector in shared memory, folded exactly as acc_pair<kCos> folds a2 (lane 0 computes, all lanes get it)
__device__ __forceinline__ float fold_a2(const float *v, uint32_t n) {
    float a2 = 0.0f;
    if (lane_id() == 0) {
        for (uint32_t i = 0; i < n; i += 2) {
            const float2 x = make_float2(v[i], v[i + 1]);
            const float2 p = mul2(x, x);
            a2 = __fadd_rn(__fadd_rn(a2, p.x), p.y);
        }
    }
    return __shfl_sync(kFull, a2, 0);
}
// d(query in shared memory, row `my_slot`) for every active lane; dpad is a multiple of 8 floats
template <int METRIC>
__device__ __noinline__ float direct_distance(const float *vectors, uint32_t dpad, const float *qs, uint32_t my_slot, bool active) {
    Acc acc = {0.0f, 0.0f, 0.0f};
    if (active) {
        const float *row = vectors + (size_t)my_slot * dpad;
        const uint32_t n8 = dpad / 8;
#pragma unroll 1
        for (uint32_t base = 0; base < n8; base += 4) { // 4 sectors (one 128-byte line) in flight per lane
            float v[4][8];
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (base + j < n8) ldg256(row + (base + j) * 8, v[j]);
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if (base + j < n8) {
                    const float4 q0 = *reinterpret_cast<const float4 *>(qs + (base + j) * 8);
                    const float4 q1 = *reinterpret_cast<const float4 *>(qs + (base + j) * 8 + 4);
                    acc_pair<METRIC>(acc, make_float2(q0.x, q0.y), make_float2(v[j][0], v[j][1]));
                    acc_pair<METRIC>(acc, make_float2(q0.z, q0.w), make_float2(v[j][2], v[j][3]));
                    acc_pair<METRIC>(acc, make_float2(q1.x, q1.y), make_float2(v[j][4], v[j][5]));
                    acc_pair<METRIC>(acc, make_float2(q1.z, q1.w), make_float2(v[j][6], v[j][7]));
                }
            }
        }
    }
    return acc_finish<METRIC>(acc);
}

// Query-to-node distances with gather4 staging: lane i <-> row i. The 4 KB stage holds 32 rows x one 128-byte line, or —
// when a batch has at most 16 / 8 rows — 16 rows x 2 lines / 8 rows x 4 lines, so that small batches need fewer
// fetch-wait-compute steps. Sub-tile k (line ch0 + k of the rows) starts at stage + k * rows_cap * 128 B; lane L with
// L % 4 == 0 issues the copy for sub-tile L / rows_cap, rows (L % rows_cap) .. +3 (inactive lanes name a row outside the
// tensor: zero fill, no traffic), which lands at stage + L * 128 B. Reads are conflict-free: in each 8-lane phase of an
// LDS.128 the TMA swizzle (granule ^ row % 8) spreads the lanes over all 8 granules of a line.
constexpr int kG4Chunk = 32;
constexpr int kG4OobRow = 0x7FFFFFFF;
// NST = 4 KB sub-stages filled per step (all behind one mbarrier phase): 1 for the throughput kernel (shared memory buys
// resident warps there), 4 for the latency kernel (small batches: a 32-row batch of 512-byte rows is ONE fetch-wait-compute
// round instead of four).
template <int METRIC, int NST = 1>
__device__ __forceinline__ float warp_distances_g4(const GraphView &g, WarpWs &w, uint32_t my_slot, bool active) {
    const uint32_t lane = lane_id();
    const uint32_t amask = __ballot_sync(kFull, active);
    if (!amask) return 0.0f;
    const uint32_t n = 32u - (uint32_t)__clz(amask);          // rows 0..n-1 (callers use a prefix of the lanes)
    const uint32_t rows_cap = n <= 8u ? 8u : (n <= 16u ? 16u : 32u);
    const uint32_t cpt = 32u / rows_cap;                        // 128-byte lines of a row per step
    const uint32_t nchunks = (g.dpad + kG4Chunk - 1) / kG4Chunk;
    const int r0 = active ? (int)my_slot : kG4OobRow;
    const uint32_t grow = lane & (rows_cap - 1u), gk = lane / rows_cap; // as issuer: first row of my group, my sub-tile
    const int ra = __shfl_sync(kFull, r0, grow), rb = __shfl_sync(kFull, r0, grow + 1u), rc = __shfl_sync(kFull, r0, grow + 2u),
              rd = __shfl_sync(kFull, r0, grow + 3u);
    const bool issuer_row = (lane & 3u) == 0u && ((amask >> grow) & 0xFu) != 0u;
    float *sbuf = w.stage[0];
    const uint32_t sw = lane & 7u;
    Acc acc = {0.0f, 0.0f, 0.0f};
    for (uint32_t ch0 = 0; ch0 < nchunks; ch0 += cpt * NST) {
        OPCLK(t0);
        uint32_t nissue = 0;
#pragma unroll
        for (uint32_t st = 0; st < NST; st++) nissue += __popc(__ballot_sync(kFull, issuer_row && ch0 + st * cpt + gk < nchunks));
        if (lane == 0) mbar_expect_tx(w.mbar, nissue * 4u * kG4Chunk * 4u);
        __syncwarp();
#pragma unroll
        for (uint32_t st = 0; st < NST; st++)
            if (issuer_row && ch0 + st * cpt + gk < nchunks)
                tma_gather4(sbuf + st * 32u * kG4Chunk + lane * kG4Chunk, w.vmap, (int)((ch0 + st * cpt + gk) * kG4Chunk), ra, rb, rc, rd, w.mbar);
        OPCLK(t1);
        mbar_wait(w.mbar, w.phase[0]);
        OPCLK(t2);
        w.phase[0] ^= 1u;
        if (active) {
            for (uint32_t kk = 0; kk < cpt * NST && ch0 + kk < nchunks; kk++) {
                const uint32_t ch = ch0 + kk, st = kk / cpt, k = kk - st * cpt;
                const uint32_t nfl = min((uint32_t)kG4Chunk, g.dpad - ch * kG4Chunk);
                const float4 *a = reinterpret_cast<const float4 *>(w.q + ch * kG4Chunk);
                const float4 *myrow = reinterpret_cast<const float4 *>(sbuf + st * 32u * kG4Chunk + (k * rows_cap + lane) * kG4Chunk);
                if (nfl == kG4Chunk) {
#pragma unroll
                    for (uint32_t i = 0; i < kG4Chunk / 4; i++) {
                        const float4 x = a[i], y = myrow[i ^ sw];
                        acc_pair_q<METRIC, true>(acc, make_float2(x.x, x.y), make_float2(y.x, y.y));
                        acc_pair_q<METRIC, true>(acc, make_float2(x.z, x.w), make_float2(y.z, y.w));
                    }
                } else {
                    for (uint32_t i = 0; i < nfl / 4; i++) {
                        const float4 x = a[i], y = myrow[i ^ sw];
                        acc_pair_q<METRIC, true>(acc, make_float2(x.x, x.y), make_float2(y.x, y.y));
                        acc_pair_q<METRIC, true>(acc, make_float2(x.z, x.w), make_float2(y.z, y.w));
                    }
                }
            }
        }
        __syncwarp(); // all lanes are done reading the buffer before it is refilled
        OPCLK(t3);
        OPC(kOpClkStage, t1 - t0); OPC(kOpClkWait, t2 - t1); OPC(kOpClkMath, t3 - t2); OPC(kOpSteps, 1);
    }
    if (METRIC == kCos) acc.a2 = w.q_a2;
    return active ? acc_finish<METRIC>(acc) : 0.0f;
}

// Where the left-hand vector of a distance batch comes from.
struct LeftSrc {
    const float *smem; // whole row resident in shared memory (the query), or nullptr
    uint32_t slot;     // else: row `slot` of the index, streamed chunk by chunk next to the right-hand rows
};

// Distances from one left-hand vector to up to 32 rows: lane i (if `active`) gets d(left, vectors[my_slot]).
// `staged` remembers which row sits in stage[0][lane] so that back-to-back calls over the same rows
// (neighbour-selection steps) do not touch HBM/L2 again when a row is a single chunk.
// MODE = 1: the production search configuration is fixed at compile time — query-to-node distances through gather4
// staging, whole-row and visited-window prefetches on, list-as-queue allowed — so none of the alternative paths (kept for
// A/B measurements and for the build) takes part in k_search's register allocation. MODE = 2: the same with a 16 KB stage
// (latency kernel for batches that fit the GPU at once). MODE = 0: configuration from WarpWs.
template <int METRIC, int MODE = 0>
__device__ __forceinline__ float warp_distances(const GraphView &g, WarpWs &w, LeftSrc left, uint32_t my_slot,
                                                bool active, uint32_t &staged) {
    if (MODE == 1) return warp_distances_g4<METRIC, 1>(g, w, my_slot, active);
    if (MODE == 2) return warp_distances_g4<METRIC, 4>(g, w, my_slot, active); // latency kernel: 16 KB stage
    const bool use_tma = MODE == 3 ? true : w.use_tma; // MODE 3: k_build's production configuration (1-D bulk copies, no A/B paths)
    const uint32_t lane = lane_id();
    const uint32_t chunk = w.chunk;
    const uint32_t nchunks = (g.dpad + chunk - 1) / chunk;
    const bool dbuf = w.nbuf == 2;
    const uint32_t rows = MODE == 3 ? 32u : w.stage_rows; // MODE 3: lane <-> row, one pass
    const bool cacheable = nchunks == 1 && rows == 32;
    const bool stream_left = left.smem == nullptr;
    if (MODE != 3 && w.g4 && left.smem == w.q) return warp_distances_g4<METRIC>(g, w, my_slot, active); // gather4 staging serves query-to-node distances only
#if GVSS_DIRECT_PATH
    if (MODE != 3 && w.direct && !stream_left) { // no staging at all: all 32 lanes at once
        const float d = direct_distance<METRIC>(g.vectors, g.dpad, left.smem, my_slot, active);
        __syncwarp();
        return d;
    }
#endif
    const uint32_t row = lane & (rows - 1);
    float result = 0.0f;

    // with a 16-row stage the 32 lanes are served in two passes (callers on the search path only use lanes 0..15)
    for (uint32_t pass = 0; pass * rows < 32; pass++) {
        const bool act = active && (lane / rows == pass);
        if (!__ballot_sync(kFull, act)) continue;
        Acc acc = {0.0f, 0.0f, 0.0f};

        auto issue = [&](uint32_t ch) {
            const uint32_t buf = dbuf ? (ch & 1u) : 0u;
            const uint32_t off = ch * chunk;
            const uint32_t nfl = min(chunk, g.dpad - off);
            const bool need = act && !(cacheable && staged == my_slot);
            const uint32_t need_mask = __ballot_sync(kFull, need);
            const uint32_t ncopy = __popc(need_mask) + (stream_left ? 1u : 0u);
            if (ncopy == 0) return false;
            float *sbuf = buf ? w.stage[1] : w.stage[0];
            if (use_tma) {
                uint64_t *bar = w.mbar + buf;
                if (lane == 0) mbar_expect_tx(bar, ncopy * nfl * 4u);
                __syncwarp();
                if (need) tma_row(sbuf + row * w.rstride, g.vectors + (size_t)my_slot * g.dpad + off, nfl * 4u, bar);
                if (stream_left && lane == 0)
                    tma_row(w.c + buf * kChunk, g.vectors + (size_t)left.slot * g.dpad + off, nfl * 4u, bar);
            } else {
                // coalesced 16-byte cp.async (LDGSTS): one instruction moves 32 granules = 32*4/nfl rows of this chunk
                const uint32_t gpr = nfl / 4;                 // 16-byte granules per row chunk (<= 32)
                const uint32_t rpi = 32u / gpr;               // rows per instruction
                const uint32_t sub = lane / gpr, gi = lane % gpr;
                const bool prefix = (need_mask & (need_mask + 1u)) == 0u && (need_mask & 1u); // lanes 0..cnt-1 (pass 0)
                if (prefix) {
                    const uint32_t cnt = __popc(need_mask);
                    for (uint32_t r0 = 0; r0 < cnt; r0 += rpi) {
                        const uint32_t src = r0 + sub;        // lane that owns the row this lane helps to copy
                        const bool ok = sub < rpi && src < cnt;
                        const uint32_t s = __shfl_sync(kFull, my_slot, ok ? src : 0);
                        if (ok) cp_async16(sbuf + src * w.rstride + gi * 4, g.vectors + (size_t)s * g.dpad + off + gi * 4);
                    }
                } else {
                    for (uint32_t m = need_mask; m; m &= m - 1) {
                        const uint32_t src = __ffs(m) - 1;
                        const uint32_t s = __shfl_sync(kFull, my_slot, src);
                        if (lane < gpr) cp_async16(sbuf + (src & (rows - 1)) * w.rstride + lane * 4, g.vectors + (size_t)s * g.dpad + off + lane * 4);
                    }
                }
                if (stream_left && lane < gpr)
                    cp_async16(w.c + buf * kChunk + lane * 4, g.vectors + (size_t)left.slot * g.dpad + off + lane * 4);
                cp_async_commit();
            }
            return true;
        };

        // chunk pipeline with ONE issue site (the routine is inlined at many places; k_build is large enough to stall on
        // instruction fetch): chunk ch is awaited with chunk ch + 1 already in flight when the stage is double-buffered
        uint32_t issued = 0;
        bool pend0 = false, pend1 = false; // by chunk parity
#pragma unroll 1
        for (uint32_t ch = 0; ch < nchunks; ch++) {
            const uint32_t buf = dbuf ? (ch & 1u) : 0u;
            const uint32_t upto = min(nchunks, ch + (dbuf ? 2u : 1u));
#pragma unroll 1
            while (issued < upto) {
                const bool any = issue(issued);
                if (issued & 1u) pend1 = any; else pend0 = any;
                issued++;
            }
            if ((ch & 1u) ? pend1 : pend0) {
                if (use_tma) {
                    if (buf) { mbar_wait(w.mbar + 1, w.phase[1]); w.phase[1] ^= 1u; }
                    else { mbar_wait(w.mbar, w.phase[0]); w.phase[0] ^= 1u; }
                } else {
                    if (issued > ch + 1u && ((ch & 1u) ? pend0 : pend1)) cp_async_wait<1>(); else cp_async_wait<0>();
                    __syncwarp();
                }
            }
            const uint32_t off = ch * chunk;
            const uint32_t nfl = min(chunk, g.dpad - off);
            if (act) {
                const float4 *a = reinterpret_cast<const float4 *>(stream_left ? (w.c + buf * kChunk) : (left.smem + off));
                const float4 *b = reinterpret_cast<const float4 *>((buf ? w.stage[1] : w.stage[0]) + row * w.rstride);
                acc_chunk<METRIC>(acc, a, b, (int)(nfl / 4));
            }
            __syncwarp(); // all lanes are done reading this buffer before it is refilled
        }
        if (act) {
            staged = cacheable ? my_slot : kInvalid;
            result = acc_finish<METRIC>(acc);
        }
    }
    return result;
}

// ---------------------------------------------------------------------------------------------------------
// visited set: open addressing in global scratch, one table per warp (any exact set is equivalent to
// usearch's growing_hash_set_gt, index.hpp:852-970 — only membership matters)
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void vis_clear(WarpWs &w) {
    if (w.use_bitmap) {
        uint4 *p = reinterpret_cast<uint4 *>(w.bitmap);
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        for (uint32_t i = lane_id(); i < w.bitmap_words / 4; i += 32) p[i] = z;
    } else {
        uint4 *p = reinterpret_cast<uint4 *>(w.vis);
        const uint32_t n4 = (w.vis_mask + 1) / 4;
        const uint4 e = make_uint4(kInvalid, kInvalid, kInvalid, kInvalid);
        for (uint32_t i = lane_id(); i < n4; i += 32) p[i] = e;
    }
    __syncwarp();
}
constexpr uint32_t kVisWindow = 8; // slots per probe window = one 32-byte sector

// returns true when `key` was not yet present (and is now). Single-lane form (atomicCAS over the probe windows).
__device__ __forceinline__ bool vis_insert(WarpWs &w, uint32_t key) {
    if (w.use_bitmap) {
        const uint32_t bit = 1u << (key & 31u);
        return (atomicOr(&w.bitmap[key >> 5], bit) & bit) == 0u;
    }
    uint32_t h = ((key * 2654435761u) >> 7) & w.vis_mask & ~(kVisWindow - 1);
    for (;;) {
        for (uint32_t j = 0; j < kVisWindow; j++) {
            const uint32_t old = atomicCAS(&w.vis[h + j], kInvalid, key);
            if (old == kInvalid) return true;
            if (old == key) return false;
        }
        h = (h + kVisWindow) & w.vis_mask;
    }
}

// Warp-wide form: every lane with `want` tests-and-sets its own key; keys of wanting lanes are distinct.
// The table belongs to this warp alone, so no atomics are needed: each lane reads its 8-slot window (one 32-byte
// sector, one L2 round trip), lanes that target the same empty slot are ordered by lane id in registers
// (__match_any_sync), winners write with a plain store; a lane whose window is full — or was filled by the winners —
// moves to the next window of its probe sequence. A key lives in the first window of its sequence that had room when
// it was inserted, so a lookup ends at the first window with an empty slot.
__device__ __forceinline__ bool vis_insert_warp(WarpWs &w, uint32_t key, bool want) {
    if (w.use_bitmap) return want ? vis_insert(w, key) : false;
    const uint32_t lane = lane_id();
    uint32_t base = ((key * 2654435761u) >> 7) & w.vis_mask & ~(kVisWindow - 1);
    bool pending = want, isnew = false;
    while (__ballot_sync(kFull, pending)) {
        OPC(kOpVisRounds, 1);
        uint4 a = make_uint4(0u, 0u, 0u, 0u), b = a;
        if (pending) {
            a = __ldcg(reinterpret_cast<const uint4 *>(w.vis + base));
            b = __ldcg(reinterpret_cast<const uint4 *>(w.vis + base + 4));
        }
        const uint32_t e[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        bool found = false;
        uint32_t empties = 0; // bit j: slot j of the window is empty
#pragma unroll
        for (int j = 0; j < 8; j++) {
            found |= e[j] == key;
            empties |= (e[j] == kInvalid) ? (1u << j) : 0u;
        }
        if (pending && found) pending = false;
        bool claiming = pending && empties != 0u;
        for (;;) { // settle contested slots: the lowest lane keeps it, the others move to their next empty slot
            const uint32_t pos = claiming ? base + (uint32_t)(__ffs(empties) - 1) : (0x80000000u | lane);
            const uint32_t peers = __match_any_sync(kFull, pos);
            const bool lost = claiming && (uint32_t)(__ffs(peers) - 1) != lane;
            if (!__ballot_sync(kFull, lost)) {
                if (claiming) { __stcg(w.vis + pos, key); isnew = true; pending = false; }
                break;
            }
            if (lost) {
                empties &= empties - 1;
                if (!empties) claiming = false; // every empty slot went to a lower lane: window is full now
            }
        }
        __syncwarp();
        if (pending) base = (base + kVisWindow) & w.vis_mask;
    }
    return isnew;
}

// ---------------------------------------------------------------------------------------------------------
// candidate queue: binary heap with usearch's shift_up / shift_down (index.hpp:624-644). The reference keeps
// {-distance, slot} in a max-heap; here +distance in a min-heap with every comparison mirrored. Lane 0 is the single
// writer (lanes of a warp are not guaranteed to run in lockstep, so redundant all-lane writes could land out of order);
// the other lanes see the result after the __syncwarp() that ends each operation.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ Cand heap_get(const WarpWs &w, uint32_t i) {
    return i < w.heap_smem ? w.heap[i] : w.heap_ovf[i - w.heap_smem];
}
__device__ __forceinline__ void heap_set(WarpWs &w, uint32_t i, Cand v) {
    if (i < w.heap_smem) w.heap[i] = v; else w.heap_ovf[i - w.heap_smem] = v;
}
// Drops queue entries that can never be expanded any more (distance above the current radius while the result list is
// full: the radius only shrinks from there, and such an entry would only end the search when popped). Called when the
// shared-memory part of the queue is full, instead of letting the queue grow into global memory. The survivors are
// re-inserted in index order; among entries with EQUAL distance the pop order may then differ from the reference's heap
// (no effect unless two candidates tie exactly, and never on queues that fit shared memory).
__device__ __forceinline__ void heap_prune(WarpWs &w, uint32_t &size, float radius) {
    const uint32_t lane = lane_id();
    uint32_t kept = 0;
    for (uint32_t base = 0; base < size; base += 32) {
        const uint32_t i = base + lane;
        Cand e = {0.f, 0u};
        bool keep = false;
        if (i < size) { e = heap_get(w, i); keep = !(e.d > radius); }
        const uint32_t km = __ballot_sync(kFull, keep);
        if (keep) w.heap_ovf[kept + __popc(km & ((1u << lane) - 1u))] = e;
        kept += __popc(km);
    }
    __syncwarp();
    if (lane == 0) {
        for (uint32_t j = 0; j < kept; j++) {
            const Cand e = w.heap_ovf[j];
            uint32_t i = j;
            while (i) {
                const uint32_t p = (i - 1) >> 1;
                const Cand pe = w.heap[p];
                if (!(pe.d > e.d)) break;
                w.heap[i] = pe;
                i = p;
            }
            w.heap[i] = e;
        }
    }
    size = kept;
    __syncwarp();
}

__device__ __forceinline__ bool heap_push(WarpWs &w, uint32_t &size, float d, uint32_t s) {
    if (size >= w.heap_total) return false;
    if (lane_id() == 0) {
        uint32_t i = size;
        while (i) {
            const uint32_t p = (i - 1) >> 1;
            const Cand pe = heap_get(w, p);
            if (!(pe.d > d)) break; // less(parent, child) on negated distances
            heap_set(w, i, pe);
            i = p;
            OPC(kOpPushLevels, 1);
        }
        heap_set(w, i, Cand{d, s});
    }
    size++;
    __syncwarp();
    return true;
}
__device__ __forceinline__ void heap_pop(WarpWs &w, uint32_t &size) {
    size--;
    if (size && lane_id() == 0) {
        const Cand x = heap_get(w, size);
        uint32_t i = 0;
        for (;;) {
            uint32_t mx = i;
            float best = x.d;
            const uint32_t l = 2 * i + 1, r = 2 * i + 2;
            Cand le = x, re = x;
            if (l < size) { le = heap_get(w, l); if (best > le.d) { mx = l; best = le.d; } }
            if (r < size) { re = heap_get(w, r); if (best > re.d) { mx = r; } }
            if (mx == i) break;
            heap_set(w, i, mx == l ? le : re);
            i = mx;
            OPC(kOpPopLevels, 1);
        }
        heap_set(w, i, x);
    }
    __syncwarp();
}

// ---------------------------------------------------------------------------------------------------------
// result list: sorted_buffer_gt (index.hpp:725-749). lower_bound => an element equal to existing ones goes
// BEFORE them; at the limit the worst element is dropped; position == limit rejects.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool top_insert(WarpWs &w, uint32_t &size, uint32_t limit, float d, uint32_t s) {
    const uint32_t lane = lane_id();
    uint32_t cnt = 0;
    for (uint32_t i = lane; i < size; i += 32) cnt += (w.top_d[i] < d) ? 1u : 0u;
    const uint32_t pos = __reduce_add_sync(kFull, cnt);
    if (pos == limit) return false;
    const bool full = size == limit;
    const uint32_t end = full ? size - 1 : size; // elements [pos, end) move up by one
    // move in 32-wide windows from the top so that no element is overwritten before it is read
    for (int32_t hi = (int32_t)end; hi > (int32_t)pos; hi -= 32) {
        const int32_t i = hi - 1 - (int32_t)lane;
        float td = 0.f;
        uint32_t ts = 0;
        const bool mv = i >= (int32_t)pos;
        if (mv) { td = w.top_d[i]; ts = w.top_s[i]; }
        __syncwarp();
        if (mv) { w.top_d[i + 1] = td; w.top_s[i + 1] = ts; }
        __syncwarp();
    }
    if (lane == 0) { w.top_d[pos] = d; w.top_s[pos] = s; }
    __syncwarp();
    if (!full) size++;
    return true;
}

// ---------------------------------------------------------------------------------------------------------
// adjacency access
// ---------------------------------------------------------------------------------------------------------
template <bool VOLATILE> __device__ __forceinline__ uint32_t ld_adj(const uint32_t *p) {
    return VOLATILE ? __ldcg(p) : __ldg(p);
}
__device__ __forceinline__ uint32_t *list_ptr(const GraphView &g, uint32_t slot, int level, uint32_t &cap) {
    if (level == 0) { cap = g.M0; return g.nbr0 + (size_t)slot * g.stride0; }
    cap = g.M;
    return g.upper + ((size_t)__ldg(&g.upper_off[slot]) + (uint32_t)(level - 1)) * g.strideU;
}

// search_for_one_ (index.hpp:3384-3422): greedy descent from `closest` through levels begin..end+1.
template <int METRIC, bool VOLATILE, int MODE = 0>
__device__ __forceinline__ uint32_t greedy_descent(const GraphView &g, WarpWs &w, uint32_t closest, int begin_level,
                                                   int end_level, Counters &cnt, uint32_t &staged, float &closest_d) {
    const uint32_t lane = lane_id();
    LeftSrc q{w.q, 0};
    float cd = warp_distances<METRIC, MODE>(g, w, q, closest, lane == 0, staged);
    cd = __shfl_sync(kFull, cd, 0);
    cnt.computed++;
    for (int level = begin_level; level > end_level; --level) {
        bool changed;
        do {
            changed = false;
            uint32_t cap;
            const uint32_t *list = list_ptr(g, closest, level, cap);
            for (uint32_t base = 0; base < cap; base += 32) {
                const uint32_t idx = base + lane;
                const uint32_t s = idx < cap ? ld_adj<VOLATILE>(list + idx) : kInvalid;
                const bool valid = s != kInvalid;
                const uint32_t vmask = __ballot_sync(kFull, valid);
                if (!vmask) break;
                cnt.computed += __popc(vmask);
                const float d = warp_distances<METRIC, MODE>(g, w, q, s, valid, staged);
                // sequential scan with a running minimum == first strict minimum in list order
                float best = valid ? d : CUDART_INF_F;
                uint32_t bl = lane;
                bool ok = valid && (d < cd);
                uint32_t okmask = __ballot_sync(kFull, ok);
                if (okmask) {
                    // among lanes that beat cd, the strict running minimum picks the smallest value, earliest lane on ties
                    if (!ok) best = CUDART_INF_F;
#pragma unroll
                    for (int o = 16; o; o >>= 1) {
                        const float ob = __shfl_xor_sync(kFull, best, o);
                        const uint32_t ol = __shfl_xor_sync(kFull, bl, o);
                        if (ob < best || (ob == best && ol < bl)) { best = ob; bl = ol; }
                    }
                    cd = best;
                    closest = __shfl_sync(kFull, s, bl);
                    changed = true;
                }
            }
            cnt.cycles++;
        } while (changed);
    }
    closest_d = cd; // d(query, closest): the beam search that starts there need not fetch the row again
    return closest;
}

// per-lane binary searches over the ascending result list
__device__ __forceinline__ uint32_t top_lower_bound(const float *td, uint32_t n, float d) { // #{x : x < d}
    uint32_t lo = 0, hi = n;
    while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (td[mid] < d) lo = mid + 1; else hi = mid; }
    return lo;
}
__device__ __forceinline__ uint32_t top_upper_bound(const float *td, uint32_t n, float d) { // #{x : x <= d}
    uint32_t lo = 0, hi = n;
    while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (td[mid] <= d) lo = mid + 1; else hi = mid; }
    return lo;
}

// search_to_find_in_base_ (INSERT=false, index.hpp:3504-3573) / search_to_insert_ (INSERT=true, :3430-3496).
// On return `top` (w.top_d/top_s, ascending) holds `top_size` results.
//
// One expansion handles all unvisited neighbours of the popped node as a batch, with exactly the outcome of the
// reference's one-by-one loop (`if (top.size() < limit || d < radius) { next.insert; top.insert; radius = top.top() }`):
//   * radius before candidate j = limit-th smallest of (old top ∪ allowed candidates before j) — a rejected candidate is
//     never below the radius, so it cannot change that order statistic — hence
//       accepted_j  <=>  #{x in old top : x <= d_j} + #{i < j allowed : d_i <= d_j} < limit ;
//   * accepted candidates enter the queue in list order (same heap operations as the reference);
//   * the new result list is the merge of the old one with the accepted allowed candidates, an element equal to
//     existing ones going BEFORE them and later candidates before earlier equal ones (sorted_buffer_gt::insert uses
//     lower_bound), cut at `limit`.
//
// lq = true ("list as queue"): every candidate the reference can still pop is an unexpanded member of the result list
// (an entry that fell off the list lies above the radius for good), and as long as the smallest live distance is unique
// the heap pops candidates in ascending distance — i.e. in list order. The queue is then not materialised: expanded
// members carry a flag in the top bit of their slot and the next candidate is the first unflagged entry; the queue
// operations the reference would have done (pushes of accepted candidates, one pop per expansion) are only LOGGED, 8
// bytes each, at the far end of the queue's overflow area. When two live candidates tie for the head of the queue — or
// an entry that fell off the list equals it — the pop order depends on the history of usearch's heap: the log is
// replayed on the real heap (cheap: no distance is recomputed) and the search continues with the heap as the queue
// (flag bit2). If the log ran out of room the search ends with flag bit3 and the caller repeats it with lq = false.
// Removed members (they enter the reference's queue but not its list) wait in a small min-heap of their own and are
// expanded when they are closer than every live candidate; a tie with a live candidate switches to the real heap as above.
constexpr uint32_t kExpanded = 0x80000000u;
constexpr uint32_t kSlotMask = 0x7FFFFFFFu;

template <int METRIC, bool INSERT, bool VOLATILE, int MODE = 0>
// start_d = d(query, start), already known to the caller (greedy descent, or the previous level's best link): the reference
// evaluates it again here (counted below), the bits are the same.
__device__ __forceinline__ void beam_search(const GraphView &g, WarpWs &w, uint32_t start, float start_d, uint32_t limit, int level,
                                            uint32_t new_slot, Counters &cnt, uint32_t &top_size, uint32_t &staged, bool lq = false) {
    const uint32_t lane = lane_id();
    const uint32_t lt_mask = (1u << lane) - 1u;
    const bool cfg_g4 = MODE == 3 ? false : (MODE != 0 ? true : w.g4), cfg_pf_rows = MODE == 3 ? false : (MODE != 0 ? true : w.prefetch_rows),
               cfg_pf_vis = MODE != 0 ? true : w.prefetch_vis, cfg_spec = MODE != 0 ? false : w.spec_prefetch;
    LeftSrc q{w.q, 0};
    uint32_t heap_size = 0, nvis = 1;
    uint32_t scan_from = 0; // LQ: every list entry before this index is expanded
#if GVSS_OPCOUNT
    uint32_t pred_slot = kInvalid; // the candidate the previous expansion predicted as next
#endif
    float ghost = CUDART_INF_F; // LQ: smallest distance among the live entries that fell off the list (still in the reference's queue)
    // LQ: log of the reference's queue operations, written downwards from the end of the overflow area ({d, slot} = push,
    // slot == kInvalid = pop); it may use half of the area, the replayed heap then fits below it
    const uint32_t log_end = w.heap_total - w.heap_smem, log_cap = log_end / 2;
    uint32_t logn = 0;
    bool log_ok = true;
    auto log_put = [&](uint32_t at, float d, uint32_t s) { w.heap_ovf[log_end - 1u - at] = Cand{d, s}; };
    // replays the log on usearch's heap and leaves list-as-queue mode; false (flag bit3) when the log is incomplete
    auto to_heap = [&]() -> bool {
        if (!log_ok) { cnt.flags |= 8u; return false; }
        __syncwarp();
        heap_size = 0;
        for (uint32_t base = 0; base < logn; base += 32) {
            Cand mine = {0.0f, kInvalid};
            if (base + lane < logn) mine = w.heap_ovf[log_end - 1u - (base + lane)];
            const uint32_t nb = min(32u, logn - base);
            for (uint32_t j = 0; j < nb; j++) {
                const float d = __shfl_sync(kFull, mine.d, j);
                const uint32_t s = __shfl_sync(kFull, mine.s, j);
                if (s == kInvalid) heap_pop(w, heap_size);
                else heap_push(w, heap_size, d, s);
            }
        }
        lq = false;
        cnt.flags |= 4u;
        return true;
    };
    top_size = 0;
    vis_clear(w);

    float radius = start_d;
    if (radius != radius) radius = CUDART_INF_F;
    cnt.computed++;
    if (!lq) heap_push(w, heap_size, radius, start);
    else { if (lane == 0) log_put(0, radius, start); logn = 1; }
    if (lane == 0) vis_insert(w, start);
    const bool start_ok = INSERT || !g.has_deleted || __ldg(&g.keys[start]) != INT64_MAX;
    if (start_ok) {
        if (lane == 0) { w.top_d[0] = radius; w.top_s[0] = start; }
        top_size = 1;
    } else {
        // a removed entry point is queued but not listed; in list-as-queue mode such members wait in their own min-heap
        if (lq) heap_push(w, heap_size, radius, start);
        radius = CUDART_INF_F; // reference reads top.top() of an empty buffer here (UB); +inf is the documented choice
    }
    __syncwarp();

    for (;;) {
        OPCLK(t_sel);
        Cand c;
        uint32_t next_slot = kInvalid; // most likely next candidate (prefetch hint)
        if (lq) {
            uint32_t cur = top_size, nxt = top_size;
            bool later = false; // the next unexpanded entry (if any) lies beyond the window that holds `cur`
            for (uint32_t i0 = scan_from & ~31u; i0 < top_size; i0 += 32) {
                const uint32_t i = i0 + lane;
                const bool un = i >= scan_from && i < top_size && !(w.top_s[i] & kExpanded);
                const uint32_t um = __ballot_sync(kFull, un);
                if (um) {
                    cur = i0 + (uint32_t)__ffs(um) - 1u;
                    const uint32_t um2 = um & (um - 1u);
                    if (um2) nxt = i0 + (uint32_t)__ffs(um2) - 1u;
                    else { nxt = i0 + 32u; later = true; }
                    break;
                }
            }
            // removed members (search only): the reference keeps them in its queue but not in the list. Here they wait in a
            // min-heap of their own (the heap_* storage) and are expanded when they are closer than every live candidate.
            const bool have_t = !INSERT && heap_size != 0; // (an insert-time search lists every member: nothing ever waits there)
            const Cand tm = have_t ? heap_get(w, 0) : Cand{CUDART_INF_F, kInvalid};
            bool tie, take_t = false;
            if (cur >= top_size) {
                if (have_t && !(tm.d > radius)) { take_t = true; tie = tm.d == ghost; }
                else {
                    // nothing left to expand: the reference's queue is empty or its head lies above the radius and ends the
                    // search — unless an entry that fell off the list EQUALS the radius, then the reference expands it
                    tie = top_size == limit && ghost == radius;
                    if (!tie) break;
                }
            } else {
                c.d = w.top_d[cur];
                c.s = w.top_s[cur] & kSlotMask;
                if (have_t && tm.d < c.d) { take_t = true; tie = false; }
                else {
                    // an entry outside the list, or a second candidate, at the same distance as the queue head: the
                    // reference's pop order depends on its heap's history (`later`: conservative, the equal entry may be expanded)
                    tie = c.d == ghost || (nxt < top_size && w.top_d[nxt] == c.d) || (have_t && tm.d == c.d);
                    if (!tie && nxt < top_size && !later) next_slot = w.top_s[nxt] & kSlotMask;
                }
            }
            if (take_t && !tie) { // two removed members at the same distance: ambiguous as well
                tie = (heap_size > 1 && heap_get(w, 1).d == tm.d) || (heap_size > 2 && heap_get(w, 2).d == tm.d);
            }
            if (tie) {
                if (!to_heap()) return;
                continue; // the same iteration again, with the heap as the queue
            }
            if (take_t) {
                c = tm;
                heap_pop(w, heap_size);
                if (lane == 0 && logn < log_cap) log_put(logn, 0.0f, kInvalid);
                log_ok = log_ok && logn < log_cap;
                logn++;
                __syncwarp();
            } else {
            __syncwarp();
            if (lane == 0) {
                w.top_s[cur] = c.s | kExpanded;
                if (logn < log_cap) log_put(logn, 0.0f, kInvalid);
            }
            log_ok = log_ok && logn < log_cap;
            logn++;
            __syncwarp();
            scan_from = cur + 1;
            }
        } else {
            if (!heap_size) break;
            c = heap_get(w, 0);
            if (INSERT) { if (c.d > radius && top_size == limit) break; }
            else if (c.d > radius) break;
            heap_pop(w, heap_size);
            if (heap_size) next_slot = heap_get(w, 0).s;
        }
        cnt.cycles++;
        cnt.base++;
        OPC(kOpHops, 1);
#if GVSS_OPCOUNT
        if (c.s == pred_slot) OPC(kOpPredHit, 1);
        pred_slot = next_slot;
#endif
        if ((cnt.flags & 0x80000000u) && lane == 0)
            printf("pop %u slot %u d %.9g radius %.9g heap %u top %u computed %u\n", cnt.base, c.s, c.d, radius, heap_size, top_size, cnt.computed);
        if (INSERT && c.s == new_slot) continue;
        uint32_t pcap = 0;
        const uint32_t *plist = nullptr;
        if (next_slot != kInvalid) { // pull its adjacency row towards L2
            plist = list_ptr(g, next_slot, level, pcap);
            if (lane * 32 < pcap) prefetch_l2(plist + lane * 32);
        }

        OPCLK(t_gather);
        OPC(kOpClkSelect, t_gather - t_sel);
        // 1. unvisited neighbours, compacted in list order (visits.set(): the first occurrence wins)
        uint32_t cap;
        const uint32_t *list = list_ptr(g, c.s, level, cap);
        uint32_t nnew = 0;
        if (cfg_pf_vis && !w.use_bitmap && cap > 32) {
            // the visited windows of the second block of neighbours start their trip from DRAM while the first block is tested
            const uint32_t s1 = 32 + lane < cap ? ld_adj<VOLATILE>(list + 32 + lane) : kInvalid;
            if (s1 != kInvalid) prefetch_l2(w.vis + (((s1 * 2654435761u) >> 7) & w.vis_mask & ~(kVisWindow - 1)));
        }
        for (uint32_t base = 0; base < cap; base += 32) {
            const uint32_t idx = base + lane;
            const uint32_t s = idx < cap ? ld_adj<VOLATILE>(list + idx) : kInvalid;
            if (!__ballot_sync(kFull, s != kInvalid)) break;
            OPCLK(t_v0);
            if (base == 0) OPC(kOpClkAdj, t_v0 - t_gather);
            const bool valid = s != kInvalid && s < g.count;
            if (cfg_spec && valid) { // start pulling the row towards L2 while the visited test is in flight
                const float *row = g.vectors + (size_t)s * g.dpad;
                for (uint32_t l = 0; l * 32 < g.dpad; l++) prefetch_l2(row + l * 32);
            }
            const uint32_t same = __match_any_sync(kFull, s);
            const bool leader = valid && (uint32_t)(__ffs(same) - 1) == lane;
            const bool room = w.use_bitmap || nvis < w.vis_limit;
            const bool isnew = vis_insert_warp(w, s, leader && room);
            const uint32_t nmask = __ballot_sync(kFull, isnew);
            OPCLK(t_v1);
            OPC(kOpClkVis, t_v1 - t_v0);
            if (!room && __ballot_sync(kFull, leader)) cnt.flags |= 1u;
            if (isnew) w.nb[nnew + __popc(nmask & lt_mask)] = s;
            nnew += __popc(nmask);
            nvis += __popc(nmask);
        }
        __syncwarp();
        OPC(kOpNew, nnew);
        if (nnew <= 16) OPC(kOpHopsLe16, 1);
        if (!nnew) OPC(kOpHopsNone, 1);
        if (!nnew) continue;
        cnt.computed += nnew;

        // the likely next candidate's list has reached L2 by now: read it, and after the distance phase start the visited
        // windows of ITS neighbours on their way from DRAM (a hint only: wrong guesses cost 32 bytes of traffic each)
        uint32_t sp0 = kInvalid, sp1 = kInvalid;
        if (cfg_pf_vis && !w.use_bitmap && plist) {
            if (lane < pcap) sp0 = ld_adj<VOLATILE>(plist + lane);
            if (32 + lane < pcap) sp1 = ld_adj<VOLATILE>(plist + 32 + lane);
        }
        OPCLK(t_dist);
        OPC(kOpClkGather, t_dist - t_gather);
        // 2. their distances, stage_rows rows at a time (rows of the later batches are prefetched to L2 meanwhile)
        const uint32_t bw = MODE == 3 ? 32u : (((GVSS_DIRECT_PATH && MODE == 0 && w.direct) || cfg_g4) ? 32u : w.stage_rows); // rows per distance batch
        if (cfg_g4 && cfg_pf_rows) {
            // gather4 staging takes one 128-byte line of every row per step: get ALL lines of all new rows moving towards L2
            // now, so that only the first step of the hop waits for DRAM
            const uint32_t lines = (g.dpad * 4 + 127) / 128;
            for (uint32_t i = lane; i < nnew * lines; i += 32) { // consecutive lanes <-> consecutive lines of a row (measured best)
                const uint32_t r = i / lines, l = i - r * lines;
                if (r || l) prefetch_l2(g.vectors + (size_t)w.nb[r] * g.dpad + l * 32);
            }
        } else {
            for (uint32_t i = bw * 4 + lane; i < nnew * 4; i += 32) {
                const uint32_t lines = (g.dpad * 4 + 127) / 128;
                if ((i & 3u) < lines) prefetch_l2(g.vectors + (size_t)w.nb[i >> 2] * g.dpad + (i & 3u) * 32);
            }
        }
        for (uint32_t b = 0; b < nnew; b += bw) {
            const bool act = lane < bw && b + lane < nnew;
            const uint32_t s = act ? w.nb[b + lane] : kInvalid;
            const float d = warp_distances<METRIC, MODE>(g, w, q, s, act, staged);
            // a NaN distance (NULL / NaN query rows reach the index join, hnsw_lateral_join.test:57-67) is ordered as +inf:
            // the reference only guarantees termination there, and a total order keeps the merge below well defined
            if (act) w.nd[b + lane] = (d == d) ? d : CUDART_INF_F;
        }
        __syncwarp();

        if (sp0 != kInvalid) prefetch_l2(w.vis + (((sp0 * 2654435761u) >> 7) & w.vis_mask & ~(kVisWindow - 1)));
        if (sp1 != kInvalid) prefetch_l2(w.vis + (((sp1 * 2654435761u) >> 7) & w.vis_mask & ~(kVisWindow - 1)));
        OPCLK(t_merge);
        OPC(kOpClkDist, t_merge - t_dist);
        // 3. acceptance, queue and result-list update, 32 candidates at a time
        for (uint32_t b = 0; b < nnew; b += 32) {
            const bool have = b + lane < nnew;
            const float dj = have ? w.nd[b + lane] : CUDART_INF_F;
            const uint32_t sj = have ? w.nb[b + lane] : kInvalid;
            // once the result list is full a candidate at or above the radius can neither be accepted nor influence the
            // acceptance of a later one (the radius only shrinks), so the group is thinned before any bookkeeping
            const bool maybe = have && (top_size < limit || dj < radius);
            if (!__ballot_sync(kFull, maybe)) continue;
            OPC(kOpMaybeBatches, 1);
            OPCLK(t_a0);
            const bool allowed = maybe && (INSERT || !g.has_deleted || __ldg(&g.keys[sj]) != INT64_MAX);
            const uint32_t amask = __ballot_sync(kFull, allowed);
            // list size and order statistic as the reference would see them just before candidate j
            uint32_t le_prev = 0;
            for (uint32_t m = amask; m; m &= m - 1) {
                const uint32_t i = __ffs(m) - 1;
                const float di = __shfl_sync(kFull, dj, i);
                le_prev += (i < lane && di <= dj) ? 1u : 0u;
            }
            const uint32_t size_before = top_size + __popc(amask & lt_mask);
            const uint32_t le_top = maybe ? top_upper_bound(w.top_d, top_size, dj) : 0u;
            const bool acc = maybe && (size_before < limit || le_top + le_prev < limit);
            const uint32_t accmask = __ballot_sync(kFull, acc);
            OPCLK(t_a1);
            OPC(kOpClkAccept, t_a1 - t_a0);
            if (!accmask) continue;

            OPC(kOpPushes, __popc(accmask));
            if (!lq) {
                for (uint32_t m = accmask; m; m &= m - 1) { // next.insert in list order
                    const uint32_t i = __ffs(m) - 1;
                    // (pruning rebuilds the heap, so it is only allowed where the queue was never list-shaped: heap_smem >= 256)
                    if (w.heap_smem >= 256u && heap_size == w.heap_smem && top_size == limit) heap_prune(w, heap_size, radius);
                    if (!heap_push(w, heap_size, __shfl_sync(kFull, dj, i), __shfl_sync(kFull, sj, i))) cnt.flags |= 1u;
                }
            } else { // log the pushes in list order; removed members go to their own heap
                const uint32_t at = logn + __popc(accmask & lt_mask);
                if (acc && at < log_cap) log_put(at, dj, sj);
                logn += __popc(accmask);
                log_ok = log_ok && logn <= log_cap;
                for (uint32_t m = INSERT ? 0u : (accmask & ~amask); m; m &= m - 1) {
                    const uint32_t i = __ffs(m) - 1;
                    if (!heap_push(w, heap_size, __shfl_sync(kFull, dj, i), __shfl_sync(kFull, sj, i))) cnt.flags |= 1u;
                }
            }

            const bool ins = acc && allowed;
            const uint32_t imask = __ballot_sync(kFull, ins);
            const uint32_t m_ins = __popc(imask);
            if (m_ins) {
                const uint32_t lb = ins ? top_lower_bound(w.top_d, top_size, dj) : 0xFFFFFFFFu;
                uint32_t before = 0; // inserted candidates that end up in front of this one
                for (uint32_t m = imask; m; m &= m - 1) {
                    const uint32_t i = __ffs(m) - 1;
                    const float di = __shfl_sync(kFull, dj, i);
                    before += (di < dj || (di == dj && i > lane)) ? 1u : 0u;
                }
                // LQ: smallest distance among the live entries this update pushes off the list (they stay in the
                // reference's queue; harmless there unless the queue head ever EQUALS one of them)
                float dropped = CUDART_INF_F;
                bool has_drop = false;
                if (lq) {
                    const uint32_t at = (ins && lb + before < limit) ? lb + before : 0xFFFFFFFFu;
                    scan_from = min(scan_from, __reduce_min_sync(kFull, at));
                    if (ins && lb + before >= limit) { dropped = dj; has_drop = true; }
                }
                const uint32_t pmin = __reduce_min_sync(kFull, lb);
                OPC(kOpInsBatches, 1);
                OPC(kOpInserted, m_ins);
                OPC(kOpShifted, top_size - pmin);
                // old elements [pmin, top_size) move up by the number of insertions at or before them. The insertion points go
                // to shared memory in ascending order (this batch's slice of w.nd is dead: dj/sj live in registers), then
                // aligned blocks of 128 elements are moved from the top down, 4 consecutive elements per lane: all reads of a
                // block happen before its writes, and writes only reach upwards into blocks that were already moved.
                OPCLK(t_s0);
                OPC(kOpClkMerge, t_s0 - t_a1);
#if GVSS_VSHIFT
                uint32_t *ins_lb = reinterpret_cast<uint32_t *>(w.nd + b);
                __syncwarp();
                if (ins) ins_lb[before] = lb;
                __syncwarp();
                for (int32_t blk = pmin < top_size ? (int32_t)((top_size - 1u) >> 7) : -1; blk >= (int32_t)(pmin >> 7); blk--) {
                    const uint32_t i0 = (uint32_t)blk * 128u + lane * 4u;
                    const bool any = i0 < top_size && i0 + 3u >= pmin;
                    float4 td = make_float4(0.f, 0.f, 0.f, 0.f);
                    uint4 ts = make_uint4(0u, 0u, 0u, 0u);
                    uint32_t sh0 = 0, sh1 = 0, sh2 = 0, sh3 = 0;
                    if (any) {
                        td = *reinterpret_cast<const float4 *>(w.top_d + i0);
                        ts = *reinterpret_cast<const uint4 *>(w.top_s + i0);
#pragma unroll 4
                        for (uint32_t r = 0; r < m_ins; r++) {
                            const uint32_t pa = ins_lb[r];
                            sh0 += pa <= i0 ? 1u : 0u;
                            sh1 += pa <= i0 + 1u ? 1u : 0u;
                            sh2 += pa <= i0 + 2u ? 1u : 0u;
                            sh3 += pa <= i0 + 3u ? 1u : 0u;
                        }
                    }
                    __syncwarp();
                    if (any) {
                        const float tdv[4] = {td.x, td.y, td.z, td.w};
                        const uint32_t tsv[4] = {ts.x, ts.y, ts.z, ts.w};
                        const uint32_t shv[4] = {sh0, sh1, sh2, sh3};
#pragma unroll
                        for (uint32_t e = 0; e < 4; e++) {
                            const uint32_t i = i0 + e;
                            if (i >= pmin && i < top_size) {
                                if (i + shv[e] < limit) { w.top_d[i + shv[e]] = tdv[e]; w.top_s[i + shv[e]] = tsv[e]; }
                                else if (lq) {
                                    if (!(tsv[e] & kExpanded)) { dropped = fminf(dropped, tdv[e]); has_drop = true; }
                                }
                            }
                        }
                    }
                    __syncwarp();
                }
#else
                // old elements [pmin, top_size) move up by the number of insertions at or before them; windows from the top
                for (int32_t hi = (int32_t)top_size; hi > (int32_t)pmin; hi -= 32) {
                    const int32_t i = hi - 1 - (int32_t)lane;
                    const bool mv = i >= (int32_t)pmin;
                    float td = 0.f;
                    uint32_t ts = 0, sh = 0;
                    if (mv) { td = w.top_d[i]; ts = w.top_s[i]; }
                    for (uint32_t m = imask; m; m &= m - 1) {
                        const uint32_t pa = __shfl_sync(kFull, lb, __ffs(m) - 1);
                        sh += ((int32_t)pa <= i) ? 1u : 0u;
                    }
                    __syncwarp();
                    if (mv && (uint32_t)i + sh < limit) { w.top_d[i + sh] = td; w.top_s[i + sh] = ts; }
                    if (lq) {
                        if (mv && (uint32_t)i + sh >= limit && !(ts & kExpanded)) { dropped = fminf(dropped, td); has_drop = true; }
                    }
                    __syncwarp();
                }
#endif
                OPCLK(t_s1);
                OPC(kOpClkShiftLoop, t_s1 - t_s0);
                if (ins && lb + before < limit) { w.top_d[lb + before] = dj; w.top_s[lb + before] = sj; }
                top_size = min(limit, top_size + m_ins);
                __syncwarp();
                radius = w.top_d[top_size - 1];
                if (lq) {
                    if (__ballot_sync(kFull, has_drop)) {
                        uint32_t u = __float_as_uint(dropped); // order-preserving key for the warp-wide minimum
                        u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
                        u = __reduce_min_sync(kFull, u);
                        u = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u;
                        ghost = fminf(ghost, __uint_as_float(u));
                    }
                }
                OPCLK(t_a2);
                OPC(kOpClkShift, t_a2 - t_a1);
            }
        }
    }
}

} // namespace gvss
