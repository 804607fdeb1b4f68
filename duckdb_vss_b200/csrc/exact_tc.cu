// K5 (tensor-core variant) + K6 — exact top-k through error-compensated TF32 distance tiles and an exact re-rank.
//
// North star (4): "batched query-by-base distance tiles for brute-force top-k and the index join run on tensor cores,
// but only through an error-compensated split (3xTF32) that holds fp32 tolerance". The pipeline:
//   k_split      x -> (hi, lo): hi = x with the 13 low mantissa bits cleared (exactly a TF32 value), lo = TF32(x - hi);
//                plus ||x||^2 per row. Once per base set / query batch.
//   k_tiles_tc   CTA = one tile of 128 queries, loops over 128-row base tiles of its split. Per 32-float K block the four
//                operand tiles (Q_hi, Q_lo, B_hi, B_lo) arrive by 2-D TMA (128-byte swizzle) in a 3-stage mbarrier ring;
//                one elected thread issues tcgen05.mma kind::tf32 three times per 8-wide K step
//                (hi·hi + lo·hi + hi·lo, fp32 accumulation in TMEM, two accumulator buffers); four epilogue warps read
//                the accumulators with tcgen05.ld (thread <-> query row), form the score (||b||^2 - 2 q·b, -q·b or the
//                cosine form) and keep the 32 best rows per query in registers.
//   k_merge<0>   merges the splits (exact.cu).
//   k_rerank     recomputes the 32 candidates per query with DuckDB's folds bit for bit (array_kernels.hpp:11-105, same
//                code path as k_exact), orders them by (distance, row) and checks that no row outside the candidate set
//                can belong to the top k: every discarded row has an approximate score >= the worst kept one, so if
//                that bound minus the error budget still exceeds the exact k-th distance the result is exact; otherwise
//                the query is flagged and recomputed by k_exact (k_exact_flagged).
// Ids and distances are therefore bit-identical to k_exact / DuckDB; the tensor cores only choose the candidates.
#include "kernels.h"
#include "exact_common.cuh"

#include <cuda.h>
#include <algorithm>
#include <cstdio>
#include <type_traits>

namespace gvss {

constexpr int kTM = 128;     // queries per tile (UMMA M)
constexpr int kTN = 128;     // base rows per tile (UMMA N)
constexpr int kKB = 32;      // floats per K block = one 128-byte swizzle row
constexpr int kStages = 3;        // ring depth of the query-resident variant (32 KB stages)
#ifndef GVSS_TC_PROF
#define GVSS_TC_PROF 0 // per-role clocks of one CTA on stdout (tuning aid)
#endif
#ifndef GVSS_TC_WAIT_HINT
#define GVSS_TC_WAIT_HINT 2000
#endif
constexpr int kStagesStream = 2;  // ring depth when the query tile streams too (64 KB stages)
constexpr int kCand = 32;    // candidates kept per query
constexpr uint32_t kTileBytes = kTM * kKB * 4; // 16 KB per operand tile
constexpr uint32_t kStageBytes = 4 * kTileBytes;
constexpr int kScStride = 36;       // floats per query in the score staging area: 9 x 16 B -> conflict-free 16-byte stores
constexpr uint32_t kTcThreads = 192; // warp 0: TMA, warp 1: MMA + TMEM alloc, warps 2-5: epilogue

__device__ __forceinline__ void mbar_arrive_expect(uint64_t *bar, uint32_t bytes) { mbar_expect_tx(bar, bytes); }
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_tile_2d(void *dst, const CUtensorMap *map, int c0, int c1, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
                     smem_u32(dst)),
                 "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ uint64_t umma_desc_sw128(const void *tile) {
    // K-major operand tile, rows of 128 B, 8-row swizzle atoms of 1024 B (SBO), SWIZZLE_128B, descriptor version 1 (sm_100)
    const uint64_t addr = (uint64_t)(smem_u32(tile) >> 4) & 0x3FFFull;
    return addr | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// wait with a suspend-time hint: the epilogue warps sleep in hardware instead of spinning on issue slots
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAITR_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
        "@p bra DONER_%=;\n"
        "bra WAITR_%=;\n"
        "DONER_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity), "r"((uint32_t)GVSS_TC_WAIT_HINT)
        : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        "tcgen05.wait::ld.sync.aligned;\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
#pragma unroll
    for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]);
}

// ---- k_split: hi / lo planes (row stride dk = d rounded up to 32 floats, zero padded) and squared norms ------------------
__global__ void k_split(const float *__restrict__ x, uint64_t rows, uint32_t d, uint32_t dk, float *__restrict__ hi,
                        float *__restrict__ lo, float *__restrict__ norm2) {
    const uint64_t row = (uint64_t)blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const uint32_t lane = threadIdx.x & 31u;
    float s = 0.f;
    for (uint32_t c = lane; c < dk; c += 32) {
        const float v = c < d ? x[row * d + c] : 0.f;
        const float h = __uint_as_float(__float_as_uint(v) & 0xFFFFE000u);
        const float l = __uint_as_float(__float_as_uint(__fsub_rn(v, h)) & 0xFFFFE000u);
        hi[row * dk + c] = h;
        lo[row * dk + c] = l;
        s = fmaf(v, v, s);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(kFull, s, o);
    if (lane == 0) norm2[row] = s;
}

struct TcParams {
    uint64_t n, nq, rows_per_split;
    uint32_t dk;      // padded dimension (multiple of 32)
    int metric;       // 'e' / 'c' / 'i'
    const float *bnorm2, *qnorm2;
    unsigned long long *part_key; // [split][nq][kCand] (order-key << 32 | row)
    float *part_score;            // [split][nq][kCand]
    float *dbg_dots;              // optional [nq][n] raw q·b (validation of the tensor-core path)
    uint32_t *shared_worst;       // [nq] order key of the smallest 'worst kept score' any CTA has reached for the query (0xFFFFFFFF = none)
};

// MODE 0: every ring stage carries (Q_hi, Q_lo, B_hi, B_lo) of one K block — 64 KB per stage, any dk.
// MODE 1 (dk <= 64): the query tile (hi + lo planes of every K block) stays resident in shared memory for the whole CTA and
//         only (B_hi, B_lo) stream through the ring — half the L2->SM traffic per tile.
// MODE 2 (dk <= 128): only the hi plane of the query tile is resident (64 KB at dk = 128); a stage carries (Q_lo, B_hi, B_lo) —
//         three quarters of the streamed bytes of MODE 0, which is what bounds the kernel (L2->SM operand stream).
template <int MODE>
__global__ void __launch_bounds__(kTcThreads, 1)
k_tiles_tc(const __grid_constant__ CUtensorMap map_q_hi, const __grid_constant__ CUtensorMap map_q_lo,
           const __grid_constant__ CUtensorMap map_b_hi, const __grid_constant__ CUtensorMap map_b_lo, TcParams p) {
    constexpr bool QRES = MODE == 1, HIRES = MODE == 2;
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char *tiles = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(smem) + 1023) & ~uintptr_t(1023));
    constexpr uint32_t stage_bytes = QRES ? 2 * kTileBytes : (HIRES ? 3 * kTileBytes : kStageBytes);
    constexpr uint32_t b_off = QRES ? 0u : (HIRES ? kTileBytes : 2 * kTileBytes);   // offset of B_hi inside a stage
    constexpr uint32_t NST = QRES ? kStages : kStagesStream;
    constexpr uint32_t res_tiles = QRES ? 2u : (HIRES ? 1u : 0u);              // resident query tiles per K block
    unsigned char *qres = tiles + NST * stage_bytes;                            // resident: [kblock][hi(, lo)] tiles
    unsigned char *after = qres + (p.dk / kKB) * res_tiles * kTileBytes;
    float *cl_s = reinterpret_cast<float *>(after);                          // [128 queries][kCand] candidate scores, ascending
    uint32_t *cl_r = reinterpret_cast<uint32_t *>(after + kTM * kCand * 4);  // ... and rows
    float *sc = reinterpret_cast<float *>(after + 2 * kTM * kCand * 4);      // [128 queries][36] scores of the current 32-column chunk
    float *bn_s = sc + kTM * kScStride;                                      // [4 epilogue warps][128] norms of the current base tile
    uint64_t *bars = reinterpret_cast<uint64_t *>(after + 2 * kTM * kCand * 4 + kTM * kScStride * 4 + 4 * kTN * 4);
    uint64_t *full = bars, *empty = bars + NST, *acc_full = bars + 2 * NST, *acc_empty = bars + 2 * NST + 2;
    uint64_t *q_full = bars + 2 * NST + 4;
    uint32_t *tmem_base_smem = reinterpret_cast<uint32_t *>(bars + 2 * NST + 5);

    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint32_t kblocks = p.dk / kKB;
    const uint64_t q0 = (uint64_t)blockIdx.x * kTM;
    const uint64_t r_begin = (uint64_t)blockIdx.y * p.rows_per_split;
    const uint64_t r_end = min(p.n, r_begin + p.rows_per_split);
    const uint32_t ntiles = r_end > r_begin ? (uint32_t)((r_end - r_begin + kTN - 1) / kTN) : 0;

    if (threadIdx.x == 0) {
        for (uint32_t i = 0; i < NST; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4 * 32); }
        mbar_init(q_full, 1);
        mbar_fence_init();
    }
    if (warp == 1) { // TMEM: 2 accumulator buffers x 128 fp32 columns
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_smem)), "n"(2 * kTN));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_base_smem;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            if (res_tiles && ntiles) {
                mbar_arrive_expect(q_full, kblocks * res_tiles * kTileBytes);
                for (uint32_t kb = 0; kb < kblocks; kb++) {
                    tma_tile_2d(qres + (kb * res_tiles + 0) * kTileBytes, &map_q_hi, (int)(kb * kKB), (int)q0, q_full);
                    if (QRES) tma_tile_2d(qres + (kb * 2 + 1) * kTileBytes, &map_q_lo, (int)(kb * kKB), (int)q0, q_full);
                }
            }
            uint32_t it = 0;
            for (uint32_t t = 0; t < ntiles; t++) {
                const int row0 = (int)(r_begin + (uint64_t)t * kTN);
                for (uint32_t kb = 0; kb < kblocks; kb++, it++) {
                    const uint32_t s = it % NST, ph = (it / NST) & 1u;
                    mbar_wait_relaxed(&empty[s], ph ^ 1u);
                    unsigned char *st = tiles + s * stage_bytes;
                    mbar_arrive_expect(&full[s], stage_bytes);
                    if (MODE == 0) {
                        tma_tile_2d(st + 0 * kTileBytes, &map_q_hi, (int)(kb * kKB), (int)q0, &full[s]);
                        tma_tile_2d(st + 1 * kTileBytes, &map_q_lo, (int)(kb * kKB), (int)q0, &full[s]);
                    } else if (HIRES) {
                        tma_tile_2d(st, &map_q_lo, (int)(kb * kKB), (int)q0, &full[s]);
                    }
                    tma_tile_2d(st + b_off, &map_b_hi, (int)(kb * kKB), row0, &full[s]);
                    tma_tile_2d(st + b_off + kTileBytes, &map_b_lo, (int)(kb * kKB), row0, &full[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread) =====
        if (lane == 0) {
            // kind::tf32, D = F32, A/B = TF32, both K-major, N = 128, M = 128
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kTN >> 3) << 17) | ((uint32_t)(kTM >> 4) << 24);
            uint32_t it = 0;
            if (res_tiles && ntiles) mbar_wait(q_full, 0);
#if GVSS_TC_PROF
            long long pw_acc = 0, pw_full = 0, p_t0 = clock64();
#endif
            for (uint32_t t = 0; t < ntiles; t++) {
                const uint32_t buf = t & 1u, aph = (t >> 1) & 1u;
#if GVSS_TC_PROF
                const long long a0 = clock64();
#endif
                mbar_wait_relaxed(&acc_empty[buf], aph ^ 1u);
#if GVSS_TC_PROF
                pw_acc += clock64() - a0;
#endif
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + buf * kTN;
                for (uint32_t kb = 0; kb < kblocks; kb++, it++) {
                    const uint32_t s = it % NST, ph = (it / NST) & 1u;
#if GVSS_TC_PROF
                    const long long f0 = clock64();
#endif
                    mbar_wait_relaxed(&full[s], ph);
#if GVSS_TC_PROF
                    pw_full += clock64() - f0;
#endif
                    tc_fence_after();
                    unsigned char *st = tiles + s * stage_bytes;
                    const unsigned char *qht = res_tiles ? qres + kb * res_tiles * kTileBytes : st;
                    const unsigned char *qlt = QRES ? qht + kTileBytes : (HIRES ? st : st + kTileBytes);
                    const uint64_t qh = umma_desc_sw128(qht), ql = umma_desc_sw128(qlt);
                    const uint64_t bh = umma_desc_sw128(st + b_off), bl = umma_desc_sw128(st + b_off + kTileBytes);
#pragma unroll
                    for (uint32_t k = 0; k < kKB / 8; k++) { // UMMA_K = 8 for TF32: 32 bytes along K inside the swizzle row
                        const uint64_t adv = (uint64_t)(k * 32 >> 4);
                        umma_tf32(d_tmem, qh + adv, bh + adv, idesc, (kb | k) != 0u);
                        umma_tf32(d_tmem, ql + adv, bh + adv, idesc, 1u);
                        umma_tf32(d_tmem, qh + adv, bl + adv, idesc, 1u);
                    }
                    umma_commit(&empty[s]); // frees the stage once these MMAs have read it
                }
                umma_commit(&acc_full[buf]);
            }
#if GVSS_TC_PROF
            if (blockIdx.x == 3 && blockIdx.y == 2) printf("MMA thread: tiles %u total %lld wait_acc_empty %lld wait_full %lld (cycles)\n", ntiles, clock64() - p_t0, pw_acc, pw_full);
#endif
        }
    } else {
        // ===== epilogue: thread <-> query row (TMEM lane) for the scores, warp <-> query for the candidate lists =====
        const uint32_t quarter = warp & 3u;              // tcgen05.ld: a warp may touch lanes [32*(warp%4), +32)
        const uint32_t qrow = quarter * 32 + lane;
        const uint64_t qi = q0 + qrow;
        const bool qvalid = qi < p.nq;
        const float qn2 = qvalid ? p.qnorm2[qi] : 0.f;
        // the 32 candidate lists of this warp's queries live in shared memory, one entry per lane
        for (uint32_t qq = 0; qq < 32; qq++) {
            cl_s[(quarter * 32 + qq) * kCand + lane] = CUDART_INF_F;
            cl_r[(quarter * 32 + qq) * kCand + lane] = 0xFFFFFFFFu;
        }
        __syncwarp();
        float worst = CUDART_INF_F; // this thread's query: score of its current worst kept candidate
#if GVSS_TC_PROF
        long long ew_acc = 0, e_ld = 0, e_p1 = 0, e_p2 = 0, e_t0 = clock64();
#endif
        for (uint32_t t = 0; t < ntiles; t++) {
            const uint32_t buf = t & 1u, aph = (t >> 1) & 1u;
            const uint64_t row0 = r_begin + (uint64_t)t * kTN;
            // The CTAs that scan other row ranges for the same queries publish the worst score of their (full) lists: a row
            // at or above ANY full list's worst score cannot be among the best kCand of the union, so the tightest published
            // value is a valid threshold here too — it cuts the list insertions (the critical path of this kernel) by about
            // the number of splits. A stale value only prunes less.
            if (qvalid) {
                const uint32_t wk = __ldcg(p.shared_worst + qi);
                if (wk != 0xFFFFFFFFu) {
                    const float wv = __uint_as_float((wk & 0x80000000u) ? (wk & 0x7FFFFFFFu) : ~wk); // inverse of order_key
                    worst = fminf(worst, wv);
                }
            }
            // norms of this tile's 128 base rows: loaded before the wait on the accumulator so that the L2 round trip
            // overlaps it, staged once per warp (zero beyond the end: those columns never pass)
            float bnr[kTN / 32];
#pragma unroll
            for (int c = 0; c < kTN / 32; c++) bnr[c] = row0 + c * 32 + lane < r_end ? __ldg(&p.bnorm2[row0 + c * 32 + lane]) : 0.f;
#if GVSS_TC_PROF
            const long long w0 = clock64();
#endif
            mbar_wait_relaxed(&acc_full[buf], aph);
#if GVSS_TC_PROF
            ew_acc += clock64() - w0;
#endif
            tc_fence_after();
            float *bnw = bn_s + quarter * kTN;
#pragma unroll
            for (int c = 0; c < kTN / 32; c++) bnw[c * 32 + lane] = bnr[c];
            __syncwarp();
            const uint32_t valid_cols = (uint32_t)min((uint64_t)kTN, r_end - row0);
            for (uint32_t c0 = 0; c0 < kTN; c0 += 32) {
                float acc[32];
#if GVSS_TC_PROF
                const long long l0 = clock64();
#endif
                tmem_ld32(tmem_base + ((quarter * 32u) << 16) + buf * kTN + c0, acc);
#if GVSS_TC_PROF
                const long long l1 = clock64();
                e_ld += l1 - l0;
#endif
                if (p.dbg_dots && qvalid) {
#pragma unroll
                    for (int j = 0; j < 32; j++) if (c0 + j < valid_cols) p.dbg_dots[qi * p.n + row0 + c0 + j] = acc[j];
                }
                // phase 1 (straight-line, thread <-> query): scores of 32 columns into shared memory, bit mask of the
                // columns that beat this query's threshold
                uint32_t pass = 0;
                auto phase1 = [&](auto metric_tag) { // metric fixed at compile time inside the straight-line block
                    constexpr int MT = decltype(metric_tag)::value;
                    uint32_t pmask[8];
#pragma unroll
                    for (int j4 = 0; j4 < 8; j4++) {
                        const float4 bn = *reinterpret_cast<const float4 *>(bnw + c0 + j4 * 4);
                        const float b4[4] = {bn.x, bn.y, bn.z, bn.w};
                        float s4[4];
                        uint32_t m4 = 0; // independent per group of four: no 32-long dependent chain on `pass`
#pragma unroll
                        for (int e = 0; e < 4; e++) {
                            const float a = acc[j4 * 4 + e];
                            if (MT == 'e') s4[e] = fmaf(-2.f, a, b4[e]);
                            else if (MT == 'i') s4[e] = -a;
                            else s4[e] = -a * rsqrtf(fmaxf(qn2 * b4[e], 1e-30f));
                            m4 |= (s4[e] < worst) ? (1u << e) : 0u;
                        }
                        pmask[j4] = m4;
                        *reinterpret_cast<float4 *>(sc + qrow * kScStride + j4 * 4) = make_float4(s4[0], s4[1], s4[2], s4[3]);
                    }
                    pass = (pmask[0] | (pmask[1] << 4)) | ((pmask[2] << 8) | (pmask[3] << 12)) |
                           (((pmask[4] << 16) | (pmask[5] << 20)) | ((pmask[6] << 24) | (pmask[7] << 28)));
                };
                if (p.metric == 'e') phase1(std::integral_constant<int, 'e'>{});
                else if (p.metric == 'i') phase1(std::integral_constant<int, 'i'>{});
                else phase1(std::integral_constant<int, 'c'>{});
                const uint32_t colmask = valid_cols >= c0 + 32 ? 0xFFFFFFFFu : (valid_cols > c0 ? (1u << (valid_cols - c0)) - 1u : 0u);
                pass = qvalid ? (pass & colmask) : 0u;
                __syncwarp();
#if GVSS_TC_PROF
                const long long l2 = clock64();
                e_p1 += l2 - l1;
#endif
                // phase 2 (warp <-> query): sorted insertion with one ballot and one shuffle per candidate; equal scores
                // keep the earlier row first
                uint32_t qmask = __ballot_sync(kFull, pass != 0u);
                while (qmask) {
                    const uint32_t qq = __ffs(qmask) - 1;
                    qmask &= qmask - 1;
                    uint32_t pm = __shfl_sync(kFull, pass, qq);
                    float *ls = cl_s + (quarter * 32 + qq) * kCand;
                    uint32_t *lr = cl_r + (quarter * 32 + qq) * kCand;
                    float e = ls[lane];
                    uint32_t r = lr[lane];
                    while (pm) {
                        const uint32_t j = __ffs(pm) - 1;
                        pm &= pm - 1;
                        const float s = sc[(quarter * 32 + qq) * kScStride + j];
                        const uint32_t pos = __popc(__ballot_sync(kFull, e <= s));
                        if (pos >= (uint32_t)kCand) continue;
                        const float eu = __shfl_up_sync(kFull, e, 1);
                        const uint32_t ru = __shfl_up_sync(kFull, r, 1);
                        if (lane > pos) { e = eu; r = ru; }
                        else if (lane == pos) { e = s; r = (uint32_t)(row0 + c0 + j); }
                    }
                    ls[lane] = e;
                    lr[lane] = r;
                    const float w31 = __shfl_sync(kFull, e, kCand - 1);
                    if (lane == qq && w31 < worst) {
                        worst = w31;
                        atomicMin(p.shared_worst + qi, order_key_f(w31));
                    }
                }
                __syncwarp();
#if GVSS_TC_PROF
                e_p2 += clock64() - l2;
#endif
            }
            tc_fence_before();
            mbar_arrive(&acc_empty[buf]);
        }
#if GVSS_TC_PROF
        if (blockIdx.x == 3 && blockIdx.y == 2 && warp == 2 && lane == 0)
            printf("epilogue warp 2: total %lld wait_acc_full %lld tmem_ld %lld phase1 %lld phase2 %lld (cycles)\n", clock64() - e_t0, ew_acc, e_ld, e_p1, e_p2);
#endif
        for (uint32_t qq = 0; qq < 32; qq++) { // lists out: warp-wide, one entry per lane
            const uint64_t q = q0 + quarter * 32 + qq;
            if (q >= p.nq) break;
            const size_t o = ((size_t)blockIdx.y * p.nq + q) * kCand;
            const float e = cl_s[(quarter * 32 + qq) * kCand + lane];
            const uint32_t r = cl_r[(quarter * 32 + qq) * kCand + lane];
            const bool have = r != 0xFFFFFFFFu;
            p.part_key[o + lane] = have ? (((unsigned long long)order_key_f(e) << 32) | r) : ~0ull;
            p.part_score[o + lane] = have ? e : CUDART_INF_F;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(2 * kTN));
    }
}

// ---- k_rerank: exact folds on the candidates, final order, exactness check ------------------------------------------------
template <char METRIC> __device__ __forceinline__ float fold_seq(const float *a, const float *b, uint32_t d) {
    float ab = 0.f, a2 = 0.f, b2 = 0.f;
    for (uint32_t i = 0; i < d; i++) {
        const float x = a[i], y = b[i];
        if (METRIC == 'e') { const float t = __fsub_rn(x, y); ab = __fadd_rn(ab, __fmul_rn(t, t)); }
        else {
            ab = __fadd_rn(ab, __fmul_rn(x, y));
            if (METRIC == 'c') { a2 = __fadd_rn(a2, __fmul_rn(x, x)); b2 = __fadd_rn(b2, __fmul_rn(y, y)); }
        }
    }
    return fold_finish<METRIC>(ab, a2, b2);
}

// one warp per query; lane <-> candidate. cand_key/cand_score: merged [nq][kCand] lists (row index in the low 32 bits).
template <char METRIC>
__global__ void k_rerank(const float *__restrict__ base, uint64_t n, const int64_t *__restrict__ keys, const float *__restrict__ queries,
                         uint64_t nq, uint32_t d, uint32_t k, const int64_t *__restrict__ cand_row, const float *__restrict__ cand_score,
                         const float *__restrict__ bnorm2, const float *__restrict__ qnorm2, float bnorm2_max, int64_t *__restrict__ ids_out,
                         float *__restrict__ dist_out, uint32_t *__restrict__ flags) {
    const uint64_t q = (uint64_t)blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
    if (q >= nq) return;
    const uint32_t lane = threadIdx.x & 31u;
    const int64_t row = cand_row[q * kCand + lane];
    const bool have = row >= 0;
    float dist = CUDART_INF_F;
    if (have) dist = fold_seq<METRIC>(base + (size_t)row * d, queries + q * d, d); // lhs = table row, rhs = query
    const unsigned long long key = have ? (((unsigned long long)order_key_f(dist) << 32) | (uint32_t)row) : ~0ull;
    uint32_t rank = 0;
    for (int i = 0; i < 32; i++) {
        const unsigned long long o = __shfl_sync(kFull, key, i);
        rank += (o < key || (o == key && i < (int)lane)) ? 1u : 0u;
    }
    if (rank < k) {
        ids_out[q * k + rank] = have ? (keys ? keys[row] : row) : -1;
        if (dist_out) dist_out[q * k + rank] = dist;
    }
    // Exactness: every row outside the candidate set has approximate score >= the worst kept score s_max. In distance
    // units (l2: + ||q||^2; ip: as is; cos: + 1) its exact fold is >= that bound - eps. The result is proven exact when
    // the bound minus eps is still above the exact k-th distance (compared before DuckDB's final sqrt for l2).
    const uint32_t ncand = __popc(__ballot_sync(kFull, have));
    const float s_max = __shfl_sync(kFull, cand_score[q * kCand + (ncand ? ncand - 1 : 0)], 0);
    float kth = CUDART_INF_F; // exact distance of rank k-1
    const uint32_t kth_lane = __ffs(__ballot_sync(kFull, rank == k - 1)) - 1;
    if (kth_lane < 32) kth = __shfl_sync(kFull, dist, kth_lane);
    if (lane == 0) {
        bool ok;
        if (ncand < kCand) ok = (uint64_t)ncand >= min((uint64_t)kCand, n); // fewer candidates than slots: nothing was discarded
        else {
            const float qn2 = qnorm2[q];
            float bound, kth_cmp, eps;
            // error budget: the sequential fp32 fold is within ~d * 2^-24 of the real value relative to the sum of |terms|,
            // the 3xTF32 product within ~2^-21; (d + 64) * 2^-22 covers both with a factor > 3 to spare
            const float rel = (float)(d + 64) * 2.4e-7f;
            if (METRIC == 'e') { bound = s_max + qn2; kth_cmp = kth * kth; eps = rel * (qn2 + bnorm2_max); }
            else if (METRIC == 'i') { bound = s_max; kth_cmp = kth; eps = rel * sqrtf(qn2 * bnorm2_max); }
            else { bound = s_max + 1.0f; kth_cmp = kth; eps = rel * 4.0f; }
            ok = (bound - eps > kth_cmp + eps) && (kth == kth);
        }
        if (n <= (uint64_t)kCand) ok = ncand == n;
        flags[q] = ok ? 0u : 1u;
    }
}

// ---- host side ---------------------------------------------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static encode_tiled_fn get_encode() {
    static encode_tiled_fn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess || qres != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (encode_tiled_fn)p;
    }();
    return fn;
}

static bool make_map(CUtensorMap *m, const float *ptr, uint64_t rows, uint32_t dk) {
    encode_tiled_fn enc = get_encode();
    if (!enc) return false;
    const cuuint64_t dims[2] = {dk, rows};
    const cuuint64_t strides[1] = {(cuuint64_t)dk * 4};
    const cuuint32_t box[2] = {(cuuint32_t)kKB, (cuuint32_t)kTM};
    const cuuint32_t estr[2] = {1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// k_search's gather4 staging: box = 32 floats (128 B) x 1 row; one instruction fetches that box from 4 arbitrary rows
bool encode_row_gather_map(CUtensorMap *m, const float *vectors, uint64_t rows, uint32_t dpad) {
    encode_tiled_fn enc = get_encode();
    if (!enc || dpad < 32 || rows == 0) return false;
    const cuuint64_t dims[2] = {dpad, rows};
    const cuuint64_t strides[1] = {(cuuint64_t)dpad * 4};
    const cuuint32_t box[2] = {32, 1};
    const cuuint32_t estr[2] = {1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)vectors, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
struct DevBuf {
    void *p = nullptr;
    cudaStream_t st;
    explicit DevBuf(cudaStream_t s) : st(s) {}
    cudaError_t alloc(size_t bytes) { return cudaMallocAsync(&p, bytes ? bytes : 16, st); }
    ~DevBuf() { if (p) cudaFreeAsync(p, st); }
    template <typename T> T *as() { return static_cast<T *>(p); }
};

template <char METRIC>
static cudaError_t rerank_t(const float *base, uint64_t n, const int64_t *keys, const float *queries, uint64_t nq, uint32_t d, uint32_t k,
                            const int64_t *cand_row, const float *cand_score, const float *bn2, const float *qn2, float bn2max,
                            int64_t *ids_out, float *dist_out, uint32_t *flags, cudaStream_t st) {
    k_rerank<METRIC><<<(unsigned)((nq + 3) / 4), 128, 0, st>>>(base, n, keys, queries, nq, d, k, cand_row, cand_score, bn2, qn2, bn2max,
                                                             ids_out, dist_out, flags);
    count_launch();
    return cudaGetLastError();
}

// maximum of non-negative floats (squared norms): their bit patterns order like unsigned integers; *out starts at 0
__global__ void k_max(const float *x, uint64_t n, float *out) {
    float m = 0.f;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) m = fmaxf(m, x[i]);
#pragma unroll
    for (int o = 16; o; o >>= 1) m = fmaxf(m, __shfl_xor_sync(kFull, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(reinterpret_cast<unsigned int *>(out), __float_as_uint(m));
}

// Returns cudaErrorNotSupported when the shape does not fit this path (caller then uses k_exact alone).
cudaError_t launch_exact_topk_tc(char metric, const float *base, uint64_t n, const int64_t *keys, const float *queries, uint64_t nq,
                                 uint32_t d, uint32_t k, int64_t *ids_out, float *dist_out, float *dbg_dots, uint64_t *flagged_out,
                                 cudaStream_t st) {
    if (k == 0 || k > 16 || n < 1 || d < 4 || n >= 0xFFFFFFFFull || !get_encode()) return cudaErrorNotSupported;
    const uint32_t dk = (d + kKB - 1) / kKB * kKB;
    {   // keep the stream-ordered pool's memory between calls (the planes below are > 1 GB at 1M x 128)
        static bool pool_set[64];
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev >= 0 && dev < 64 && !pool_set[dev]) {
            cudaMemPool_t pool;
            if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
                uint64_t keep = ~0ull;
                cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
            }
            pool_set[dev] = true;
        }
    }
    DevBuf b_hi(st), b_lo(st), q_hi(st), q_lo(st), bn2(st), qn2(st), pk(st), ps(st), ck(st), cs(st), fl(st), mx(st), sw(st);
    cudaError_t e;
#define TRY(x) do { e = (x); if (e != cudaSuccess) return e; } while (0)
    TRY(b_hi.alloc(n * dk * 4)); TRY(b_lo.alloc(n * dk * 4)); TRY(q_hi.alloc(nq * dk * 4)); TRY(q_lo.alloc(nq * dk * 4));
    TRY(bn2.alloc(n * 4)); TRY(qn2.alloc(nq * 4)); TRY(mx.alloc(4));
    k_split<<<(unsigned)((n + 7) / 8), 256, 0, st>>>(base, n, d, dk, b_hi.as<float>(), b_lo.as<float>(), bn2.as<float>());
    k_split<<<(unsigned)((nq + 7) / 8), 256, 0, st>>>(queries, nq, d, dk, q_hi.as<float>(), q_lo.as<float>(), qn2.as<float>());
    cudaMemsetAsync(mx.p, 0, 4, st);
    k_max<<<(unsigned)std::min<uint64_t>(592, (n + 1023) / 1024), 256, 0, st>>>(bn2.as<float>(), n, mx.as<float>());
    count_launch(3);
    CUtensorMap mqh, mql, mbh, mbl;
    if (!make_map(&mqh, q_hi.as<float>(), nq, dk) || !make_map(&mql, q_lo.as<float>(), nq, dk) || !make_map(&mbh, b_hi.as<float>(), n, dk) ||
        !make_map(&mbl, b_lo.as<float>(), n, dk))
        return cudaErrorNotSupported;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const uint64_t qtiles = (nq + kTM - 1) / kTM, tiles = (n + kTN - 1) / kTN;
    // one CTA per SM (shared memory): pick the split count whose last wave wastes the least — minimise
    // waves(qtiles * splits) / splits, the time in units of "one CTA scans the whole base"
    uint64_t splits = 1;
    double best = 1e30;
    for (uint64_t sp = 1; sp <= 64 && sp <= tiles; sp++) {
        const uint64_t waves = (qtiles * sp + sms - 1) / sms;
        const double cost = (double)waves / (double)sp + 0.002 * (double)sp; // small bias towards fewer partial lists
        if (cost < best) { best = cost; splits = sp; }
    }
    const uint64_t rows_per_split = ((tiles + splits - 1) / splits) * kTN;
    splits = (n + rows_per_split - 1) / rows_per_split;
    TRY(pk.alloc(splits * nq * kCand * 8)); TRY(ps.alloc(splits * nq * kCand * 4));
    TRY(ck.alloc(nq * kCand * 8)); TRY(cs.alloc(nq * kCand * 4)); TRY(fl.alloc(nq * 4));
    TRY(sw.alloc(nq * 4));
    TRY(cudaMemsetAsync(sw.p, 0xFF, nq * 4, st));
    TcParams p{};
    p.n = n; p.nq = nq; p.rows_per_split = rows_per_split; p.dk = dk; p.metric = metric;
    p.bnorm2 = bn2.as<float>(); p.qnorm2 = qn2.as<float>();
    p.part_key = pk.as<unsigned long long>(); p.part_score = ps.as<float>(); p.dbg_dots = dbg_dots;
    p.shared_worst = sw.as<uint32_t>();
    // operand placement (see k_tiles_tc): dk <= 64 whole query tile resident, dk <= 128 its hi plane, else everything streams
    static const int forced = [] { const char *e = getenv("GVSS_TC_MODE"); return e ? atoi(e) : -1; }();
    int mode = dk <= 64 ? 1 : (dk <= 128 ? 2 : 0);
    if (forced == 0) mode = 0;
    const size_t lists = 2 * (size_t)kTM * kCand * 4 + (size_t)kTM * kScStride * 4 + 4 * (size_t)kTN * 4; // lists + scores + norms
    const size_t ring = mode == 1 ? (size_t)kStages * 2 * kTileBytes : (mode == 2 ? (size_t)kStagesStream * 3 * kTileBytes : (size_t)kStagesStream * kStageBytes);
    const size_t resident = (size_t)(dk / kKB) * (mode == 1 ? 2 : (mode == 2 ? 1 : 0)) * kTileBytes;
    const size_t smem = ring + resident + lists + 1024 + 256;
    const dim3 grid((unsigned)qtiles, (unsigned)splits);
    if (mode == 1) {
        TRY(cudaFuncSetAttribute(k_tiles_tc<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_tiles_tc<1><<<grid, kTcThreads, smem, st>>>(mqh, mql, mbh, mbl, p);
    } else if (mode == 2) {
        TRY(cudaFuncSetAttribute(k_tiles_tc<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_tiles_tc<2><<<grid, kTcThreads, smem, st>>>(mqh, mql, mbh, mbl, p);
    } else {
        TRY(cudaFuncSetAttribute(k_tiles_tc<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k_tiles_tc<0><<<grid, kTcThreads, smem, st>>>(mqh, mql, mbh, mbl, p);
    }
    count_launch();
    TRY(cudaGetLastError());
    // merge the splits into one candidate list per query (rows as ids, scores as "distances")
    TRY(launch_merge_partials(pk.as<unsigned long long>(), ps.as<float>(), (uint32_t)splits, nq, kCand, ck.as<int64_t>(), cs.as<float>(), st));
    float bn2max = 0.f;
    TRY(cudaMemcpyAsync(&bn2max, mx.p, 4, cudaMemcpyDeviceToHost, st));
    TRY(cudaStreamSynchronize(st));
    switch (metric) {
    case 'c': TRY(rerank_t<'c'>(base, n, keys, queries, nq, d, k, ck.as<int64_t>(), cs.as<float>(), bn2.as<float>(), qn2.as<float>(), bn2max, ids_out, dist_out, fl.as<uint32_t>(), st)); break;
    case 'i': TRY(rerank_t<'i'>(base, n, keys, queries, nq, d, k, ck.as<int64_t>(), cs.as<float>(), bn2.as<float>(), qn2.as<float>(), bn2max, ids_out, dist_out, fl.as<uint32_t>(), st)); break;
    default: TRY(rerank_t<'e'>(base, n, keys, queries, nq, d, k, ck.as<int64_t>(), cs.as<float>(), bn2.as<float>(), qn2.as<float>(), bn2max, ids_out, dist_out, fl.as<uint32_t>(), st)); break;
    }
    // queries whose candidate set could not be proven sufficient are recomputed by the all-pairs exact kernel
    TRY(launch_exact_topk_flagged(metric, base, n, keys, queries, nq, d, k, ids_out, dist_out, fl.as<uint32_t>(), flagged_out, st));
#undef TRY
    return cudaSuccess;
}

} // namespace gvss
