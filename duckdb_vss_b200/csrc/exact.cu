// K5/K6 — exact (brute-force) top-k with DuckDB's distance folds, bit for bit, and K7 — top-k list merge.
// Restates DistanceOp / CosineDistanceOp / NegativeInnerProductOp (duckdb/extension/core_functions/include/
// core_functions/array_kernels.hpp:11-105) applied by ArrayGenericFold (array_functions.cpp:145-200) under
// `ORDER BY array_distance(v, q), rowid LIMIT k` and the vss_join / vss_match macros (hnsw_index_macros.cpp:10-78).
//
// Every (row, query) distance is a strict left-to-right fp32 sum with separate multiply and add owned by one thread,
// so values — and therefore the order, including ties — equal DuckDB's on x86-64. This is the FFMA-class variant the
// north star allows ("otherwise they use FFMA"); it is FP32-pipe bound, not HBM bound: a 128-row tile of the base is
// shared through shared memory by QT queries per thread.
#include "kernels.h"
#include "exact_common.cuh"
#include <cstdlib>
#include <vector>

namespace gvss {

constexpr int kRows = 128;   // base rows per tile = threads per CTA
constexpr int kDimChunk = 32; // floats per staged chunk
constexpr int kRowStride = kDimChunk + 4; // 9 x 16 B: conflict-free per-thread float4 reads

// sorted insert of (key, dist) into a k-entry list owned by one warp
__device__ __forceinline__ void list_insert(unsigned long long *lk, float *ld, uint32_t &size, uint32_t k,
                                            unsigned long long key, float d) {
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t cnt = 0;
    for (uint32_t i = lane; i < size; i += 32) cnt += lk[i] < key ? 1u : 0u;
    const uint32_t pos = __reduce_add_sync(kFull, cnt);
    if (pos >= k) return;
    const bool full = size == k;
    const uint32_t end = full ? size - 1 : size;
    for (int32_t hi = (int32_t)end; hi > (int32_t)pos; hi -= 32) {
        const int32_t i = hi - 1 - (int32_t)lane;
        unsigned long long tk = 0; float td = 0.f;
        const bool mv = i >= (int32_t)pos;
        if (mv) { tk = lk[i]; td = ld[i]; }
        __syncwarp();
        if (mv) { lk[i + 1] = tk; ld[i + 1] = td; }
        __syncwarp();
    }
    if (lane == 0) { lk[pos] = key; ld[pos] = d; }
    __syncwarp();
    if (!full) size++;
}

// grid.x = query tiles, grid.y = base splits. Output: partial lists [split][nq][k] of (dist, row index).
template <char METRIC, int QT>
__global__ void __launch_bounds__(kRows) k_exact(const float *__restrict__ base, uint64_t n, const float *__restrict__ queries,
                                                uint64_t nq, uint32_t d, uint32_t k, uint64_t rows_per_split,
                                                unsigned long long *__restrict__ part_key, float *__restrict__ part_dist) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t dpad = (d + 3) / 4 * 4;
    float *sq = reinterpret_cast<float *>(smem_raw);                 // QT × dpad
    float *sb = sq + QT * dpad;                                      // 2 × kRows × kRowStride
    float *sd = sb + 2 * kRows * kRowStride;                         // QT × kRows distances of the current tile
    unsigned long long *lk = reinterpret_cast<unsigned long long *>(sd + QT * kRows); // QT × k
    float *ld = reinterpret_cast<float *>(lk + QT * k);              // QT × k
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint64_t q0 = (uint64_t)blockIdx.x * QT;
    const uint64_t r_begin = (uint64_t)blockIdx.y * rows_per_split;
    const uint64_t r_end = min(n, r_begin + rows_per_split);

    for (uint32_t i = tid; i < QT * dpad; i += kRows) {
        const uint32_t j = i / dpad, c = i % dpad;
        sq[i] = (q0 + j < nq && c < d) ? queries[(q0 + j) * d + c] : 0.0f;
    }
    uint32_t lsize[(QT + 3) / 4]; // list sizes of the queries this warp owns (queries warp, warp+4, ...)
#pragma unroll
    for (int i = 0; i < (QT + 3) / 4; i++) lsize[i] = 0;
    __syncthreads();

    const uint32_t nchunk = (dpad + kDimChunk - 1) / kDimChunk;
    const bool vec4 = (d % 4 == 0);
    for (uint64_t t0 = r_begin; t0 < r_end; t0 += kRows) {
        float ab[QT], a2[QT], b2[QT];
#pragma unroll
        for (int j = 0; j < QT; j++) { ab[j] = 0.f; a2[j] = 0.f; b2[j] = 0.f; }
        // stage chunk 0
        auto load_chunk = [&](uint32_t ch, float4 *regs) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const uint32_t g = tid + kRows * i;      // granule: row = g / 8, 16-byte column = g % 8
                const uint32_t r = g >> 3, c = (g & 7u) * 4 + ch * kDimChunk;
                const uint64_t row = t0 + r;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (row < r_end && c < d) {
                    const float *src = base + row * d + c;
                    if (vec4) v = __ldg(reinterpret_cast<const float4 *>(src));
                    else { v.x = __ldg(src); if (c + 1 < d) v.y = __ldg(src + 1); if (c + 2 < d) v.z = __ldg(src + 2); if (c + 3 < d) v.w = __ldg(src + 3); }
                }
                regs[i] = v;
            }
        };
        auto store_chunk = [&](uint32_t buf, const float4 *regs) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const uint32_t g = tid + kRows * i;
                const uint32_t r = g >> 3, c4 = g & 7u;
                *reinterpret_cast<float4 *>(sb + (buf * kRows + r) * kRowStride + c4 * 4) = regs[i];
            }
        };
        float4 regs[8];
        load_chunk(0, regs);
        store_chunk(0, regs);
        __syncthreads();
        for (uint32_t ch = 0; ch < nchunk; ch++) {
            const uint32_t buf = ch & 1u;
            if (ch + 1 < nchunk) load_chunk(ch + 1, regs);
            const float4 *brow = reinterpret_cast<const float4 *>(sb + (buf * kRows + tid) * kRowStride);
            const uint32_t nf4 = min((uint32_t)kDimChunk, dpad - ch * kDimChunk) / 4;
            for (uint32_t c4 = 0; c4 < nf4; c4++) {
                const float4 x = brow[c4];
                const float2 x01 = make_float2(x.x, x.y), x23 = make_float2(x.z, x.w);
                // differences / products of two neighbouring dimensions with packed FP32x2 (FADD2 / FMUL2, RN per element);
                // the sums stay strict left-to-right chains
                float2 xx01 = x01, xx23 = x23;
                if (METRIC == 'c') { xx01 = mul2(x01, x01); xx23 = mul2(x23, x23); }
#pragma unroll
                for (int j = 0; j < QT; j++) {
                    const float4 y = *reinterpret_cast<const float4 *>(sq + j * dpad + ch * kDimChunk + c4 * 4);
                    const float2 y01 = make_float2(y.x, y.y), y23 = make_float2(y.z, y.w);
                    if (METRIC == 'e') {
                        const float2 t01 = sub2(x01, y01), t23 = sub2(x23, y23);
                        const float2 p01 = mul2(t01, t01), p23 = mul2(t23, t23);
                        ab[j] = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(ab[j], p01.x), p01.y), p23.x), p23.y);
                    } else {
                        const float2 p01 = mul2(x01, y01), p23 = mul2(x23, y23);
                        ab[j] = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(ab[j], p01.x), p01.y), p23.x), p23.y);
                        if (METRIC == 'c') {
                            const float2 q01 = mul2(y01, y01), q23 = mul2(y23, y23);
                            a2[j] = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(a2[j], xx01.x), xx01.y), xx23.x), xx23.y);
                            b2[j] = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(b2[j], q01.x), q01.y), q23.x), q23.y);
                        }
                    }
                }
            }
            if (ch + 1 < nchunk) store_chunk(buf ^ 1u, regs);
            __syncthreads();
        }
#pragma unroll
        for (int j = 0; j < QT; j++) sd[j * kRows + tid] = fold_finish<METRIC>(ab[j], a2[j], b2[j]);
        __syncthreads();
        // warp w owns queries w, w+4, ...: filter the tile's 128 distances against the current k-th and insert
#pragma unroll
        for (int jj = 0; jj < (QT + 3) / 4; jj++) {
            const uint32_t j = warp + 4 * jj;
            if (j >= QT || q0 + j >= nq) continue;
            unsigned long long *mk = lk + (size_t)j * k;
            float *md = ld + (size_t)j * k;
            for (uint32_t rb = 0; rb < kRows; rb += 32) {
                const uint64_t row = t0 + rb + lane;
                const float dv = sd[j * kRows + rb + lane];
                const unsigned long long key = ((unsigned long long)order_key(dv) << 32) | (uint32_t)row;
                const unsigned long long worst = lsize[jj] == k ? mk[k - 1] : ~0ull;
                uint32_t m = __ballot_sync(kFull, row < r_end && key < worst);
                while (m) {
                    const uint32_t src = __ffs(m) - 1;
                    m &= m - 1;
                    const unsigned long long kj = __shfl_sync(kFull, key, src);
                    const float dj = __shfl_sync(kFull, dv, src);
                    list_insert(mk, md, lsize[jj], k, kj, dj);
                }
            }
        }
        __syncthreads();
    }
    // write this split's lists
#pragma unroll
    for (int jj = 0; jj < (QT + 3) / 4; jj++) {
        const uint32_t j = warp + 4 * jj;
        if (j >= QT || q0 + j >= nq) continue;
        const size_t o = ((size_t)blockIdx.y * nq + (q0 + j)) * k;
        for (uint32_t i = lane; i < k; i += 32) {
            const bool have = i < lsize[jj];
            part_key[o + i] = have ? lk[(size_t)j * k + i] : ~0ull;
            part_dist[o + i] = have ? ld[(size_t)j * k + i] : CUDART_INF_F;
        }
    }
}

// merge of `shards` ascending lists per query. One warp per query; rank sort over shards×k elements.
// MODE 0: inputs are (order-key|row, dist) partials of k_exact (unique keys; row -> optional keys[] lookup)
// MODE 1: inputs are (key i64, dist) lists from different GPUs; order (dist, key), padding entries (key -1) last
template <int MODE>
__device__ __forceinline__ void merge_entry(const unsigned long long *__restrict__ in_key, const float *__restrict__ in_dist, uint64_t nq,
                                            uint64_t q, uint32_t k, uint32_t e, unsigned long long &hi, unsigned long long &lo, float &dv) {
    const uint32_t s = e / k, i = e % k;
    const size_t o = ((size_t)s * nq + q) * k + i;
    dv = in_dist[o];
    if (MODE == 0) { hi = in_key[o]; lo = e; }
    else {
        const long long key = (long long)in_key[o];
        const bool pad = key < 0 && !(dv < CUDART_INF_F);
        hi = ((unsigned long long)(pad ? 0xFFFFFFFFu : order_key(dv)) << 32) | (pad ? 0xFFFFFFFFu : 0u);
        lo = (unsigned long long)key ^ 0x8000000000000000ull; // signed order; (dist, key) lexicographic
    }
}

// STAGED: the query's shards*k sort keys are staged in shared memory once (16 B each) and every lane ranks its own
// elements (held in registers, at most kMergePerLane) against one broadcast element per step; otherwise (lists too long
// for shared memory) every comparison re-reads global memory.
constexpr uint32_t kMergePerLane = 16;
template <int MODE, bool STAGED>
__global__ void k_merge(const unsigned long long *__restrict__ in_key, const float *__restrict__ in_dist, uint32_t shards,
                        uint64_t nq, uint32_t k, const int64_t *__restrict__ id_map, int64_t *__restrict__ keys_out,
                        float *__restrict__ dist_out) {
    extern __shared__ __align__(16) unsigned long long merge_smem[];
    const uint64_t q = (uint64_t)blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
    if (q >= nq) return;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t total = shards * k;
    auto emit = [&](uint32_t e, uint32_t rank, float dv) {
        if (rank >= k) return;
        const unsigned long long kk = in_key[((size_t)(e / k) * nq + q) * k + (e % k)];
        int64_t id;
        if (MODE == 0) {
            if (kk == ~0ull) id = -1;
            else { const uint32_t row = (uint32_t)(kk & 0xFFFFFFFFull); id = id_map ? id_map[row] : (int64_t)row; }
        } else id = (int64_t)kk;
        keys_out[q * k + rank] = id;
        if (dist_out) dist_out[q * k + rank] = dv;
    };
    if (STAGED) {
        unsigned long long *sh = merge_smem + (size_t)(threadIdx.x >> 5) * total * 2; // (hi, lo) pairs
        unsigned long long mhi[kMergePerLane], mlo[kMergePerLane];
        float mdv[kMergePerLane];
        uint32_t rank[kMergePerLane];
#pragma unroll
        for (uint32_t j = 0; j < kMergePerLane; j++) {
            const uint32_t e = lane + 32 * j;
            rank[j] = 0;
            mhi[j] = ~0ull; mlo[j] = ~0ull; mdv[j] = 0.f;
            if (e < total) {
                merge_entry<MODE>(in_key, in_dist, nq, q, k, e, mhi[j], mlo[j], mdv[j]);
                sh[2 * e] = mhi[j];
                sh[2 * e + 1] = mlo[j];
            }
        }
        __syncwarp();
        for (uint32_t o = 0; o < total; o++) {
            const ulonglong2 x = *reinterpret_cast<const ulonglong2 *>(sh + 2 * o); // broadcast
#pragma unroll
            for (uint32_t j = 0; j < kMergePerLane; j++) {
                const uint32_t e = lane + 32 * j;
                rank[j] += (x.x < mhi[j] || (x.x == mhi[j] && (x.y < mlo[j] || (x.y == mlo[j] && o < e)))) ? 1u : 0u;
            }
        }
#pragma unroll
        for (uint32_t j = 0; j < kMergePerLane; j++)
            if (lane + 32 * j < total) emit(lane + 32 * j, rank[j], mdv[j]);
        return;
    }
    for (uint32_t e = lane; e < total; e += 32) {
        unsigned long long hi, lo; float dv;
        merge_entry<MODE>(in_key, in_dist, nq, q, k, e, hi, lo, dv);
        uint32_t rank = 0;
        for (uint32_t o = 0; o < total; o++) {
            unsigned long long h2, l2; float d2;
            merge_entry<MODE>(in_key, in_dist, nq, q, k, o, h2, l2, d2);
            rank += (h2 < hi || (h2 == hi && (l2 < lo || (l2 == lo && o < e)))) ? 1u : 0u;
        }
        emit(e, rank, dv);
    }
}

// Merge of `shards` lists that are each ASCENDING in their 64-bit sort key (MODE 0 partials: order key << 32 | row, padding
// ~0 at the end), k <= 32 entries per list: a tree of pairwise merges cut at k, lane i holding entry i of both lists and
// ranking it in the other list by binary search (5 steps) — about twenty times fewer comparisons than the rank sort over
// shards * k entries. A query with a list that is not ascending (NaN scores are inserted in front by the tensor-core
// kernel) is ranked by the generic rule instead; results are identical either way.
__global__ void k_merge_sorted(const unsigned long long *__restrict__ in_key, const float *__restrict__ in_dist, uint32_t shards,
                               uint64_t nq, uint32_t k, const int64_t *__restrict__ id_map, int64_t *__restrict__ keys_out,
                               float *__restrict__ dist_out) {
    extern __shared__ __align__(16) unsigned long long merge_smem[];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint64_t q = (uint64_t)blockIdx.x * (blockDim.x / 32) + warp;
    if (q >= nq) return;
    const uint32_t half = (shards + 1) / 2;
    // per warp: keys A [shards][32], keys B [half][32], origins A, origins B (entry index inside the input)
    unsigned long long *base = merge_smem + (size_t)warp * ((size_t)(shards + half) * 32 + (size_t)(shards + half) * 16);
    unsigned long long *kA = base, *kB = base + (size_t)shards * 32;
    uint32_t *oA = reinterpret_cast<uint32_t *>(kB + (size_t)half * 32), *oB = oA + (size_t)shards * 32;
    bool sorted = true;
    for (uint32_t l = 0; l < shards; l++) {
        const unsigned long long key = lane < k ? in_key[((size_t)l * nq + q) * k + lane] : ~0ull;
        kA[l * 32 + lane] = key;
        oA[l * 32 + lane] = l * k + lane;
        const unsigned long long nxt = __shfl_down_sync(0xFFFFFFFFu, key, 1);
        sorted &= lane == 31 || key <= nxt;
    }
    __syncwarp();
    if (__ballot_sync(0xFFFFFFFFu, !sorted)) { // generic rank sort over all entries (rare)
        const uint32_t total = shards * 32;
        for (uint32_t e = lane; e < total; e += 32) {
            const unsigned long long key = kA[e];
            uint32_t rank = 0;
            for (uint32_t o = 0; o < total; o++) { const unsigned long long x = kA[o]; rank += (x < key || (x == key && o < e)) ? 1u : 0u; }
            if (rank < k) {
                const uint32_t org = oA[e];
                int64_t id = -1;
                if (key != ~0ull) { const uint32_t row = (uint32_t)(key & 0xFFFFFFFFull); id = id_map ? id_map[row] : (int64_t)row; }
                keys_out[q * k + rank] = id;
                if (dist_out) dist_out[q * k + rank] = key != ~0ull ? in_dist[((size_t)(org / k) * nq + q) * k + org % k] : CUDART_INF_F;
            }
        }
        return;
    }
    uint32_t n = shards;
    while (n > 1) {
        const uint32_t pairs = n / 2;
        for (uint32_t p = 0; p < pairs; p++) {
            const unsigned long long *X = kA + (2 * p) * 32, *Y = kA + (2 * p + 1) * 32;
            const unsigned long long x = X[lane], y = Y[lane];
            uint32_t lo = 0, hi = 32; // #{Y < x}
            while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (Y[mid] < x) lo = mid + 1; else hi = mid; }
            const uint32_t rx = lane + lo;
            lo = 0; hi = 32;          // #{X <= y}
            while (lo < hi) { const uint32_t mid = (lo + hi) >> 1; if (X[mid] <= y) lo = mid + 1; else hi = mid; }
            const uint32_t ry = lane + lo;
            if (rx < 32) { kB[p * 32 + rx] = x; oB[p * 32 + rx] = oA[(2 * p) * 32 + lane]; }
            if (ry < 32) { kB[p * 32 + ry] = y; oB[p * 32 + ry] = oA[(2 * p + 1) * 32 + lane]; }
        }
        if (n & 1u) { kB[pairs * 32 + lane] = kA[(n - 1) * 32 + lane]; oB[pairs * 32 + lane] = oA[(n - 1) * 32 + lane]; }
        __syncwarp();
        n = (n + 1) / 2;
        for (uint32_t l = 0; l < n; l++) { kA[l * 32 + lane] = kB[l * 32 + lane]; oA[l * 32 + lane] = oB[l * 32 + lane]; }
        __syncwarp();
    }
    if (lane < k) {
        const unsigned long long key = kA[lane];
        const uint32_t org = oA[lane];
        int64_t id = -1;
        if (key != ~0ull) { const uint32_t row = (uint32_t)(key & 0xFFFFFFFFull); id = id_map ? id_map[row] : (int64_t)row; }
        keys_out[q * k + lane] = id;
        if (dist_out) dist_out[q * k + lane] = key != ~0ull ? in_dist[((size_t)(org / k) * nq + q) * k + org % k] : CUDART_INF_F;
    }
}

template <int MODE>
static cudaError_t launch_merge(const unsigned long long *in_key, const float *in_dist, uint32_t shards, uint64_t nq, uint32_t k,
                                const int64_t *id_map, int64_t *keys_out, float *dist_out, cudaStream_t st) {
    const uint32_t wpb = 4, total = shards * k;
    const unsigned grid = (unsigned)((nq + wpb - 1) / wpb);
    if (MODE == 0 && k <= 32 && shards >= 2 && shards <= 64) { // partial lists of k_exact / k_tiles_tc: ascending
        const uint32_t half = (shards + 1) / 2;
        const size_t per_warp = ((size_t)(shards + half) * 32 + (size_t)(shards + half) * 16) * 8;
        if (per_warp * wpb <= 96 * 1024) {
            if (per_warp * wpb > 48 * 1024) cudaFuncSetAttribute(k_merge_sorted, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
            k_merge_sorted<<<grid, wpb * 32, per_warp * wpb, st>>>(in_key, in_dist, shards, nq, k, id_map, keys_out, dist_out);
            count_launch();
            return cudaGetLastError();
        }
    }
    const size_t smem = (size_t)wpb * total * 16;
    if (total <= 32 * kMergePerLane && smem <= 48 * 1024) k_merge<MODE, true><<<grid, wpb * 32, smem, st>>>(in_key, in_dist, shards, nq, k, id_map, keys_out, dist_out);
    else k_merge<MODE, false><<<grid, wpb * 32, 0, st>>>(in_key, in_dist, shards, nq, k, id_map, keys_out, dist_out);
    count_launch();
    return cudaGetLastError();
}

template <char METRIC, int QT>
static cudaError_t run_exact(const float *base, uint64_t n, const int64_t *keys, const float *queries, uint64_t nq, uint32_t d,
                             uint32_t k, int64_t *ids_out, float *dist_out, cudaStream_t st) {
    const uint32_t dpad = (d + 3) / 4 * 4;
    const size_t smem = (size_t)QT * dpad * 4 + 2ull * kRows * kRowStride * 4 + (size_t)QT * kRows * 4 + (size_t)QT * k * 12 + 16;
    if (smem > 220 * 1024) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(k_exact<METRIC, QT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const uint64_t qtiles = (nq + QT - 1) / QT;
    uint64_t splits = 1;
    if (qtiles < 2ull * sms) splits = (2ull * sms + qtiles - 1) / qtiles;
    const uint64_t tiles = (n + kRows - 1) / kRows;
    if (splits > tiles) splits = tiles ? tiles : 1;
    if (splits > 1024) splits = 1024;
    const uint64_t rows_per_split = ((tiles + splits - 1) / splits) * kRows;
    splits = n ? (n + rows_per_split - 1) / rows_per_split : 1;
    unsigned long long *pk = nullptr; float *pd = nullptr;
    e = cudaMallocAsync(&pk, splits * nq * k * 8, st);
    if (e != cudaSuccess) return e;
    e = cudaMallocAsync(&pd, splits * nq * k * 4, st);
    if (e != cudaSuccess) { cudaFreeAsync(pk, st); return e; }
    dim3 grid((unsigned)qtiles, (unsigned)splits);
    k_exact<METRIC, QT><<<grid, kRows, smem, st>>>(base, n, queries, nq, d, k, rows_per_split, pk, pd);
    count_launch();
    e = launch_merge<0>(pk, pd, (uint32_t)splits, nq, k, keys, ids_out, dist_out, st);
    cudaFreeAsync(pk, st);
    cudaFreeAsync(pd, st);
    return e;
}

cudaError_t launch_exact_topk(char metric, const float *base, uint64_t n, const int64_t *keys, const float *queries, uint64_t nq,
                              uint32_t d, uint32_t k, int64_t *ids_out, float *dist_out, cudaStream_t st) {
    if (!nq) return cudaSuccess;
    switch (metric) {
    case 'c': return run_exact<'c', 8>(base, n, keys, queries, nq, d, k, ids_out, dist_out, st);
    case 'i': return run_exact<'i', 16>(base, n, keys, queries, nq, d, k, ids_out, dist_out, st);
    default: return run_exact<'e', 16>(base, n, keys, queries, nq, d, k, ids_out, dist_out, st);
    }
}

// merge of per-split candidate lists of the tensor-core path: ids = row indices (-1 = empty), "distances" = scores
cudaError_t launch_merge_partials(const unsigned long long *pk, const float *ps, uint32_t splits, uint64_t nq, uint32_t kc, int64_t *rows_out,
                                  float *scores_out, cudaStream_t st) {
    if (!nq) return cudaSuccess;
    return launch_merge<0>(pk, ps, splits, nq, kc, nullptr, rows_out, scores_out, st);
}

__global__ void k_gather_rows(const float *__restrict__ src, const uint32_t *__restrict__ idx, uint64_t m, uint32_t d, float *__restrict__ dst) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m * d) dst[i] = src[(uint64_t)idx[i / d] * d + i % d];
}
__global__ void k_scatter_results(const int64_t *__restrict__ ids, const float *__restrict__ dist, const uint32_t *__restrict__ idx, uint64_t m,
                                  uint32_t k, int64_t *__restrict__ ids_out, float *__restrict__ dist_out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m * k) return;
    const uint64_t o = (uint64_t)idx[i / k] * k + i % k;
    ids_out[o] = ids[i];
    if (dist_out) dist_out[o] = dist[i];
}

// Recomputes the queries whose flag is set with the all-pairs exact kernel (synchronises the stream).
cudaError_t launch_exact_topk_flagged(char metric, const float *base, uint64_t n, const int64_t *keys, const float *queries, uint64_t nq,
                                      uint32_t d, uint32_t k, int64_t *ids_out, float *dist_out, const uint32_t *flags, uint64_t *flagged_out,
                                      cudaStream_t st) {
    std::vector<uint32_t> h(nq);
    cudaError_t e = cudaMemcpyAsync(h.data(), flags, nq * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return e;
    std::vector<uint32_t> idx;
    for (uint64_t i = 0; i < nq; i++) if (h[i]) idx.push_back((uint32_t)i);
    if (flagged_out) *flagged_out = idx.size();
    if (idx.empty()) return cudaSuccess;
    const uint64_t m = idx.size();
    uint32_t *d_idx = nullptr; float *qtmp = nullptr, *dtmp = nullptr; int64_t *itmp = nullptr;
    auto cleanup = [&]() { cudaFreeAsync(d_idx, st); cudaFreeAsync(qtmp, st); cudaFreeAsync(dtmp, st); cudaFreeAsync(itmp, st); };
    if ((e = cudaMallocAsync(&d_idx, m * 4, st)) != cudaSuccess || (e = cudaMallocAsync(&qtmp, m * d * 4, st)) != cudaSuccess ||
        (e = cudaMallocAsync(&dtmp, m * k * 4, st)) != cudaSuccess || (e = cudaMallocAsync(&itmp, m * k * 8, st)) != cudaSuccess) { cleanup(); return e; }
    e = cudaMemcpyAsync(d_idx, idx.data(), m * 4, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) {
        k_gather_rows<<<(unsigned)((m * d + 255) / 256), 256, 0, st>>>(queries, d_idx, m, d, qtmp);
        count_launch();
        e = launch_exact_topk(metric, base, n, keys, qtmp, m, d, k, itmp, dtmp, st);
    }
    if (e == cudaSuccess) {
        k_scatter_results<<<(unsigned)((m * k + 255) / 256), 256, 0, st>>>(itmp, dtmp, d_idx, m, k, ids_out, dist_out);
        count_launch();
        e = cudaStreamSynchronize(st); // idx must outlive the copies above
    }
    cleanup();
    return e;
}

cudaError_t launch_merge_topk(const int64_t *gk, const float *gd, uint32_t shards, uint64_t nq, uint32_t k, int64_t *keys_out,
                              float *dist_out, cudaStream_t st) {
    if (!nq) return cudaSuccess;
    return launch_merge<1>(reinterpret_cast<const unsigned long long *>(gk), gd, shards, nq, k, nullptr, keys_out, dist_out, st);
}

// ---------------------------------------------------------------------------------------------------------------
// Multi-GPU exchange without a collective library: every rank's merge kernel reads the peers' per-shard lists straight
// out of THEIR memory over NVLink (the host maps the peers' buffers into this process, e.g. through CUDA IPC / symmetric
// memory) and the ranks synchronise through one 32-bit flag per (reader, writer) pair:
//   1. this kernel follows the rank's own search in stream order, so when it starts the rank's lists are complete; CTA 0
//      publishes `step` into slot [rank] of every peer's flag array (release, system scope);
//   2. every warp waits until all slots of its OWN flag array have reached `step` (acquire, system scope) — the peers' lists
//      of this step are then complete and visible;
//   3. the lists are merged by (distance, key) like k_merge<1>.
// Steps must increase by one per call on every rank and result buffers must be double-buffered by the caller (a rank may
// start step s+1's search while a slow peer still merges step s). A peer that never arrives trips the time-out
// (GVSS_PEER_TIMEOUT_MS, default 20 s): the merge writes nothing and records the late rank in an error word the host reads
// with gvss_peer_exchange_status — no trap, the context survives.
// ---------------------------------------------------------------------------------------------------------------
constexpr uint32_t kMaxPeers = 16;
struct PeerLists {
    const unsigned long long *keys[kMaxPeers];
    const float *dist[kMaxPeers];
    uint32_t *flags[kMaxPeers]; // flags[r] = rank r's flag array (kMaxPeers u32), mapped here
};

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ld_relaxed_sys_f32(const float *p) {
    float v;
    asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}

static __device__ unsigned int g_peer_timeouts; // bit r: a merge on this device gave up waiting for rank r

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__global__ void k_merge_peers(PeerLists pl, uint32_t rank, uint32_t shards, uint32_t step, uint64_t nq, uint32_t k,
                              int64_t *__restrict__ keys_out, float *__restrict__ dist_out, unsigned long long timeout_ns) {
    const uint32_t lane = threadIdx.x & 31u;
    // step == 0: one process drives all GPUs and has ordered this launch behind the shards' searches with CUDA events
    // (router.cu): no handshake, the lists are simply read where they are
    if (step && blockIdx.x == 0 && threadIdx.x < shards) {
        __threadfence_system();
        st_release_sys(pl.flags[threadIdx.x] + rank, step);
    }
    // Wait for the PEERS only. This rank's own lists are complete by stream order, and its own flag slot is written by CTA 0
    // of this very grid: waiting on it would make every CTA depend on CTA 0 having been scheduled, which CUDA does not
    // promise once the grid exceeds what the GPU holds at once.
    bool late = false;
    if (step && lane < shards && lane != rank) {
        const uint32_t *mine = pl.flags[rank] + lane;
        const unsigned long long t0 = global_ns();
        while ((int32_t)(ld_acquire_sys(mine) - step) < 0) { // wrap-safe "reached step"
            if (global_ns() - t0 > timeout_ns) { late = true; break; }
            __nanosleep(200);
        }
        if (late) atomicOr(&g_peer_timeouts, 1u << lane);
    }
    // a peer that never arrived: no result is written (its list would be stale); the host learns it from
    // gvss_peer_exchange_status instead of losing the context to a trap
    if (__ballot_sync(kFull, late)) return;
    const uint64_t q = (uint64_t)blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);
    if (q >= nq) return;
    const uint32_t total = shards * k;
    auto get = [&](uint32_t e, unsigned long long &hi, unsigned long long &lo, float &dv, long long &key) {
        const uint32_t s = e / k, i = e % k;
        key = (long long)ld_relaxed_sys_u64(pl.keys[s] + q * k + i);
        dv = ld_relaxed_sys_f32(pl.dist[s] + q * k + i);
        const bool pad = key < 0 && !(dv < CUDART_INF_F);
        hi = ((unsigned long long)(pad ? 0xFFFFFFFFu : order_key(dv)) << 32) | (pad ? 0xFFFFFFFFu : 0u);
        lo = (unsigned long long)key ^ 0x8000000000000000ull;
    };
    // each lane owns up to 8 elements (shards * k <= 256); the others are broadcast from the owners' registers
    unsigned long long mhi[8], mlo[8];
    float mdv[8];
    long long mkey[8];
    uint32_t rk[8];
#pragma unroll
    for (uint32_t j = 0; j < 8; j++) {
        rk[j] = 0; mhi[j] = ~0ull; mlo[j] = ~0ull; mdv[j] = 0.f; mkey[j] = -1;
        if (lane + 32 * j < total) get(lane + 32 * j, mhi[j], mlo[j], mdv[j], mkey[j]);
    }
#pragma unroll
    for (uint32_t jo = 0; jo < 8; jo++) {
        if (jo * 32 >= total) break;
        for (uint32_t src = 0; src < 32 && jo * 32 + src < total; src++) {
            const unsigned long long xh = __shfl_sync(kFull, mhi[jo], src), xl = __shfl_sync(kFull, mlo[jo], src);
            const uint32_t o = jo * 32 + src;
#pragma unroll
            for (uint32_t j = 0; j < 8; j++) {
                const uint32_t e = lane + 32 * j;
                rk[j] += (xh < mhi[j] || (xh == mhi[j] && (xl < mlo[j] || (xl == mlo[j] && o < e)))) ? 1u : 0u;
            }
        }
    }
#pragma unroll
    for (uint32_t j = 0; j < 8; j++) {
        if (lane + 32 * j < total && rk[j] < k) {
            keys_out[q * k + rk[j]] = mkey[j];
            if (dist_out) dist_out[q * k + rk[j]] = mdv[j];
        }
    }
}

cudaError_t launch_merge_topk_peers(const int64_t *const *peer_keys, const float *const *peer_dist, uint32_t *const *peer_flags,
                                    uint32_t rank, uint32_t shards, uint32_t step, uint64_t nq, uint32_t k, int64_t *keys_out,
                                    float *dist_out, cudaStream_t st) {
    if (!shards || shards > kMaxPeers || rank >= shards || (uint64_t)shards * k > 256) return cudaErrorInvalidValue;
    PeerLists pl{};
    for (uint32_t s = 0; s < shards; s++) {
        pl.keys[s] = reinterpret_cast<const unsigned long long *>(peer_keys[s]);
        pl.dist[s] = peer_dist[s];
        pl.flags[s] = peer_flags ? peer_flags[s] : nullptr;
    }
    const uint32_t wpb = 4;
    const unsigned grid = (unsigned)(((nq ? nq : 1) + wpb - 1) / wpb); // at least CTA 0: the flags must be published even for an empty batch
    // how long a merge waits for a peer before it gives up (GVSS_PEER_TIMEOUT_MS, default 20 s: a slow shard, a first-call
    // module load or a reallocation on the peer must not look like a dead rank)
    static const unsigned long long timeout_ns = [] {
        const char *e = getenv("GVSS_PEER_TIMEOUT_MS");
        const unsigned long long ms = e ? strtoull(e, nullptr, 10) : 20000ull;
        return (ms ? ms : 1ull) * 1000000ull;
    }();
    k_merge_peers<<<grid, wpb * 32, 0, st>>>(pl, rank, shards, step, nq, k, keys_out, dist_out, timeout_ns);
    count_launch();
    return cudaGetLastError();
}

cudaError_t read_peer_timeouts(uint32_t *out) {
    unsigned int v = 0;
    cudaError_t e = cudaMemcpyFromSymbol(&v, g_peer_timeouts, sizeof v);
    if (e != cudaSuccess) return e;
    if (v) {
        const unsigned int z = 0;
        e = cudaMemcpyToSymbol(g_peer_timeouts, &z, sizeof z);
    }
    *out = v;
    return e;
}

} // namespace gvss
