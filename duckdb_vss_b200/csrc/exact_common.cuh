// Device helpers shared by the exact top-k kernels (exact.cu: all-pairs FP32 folds; exact_tc.cu: tensor-core candidates + re-rank).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace gvss {

__device__ __forceinline__ uint32_t order_key(float d) {
    if (d != d) return 0xFFFFFFFFu;           // NaN sorts last (DuckDB orders NaN as the greatest value)
    if (d == 0.0f) return 0x80000000u;        // -0 == +0
    const uint32_t u = __float_as_uint(d);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

template <char METRIC> __device__ __forceinline__ float fold_finish(float ab, float a2, float b2) {
    if (METRIC == 'e') return __fsqrt_rn(ab);
    if (METRIC == 'i') return -ab;
    const float sim = __fdiv_rn(ab, __fsqrt_rn(__fmul_rn(a2, b2)));
    const float mn = (1.0f < sim) ? 1.0f : sim;   // std::min(sim, 1)
    const float mx = (-1.0f < mn) ? mn : -1.0f;   // std::max(-1, .)  (a NaN similarity ends as -1, like DuckDB)
    return __fsub_rn(1.0f, mx);
}

__device__ __forceinline__ uint32_t order_key_f(float d) { return order_key(d); }

} // namespace gvss
