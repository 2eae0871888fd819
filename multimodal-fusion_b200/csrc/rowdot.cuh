// One summation order for every FP32 row reduction whose result is compared or reused across
// kernels (squared norms from split.cu, the exact distances / dot products of rerank.cu, the cosine
// edge weights of graph_ops.cu).  Because all of them go through these helpers, a cosine computed as
// a by-product of the re-rank is bit-identical to the one hg_edge_weights_cosine computes later.
//
// The order, for a row of d floats, is that of ONE warp:
//   * with 16-byte-aligned rows and d % 4 == 0, lane l owns the float4 groups l, l + 32, ... and keeps
//     two accumulators, one fed by the x then z components and one by y then w -- exactly what two
//     packed fma.rn.f32x2 on (x, y) and (z, w) do -- added at the end (lo + hi); otherwise it owns
//     the scalars l, l + 32, ... in one accumulator;
//   * a 5-step xor butterfly over the 32 lanes.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace hg {

// ---- packed FP32 pairs (sm_100: fma.rn.f32x2 issues two FMAs per instruction) -------------------
__device__ __forceinline__ uint64_t pk_fma(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ uint64_t pk_sub(uint64_t a, uint64_t b) {          // a - b, one rounding per half
    constexpr uint64_t kMinusOne = 0xBF800000BF800000ull;
    return pk_fma(b, kMinusOne, a);
}
__device__ __forceinline__ float pk_hsum(uint64_t v) {                         // lo + hi
    return __uint_as_float(uint32_t(v)) + __uint_as_float(uint32_t(v >> 32));
}

__device__ __forceinline__ float warp_butterfly(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ bool rows_vectorisable(const float* base, int64_t d) {
    return (d & 3) == 0 && (reinterpret_cast<uintptr_t>(base) & 15) == 0;
}

// dot(a, b), |a|^2, |b|^2 of two rows by one warp, in the canonical order (all lanes get the result)
__device__ __forceinline__ void warp_dot_norms(const float* __restrict__ a, const float* __restrict__ b, int64_t d,
                                               int lane, bool vec, float& dot, float& na, float& nb) {
    if (vec) {
        uint64_t ad = 0, aa = 0, ab = 0;
        for (int64_t v = lane; v < (d >> 2); v += 32) {
            const ulonglong2 x = __ldg(reinterpret_cast<const ulonglong2*>(a) + v);
            const ulonglong2 y = __ldg(reinterpret_cast<const ulonglong2*>(b) + v);
            ad = pk_fma(x.x, y.x, ad); ad = pk_fma(x.y, y.y, ad);
            aa = pk_fma(x.x, x.x, aa); aa = pk_fma(x.y, x.y, aa);
            ab = pk_fma(y.x, y.x, ab); ab = pk_fma(y.y, y.y, ab);
        }
        dot = pk_hsum(ad); na = pk_hsum(aa); nb = pk_hsum(ab);
    } else {
        dot = 0.f; na = 0.f; nb = 0.f;
        for (int64_t c = lane; c < d; c += 32) {
            const float x = __ldg(a + c), y = __ldg(b + c);
            dot = fmaf(x, y, dot); na = fmaf(x, x, na); nb = fmaf(y, y, nb);
        }
    }
    dot = warp_butterfly(dot);
    na = warp_butterfly(na);
    nb = warp_butterfly(nb);
}

// |a|^2 of one row by one warp, canonical order
__device__ __forceinline__ float warp_sqnorm(const float* __restrict__ a, int64_t d, int lane, bool vec) {
    float na;
    if (vec) {
        uint64_t aa = 0;
        for (int64_t v = lane; v < (d >> 2); v += 32) {
            const ulonglong2 x = __ldg(reinterpret_cast<const ulonglong2*>(a) + v);
            aa = pk_fma(x.x, x.x, aa); aa = pk_fma(x.y, x.y, aa);
        }
        na = pk_hsum(aa);
    } else {
        na = 0.f;
        for (int64_t c = lane; c < d; c += 32) { const float x = __ldg(a + c); na = fmaf(x, x, na); }
    }
    return warp_butterfly(na);
}

// max(0, cos) with the reference's clamp (F.cosine_similarity eps = 1e-8, preprocess_hypergraph.py:419)
__device__ __forceinline__ float cosine_from(float dot, float na, float nb) {
    return dot / (fmaxf(sqrtf(na), 1e-8f) * fmaxf(sqrtf(nb), 1e-8f));
}

}  // namespace hg
