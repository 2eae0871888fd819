// Operand preparation for the tensor-core Gram product.
//
// FP32 row x  ->  s = op_scale(|x|^2), a power of two with |s x| in [2^5, 2^6) (knn_tc.cuh);
//                 hi = fp16(s x), lo = fp16(s x - hi)  (s x = hi + lo up to 2^-22 |s x| + 2^-25 per element)
// plus the FP32 squared row norm taken from the ORIGINAL data.  Replaces
// `torch.sum(features ** 2, dim=1)` (reference similarity_kernel.py:43) and
// prepares the inputs of the MMA that replaces `torch.mm` (:46).
//
// Layout written: ops[row][col / 32][half][col % 32], half 0 = hi, 1 = lo: every 128-byte run of a
// row holds one 32-wide k-block as 32 hi | 32 lo values, which is one TMA box row and one 128-byte
// swizzled shared-memory row of the fused kernel.  Rows are padded to a multiple of 256 and cols to
// a multiple of 64 with zeros, so that TMA boxes never straddle undefined memory and padded rows
// contribute dot = 0.
//
// HBM-bound: reads 4 B and writes 4 B per element; one warp per row (4 rows per CTA),
// 128-bit loads, 64-bit stores.  Two passes over the row: the norm (which fixes the scale), then the split --
// the second pass finds the row in L1.
#include "knn_tc.cuh"
#include "rowdot.cuh"
#include <cuda_fp16.h>
#include <stdlib.h>

namespace hg {

constexpr int SPLIT_WARPS = 4;          // rows per CTA, one warp each

// Row pitch of the operand buffer.  HG_B200_OPS_PAD_BYTES (multiple of 16) overrides the pad.
int64_t ops_row_elems(int64_t d_pad) {
    static const int pad_bytes = [] {
        if (const char* e = getenv("HG_B200_OPS_PAD_BYTES")) {
            const int v = atoi(e);
            if (v >= 0 && v <= 4096 && v % 16 == 0) return v;
        }
        return 0;
    }();
    return 2 * d_pad + pad_bytes / 2;
}

// hi half of a scaled element: FP16 round to nearest, or (hi8, dense-output launches) rounded to 8 significant bits
// first.  The tensor core aligns the 16 products of a k-step to the largest and truncates what falls below its
// adder: products of two 11-bit values are 22 bits wide and lose their tails there (measured: 1.2e-5 on exp(-d2) at
// D = 1536 against 6e-6), products of two 8-bit values go through whole -- and lo = fp16(x - hi) still holds the next
// 11 bits, so hi + lo carries 19 bits; the dense mode adds lo*lo as well.
__device__ __forceinline__ __half hi_half(float x, int hi8) {
    return hi8 ? __float2half_rn(__bfloat162float(__float2bfloat16_rn(x))) : __float2half_rn(x);
}

template <bool VEC4>
__global__ void __launch_bounds__(SPLIT_WARPS * 32)
split_rows_kernel(const float* __restrict__ feats, int64_t n, int64_t d, int64_t n_pad, int64_t d_pad,
                  int64_t row_elems, op16_t* __restrict__ ops, float* __restrict__ sq_norms,
                  const RrSlide* __restrict__ g_slides, const int32_t* __restrict__ g_blk2slide, int hi8,
                  float* __restrict__ rho_max) {
    const int lane = threadIdx.x & 31;
    int64_t row = int64_t(blockIdx.x) * SPLIT_WARPS + (threadIdx.x >> 5);
    if (row >= n_pad) return;
    op16_t* out_row = ops + row * row_elems;
    float* norm_out = sq_norms != nullptr ? sq_norms + row : nullptr;
    if (g_slides != nullptr) {                   // grouped launch: operand row `row` is row (row - off) of its slide
        const RrSlide sl = g_slides[g_blk2slide[row >> 8]];
        feats = sl.feats; n = sl.n; row -= sl.off;
    }
    // column c lives at (c / 32) * 64 + (c % 32) (hi) and 32 further (lo)
    auto hi_at = [&](int64_t c) { return out_row + ((c >> 5) << 6) + (c & 31); };
    if (row >= n) {                              // padding rows are zero in both halves
        for (int64_t v = lane; v < (d_pad >> 1); v += 32) reinterpret_cast<uint2*>(out_row)[v] = make_uint2(0u, 0u);
        return;
    }
    const float* src = feats + row * d;
    // one warp per row, lane l owns the float4 groups l, l + 32, ...: the squared norm comes out in
    // the canonical order of rowdot.cuh in the same pass
    float norm, res2 = 0.f;                      // res2: |s x - hi|^2, what a product that skips this row's lo half leaves out
    if (VEC4) {
        const int64_t nvec = d >> 2;
        float lo = 0.f, hi = 0.f;               // the two halves of the packed accumulator
        for (int64_t v = lane; v < (d_pad >> 2); v += 32) {
            float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
            if (v < nvec) x = __ldg(reinterpret_cast<const float4*>(src) + v);
            lo = fmaf(x.x, x.x, lo); hi = fmaf(x.y, x.y, hi);
            lo = fmaf(x.z, x.z, lo); hi = fmaf(x.w, x.w, hi);
        }
        norm = warp_butterfly(lo + hi);
        const float s = op_scale(norm);
        for (int64_t v = lane; v < (d_pad >> 2); v += 32) {
            float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
            if (v < nvec) x = __ldg(reinterpret_cast<const float4*>(src) + v);
            x.x *= s; x.y *= s; x.z *= s; x.w *= s;
            const __half h0 = hi_half(x.x, hi8), h1 = hi_half(x.y, hi8);
            const __half h2 = hi_half(x.z, hi8), h3 = hi_half(x.w, hi8);
            const float r0 = x.x - __half2float(h0), r1 = x.y - __half2float(h1);     // exact: the residue of a rounding
            const float r2 = x.z - __half2float(h2), r3 = x.w - __half2float(h3);
            res2 = fmaf(r0, r0, fmaf(r1, r1, fmaf(r2, r2, fmaf(r3, r3, res2))));
            const __half l0 = __float2half_rn(r0), l1 = __float2half_rn(r1);
            const __half l2 = __float2half_rn(r2), l3 = __float2half_rn(r3);
            uint2 hv, lv;
            hv.x = uint32_t(__half_as_ushort(h0)) | (uint32_t(__half_as_ushort(h1)) << 16);
            hv.y = uint32_t(__half_as_ushort(h2)) | (uint32_t(__half_as_ushort(h3)) << 16);
            lv.x = uint32_t(__half_as_ushort(l0)) | (uint32_t(__half_as_ushort(l1)) << 16);
            lv.y = uint32_t(__half_as_ushort(l2)) | (uint32_t(__half_as_ushort(l3)) << 16);
            op16_t* dst = hi_at(v << 2);
            *reinterpret_cast<uint2*>(dst) = hv;
            *reinterpret_cast<uint2*>(dst + 32) = lv;
        }
    } else {
        norm = 0.f;
        for (int64_t c = lane; c < d_pad; c += 32) {
            const float x = (c < d) ? __ldg(src + c) : 0.f;
            norm = fmaf(x, x, norm);
        }
        norm = warp_butterfly(norm);
        const float s = op_scale(norm);
        for (int64_t c = lane; c < d_pad; c += 32) {
            const float x = ((c < d) ? __ldg(src + c) : 0.f) * s;
            const __half h = hi_half(x, hi8);
            const float r = x - __half2float(h);
            res2 = fmaf(r, r, res2);
            unsigned short* dst = reinterpret_cast<unsigned short*>(hi_at(c));
            dst[0] = __half_as_ushort(h);
            dst[32] = __half_as_ushort(__float2half_rn(r));
        }
    }
    if (lane == 0 && norm_out != nullptr) *norm_out = norm;
    if (rho_max != nullptr) {
        // rho = |s x - hi| / |s x| (<= 2^-11 up to the subnormal grid); the largest over the rows bounds the error of a
        // product that reads only hi halves of one operand (tc_eps, knn_tc.cuh).  Non-negative floats order like ints.
        res2 = warp_butterfly(res2);
        const float sn = op_scale(norm) * sqrtf(norm);
        if (lane == 0 && sn > 0.f) atomicMax(reinterpret_cast<int*>(rho_max), __float_as_int(sqrtf(res2) / sn));
    }
}

int launch_split(const float* feats, int64_t n, int64_t d, op16_t* ops, float* sq_norms,
                 cudaStream_t st, bool dense, float* rho_max) {
    const int64_t n_pad = ops_rows(n), d_pad = ops_cols(d);
    const bool vec = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(feats) & 15) == 0);
    if (vec)
        split_rows_kernel<true><<<(unsigned)ceil_div(n_pad, SPLIT_WARPS), SPLIT_WARPS * 32, 0, st>>>(feats, n, d, n_pad, d_pad, ops_row_elems(d_pad), ops, sq_norms, nullptr, nullptr, dense ? 1 : 0, rho_max);
    else
        split_rows_kernel<false><<<(unsigned)ceil_div(n_pad, SPLIT_WARPS), SPLIT_WARPS * 32, 0, st>>>(feats, n, d, n_pad, d_pad, ops_row_elems(d_pad), ops, sq_norms, nullptr, nullptr, dense ? 1 : 0, rho_max);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

// the rows of a whole group of slides (device tables, knn_tc.cuh) in one launch; n_pad = the group's operand rows
int launch_split_group(const RrSlide* g_slides, const int32_t* g_blk2slide, int64_t n_pad, int64_t d, bool vec,
                       op16_t* ops, float* sq_norms, cudaStream_t st, float* rho_max) {
    const int64_t d_pad = ops_cols(d);
    const unsigned blocks = (unsigned)ceil_div(n_pad, SPLIT_WARPS);
    if (vec)
        split_rows_kernel<true><<<blocks, SPLIT_WARPS * 32, 0, st>>>(nullptr, 0, d, n_pad, d_pad, ops_row_elems(d_pad), ops, sq_norms, g_slides, g_blk2slide, 0, rho_max);
    else
        split_rows_kernel<false><<<blocks, SPLIT_WARPS * 32, 0, st>>>(nullptr, 0, d, n_pad, d_pad, ops_row_elems(d_pad), ops, sq_norms, g_slides, g_blk2slide, 0, rho_max);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

}  // namespace hg

extern "C" size_t hg_operand_bytes(int64_t n, int64_t d) {
    if (n <= 0 || d <= 0) return 0;
    return size_t(hg::ops_rows(n)) * size_t(hg::ops_row_elems(hg::ops_cols(d))) * sizeof(hg::op16_t);
}

extern "C" int hg_split_operands(const float* feats, int64_t n, int64_t d, void* ops, float* sq_norms,
                                 void* stream) {
    HG_REQUIRE(feats && ops && n > 0 && d > 0, HG_ERR_BAD_ARG, "hg_split_operands: bad argument");
    return hg::launch_split(feats, n, d, static_cast<hg::op16_t*>(ops), sq_norms,
                            static_cast<cudaStream_t>(stream), false, nullptr);
}
