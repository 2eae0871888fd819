// FP32 SIMT tile kernel: dense Gaussian similarity (A1/A2/A7) and the plain-FP32
// variant of the fused similarity + top-k (precision = 1 in hg_knn_feature).
//
// Reference semantics:
//   expansion form  |a|^2 + |b|^2 - 2 a.b, un-clamped     similarity_kernel.py:43-52, 79-84
//   direct form     sum (a-b)^2                            preprocess_hypergraph.py:254-256
//   product K_h*K_g folded into one exp                    similarity_kernel.py:122
//
// 64 x 64 output tile per 256-thread CTA, 4 x 4 micro-tile per thread, BK = 16,
// operands transposed into padded shared memory so the inner loop is two
// conflict-free LDS.128 + 16 FFMA.  The dense path is for small N (N x N fits);
// large N goes through the tcgen05 kernel in knn_tc.cu which never writes N x N.
#include "sim_simt.cuh"

namespace hg {

constexpr int BM = 64, BN = 64, BK = 16, PADT = 4;

enum : int { MODE_DENSE = 0, MODE_TOPK = 1 };

template <bool VEC>
__device__ __forceinline__ void load4(const float* base, int64_t rows, int64_t d, int64_t r, int64_t c,
                                      float (&v)[4]) {
    if (VEC) {
        if (r < rows && c + 3 < d) {
            float4 t = __ldg(reinterpret_cast<const float4*>(base + r * d + c));
            v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
            return;
        }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) v[q] = (r < rows && c + q < d) ? __ldg(base + r * d + c + q) : 0.f;
}

template <int MODE, int FORM, bool VEC, int KMAX>
__global__ void __launch_bounds__(256)
sim_tile_kernel(SimArgs p) {
    __shared__ __align__(16) float As[BK][BM + PADT];
    __shared__ __align__(16) float Bs[BK][BN + PADT];
    __shared__ float S[MODE == MODE_TOPK ? BM : 1][MODE == MODE_TOPK ? BN + 1 : 1];

    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int lr = tid >> 2, lk = (tid & 3) * 4;     // loader: row within tile, k offset
    const int64_t row0 = (MODE == MODE_TOPK ? p.row_begin : 0) + int64_t(blockIdx.y) * BM;

    // running top-k of row (row0 + tid) for tid < BM, ascending (score, index)
    float lv[KMAX];
    int li[KMAX];
    if (MODE == MODE_TOPK) {
#pragma unroll
        for (int q = 0; q < KMAX; ++q) { lv[q] = INFINITY; li[q] = -1; }
    }

    const int64_t ct_begin = (MODE == MODE_TOPK) ? 0 : blockIdx.x;
    const int64_t ct_end = (MODE == MODE_TOPK) ? ceil_div(p.n, BN) : blockIdx.x + 1;
    for (int64_t ct = ct_begin; ct < ct_end; ++ct) {
        const int64_t col0 = ct * BN;
        float acc[4][4] = {};
        float na[4] = {}, nb[4] = {};
        for (int64_t k0 = 0; k0 < p.d; k0 += BK) {
            float va[4], vb[4];
            load4<VEC>(p.a, p.m, p.d, row0 + lr, k0 + lk, va);
            load4<VEC>(p.b, p.n, p.d, col0 + lr, k0 + lk, vb);
            __syncthreads();
#pragma unroll
            for (int q = 0; q < 4; ++q) { As[lk + q][lr] = va[q]; Bs[lk + q][lr] = vb[q]; }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < BK; ++kk) {
                const float4 a4 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
                const float4 b4 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
                const float av[4] = {a4.x, a4.y, a4.z, a4.w};
                const float bv[4] = {b4.x, b4.y, b4.z, b4.w};
                if (FORM == HG_FORM_DIRECT) {
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float df = av[i] - bv[j];
                            acc[i][j] = fmaf(df, df, acc[i][j]);
                        }
                } else {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        na[i] = fmaf(av[i], av[i], na[i]);
                        nb[i] = fmaf(bv[i], bv[i], nb[i]);
#pragma unroll
                        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
                    }
                }
            }
        }
        // ---- epilogue -----------------------------------------------------------
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int64_t r = row0 + ty * 4 + i;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t c = col0 + tx * 4 + j;
                float d2 = (FORM == HG_FORM_DIRECT) ? acc[i][j] : (na[i] + nb[j]) - 2.f * acc[i][j];
                float d2g = 0.f;
                if (p.pa != nullptr && r < p.m && c < p.n) {
                    for (int q = 0; q < p.pos_dim; ++q) {
                        const float df = __ldg(p.pa + r * p.pos_dim + q) - __ldg(p.pb + c * p.pos_dim + q);
                        d2g = fmaf(df, df, d2g);
                    }
                }
                if (MODE == MODE_DENSE) {
                    if (r < p.m && c < p.n)
                        p.out[r * p.ld_out + c] = expf(-(p.lambda_h * d2 + p.lambda_g * d2g));
                } else {
                    float s;
                    if (p.metric == HG_METRIC_EUCLIDEAN) s = fmaxf(d2, 0.f);
                    else if (p.metric == HG_METRIC_GAUSSIAN) s = p.lambda_h * d2 + p.lambda_g * d2g;
                    else s = -acc[i][j] / (fmaxf(sqrtf(na[i]), 1e-8f) * fmaxf(sqrtf(nb[j]), 1e-8f));
                    if (c >= p.n || (p.exclude_self && c == r)) s = INFINITY;
                    S[ty * 4 + i][tx * 4 + j] = s;
                }
            }
        }
        if (MODE == MODE_TOPK) {
            __syncthreads();
            if (tid < BM) {
                for (int j = 0; j < BN; ++j) {
                    const float s = S[tid][j];
                    if (s < lv[p.k - 1]) {
                        int q = p.k - 1;
                        while (q > 0 && lv[q - 1] > s) { lv[q] = lv[q - 1]; li[q] = li[q - 1]; --q; }
                        lv[q] = s; li[q] = int(col0 + j);
                    }
                }
            }
            // S is rewritten only after the two __syncthreads() of the next k-loop
        }
    }
    if (MODE == MODE_TOPK && tid < BM) {
        const int64_t r = row0 + tid;
        if (r < p.row_end) {
            const int64_t o = (r - p.row_begin) * p.k;
            for (int q = 0; q < p.k; ++q) {
                float v = lv[q];
                if (p.metric == HG_METRIC_GAUSSIAN) v = expf(-v);
                else if (p.metric == HG_METRIC_COSINE) v = -v;
                p.out_idx[o + q] = li[q];
                p.out_val[o + q] = (li[q] < 0) ? NAN : v;
            }
        }
    }
}

// Dense expansion-form kernel for larger outputs: 128 x 128 tile per 256-thread CTA, 8 x 8 outputs per thread (four
// 4 x 4 blocks, 64 apart, so the shared-memory reads stay conflict-free LDS.128), BK = 16, two shared-memory
// buffers with the next k-block prefetched into registers while the current one is multiplied.  The squared norms
// are not re-accumulated by every thread as in the 64 x 64 kernel (8 extra FMAs per 16 there): thread t < 128 owns
// row t's norm, thread 128 + t column t's, one FMA per k.  Every accumulator is the same k-ordered FMA chain as
// in sim_tile_kernel, so both kernels return identical bits.
constexpr int BM2 = 128, BN2 = 128;

template <bool VEC>
__global__ void __launch_bounds__(256, 2)
sim_tile128_kernel(SimArgs p) {
    __shared__ __align__(16) float As[2][BK][BM2 + PADT];
    __shared__ __align__(16) float Bs[2][BK][BN2 + PADT];
    __shared__ float sna[BM2], snb[BN2];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int lr = tid >> 2, lk = (tid & 3) * 4;     // loader: rows lr and lr + 64 of the tile, k offset lk
    const int64_t row0 = int64_t(blockIdx.y) * BM2, col0 = int64_t(blockIdx.x) * BN2;

    float acc[8][8] = {};
    float nrm = 0.f;                                  // |a_row|^2 (tid < 128) or |b_col|^2 (tid >= 128)
    float ra[2][4], rb[2][4];
    auto fetch = [&](int64_t k0) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            load4<VEC>(p.a, p.m, p.d, row0 + lr + 64 * h, k0 + lk, ra[h]);
            load4<VEC>(p.b, p.n, p.d, col0 + lr + 64 * h, k0 + lk, rb[h]);
        }
    };
    auto stash = [&](int buf) {
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int q = 0; q < 4; ++q) { As[buf][lk + q][lr + 64 * h] = ra[h][q]; Bs[buf][lk + q][lr + 64 * h] = rb[h][q]; }
    };
    fetch(0);
    stash(0);
    __syncthreads();
    int buf = 0;
    for (int64_t k0 = 0; k0 < p.d; k0 += BK) {
        const bool more = k0 + BK < p.d;
        if (more) fetch(k0 + BK);                      // global loads in flight during the multiply
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + ty * 4]);
            const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
            const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][64 + tx * 4]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
            const float v = (tid < BM2) ? As[buf][kk][tid] : Bs[buf][kk][tid - BM2];
            nrm = fmaf(v, v, nrm);
        }
        if (more) {
            stash(buf ^ 1);                            // nobody reads buf ^ 1 any more: its readers passed the barrier below
            __syncthreads();
            buf ^= 1;
        }
    }
    if (tid < BM2) sna[tid] = nrm; else snb[tid - BM2] = nrm;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int li = (i < 4) ? ty * 4 + i : 64 + ty * 4 + (i - 4);
        const int64_t r = row0 + li;
        if (r >= p.m) continue;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int lj = (j < 4) ? tx * 4 + j : 64 + tx * 4 + (j - 4);
            const int64_t c = col0 + lj;
            if (c >= p.n) continue;
            const float d2 = (sna[li] + snb[lj]) - 2.f * acc[i][j];
            float d2g = 0.f;
            if (p.pa != nullptr) {
                for (int q = 0; q < p.pos_dim; ++q) {
                    const float df = __ldg(p.pa + r * p.pos_dim + q) - __ldg(p.pb + c * p.pos_dim + q);
                    d2g = fmaf(df, df, d2g);
                }
            }
            p.out[r * p.ld_out + c] = expf(-(p.lambda_h * d2 + p.lambda_g * d2g));
        }
    }
}

static bool vec_ok(const float* a, const float* b, int64_t d) {
    return d % 4 == 0 && ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0;
}

int launch_dense(const SimArgs& p, int form, cudaStream_t st) {
    dim3 grid((unsigned)ceil_div(p.n, BN), (unsigned)ceil_div(p.m, BM));
    const bool v = vec_ok(p.a, p.b, p.d);
    if (form != HG_FORM_DIRECT && p.m >= 512 && p.n >= 512) {          // enough 128 x 128 tiles to fill the SMs
        dim3 grid2((unsigned)ceil_div(p.n, BN2), (unsigned)ceil_div(p.m, BM2));
        if (v) sim_tile128_kernel<true><<<grid2, 256, 0, st>>>(p);
        else   sim_tile128_kernel<false><<<grid2, 256, 0, st>>>(p);
    } else if (form == HG_FORM_DIRECT) {
        if (v) sim_tile_kernel<MODE_DENSE, HG_FORM_DIRECT, true, 1><<<grid, 256, 0, st>>>(p);
        else   sim_tile_kernel<MODE_DENSE, HG_FORM_DIRECT, false, 1><<<grid, 256, 0, st>>>(p);
    } else {
        if (v) sim_tile_kernel<MODE_DENSE, HG_FORM_EXPANSION, true, 1><<<grid, 256, 0, st>>>(p);
        else   sim_tile_kernel<MODE_DENSE, HG_FORM_EXPANSION, false, 1><<<grid, 256, 0, st>>>(p);
    }
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

int launch_topk_fp32(const SimArgs& p, cudaStream_t st) {
    dim3 grid(1, (unsigned)ceil_div(p.row_end - p.row_begin, BM));
    const bool v = vec_ok(p.a, p.b, p.d);
    HG_REQUIRE(p.k >= 1 && p.k <= 64, HG_ERR_UNSUPPORTED, "fp32 top-k supports 1 <= k <= 64 (got %d)", p.k);
    if (p.k <= 16) {
        if (v) sim_tile_kernel<MODE_TOPK, HG_FORM_EXPANSION, true, 16><<<grid, 256, 0, st>>>(p);
        else   sim_tile_kernel<MODE_TOPK, HG_FORM_EXPANSION, false, 16><<<grid, 256, 0, st>>>(p);
    } else {
        if (v) sim_tile_kernel<MODE_TOPK, HG_FORM_EXPANSION, true, 64><<<grid, 256, 0, st>>>(p);
        else   sim_tile_kernel<MODE_TOPK, HG_FORM_EXPANSION, false, 64><<<grid, 256, 0, st>>>(p);
    }
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

}  // namespace hg

extern "C" int hg_similarity_dense(const float* a, int64_t m, const float* b, int64_t n, int64_t d,
                                   const float* pos_a, const float* pos_b, int pos_dim,
                                   float lambda_h, float lambda_g, int form,
                                   float* out, int64_t ld_out, void* stream) {
    HG_REQUIRE(a && out && m > 0 && d > 0, HG_ERR_BAD_ARG, "hg_similarity_dense: bad argument");
    if (b == nullptr) { b = a; n = m; if (pos_b == nullptr) pos_b = pos_a; }
    HG_REQUIRE(n > 0 && ld_out >= n, HG_ERR_BAD_ARG, "hg_similarity_dense: bad n / ld_out");
    HG_REQUIRE((pos_a == nullptr) == (pos_b == nullptr), HG_ERR_BAD_ARG, "positions must be given for both sides");
    HG_REQUIRE(pos_a == nullptr || (pos_dim >= 1 && pos_dim <= 3), HG_ERR_BAD_ARG, "pos_dim must be 1..3");
    HG_REQUIRE(form == HG_FORM_EXPANSION || form == HG_FORM_DIRECT, HG_ERR_BAD_ARG, "unknown form %d", form);
    hg::SimArgs p{};
    p.a = a; p.b = b; p.m = m; p.n = n; p.d = d; p.pa = pos_a; p.pb = pos_b; p.pos_dim = pos_dim;
    p.lambda_h = lambda_h; p.lambda_g = lambda_g; p.out = out; p.ld_out = ld_out;
    return hg::launch_dense(p, form, static_cast<cudaStream_t>(stream));
}
