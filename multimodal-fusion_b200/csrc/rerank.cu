// Merge of the per-split candidate lists + exact FP32 re-rank, and the per-column
// epilogue parameters of the fused kernel.
//
// The tensor-core pass (knn_tc.cu) ranks with 3xBF16 dot products; this pass
// recomputes the distance of every surviving candidate in direct-difference
// FP32 -- sum (x_i - x_j)^2, the more accurate of the two forms the reference
// uses (preprocess_hypergraph.py:254-255 vs similarity_kernel.py:49) -- and
// orders the final k by (exact value, index).  One warp per query row, 128-bit
// loads; the query row stays in L1 while its ~KL candidates stream from L2.
// Algorithmic bytes per row: KL * 4 D (candidate rows) + 4 D (query) + 8 k out.
#include "knn_tc.cuh"
#include "rowdot.cuh"

namespace hg {

__global__ void colparams_kernel(const float* __restrict__ sq_norms, const float* __restrict__ pos, int pos_dim,
                                 int64_t n, int64_t n_pad, int metric, float2* __restrict__ colp,
                                 float4* __restrict__ pos4, float* __restrict__ thr, int32_t* __restrict__ bkt_cnt,
                                 float* __restrict__ norm_max) {
    const int64_t j = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (j >= n_pad) return;
    if (thr != nullptr) { thr[j] = __int_as_float(0x7fc00000); bkt_cnt[j] = 0; }
    if (norm_max != nullptr) {
        // non-negative floats order like their bit patterns; one atomic per warp
        float m = (j < n) ? sq_norms[j] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<int*>(norm_max), __float_as_int(m));
    }
    float2 cp = make_float2(INFINITY, 0.f);          // padded columns can never be selected
    if (j < n) {
        const float nj = sq_norms[j];
        if (metric == HG_METRIC_COSINE) cp = make_float2(0.f, -1.f / fmaxf(sqrtf(nj), 1e-8f));
        else cp = make_float2(nj, -2.f);             // n_j - 2 dot (+ the row constant n_i, irrelevant to ranking)
    }
    colp[j] = cp;
    if (pos4 != nullptr) {
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        if (j < n) {
            q.x = pos[j * pos_dim];
            if (pos_dim > 1) q.y = pos[j * pos_dim + 1];
            if (pos_dim > 2) q.z = pos[j * pos_dim + 2];
        }
        pos4[j] = q;
    }
}

int launch_colparams(const float* sq_norms, const float* pos, int pos_dim, int64_t n, int64_t n_pad, int metric,
                     float2* colp, float4* pos4, cudaStream_t st, float* thr, int32_t* bkt_cnt, float* norm_max) {
    colparams_kernel<<<(unsigned)ceil_div(n_pad, 256), 256, 0, st>>>(sq_norms, pos, pos_dim, n, n_pad, metric, colp, pos4,
                                                                     thr, bkt_cnt, norm_max);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

__device__ __forceinline__ bool pair_less(float va, int ia, float vb, int ib) {
    return va < vb || (va == vb && ia < ib);
}

constexpr int RR_WARPS = 8;
constexpr int RR_MAX_LIST = TC_MAX_SPLITS * 32;      // 256 list candidates per row at most
constexpr int RR_MAX_CAND = RR_MAX_LIST + TC_BKT_CAP; // + the row's bucket (symmetric mode)

enum { RR_D2 = 0, RR_D2_DOT = 1, RR_DOT = 2 };       // what the exact pass accumulates per candidate

template <int MODE>
__global__ void __launch_bounds__(RR_WARPS * 32, 3)
rerank_kernel(const RerankParams p) {
    const int lane = threadIdx.x & 31;
    const int64_t r = int64_t(blockIdx.x) * RR_WARPS + (threadIdx.x >> 5);
    if (r >= p.n_rows) return;
    const int64_t row = p.row_begin + r;
    // symmetric mode unless the bucket overflow flag made the classic kernel re-run
    const bool sym = p.ovf != nullptr && *p.ovf == 0;
    const int splits = sym ? p.sym_plan.splits : p.splits;
    const int total = splits * p.kl;
    const int nbkt = sym ? min(p.bkt_cnt[row], p.bkt_cap) : 0;

    // ---- 1. merge: the kl best approximate candidates, lane t ends up holding the t-th -------------
    // slots [0, total) = the split lists, [total, ncand) = the bucket
    const int ncand = total + nbkt;
    float cv[RR_MAX_CAND / 32];
    int ci[RR_MAX_CAND / 32];
#pragma unroll
    for (int q = 0; q < RR_MAX_CAND / 32; ++q) {
        const int t = lane + 32 * q;
        cv[q] = INFINITY; ci[q] = 0x7fffffff;
        if (t < total) {
            const int s = t / p.kl, w = t - s * p.kl;
            const int64_t o = (int64_t(s) * p.n_rows + r) * p.kl + w;
            const int idx = p.part_idx[o];
            if (idx >= 0) { cv[q] = p.part_val[o]; ci[q] = idx; }
        } else if (t < ncand) {
            const uint2 b = p.bkt[row * p.bkt_cap + (t - total)];
            cv[q] = __uint_as_float(b.x); ci[q] = int(b.y);
        }
    }
    // Selection stops once a candidate's approximate score exceeds the k-th best by more than the
    // guaranteed 3xBF16 error slack (both scores off by <= eps in opposite directions): such a
    // candidate cannot enter the exact top-k, and neither can any later one.
    const float ni = p.sq_norms[row];
    const float slack = 4.8828125e-4f * ((p.metric == HG_METRIC_COSINE) ? sqrtf(ni) : ni + __ldg(p.norm_max));
    int sel = -1;                                    // candidate column held by this lane (lane < kl)
    int nsel = 0;
    float kth = INFINITY;
    for (int t = 0; t < p.kl; ++t) {
        float bv = INFINITY; int bi = 0x7fffffff;
#pragma unroll
        for (int q = 0; q < RR_MAX_CAND / 32; ++q) {
            if (q * 32 >= ncand) break;              // warp-uniform
            if (pair_less(cv[q], ci[q], bv, bi)) { bv = cv[q]; bi = ci[q]; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (pair_less(ov, oi, bv, bi)) { bv = ov; bi = oi; }
        }
        if (bi == 0x7fffffff) break;                 // fewer than kl real candidates
        if (t == p.k - 1) kth = bv;
        if (t >= p.k && bv > kth + slack) break;
#pragma unroll
        for (int q = 0; q < RR_MAX_CAND / 32; ++q) {
            if (q * 32 >= ncand) break;
            if (ci[q] == bi) { cv[q] = INFINITY; ci[q] = 0x7fffffff; }   // column ids are unique per row
        }
        if (lane == t) sel = bi;
        nsel = t + 1;
    }

    // ---- 2. exact values: sum (x_i - x_j)^2 and / or x_i . x_j in the canonical order of rowdot.cuh ----
    const float* xi = p.feats + row * p.d;
    const bool vec = rows_vectorisable(p.feats, p.d);
    constexpr bool WANT_D2 = MODE != RR_DOT, WANT_DOT = MODE != RR_D2;
    float my_val = INFINITY;                         // ranking key (ascending) of this lane's candidate
    float my_out = NAN, my_cos = NAN;
    // four candidates per pass: one read of the query row serves four rows of sums, and the four
    // reductions overlap.  Lane l owns the float4 groups l, l + 32, ... (rowdot.cuh order).
    constexpr int CB = 4;
    for (int t0 = 0; t0 < nsel; t0 += CB) {
        int j[CB];
        const float* xj[CB];
#pragma unroll
        for (int q = 0; q < CB; ++q) {
            j[q] = __shfl_sync(0xffffffffu, sel, t0 + q);
            xj[q] = p.feats + int64_t(j[q] >= 0 ? j[q] : row) * p.d;     // dummy read of the own row if absent
        }
        float d2[CB], dt[CB];
        if (vec) {
            uint64_t ad[CB], at[CB];
#pragma unroll
            for (int q = 0; q < CB; ++q) { ad[q] = 0; at[q] = 0; }
#pragma unroll 2
            for (int64_t v = lane; v < (p.d >> 2); v += 32) {
                const ulonglong2 x = __ldg(reinterpret_cast<const ulonglong2*>(xi) + v);
                ulonglong2 y[CB];
#pragma unroll
                for (int q = 0; q < CB; ++q) y[q] = __ldg(reinterpret_cast<const ulonglong2*>(xj[q]) + v);
#pragma unroll
                for (int q = 0; q < CB; ++q) {
                    if (WANT_D2) {
                        const uint64_t e0 = pk_sub(x.x, y[q].x), e1 = pk_sub(x.y, y[q].y);
                        ad[q] = pk_fma(e0, e0, ad[q]); ad[q] = pk_fma(e1, e1, ad[q]);
                    }
                    if (WANT_DOT) { at[q] = pk_fma(x.x, y[q].x, at[q]); at[q] = pk_fma(x.y, y[q].y, at[q]); }
                }
            }
#pragma unroll
            for (int q = 0; q < CB; ++q) { d2[q] = pk_hsum(ad[q]); dt[q] = pk_hsum(at[q]); }
        } else {
#pragma unroll
            for (int q = 0; q < CB; ++q) { d2[q] = 0.f; dt[q] = 0.f; }
            for (int64_t c = lane; c < p.d; c += 32) {
                const float x = __ldg(xi + c);
#pragma unroll
                for (int q = 0; q < CB; ++q) {
                    const float y = __ldg(xj[q] + c);
                    if (WANT_D2) { const float df = x - y; d2[q] = fmaf(df, df, d2[q]); }
                    if (WANT_DOT) dt[q] = fmaf(x, y, dt[q]);
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int q = 0; q < CB; ++q) {
                if (WANT_D2) d2[q] += __shfl_xor_sync(0xffffffffu, d2[q], o);
                if (WANT_DOT) dt[q] += __shfl_xor_sync(0xffffffffu, dt[q], o);
            }
        }
#pragma unroll
        for (int q = 0; q < CB; ++q) {
            if (lane != t0 + q || j[q] < 0) continue;
            const float cs = WANT_DOT ? cosine_from(dt[q], ni, __ldg(p.sq_norms + j[q])) : 0.f;
            my_cos = cs;
            if (MODE == RR_DOT) {
                my_val = -cs; my_out = cs;
            } else if (p.metric == HG_METRIC_EUCLIDEAN) {
                my_val = d2[q]; my_out = d2[q];
            } else {
                float d2g = 0.f;
                if (p.pos != nullptr) {
                    for (int c = 0; c < p.pos_dim; ++c) {
                        const float df = __ldg(p.pos + row * p.pos_dim + c) - __ldg(p.pos + int64_t(j[q]) * p.pos_dim + c);
                        d2g = fmaf(df, df, d2g);
                    }
                }
                my_val = p.lambda_h * d2[q] + p.lambda_g * d2g;
                my_out = expf(-my_val);
            }
        }
    }

    // ---- 3. final order by (exact value, index); best k written out ----------------------------------
    const int my_idx = (sel >= 0) ? sel : 0x7fffffff;
    int rank = 0;
    for (int t = 0; t < nsel; ++t) {
        const float ov = __shfl_sync(0xffffffffu, my_val, t);
        const int oi = __shfl_sync(0xffffffffu, my_idx, t);
        rank += pair_less(ov, oi, my_val, my_idx) ? 1 : 0;
    }
    if (lane < nsel && sel >= 0 && rank < p.k) {
        p.out_idx[r * p.k + rank] = sel;
        p.out_val[r * p.k + rank] = my_out;
        if (p.out_cos != nullptr) p.out_cos[r * p.k + rank] = my_cos;
    }
    // rows with fewer than k candidates: pad with (-1, NaN)
    for (int t = nsel + lane; t < p.k; t += 32) {
        p.out_idx[r * p.k + t] = -1; p.out_val[r * p.k + t] = NAN;
        if (p.out_cos != nullptr) p.out_cos[r * p.k + t] = NAN;
    }
}

// Pre-pass run with several column splits: thr[row] = the rank-th smallest score over the row's S
// ascending lists (an upper bound of the row's final k-th best) plus the error slack the fused
// kernel adds in its single-split path.  One thread per row, S-way merge by list heads.
__global__ void thr_merge_kernel(const float* __restrict__ part_val, int splits, int64_t n, int kl, int rank,
                                 const float2* __restrict__ colp, const float* __restrict__ norm_max, int cosine,
                                 float slack, float* __restrict__ thr) {
    const int64_t row = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (row >= n) return;
    int head[TC_MAX_SPLITS];
#pragma unroll
    for (int s = 0; s < TC_MAX_SPLITS; ++s) head[s] = 0;
    float t = INFINITY;
    for (int step = 0; step <= rank; ++step) {
        float best = INFINITY; int bs = -1;
#pragma unroll
        for (int s = 0; s < TC_MAX_SPLITS; ++s) {
            if (s < splits && head[s] < kl) {
                const float v = part_val[(int64_t(s) * n + row) * kl + head[s]];
                if (v < best) { best = v; bs = s; }
            }
        }
        t = best;
        if (bs < 0) break;                            // fewer than rank + 1 finite scores: no bound
#pragma unroll
        for (int s = 0; s < TC_MAX_SPLITS; ++s) head[s] += (s == bs) ? 1 : 0;
    }
    const float2 me = colp[row];
    const float scale = cosine ? 1.f / fmaxf(fabsf(me.y), 1e-30f) : me.x + *norm_max;
    thr[row] = t + slack * scale;
}

int thr_merge_launch(const float* part_val, int splits, int64_t n, int kl, int rank, const float2* colp,
                     const float* norm_max, int cosine, float slack, float* thr, cudaStream_t st) {
    thr_merge_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, st>>>(part_val, splits, n, kl, rank, colp, norm_max, cosine,
                                                                 slack, thr);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

int rerank_launch(const RerankParams& p, cudaStream_t st) {
    // without buckets (classic / staged lists only) the bucket slots of the merge hold list entries as well
    const int list_cap = (p.ovf != nullptr) ? RR_MAX_LIST : RR_MAX_CAND;
    HG_REQUIRE(p.kl <= 32 && p.splits * p.kl <= list_cap && p.sym_plan.splits * p.kl <= RR_MAX_LIST &&
               p.bkt_cap <= TC_BKT_CAP, HG_ERR_UNSUPPORTED, "rerank: too many candidates per row");
    HG_REQUIRE(p.sq_norms && p.norm_max, HG_ERR_BAD_ARG, "rerank: row norms missing");
    const unsigned blocks = (unsigned)ceil_div(p.n_rows, RR_WARPS);
    if (p.metric == HG_METRIC_COSINE) rerank_kernel<RR_DOT><<<blocks, RR_WARPS * 32, 0, st>>>(p);
    else if (p.out_cos != nullptr) rerank_kernel<RR_D2_DOT><<<blocks, RR_WARPS * 32, 0, st>>>(p);
    else rerank_kernel<RR_D2><<<blocks, RR_WARPS * 32, 0, st>>>(p);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

}  // namespace hg
