// Merge of the per-split candidate lists + exact FP32 re-rank, and the per-column
// epilogue parameters of the fused kernel.
//
// The tensor-core pass (knn_tc.cu) ranks with split-FP16 dot products (2 or 3 products per k-step); this pass
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
                                 float* __restrict__ norm_max, const RrSlide* __restrict__ g_slides,
                                 const int32_t* __restrict__ g_blk2slide) {
    const int64_t j = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (j >= n_pad) return;
    int64_t jl = j;                                  // row of `pos` behind operand row j
    bool real = j < n;
    if (g_slides != nullptr) {                       // grouped launch: per-slide row count and positions
        const RrSlide sl = g_slides[g_blk2slide[j >> 8]];
        jl = j - sl.off; real = jl < sl.n; pos = sl.pos;
    }
    if (thr != nullptr) { thr[j] = __int_as_float(0x7fc00000); bkt_cnt[j] = 0; }
    if (norm_max != nullptr) {
        // non-negative floats order like their bit patterns; one atomic per warp
        float m = real ? sq_norms[j] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<int*>(norm_max), __float_as_int(m));
    }
    float2 cp = make_float2(INFINITY, 0.f);          // padded columns can never be selected
    if (real) {
        const float nj = sq_norms[j];
        if (metric == HG_METRIC_COSINE) cp = make_float2(0.f, -1.f / fmaxf(sqrtf(nj), 1e-8f));
        else cp = make_float2(nj, -2.f);             // n_j - 2 dot (+ the row constant n_i, irrelevant to ranking)
    }
    colp[j] = cp;
    if (pos4 != nullptr) {
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        if (real && pos != nullptr) {
            q.x = pos[jl * pos_dim];
            if (pos_dim > 1) q.y = pos[jl * pos_dim + 1];
            if (pos_dim > 2) q.z = pos[jl * pos_dim + 2];
        }
        pos4[j] = q;
    }
}

int launch_colparams(const float* sq_norms, const float* pos, int pos_dim, int64_t n, int64_t n_pad, int metric,
                     float2* colp, float4* pos4, cudaStream_t st, float* thr, int32_t* bkt_cnt, float* norm_max) {
    colparams_kernel<<<(unsigned)ceil_div(n_pad, 256), 256, 0, st>>>(sq_norms, pos, pos_dim, n, n_pad, metric, colp, pos4,
                                                                     thr, bkt_cnt, norm_max, nullptr, nullptr);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

int launch_colparams_group(const float* sq_norms, int pos_dim, int64_t n_pad, int metric, float2* colp, float4* pos4,
                           float* norm_max, const RrSlide* g_slides, const int32_t* g_blk2slide, cudaStream_t st) {
    colparams_kernel<<<(unsigned)ceil_div(n_pad, 256), 256, 0, st>>>(sq_norms, nullptr, pos_dim, n_pad, n_pad, metric, colp, pos4,
                                                                     nullptr, nullptr, norm_max, g_slides, g_blk2slide);
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


// The per-row view of the parameter block.  Plain launches: the block's own fields.  Grouped launches (several
// slides in one operand buffer, knn_tc.cuh TcUnit): row `rg` of the group belongs to slide blk2slide[rg / 256] and
// every per-slide quantity comes from that slide's table entry.
struct RowCtx {
    const float* feats; const float* pos; const float* sq_norms;
    const float* part_val; const int32_t* part_idx;
    int64_t n, n_rows;            // rows of the slide; row stride between two splits' lists
    int splits;
    int64_t cols_per_split;       // columns swept by one split of the lists (0: unknown), classic plan
    int32_t* out_idx; float* out_val; float* out_cos;
    int64_t r, row;               // r: row of the lists / outputs, row: row of the features
};
// false for the padding rows of a group.  Deliberately NOT inlined: inlined into the kernels, nvcc 12.9 replaced the
// grouped branch's c.feats by p.feats (null in a grouped parameter block) -- the PTX simply had no load of the table's
// first field; behind a call the struct is written to local memory and read back as written.
__device__ __noinline__ bool row_ctx(const RerankParams& p, int64_t rg, RowCtx& c) {
    if (p.g_slides == nullptr) {
        c.feats = p.feats; c.pos = p.pos; c.sq_norms = p.sq_norms; c.part_val = p.part_val; c.part_idx = p.part_idx;
        c.n = p.n; c.n_rows = p.n_rows; c.splits = p.splits; c.cols_per_split = p.cols_per_split;
        c.out_idx = p.out_idx; c.out_val = p.out_val; c.out_cos = p.out_cos;
        c.r = rg; c.row = p.row_begin + rg;
        return true;
    }
    const RrSlide* sl = p.g_slides + p.g_blk2slide[rg >> 8];
    const int64_t off = sl->off;
    c.n = sl->n; c.n_rows = c.n; c.splits = sl->splits; c.cols_per_split = sl->cols_per_split;
    c.r = rg - off; c.row = c.r;
    if (c.r >= c.n) return false;
    c.feats = sl->feats; c.pos = sl->pos; c.sq_norms = p.sq_norms + off;
    c.part_val = p.part_val + sl->part_off; c.part_idx = p.part_idx + sl->part_off;
    c.out_idx = sl->out_idx; c.out_val = sl->out_val; c.out_cos = sl->out_cos;
    return true;
}

// Exact sums of one query row against CB candidate rows by one warp, in the canonical order of
// rowdot.cuh (lane l owns the float4 groups l, l + 32, ...; packed FMA pairs; xor butterfly): one read
// of the query row serves CB rows of sums and the CB reductions overlap.  Shared by the re-rank and by
// the exact fallback of uncertified rows, so both produce the same bits for the same pair.
template <bool WANT_D2, bool WANT_DOT, int CB>
__device__ __forceinline__ void warp_pair_sums(const float* __restrict__ xi, const float* const (&xj)[CB], int64_t d,
                                               int lane, bool vec, float (&d2)[CB], float (&dt)[CB]) {
    if (vec) {
        uint64_t ad[CB], at[CB];
#pragma unroll
        for (int q = 0; q < CB; ++q) { ad[q] = 0; at[q] = 0; }
#pragma unroll 2
        for (int64_t v = lane; v < (d >> 2); v += 32) {
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
        for (int64_t c = lane; c < d; c += 32) {
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
}

// ranking key (ascending), returned value and cosine of the pair (row, j) from its exact sums
template <int MODE>
__device__ __forceinline__ void exact_key(const RerankParams& p, const RowCtx& c, int j, float d2, float dot, float ni,
                                          float& key, float& out, float& cs) {
    const int64_t row = c.row;
    cs = (MODE != RR_D2) ? cosine_from(dot, ni, __ldg(c.sq_norms + j)) : 0.f;
    if (MODE == RR_DOT) {
        key = -cs; out = cs;
    } else if (p.metric == HG_METRIC_EUCLIDEAN) {
        key = d2; out = d2;
    } else {
        float d2g = 0.f;
        if (c.pos != nullptr) {
            for (int q = 0; q < p.pos_dim; ++q) {
                const float df = __ldg(c.pos + row * p.pos_dim + q) - __ldg(c.pos + int64_t(j) * p.pos_dim + q);
                d2g = fmaf(df, df, d2g);
            }
        }
        key = p.lambda_h * d2 + p.lambda_g * d2g;
        out = expf(-key);
    }
}

// MULTI = false (every row): ONE selection round of up to 32 candidates, which settles all but a handful of rows;
// a row with more than 32 candidates inside the slack is appended to the `more` worklist instead.  MULTI = true
// (grid-stride over that worklist): as many rounds as it takes, merged into a running exact top-k.  Two
// instantiations because the merge state of the multi-round form costs ~45 registers: with it in the common kernel
// the re-rank ran at 2 blocks per SM instead of 3 and was 25-30 % slower (C2: 0.84 -> 1.07 ms, ncu).
template <int MODE, bool MULTI>
__global__ void __launch_bounds__(RR_WARPS * 32, MULTI ? 2 : 3)
rerank_kernel(const RerankParams p) {
    const int lane = threadIdx.x & 31;
    const int64_t first = int64_t(blockIdx.x) * RR_WARPS + (threadIdx.x >> 5);
    const int64_t work = MULTI ? int64_t(*p.more_cnt) : p.n_rows;
    const int64_t step = MULTI ? int64_t(gridDim.x) * RR_WARPS : work;        // MULTI = false: one row per warp
    for (int64_t w = first; w < work; w += step) {
    const int64_t rg = MULTI ? int64_t(p.more_rows[w]) : w;
    RowCtx c;
    if (!row_ctx(p, rg, c)) continue;
    const int64_t r = c.r, row = c.row;
    // symmetric mode unless the bucket overflow flag made the classic kernel re-run
    const bool sym = p.ovf != nullptr && *p.ovf == 0;
    const int splits = sym ? p.sym_plan.splits : c.splits;
    const int total = splits * p.kl;
    const int nbkt = sym ? min(p.bkt_cnt[row - p.bkt_row0], p.bkt_cap) : 0;

    const int ncand = total + nbkt;
    // Selection stops once a candidate's approximate score exceeds the k-th best by more than the
    // guaranteed tensor-core error slack (both scores off by <= eps in opposite directions): such a
    // candidate cannot enter the exact top-k, and neither can any later one.
    const float ni = c.sq_norms[row];
    const int prods = sym ? p.sym_prods : p.prods;       // tensor-core products behind the candidates' scores
    const float rho = p.rho_max != nullptr ? __ldg(p.rho_max) : TC_RHO_WORST;
    const float slack_abs = tc_slack(round_up(p.d, kColPad), prods, prods, rho) * ((p.metric == HG_METRIC_COSINE) ? sqrtf(ni) : ni + __ldg(p.norm_max));
    float slack = slack_abs;                             // + TC_SLACK_REL |k-th best| once that is known
    // Certification.  A list keeps the kl best of its column range; what it dropped scores at least its
    // last entry.  If a FULL list ends inside the slack of the k-th best, a dropped column could still
    // belong to the exact top-k (many near-duplicates): the row goes to the exact fallback
    // (exact_slices_kernel).  Everything else is decided here, however many candidates sit inside the
    // slack: they are taken 32 at a time (almost always one round).
    const float* xi = c.feats + row * p.d;
    const bool vec = rows_vectorisable(c.feats, p.d);
    constexpr bool WANT_D2 = MODE != RR_DOT, WANT_DOT = MODE != RR_D2;
    constexpr int CB = 4;
    float kth = INFINITY;
    int taken = 0;
    float last_v = -INFINITY;                            // the last candidate a round selected
    int last_i = -1;
    // running exact top-k, lane t = the t-th best by (key, index); materialised only from a second round on
    float bk = INFINITY, bo = NAN, bc = NAN;
    int bidx = 0x7fffffff, nbest = 0;
    for (int round = 0;; ++round) {
        // ---- 1. the next (up to 32) best approximate candidates; lane t holds the t-th of them --------
        // slots [0, total) = the split lists, [total, ncand) = the bucket.  The pool lives in registers only while
        // a round selects; a later round (rare) reloads it and drops what the earlier ones took: selection
        // runs in (score, index) order, so "taken" = everything up to the last candidate selected
        float cv[RR_MAX_CAND / 32];
        int ci[RR_MAX_CAND / 32];
#pragma unroll
        for (int q = 0; q < RR_MAX_CAND / 32; ++q) {
            const int t = lane + 32 * q;
            cv[q] = INFINITY; ci[q] = 0x7fffffff;
            if (t < total) {
                const int s = t / p.kl, w = t - s * p.kl;
                const int64_t o = (int64_t(s) * c.n_rows + r) * p.kl + w;
                const int idx = c.part_idx[o];
                if (idx >= 0) { cv[q] = c.part_val[o]; ci[q] = idx; }
            } else if (t < ncand) {
                const uint2 b = p.bkt[(row - p.bkt_row0) * p.bkt_cap + (t - total)];
                cv[q] = __uint_as_float(b.x); ci[q] = int(b.y);
            }
            if (round > 0 && !pair_less(last_v, last_i, cv[q], ci[q])) { cv[q] = INFINITY; ci[q] = 0x7fffffff; }
        }
        int sel = -1, nsel = 0;
        bool last = false;
        for (int t = 0; t < 32; ++t) {
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
            if (bi == 0x7fffffff) { last = true; break; }                  // no candidates left
            if (taken + t == p.k - 1) { kth = bv; slack = slack_abs + TC_SLACK_REL * fabsf(bv); }
            if (taken + t >= p.k && bv > kth + slack) { last = true; break; }
#pragma unroll
            for (int q = 0; q < RR_MAX_CAND / 32; ++q) {
                if (q * 32 >= ncand) break;
                if (ci[q] == bi) { cv[q] = INFINITY; ci[q] = 0x7fffffff; }   // column ids are unique per row
            }
            if (lane == t) sel = bi;
            nsel = t + 1;
            last_v = bv; last_i = bi;
        }
        taken += nsel;

        // ---- 2. exact values: sum (x_i - x_j)^2 and / or x_i . x_j in the canonical order of rowdot.cuh ----
        float my_val = INFINITY;                         // ranking key (ascending) of this lane's candidate
        float my_out = NAN, my_cos = NAN;
        for (int t0 = 0; t0 < nsel; t0 += CB) {
            int j[CB];
            const float* xj[CB];
#pragma unroll
            for (int q = 0; q < CB; ++q) {
                j[q] = __shfl_sync(0xffffffffu, sel, t0 + q);
                xj[q] = c.feats + int64_t(j[q] >= 0 ? j[q] : row) * p.d;     // dummy read of the own row if absent
            }
            float d2[CB], dt[CB];
            warp_pair_sums<WANT_D2, WANT_DOT, CB>(xi, xj, p.d, lane, vec, d2, dt);
#pragma unroll
            for (int q = 0; q < CB; ++q) {
                if (lane != t0 + q || j[q] < 0) continue;
                exact_key<MODE>(p, c, j[q], d2[q], dt[q], ni, my_val, my_out, my_cos);
            }
        }
        const int my_idx = (sel >= 0) ? sel : 0x7fffffff;

        if constexpr (!MULTI) {
            if (!last) {                                 // more than 32 candidates inside the slack: the multi-round kernel
                if (lane == 0) p.more_rows[atomicAdd(p.more_cnt, 1)] = int32_t(rg);
                nbest = -1;
                break;
            }
        }
        if (round == 0 && last) {
            // ---- 3. (the usual case) final order by (exact value, index); best k written out ----------------
            int rank = 0;
            for (int t = 0; t < nsel; ++t) {
                const float ov = __shfl_sync(0xffffffffu, my_val, t);
                const int oi = __shfl_sync(0xffffffffu, my_idx, t);
                rank += pair_less(ov, oi, my_val, my_idx) ? 1 : 0;
            }
            if (lane < nsel && sel >= 0 && rank < p.k) {
                c.out_idx[r * p.k + rank] = sel;
                c.out_val[r * p.k + rank] = my_out;
                if (c.out_cos != nullptr) c.out_cos[r * p.k + rank] = my_cos;
            }
            nbest = min(nsel, p.k);
            break;
        }
        if constexpr (MULTI) {
        // ---- 3'. more than 32 candidates inside the slack: merge this round into the running top-k ----------
        int r_old = lane, r_new = 0;
        for (int t = 0; t < nsel; ++t) {
            const float ov = __shfl_sync(0xffffffffu, my_val, t);
            const int oi = __shfl_sync(0xffffffffu, my_idx, t);
            r_old += pair_less(ov, oi, bk, bidx) ? 1 : 0;
            r_new += pair_less(ov, oi, my_val, my_idx) ? 1 : 0;
        }
        for (int t = 0; t < nbest; ++t) {
            const float ov = __shfl_sync(0xffffffffu, bk, t);
            const int oi = __shfl_sync(0xffffffffu, bidx, t);
            r_new += pair_less(ov, oi, my_val, my_idx) ? 1 : 0;
        }
        if (lane >= nbest) r_old = 1 << 20;
        if (lane >= nsel) r_new = 1 << 20;
        const int total_best = min(p.k, nbest + nsel);
        float nk = INFINITY, no = NAN, nc = NAN;
        int nidx = 0x7fffffff;
        for (int t = 0; t < total_best; ++t) {          // lane t takes the entry of rank t (indices are unique: no ties)
            const unsigned mo = __ballot_sync(0xffffffffu, r_old == t);
            const unsigned mn = __ballot_sync(0xffffffffu, r_new == t);
            const int src = __ffs(mo ? mo : mn) - 1;
            const float vk = __shfl_sync(0xffffffffu, mo ? bk : my_val, src);
            const float vo = __shfl_sync(0xffffffffu, mo ? bo : my_out, src);
            const float vc = __shfl_sync(0xffffffffu, mo ? bc : my_cos, src);
            const int vi = __shfl_sync(0xffffffffu, mo ? bidx : my_idx, src);
            if (lane == t) { nk = vk; no = vo; nc = vc; nidx = vi; }
        }
        bk = nk; bo = no; bc = nc; bidx = nidx; nbest = total_best;
        if (last) {
            if (lane < nbest) {
                c.out_idx[r * p.k + lane] = bidx;
                c.out_val[r * p.k + lane] = bo;
                if (c.out_cos != nullptr) c.out_cos[r * p.k + lane] = bc;
            }
            break;
        }
        }  // MULTI
    }
    if (nbest < 0) continue;                             // handed to the multi-round kernel: it writes and certifies the row
    // rows with fewer than k candidates: pad with (-1, NaN)
    for (int t = nbest + lane; t < p.k; t += 32) {
        c.out_idx[r * p.k + t] = -1; c.out_val[r * p.k + t] = NAN;
        if (c.out_cos != nullptr) c.out_cos[r * p.k + t] = NAN;
    }
    if (p.unc_cnt != nullptr) {
        bool full_inside = false;                        // lane s2 looks at split s2 (at most 32 splits)
        if (lane < splits) {
            const int64_t o = (int64_t(lane) * c.n_rows + r) * p.kl + (p.kl - 1);
            full_inside = c.part_idx[o] >= 0 && c.part_val[o] <= kth + slack;
        }
        const unsigned mask = __ballot_sync(0xffffffffu, full_inside);
        if (mask != 0u && lane == 0) {
            const int e = atomicAdd(p.unc_cnt, 1);
            p.unc_rows[e] = int32_t(rg);
            p.unc_rows[2 * p.n_rows + e] = int32_t(mask);            // which splits' column ranges have to be rescanned
        }
    }
    }  // rows of this warp
}

// ---- exact fallback for uncertified rows ------------------------------------------------------------
// Work item = (uncertified row e, column slice): one CTA ranks the slice's columns by the exact key (the
// same sums, in the same order, as the re-rank) and leaves its k best in unc_part[e][slice][k]; the merge
// kernel picks the k best of the RR_UNC_SLICES lists.  Cost: one pass over the features per row, shared
// by the CTAs working on it -- paid only by rows the approximate candidate lists cannot vouch for.
constexpr int RR_UNC_WARPS = 8;

// ascending (key, index) list of <= 32 entries spread over the lanes of a warp; lane t holds the t-th best
__device__ __forceinline__ void warp_list_insert(float& mv, int& mi, float v, int j, int k, int lane) {
    const bool before = pair_less(mv, mi, v, j);                     // this lane's entry stays ahead of the new one
    const int pos = __popc(__ballot_sync(0xffffffffu, before && lane < k));
    const float pv = __shfl_up_sync(0xffffffffu, mv, 1);
    const int pi = __shfl_up_sync(0xffffffffu, mi, 1);
    if (lane > pos) { mv = pv; mi = pi; }
    else if (lane == pos) { mv = v; mi = j; }
}

template <int MODE>
__global__ void __launch_bounds__(RR_UNC_WARPS * 32)
exact_slices_kernel(const RerankParams p) {
    __shared__ float s_val[RR_UNC_WARPS][32];
    __shared__ int s_idx[RR_UNC_WARPS][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cnt = *p.unc_cnt;
    constexpr bool WANT_D2 = MODE != RR_DOT, WANT_DOT = MODE != RR_D2;
    constexpr int CB = 4;
    for (int64_t item = blockIdx.x; item < int64_t(cnt) * RR_UNC_SLICES; item += gridDim.x) {
        const int e = int(item / RR_UNC_SLICES), slice = int(item % RR_UNC_SLICES);
        RowCtx c;
        row_ctx(p, p.unc_rows[e], c);
        const int64_t row = c.row;
        const bool vec = rows_vectorisable(c.feats, p.d);
        const int64_t per_slice = round_up(ceil_div(c.n, RR_UNC_SLICES), CB);
        const float* xi = c.feats + row * p.d;
        const float ni = c.sq_norms[row];
        const int64_t c0 = slice * per_slice, c1 = min64(c.n, c0 + per_slice);
        // only the column ranges of the splits whose list could have dropped a member of the exact top-k: every other
        // column inside the slack is in a list, was evaluated by the re-rank and competes through its output (merge)
        const bool sym = p.ovf != nullptr && *p.ovf == 0;
        const int64_t cps = sym ? int64_t(p.sym_plan.ct_per_split) * TC_BN : c.cols_per_split;
        const unsigned mask = cps > 0 ? unsigned(p.unc_rows[2 * p.n_rows + e]) : 0xffffffffu;
        float mv = INFINITY; int mi = 0x7fffffff;                      // this warp's list, lane t = t-th best
        for (int64_t j0 = c0 + warp * CB; j0 < c1; j0 += RR_UNC_WARPS * CB) {
            if (cps > 0 && ((mask >> int(j0 / cps)) & 1u) == 0u) continue;     // a batch of CB columns lies in one split
            const float* xj[CB];
#pragma unroll
            for (int q = 0; q < CB; ++q) xj[q] = c.feats + min64(j0 + q, c.n - 1) * p.d;
            float d2[CB], dt[CB];
            warp_pair_sums<WANT_D2, WANT_DOT, CB>(xi, xj, p.d, lane, vec, d2, dt);
#pragma unroll
            for (int q = 0; q < CB; ++q) {
                const int64_t j = j0 + q;
                if (j >= c1 || (p.exclude_self && j == row)) continue;                 // warp-uniform
                float key, out, cs;
                exact_key<MODE>(p, c, int(j), d2[q], dt[q], ni, key, out, cs);
                const float kv = __shfl_sync(0xffffffffu, mv, p.k - 1);
                const int ki = __shfl_sync(0xffffffffu, mi, p.k - 1);
                if (pair_less(key, int(j), kv, ki)) warp_list_insert(mv, mi, key, int(j), p.k, lane);
            }
        }
        s_val[warp][lane] = mv; s_idx[warp][lane] = mi;
        __syncthreads();
        if (warp == 0) {
            // k best of the RR_UNC_WARPS lists: lane holds one entry per list, k rounds of warp arg-min
            float cv[RR_UNC_WARPS]; int ci[RR_UNC_WARPS];
#pragma unroll
            for (int w = 0; w < RR_UNC_WARPS; ++w) {
                const bool ok = lane < p.k;
                cv[w] = ok ? s_val[w][lane] : INFINITY; ci[w] = ok ? s_idx[w][lane] : 0x7fffffff;
            }
            uint2* dst = p.unc_part + (int64_t(e) * RR_UNC_SLICES + slice) * p.k;
            for (int t = 0; t < p.k; ++t) {
                float bv = INFINITY; int bi = 0x7fffffff;
#pragma unroll
                for (int w = 0; w < RR_UNC_WARPS; ++w)
                    if (pair_less(cv[w], ci[w], bv, bi)) { bv = cv[w]; bi = ci[w]; }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                    if (pair_less(ov, oi, bv, bi)) { bv = ov; bi = oi; }
                }
#pragma unroll
                for (int w = 0; w < RR_UNC_WARPS; ++w)
                    if (ci[w] == bi) { cv[w] = INFINITY; ci[w] = 0x7fffffff; }
                if (lane == 0) dst[t] = make_uint2(__float_as_uint(bv), uint32_t(bi));
            }
        }
        __syncthreads();
    }
}

// one warp per uncertified row: k best of its slice lists by (exact key, index), values recomputed
template <int MODE>
__global__ void __launch_bounds__(RR_WARPS * 32)
exact_merge_kernel(const RerankParams p) {
    const int lane = threadIdx.x & 31;
    const int cnt = *p.unc_cnt;
    constexpr bool WANT_D2 = MODE != RR_DOT, WANT_DOT = MODE != RR_D2;
    for (int e = blockIdx.x * RR_WARPS + (threadIdx.x >> 5); e < cnt; e += gridDim.x * RR_WARPS) {
        RowCtx c;
        row_ctx(p, p.unc_rows[e], c);
        const int64_t r = c.r, row = c.row;
        const bool vec = rows_vectorisable(c.feats, p.d);
        const int total = RR_UNC_SLICES * p.k;                          // <= 256
        float cv[RR_UNC_SLICES]; int ci[RR_UNC_SLICES];
#pragma unroll
        for (int q = 0; q < RR_UNC_SLICES; ++q) {
            const int t = lane + 32 * q;
            cv[q] = INFINITY; ci[q] = 0x7fffffff;
            if (t < total) {
                const uint2 u = p.unc_part[int64_t(e) * total + t];
                cv[q] = __uint_as_float(u.x); ci[q] = int(u.y);
            }
        }
        // ... plus what the re-rank already wrote for this row: the exact top-k of every candidate the lists held (the
        // rescan above skipped the column ranges of the splits that were not flagged); keys recomputed, lane t = entry t
        const float* xi = c.feats + row * p.d;
        const float ni = c.sq_norms[row];
        float pv = INFINITY; int pi = 0x7fffffff;
        for (int t = 0; t < p.k; ++t) {
            const int j = c.out_idx[r * p.k + t];
            if (j < 0) break;                                        // warp-uniform: the list is padded at its end
            const float* xj[1] = {c.feats + int64_t(j) * p.d};
            float d2[1], dt[1];
            warp_pair_sums<WANT_D2, WANT_DOT, 1>(xi, xj, p.d, lane, vec, d2, dt);
            float key, out, cs;
            exact_key<MODE>(p, c, j, d2[0], dt[0], ni, key, out, cs);
            if (lane == t) { pv = key; pi = j; }
        }
        int sel = 0x7fffffff;
        int nsel = 0;
        for (int t = 0; t < p.k; ++t) {
            float bv = pv; int bi = pi;
#pragma unroll
            for (int q = 0; q < RR_UNC_SLICES; ++q)
                if (pair_less(cv[q], ci[q], bv, bi)) { bv = cv[q]; bi = ci[q]; }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (pair_less(ov, oi, bv, bi)) { bv = ov; bi = oi; }
            }
            if (bi == 0x7fffffff) break;
#pragma unroll
            for (int q = 0; q < RR_UNC_SLICES; ++q)
                if (ci[q] == bi) { cv[q] = INFINITY; ci[q] = 0x7fffffff; }
            if (pi == bi) { pv = INFINITY; pi = 0x7fffffff; }           // a column may be in a slice list AND in the output
            if (lane == t) sel = bi;
            nsel = t + 1;
        }
        // the selection order IS the (exact key, index) order; recompute value and cosine of every winner
        for (int t = 0; t < nsel; ++t) {
            const int j = __shfl_sync(0xffffffffu, sel, t);
            const float* xj[1] = {c.feats + int64_t(j) * p.d};
            float d2[1], dt[1];
            warp_pair_sums<WANT_D2, WANT_DOT, 1>(xi, xj, p.d, lane, vec, d2, dt);
            float key, out, cs;
            exact_key<MODE>(p, c, j, d2[0], dt[0], ni, key, out, cs);
            if (lane == 0) {
                c.out_idx[r * p.k + t] = j;
                c.out_val[r * p.k + t] = out;
                if (c.out_cos != nullptr) c.out_cos[r * p.k + t] = cs;
            }
        }
        for (int t = nsel + lane; t < p.k; t += 32) {
            c.out_idx[r * p.k + t] = -1; c.out_val[r * p.k + t] = NAN;
            if (c.out_cos != nullptr) c.out_cos[r * p.k + t] = NAN;
        }
    }
}

template <int MODE>
static int launch_exact_fallback(const RerankParams& p, cudaStream_t st) {
    exact_slices_kernel<MODE><<<2 * kNumSMs, RR_UNC_WARPS * 32, 0, st>>>(p);
    HG_CHECK_LAUNCH();
    exact_merge_kernel<MODE><<<kNumSMs, RR_WARPS * 32, 0, st>>>(p);
    HG_CHECK_LAUNCH();
    count_launches(2);
    return HG_OK;
}

// Pre-pass run with several column splits: thr[row] = the rank-th smallest score over the row's S
// ascending lists (an upper bound of the row's final k-th best) plus the error slack the fused
// kernel adds in its single-split path.  One thread per row, S-way merge by list heads.
__global__ void thr_merge_kernel(const float* __restrict__ part_val, int splits, int64_t n, int kl, int rank,
                                 const float2* __restrict__ colp, const float* __restrict__ norm_max, int cosine,
                                 int64_t d_pad, int prods_pre, int prods_main, const float* __restrict__ rho_max,
                                 float* __restrict__ thr) {
    const int64_t row = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (row >= n) return;
    const float slack = tc_slack(d_pad, prods_pre, prods_main, rho_max != nullptr ? __ldg(rho_max) : TC_RHO_WORST);
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
    thr[row] = t + slack * scale + TC_SLACK_REL * fabsf(t);
}

int thr_merge_launch(const float* part_val, int splits, int64_t n, int kl, int rank, const float2* colp,
                     const float* norm_max, int cosine, int64_t d_pad, int prods_pre, int prods_main, const float* rho_max,
                     float* thr, cudaStream_t st) {
    thr_merge_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, st>>>(part_val, splits, n, kl, rank, colp, norm_max, cosine,
                                                                 d_pad, prods_pre, prods_main, rho_max, thr);
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
    HG_REQUIRE((p.unc_cnt == nullptr) == (p.unc_rows == nullptr) && (p.unc_cnt == nullptr) == (p.unc_part == nullptr),
               HG_ERR_BAD_ARG, "rerank: incomplete fallback buffers");
    const unsigned blocks = (unsigned)ceil_div(p.n_rows, RR_WARPS);
    HG_REQUIRE(p.more_cnt && p.more_rows, HG_ERR_BAD_ARG, "rerank: worklist of the multi-round kernel missing");
    // every row: one selection round; the handful of rows with more than 32 candidates inside the slack: multi-round
    if (p.metric == HG_METRIC_COSINE) {
        rerank_kernel<RR_DOT, false><<<blocks, RR_WARPS * 32, 0, st>>>(p);
        rerank_kernel<RR_DOT, true><<<kNumSMs, RR_WARPS * 32, 0, st>>>(p);
    } else if (p.out_cos != nullptr) {
        rerank_kernel<RR_D2_DOT, false><<<blocks, RR_WARPS * 32, 0, st>>>(p);
        rerank_kernel<RR_D2_DOT, true><<<kNumSMs, RR_WARPS * 32, 0, st>>>(p);
    } else {
        rerank_kernel<RR_D2, false><<<blocks, RR_WARPS * 32, 0, st>>>(p);
        rerank_kernel<RR_D2, true><<<kNumSMs, RR_WARPS * 32, 0, st>>>(p);
    }
    HG_CHECK_LAUNCH();
    count_launches(2);
    if (p.unc_cnt == nullptr) return HG_OK;
    // rows the lists cannot vouch for (normally none: both kernels then exit at once)
    if (p.metric == HG_METRIC_COSINE) return launch_exact_fallback<RR_DOT>(p, st);
    if (p.out_cos != nullptr) return launch_exact_fallback<RR_D2_DOT>(p, st);
    return launch_exact_fallback<RR_D2>(p, st);
}

}  // namespace hg
