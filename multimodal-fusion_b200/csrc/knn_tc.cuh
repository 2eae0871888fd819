// Interface between the fused tensor-core kNN kernel (knn_tc.cu) and its host driver (knn_api.cu).
#pragma once
#include "common.cuh"

namespace hg {

constexpr int TC_BM = 128;      // query rows per CTA tile  (UMMA M)
constexpr int TC_BN = 256;      // candidate columns per tile (UMMA N)
constexpr int TC_BK = 32;       // K per pipeline stage: one 128-byte operand row = 32 hi | 32 lo FP16
constexpr int TC_MAX_SPLITS = 8;

struct TcPlan {
    int64_t n_rows;             // row_end - row_begin
    int num_rb;                 // ceil(n_rows / 128)
    int num_ct;                 // ceil(n / 256)
    int splits;                 // column splits S (partial lists per row), a multiple of cc
    int rc;                     // row blocks per chunk: CTAs running together cover rc row blocks ...
    int cc;                     // ... x cc column splits (rc * cc = 148): bounds the L2 working set of A
    int ct_per_split;
    int kl;                     // list length kept per (row, split): 8 / 16 / 24 / 32
    int grid;                   // CTAs launched
    int cg;                     // 1 = one CTA per 128-row tile, 2 = CTA pairs (cta_group::2) on 256-row tiles
    int mc;                     // 1 = multicast pairs: cg 1 MMAs, the two CTAs of a cluster share every B tile
    int cl;                     // CTAs per cluster = 128-row blocks per work unit (2 for cg 2 and for mc)
    // symmetric mode (sym = 1, full row range only): a checkerboard half of the tiles is computed
    // and each one also serves the transposed direction (see knn_tc.cu)
    int sym;
    int num_u;                  // major units of the work list (= num_rb)
};

// One work unit of a GROUPED launch: several (small) slides share one operand buffer, each at a 256-row aligned
// offset, and one persistent launch sweeps the units of all of them (reference loop over slides:
// preprocess_hypergraph.py:651-670).  All row / tile numbers are in group coordinates.
struct TcUnit {
    int32_t row0;               // first operand row of the unit's 128-row block
    int32_t row_end;            // one past the slide's last real row
    int32_t ct0, ct1;           // column tiles [ct0, ct1) of the slide's own columns swept by the unit
    int32_t col_off;            // the slide's first operand row: a list entry is (column - col_off)
    int32_t reserved;
    int64_t part_off;           // element offset of the unit's (slide, split) list block; local row r at + r * kl
};

struct TcParams {
    int64_t n, row_begin, row_end;
    int d_pad;
    TcPlan plan;
    const float2* colp;         // [n_pad]  score_j = colp.x + colp.y * dot   (+inf, 0 beyond n)
    const float4* pos4;         // [n_pad]  (x, y, z, 0) or null
    float gscale;               // weight of the squared spatial distance in the ranking score
    int exclude_self;
    float* part_val;            // [S][n_rows][kl]
    int32_t* part_idx;
    // dense mode (plan.kl == 0): K[i, j] = exp(-lambda_h * (colp[i].x + score_ij)) written to
    // dense_out[(i - row_begin) * dense_ld + j] instead of keeping lists (streaming statistics, A9)
    float* dense_out;
    int64_t dense_ld;
    float lambda_h;
    int dense_self;             // rows and columns are the same matrix: K[i, i] = 1 exactly
    // threshold pre-pass: the kernel's column c is operand row c * col_stride (the B tensor map is
    // strided to match); instead of candidate lists it writes thr_out[row] = the thr_rank-th best
    // score of the sample, an upper bound of the row's final k-th best.
    int64_t n_pad;
    int col_stride;
    float* thr_out;
    int thr_rank;
    // tensor-core products per k-step (see tc_eps below): 3 = hi*hi + hi*lo + lo*hi, 2 = hi*hi + lo*hi (the B rows
    // rounded to FP16, the A rows whole; symmetric passes), 1 = hi*hi (threshold pre-pass only).  0 means 3.
    int prods;
    int b_hi_only;              // prods <= 2: the B tiles carry the hi halves only (64-byte rows); set by tc_launch
    int stages;                 // pipeline stages that fit the shared-memory buffer with such tiles; set by tc_launch
    int pre_main_prods;         // pre-pass only: products of the main pass its bound will gate (slack of thr_out)
    // squared norms of the operand rows behind the query rows / the columns: every operand row was scaled by the
    // power of two op_scale(norm) before its FP16 split, the epilogue undoes both scales.  norms_b null = norms_a.
    const float* norms_a;
    const float* norms_b;
    const float* rho_max;       // measured rounding residual of the operand rows (tc_eps), or null = worst case
    int thr_margin_cos;         // error slack added to thr_out scales with |x_i| (cosine) or n_i + norm_max
    const float* norm_max;      // largest squared row norm
    // symmetric mode: transposed candidates (column j gets row i) with score < thr[j] are
    // appended to bucket j (bkt_cap entries of (score bits, row)); *ovf is raised on overflow
    const float* thr;
    int32_t* bkt_cnt;
    uint2* bkt;
    int bkt_cap;
    int* ovf;
    // symmetric mode over a ROW BLOCK of a slide whose rows are partitioned over several GPUs (C5): sblk_base =
    // row_begin / 256; the bucket of column j lives on the GPU that owns row j = peer j / peer_chunk, whose arrays
    // (mapped over NVLink, or local for the own block) are peer_cnt / peer_bkt[owner], indexed by
    // j - owner * peer_chunk.  peer_chunk = 0: one GPU, bkt_cnt / bkt / ovf above.
    int sblk_base;
    int64_t peer_chunk;
    int32_t* peer_cnt[8];
    uint2* peer_bkt[8];         // an overflow raises the PUSHING GPU's own flag (ovf): the owner may already be past its sweep
    const int* guard;           // non-null: the whole grid exits at once unless *guard != 0
    uint64_t hint_a, hint_b;    // L2 cache hints of the A / B tile loads (filled by tc_launch)
    int pair_stages;            // producer refills two 48 KB stages at a time (see tc_launch)
    // grouped launch: non-null unit table (classic list mode, single CTAs); the plan then only carries kl
    const TcUnit* unit_tab;
    int num_tab_units;
};

TcPlan tc_make_plan(int64_t n, int64_t n_rows, int k, int64_t d_pad, bool lists = true);   // lists = false: dense-output mode
TcPlan tc_make_plan_sym(int64_t n, int k, int64_t d_pad);
TcPlan tc_make_plan_prepass(int64_t n, int k, int64_t d_pad, int stride);
// symmetric mode over the row block [row_begin, row_end) of a partitioned slide (row_begin a multiple of 256)
TcPlan tc_make_plan_sym_rows(int64_t n, int64_t n_rows, int k, int64_t d_pad);
TcPlan tc_make_plan_prepass_rows(int64_t n, int64_t n_rows, int k, int64_t d_pad, int stride);
int tc_sym_stride(int64_t n, int64_t n_rows, int k);     // 0 = classic, else the pre-pass column stride
constexpr int TC_BKT_CAP = 512;
int tc_launch(const op16_t* ops, int64_t n_pad, const TcParams& p, cudaStream_t st);
int tc_launch_units(const op16_t* ops, int64_t n_pad, int k, const TcUnit* unit_tab, int num_units,
                    const TcParams& p, cudaStream_t st);
int tc_list_len(int k);
// rectangular product (kmeans.cu): rows of A against the rows of another matrix B, classic list mode
TcPlan tc_make_plan_rect(int64_t cols, int64_t n_rows, int k, int64_t d_pad);
int tc_launch_rect(const op16_t* ops_a, int64_t n_pad_a, const op16_t* ops_b, int64_t n_pad_b,
                   const TcParams& p, cudaStream_t st);

struct RrSlide {                 // one slide of a grouped launch, as the re-rank sees it
    const float* feats;          // [n, d] the slide's own FP32 rows
    const float* pos;            // [n, pos_dim] or null
    int64_t off;                 // first operand row of the slide in the group (multiple of 256)
    int64_t n;
    int64_t part_off;            // element offset of the slide's [splits][n][kl] candidate lists
    int32_t splits, reserved;
    int32_t* out_idx; float* out_val; float* out_cos;
    int64_t cols_per_split;      // columns swept by one split of the slide's lists
};

struct RerankParams {
    const float* feats; int64_t n, d;
    const float* sq_norms;       // [n] squared row norms in the canonical order (split.cu)
    const float* norm_max;       // largest of them
    const float* pos; int pos_dim;
    int metric; float lambda_h, lambda_g;
    const float* part_val; const int32_t* part_idx;
    int splits, kl; int64_t n_rows, row_begin;
    int64_t cols_per_split;      // columns swept by one split of the (classic-plan) lists; 0 = unknown: the fallback
                                 // rescans every column instead of the flagged splits' ranges
    // symmetric mode: candidates = per-(row, split) lists of the sym plan + the row's bucket;
    // if *ovf != 0 the classic kernel re-ran and (splits, kl) above describe its lists
    TcPlan sym_plan; const int32_t* bkt_cnt; const uint2* bkt; int bkt_cap; const int* ovf;
    int64_t bkt_row0;            // bkt_cnt / bkt are indexed by (row - bkt_row0): 0, or row_begin for a partitioned slide
    int k;
    int prods, sym_prods;        // tensor-core products per k-step behind the classic lists / the symmetric lists + buckets
    const float* rho_max;        // measured rounding residual of the operand rows (tc_eps), or null = worst case
    int32_t* out_idx; float* out_val;
    float* out_cos;              // optional [n_rows, k]: cos(x_row, x_idx) of every list entry (edge weights)
    // certification fallback (rerank.cu): rows whose lists cannot vouch for the exact top-k are appended to
    // unc_rows (counter *unc_cnt, zeroed by the caller) and recomputed exactly against the columns of the splits whose
    // list was full inside the slack; unc_rows holds 3 * n_rows ints (uncertified rows | rows of the multi-round
    // kernel | split masks of the uncertified rows), unc_part = [n_rows][RR_UNC_SLICES][k] scratch.  All null: no
    // certification.
    int32_t* unc_rows; int32_t* unc_cnt; uint2* unc_part;
    // rows with more than 32 candidates inside the slack: worklist of the multi-round re-rank kernel (counter zeroed
    // by the caller, list of n_rows entries)
    int32_t* more_rows; int32_t* more_cnt;
    int exclude_self;            // the fallback skips column == row (the lists already did)
    // grouped launch: n_rows = the group's operand rows, sq_norms / part_* = the group's arrays, everything else
    // per slide from g_slides[g_blk2slide[row / 256]]
    const RrSlide* g_slides; const int32_t* g_blk2slide;
};
constexpr int RR_UNC_SLICES = 8;
inline size_t rerank_unc_part_bytes(int64_t n_rows, int k) { return size_t(n_rows) * RR_UNC_SLICES * size_t(k) * 8; }
int rerank_launch(const RerankParams& p, cudaStream_t st);
int thr_merge_launch(const float* part_val, int splits, int64_t n, int kl, int rank, const float2* colp,
                     const float* norm_max, int cosine, int64_t d_pad, int prods_pre, int prods_main, const float* rho_max,
                     float* thr, cudaStream_t st);
// ---- operand format and the guaranteed error of a tensor-core score ---------------------------------------------
// An operand row x is scaled by s = op_scale(|x|^2), a power of two that puts |s x| into [2^5, 2^6), and split into
// two FP16 values per element: hi = fp16(s x), lo = fp16(s x - hi).  Every element is below 2^6 (no overflow);
// elements below 2^-14 land on FP16's subnormal grid, spacing 2^-24, an ABSOLUTE error of 2^-25 at most.
__host__ __device__ inline int op_scale_log2(float sq_norm) {
#ifdef __CUDA_ARCH__
    const int e = (__float_as_int(sq_norm) >> 23) & 0xff;
#else
    union { float f; int i; } u; u.f = sq_norm;
    const int e = (u.i >> 23) & 0xff;
#endif
    // |x|^2 in [2^(e-127), 2^(e-126))  ->  |x| < 2^h, h = ceil((e - 126) / 2);  s = 2^(6 - h)
    const int h = (e - 125) >> 1;
    return 6 - h;
}
__host__ __device__ inline float pow2f(int e) {          // 2^e, -126 <= e <= 127
#ifdef __CUDA_ARCH__
    return __int_as_float((e + 127) << 23);
#else
    union { float f; int i; } u; u.i = (e + 127) << 23; return u.f;
#endif
}
__host__ __device__ inline float op_scale(float sq_norm) { return pow2f(op_scale_log2(sq_norm)); }
__host__ __device__ inline float op_inv_scale(float sq_norm) { return pow2f(-op_scale_log2(sq_norm)); }

// Guaranteed error of ONE dot product x_i . x_j out of the tensor cores, relative to |x_i||x_j|, by the number of
// products per k-step.  With u = 2^-11 (FP16 round to nearest, 11 significant bits) and, per element,
// |x - hi| <= u |x| + 2^-25 and |x - hi - lo| <= u^2 |x| + 2^-25 (scaled units, |x_row| >= 2^5):
//   * representation, by Cauchy-Schwarz over the row (the absolute parts sum to 2^-25 sqrt(D) |y| <= 2^-30 sqrt(D)
//     |x||y| each):
//       3 products: lo.lo + r.y + x.r left out                       ->  3 u^2 + 2 * 2^-30 sqrt(D)
//       2 products: x_i . (x_j - hi_j) left out as well             ->  u + u^2 + 2 * 2^-30 sqrt(D)
//       1 product : both rows rounded                                ->  2 u + u^2 + 2 * 2^-30 sqrt(D)
//   * accumulation: prods * D_pad / 16 tensor-core steps into one FP32 accumulator whose add truncates, at most 2^-22
//     of the running magnitude (<= |x_i||x_j|) per step.  (Measured in the dense mode, one accumulator of 192 steps,
//     D = 1024: worst entry of 2.7e8 off by 7e-6 relative, an eighth of this bound.)
// A Euclidean / Gaussian score is n_j - 2 dot, so its error is 2 eps |x_i||x_j| <= eps (n_i + n_j); a cosine score
// -dot / |x_j| is off by eps |x_i|.  Two scores of one row therefore compare correctly once they differ by more than
// (eps_a + eps_b) * (n_i + n_max) resp. (eps_a + eps_b) * |x_i| -- tc_slack() -- plus the FP32 rounding of the score
// arithmetic itself (fma, spatial term), which is relative to the score: TC_SLACK_REL * |score| on top.
// `rho` = the largest ||s x - hi|| / ||s x|| over the operand rows, MEASURED by the split kernel (split.cu; at most u, and
// ~0.3 u on real rows because rounding errors do not line up): Cauchy-Schwarz with the actual residual norm instead of
// its worst case, still a guarantee.  It replaces u in the terms that come from rounding a whole operand to its hi half
// (two products: one operand; one product: both).  Pass TC_RHO_WORST when no measurement is at hand.
constexpr float TC_RHO_WORST = 4.8828125e-4f;       // u = 2^-11
__host__ __device__ inline float tc_eps(int64_t d_pad, int prods, float rho) {
    const float u2 = 2.3841858e-7f;                                    // 2^-22
    if (prods <= 0) prods = 3;
    rho = fmaxf(rho, 0.f) * 1.01f + 1e-7f;                             // FP32 rounding of the measurement itself
    const float repr = (prods >= 3 ? 3.f * u2 : prods == 2 ? rho + u2 : 2.f * rho + u2) + 1.8627e-9f * sqrtf(float(d_pad));
    return repr + float(prods * (d_pad / 16)) * u2;
}
__host__ __device__ inline float tc_slack(int64_t d_pad, int prods_a, int prods_b, float rho = TC_RHO_WORST) {
    return tc_eps(d_pad, prods_a, rho) + tc_eps(d_pad, prods_b, rho);
}
constexpr float TC_SLACK_REL = 4.8e-7f;           // 2^-21

// also resets the symmetric-mode state when given: thr = NaN, bucket counts = 0, and folds the
// largest squared norm into *norm_max (zeroed by the caller)
int launch_colparams(const float* sq_norms, const float* pos, int pos_dim, int64_t n, int64_t n_pad, int metric,
                     float2* colp, float4* pos4, cudaStream_t st,
                     float* thr = nullptr, int32_t* bkt_cnt = nullptr, float* norm_max = nullptr);

int launch_split_group(const RrSlide* g_slides, const int32_t* g_blk2slide, int64_t n_pad, int64_t d, bool vec,
                       op16_t* ops, float* sq_norms, cudaStream_t st, float* rho_max = nullptr);
int launch_colparams_group(const float* sq_norms, int pos_dim, int64_t n_pad, int metric, float2* colp, float4* pos4,
                           float* norm_max, const RrSlide* g_slides, const int32_t* g_blk2slide, cudaStream_t st);

// pinned host staging for per-group tables (knn_api.cu): acquire -> fill host -> cudaMemcpyAsync -> release(stream)
struct StageSlot { void* host = nullptr; size_t cap = 0; cudaEvent_t done = nullptr; };
int stage_acquire(size_t bytes, StageSlot** out);
int stage_release(StageSlot* s, cudaStream_t st);

// ---- grouped back half (group_back.cu): edge_index / weights / CSR of every slide of a group from its neighbour
// lists in ~22 launches per GROUP instead of ~18 per slide.  Slides of at most GB_MAX_N rows.
constexpr int64_t GB_MAX_N = 9216;
struct GbSlide {
    int64_t n;
    const float* feats;
    const int32_t* knn_idx; const float* knn_cos;                 // [n, k] lists of the slide (local indices, -1 = none)
    const int32_t* spatial_idx;                                   // [n, spatial_k] or null
    int64_t* edge_index; float* edge_weights; int64_t* num_edges;  // compact [2, E] / [E] / device count
    int64_t* rowptr; int64_t* col; int64_t* eid;                  // CSR incidence, or all null
};
size_t group_back_workspace_bytes(const GbSlide* slides, int num, int k, int spatial_k);
int group_back_enqueue(const GbSlide* slides, int num, int64_t d, int k, int spatial_k, void* workspace,
                       size_t workspace_bytes, cudaStream_t st);

// ---- grouped feature kNN (knn_api.cu): the front half of a batch of small slides in a handful of launches ----
struct KnnGroupSlide { const float* feats; const float* pos; int64_t n; int32_t* out_idx; float* out_val; float* out_cos; };
size_t knn_group_workspace_bytes(const KnnGroupSlide* slides, int num, int64_t d, int k, bool has_pos);
int knn_group_enqueue(const KnnGroupSlide* slides, int num, int64_t d, int pos_dim, int metric, float lambda_h,
                      float lambda_g, int k, int exclude_self, void* workspace, size_t workspace_bytes, cudaStream_t st,
                      cudaStream_t st_rerank = nullptr);

}  // namespace hg
