// C ABI of the feature-space kNN: operand split -> fused tcgen05 similarity/top-k ->
// merge + exact re-rank.  See include/hg_b200.h for the contract.
#include "knn_tc.cuh"
#include "sim_simt.cuh"
#include <stdlib.h>
#include <algorithm>
#include <mutex>
#include <string.h>
#include <vector>

namespace hg {

struct KnnWs {
    size_t off_ops, off_norms, off_colp, off_pos4, off_pval, off_pidx, off_thr, off_cnt, off_ovf, off_bkt, off_unc_rows,
        off_unc_part, bytes;
    TcPlan plan;                 // classic plan (the only one, or the overflow fallback of symmetric mode)
    TcPlan sym_plan, pre_plan;   // symmetric mode: main pass and threshold pre-pass
    int sym_stride;              // 0 = classic
    int64_t n_pad, d_pad;
};

// entries per row bucket; HG_B200_TC_BKT_CAP lowers it (tests force the overflow fallback with it)
static int bucket_cap() {
    if (const char* e = getenv("HG_B200_TC_BKT_CAP")) {
        const int v = atoi(e);
        if (v >= 1 && v <= TC_BKT_CAP) return v;
    }
    return TC_BKT_CAP;
}

// Tensor-core products per k-step (knn_tc.cuh tc_eps) of the symmetric passes: 2 by default -- the columns
// rounded to FP16, the query rows whole -- which issues two thirds of the MMAs and loads no lo half of a B tile, for a
// wider slack that the many short lists and the buckets of the symmetric mode hold easily; HG_B200_TC_SYM_PRODS=3
// keeps all three.  Classic, grouped and dense launches always use three.  The threshold pre-pass uses the main pass's count, or one product with
// HG_B200_TC_PRE_HI=1.
static int sym_prods(int64_t n) {
    if (const char* e = getenv("HG_B200_TC_SYM_PRODS")) { const int v = atoi(e); if (v == 2 || v == 3) return v; }
    (void)n;
    return 2;
}
static int prepass_prods(int64_t n) {
    if (const char* e = getenv("HG_B200_TC_PRE_HI")) if (atoi(e) != 0) return 1;
    return sym_prods(n);
}
// The symmetric pass over the row block of a PARTITIONED slide (the giant ones, C5: 250k rows): two products as well,
// HG_B200_TC_SYMPART_PRODS=3 keeps three.  With the WORST-CASE slack (2^-11 per rounded operand) this did not pay: the
// denser the slide, the more columns fall inside the band -- on C5's data 6-28 % of the rows had more than 32 of them
// (multi-round re-rank) and a few hundred filled a list (exact fallback): 20 ms of fused kernel per rank instead of 28,
// but 64 ms per slide instead of 36.  With the MEASURED rounding residual of the rows (tc_eps, 0.47 of the worst case)
// the band is half as wide: no row uncertified, at most 73 multi-round rows per rank, 29 ms per slide.
static int sympart_prods() {
    if (const char* e = getenv("HG_B200_TC_SYMPART_PRODS")) { const int v = atoi(e); if (v == 2 || v == 3) return v; }
    return 2;
}
// Classic launches keep ONE or two lists of 24 entries per row, so the width of the slack decides how many rows have a
// list that fills up inside it and take the exact fallback.  With the worst-case two-product slack that was 1.5-4.5 %
// of the rows of a group of small slides (C4: 111 -> 171 ms); with the measured one (tc_eps: rho) it is 60-100 rows
// of 128,000 and the batch runs in 99.5 ms (fused kernel 49 ms instead of 65).  Small slides (up to 16,384 columns:
// every grouped launch, single small slides) therefore use two products; a classic launch over more columns -- the
// row block of a large slide -- is denser per list and keeps three.  HG_B200_TC_CLASSIC_PRODS=2/3 overrides.
static int classic_prods(int64_t n_cols) {
    if (const char* e = getenv("HG_B200_TC_CLASSIC_PRODS")) { const int v = atoi(e); if (v == 2 || v == 3) return v; }
    return n_cols <= 16384 ? 2 : 3;
}

static KnnWs plan_ws(int64_t n, int64_t d, int k, int64_t r0, int64_t r1, bool has_pos) {
    KnnWs w{};
    w.n_pad = ops_rows(n);
    w.d_pad = ops_cols(d);
    w.plan = tc_make_plan(n, r1 - r0, k, w.d_pad);
    size_t off = 0;
    auto take = [&](size_t b) { size_t o = off; off += (b + 1023) & ~size_t(1023); return o; };
    w.off_ops = take(hg_operand_bytes(n, d));
    w.off_norms = take(size_t(w.n_pad) * 4);
    w.off_colp = take(size_t(w.n_pad) * 8);
    w.off_pos4 = take(has_pos ? size_t(w.n_pad) * 16 : 0);
    int max_splits = w.plan.splits;
    w.sym_stride = tc_sym_stride(n, r1 - r0, k);
    if (w.sym_stride > 0) {
        w.sym_plan = tc_make_plan_sym(n, k, w.d_pad);
        w.pre_plan = tc_make_plan_prepass(n, k, w.d_pad, w.sym_stride);
        if (w.sym_plan.splits > max_splits) max_splits = w.sym_plan.splits;
        if (w.pre_plan.splits > max_splits) max_splits = w.pre_plan.splits;
    }
    const size_t part = size_t(max_splits) * size_t(w.plan.n_rows) * size_t(w.plan.kl);
    w.off_pval = take(part * 4);
    w.off_pidx = take(part * 4);
    w.off_ovf = take(32);                       // overflow flag, largest squared norm, uncertified-row counters, measured rho
    w.off_unc_rows = take(size_t(w.plan.n_rows) * 12);       // uncertified rows | rows of the multi-round re-rank | split masks
    w.off_unc_part = take(rerank_unc_part_bytes(w.plan.n_rows, k));
    if (w.sym_stride > 0) {
        w.off_thr = take(size_t(w.n_pad) * 4);
        w.off_cnt = take(size_t(w.n_pad) * 4);
        w.off_bkt = take(size_t(w.n_pad) * TC_BKT_CAP * 8);
    }
    w.bytes = off;
    return w;
}

}  // namespace hg

using namespace hg;

static int check_knn_args(int64_t n, int64_t d, int k, int64_t r0, int64_t r1) {
    HG_REQUIRE(n > 1, HG_ERR_BAD_ARG, "number of nodes must be greater than 1, got N=%lld", (long long)n);
    HG_REQUIRE(d > 0 && k >= 1, HG_ERR_BAD_ARG, "hg_knn_feature: bad d / k");
    HG_REQUIRE(r0 >= 0 && r0 < r1 && r1 <= n, HG_ERR_BAD_ARG, "hg_knn_feature: bad row range [%lld, %lld)",
               (long long)r0, (long long)r1);
    HG_REQUIRE(n < (int64_t(1) << 31) - 512, HG_ERR_UNSUPPORTED, "hg_knn_feature: n too large");
    return HG_OK;
}

extern "C" int hg_knn_feature_workspace_bytes(int64_t n, int64_t d, int k, int64_t row_begin, int64_t row_end,
                                              size_t* bytes) {
    HG_REQUIRE(bytes, HG_ERR_BAD_ARG, "hg_knn_feature_workspace_bytes: null");
    int rc = check_knn_args(n, d, k, row_begin, row_end);
    if (rc != HG_OK) return rc;
    *bytes = plan_ws(n, d, k, row_begin, row_end, true).bytes;
    return HG_OK;
}

extern "C" int hg_knn_feature_plan(int64_t n, int64_t d, int k, int64_t row_begin, int64_t row_end,
                                   int* symmetric_stride, double* mma_flop) {
    int rc = check_knn_args(n, d, k, row_begin, row_end);
    if (rc != HG_OK) return rc;
    const KnnWs w = plan_ws(n, d, k, row_begin, row_end, false);
    auto tiles_of = [](const TcPlan& pl) {
        if (!pl.sym) return double(pl.num_rb) * double(pl.num_ct);
        double t = 0;
        for (int rb = 0; rb < pl.num_rb; ++rb)
            for (int ct = 0; ct < pl.num_ct; ++ct) {
                const int a = pl.cl == 2 ? rb : rb >> 1;
                t += (ct == a || (((a + ct) & 1) == (ct < a ? 1 : 0))) ? double(pl.cl) : 0.0;
            }
        return t;
    };
    const double per_tile = 2.0 * double(TC_BM) * double(TC_BN) * double(w.d_pad);           // one MMA sweep of a tile
    const double flop = w.sym_stride > 0 ? (tiles_of(w.sym_plan) * sym_prods(n) + tiles_of(w.pre_plan) * prepass_prods(n)) * per_tile
                                         : tiles_of(w.plan) * w.plan.cl * double(classic_prods(n)) * per_tile;
    if (symmetric_stride) *symmetric_stride = w.sym_stride;
    if (mma_flop) *mma_flop = flop;
    return HG_OK;
}

extern "C" int hg_knn_feature_products(int64_t n, int64_t d, int k, int64_t row_begin, int64_t row_end,
                                       int* main_pass, int* pre_pass) {
    int rc = check_knn_args(n, d, k, row_begin, row_end);
    if (rc != HG_OK) return rc;
    const bool sym = tc_sym_stride(n, row_end - row_begin, k) > 0;
    if (main_pass) *main_pass = sym ? sym_prods(n) : classic_prods(n);
    if (pre_pass) *pre_pass = sym ? prepass_prods(n) : 0;
    return HG_OK;
}

extern "C" int hg_knn_feature(const float* feats, int64_t n, int64_t d,
                              const float* pos, int pos_dim,
                              int metric, float lambda_h, float lambda_g,
                              int k, int64_t row_begin, int64_t row_end, int exclude_self,
                              int precision,
                              int32_t* out_idx, float* out_val,
                              void* workspace, size_t workspace_bytes, void* stream) {
    return hg_knn_feature_cos(feats, n, d, pos, pos_dim, metric, lambda_h, lambda_g, k, row_begin, row_end,
                              exclude_self, precision, out_idx, out_val, nullptr, workspace, workspace_bytes, stream);
}

extern "C" int hg_knn_feature_cos(const float* feats, int64_t n, int64_t d,
                                  const float* pos, int pos_dim,
                                  int metric, float lambda_h, float lambda_g,
                                  int k, int64_t row_begin, int64_t row_end, int exclude_self,
                                  int precision,
                                  int32_t* out_idx, float* out_val, float* out_cos,
                                  void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(feats && out_idx && out_val, HG_ERR_BAD_ARG, "hg_knn_feature: null buffer");
    int rc = check_knn_args(n, d, k, row_begin, row_end);
    if (rc != HG_OK) return rc;
    HG_REQUIRE(metric == HG_METRIC_EUCLIDEAN || metric == HG_METRIC_GAUSSIAN || metric == HG_METRIC_COSINE,
               HG_ERR_BAD_ARG, "hg_knn_feature: unknown metric %d", metric);
    if (metric != HG_METRIC_GAUSSIAN || lambda_g == 0.f) pos = nullptr;
    HG_REQUIRE(pos == nullptr || (pos_dim >= 1 && pos_dim <= 3), HG_ERR_UNSUPPORTED, "hg_knn_feature: pos_dim must be 1..3");
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    HG_REQUIRE(out_cos == nullptr || precision == 0, HG_ERR_UNSUPPORTED,
               "hg_knn_feature_cos: the cosine by-product needs the tensor-core path (precision 0)");
    if (precision == 1) {                                   // plain FP32 SIMT path
        SimArgs a{};
        a.a = feats; a.b = feats; a.m = n; a.n = n; a.d = d; a.pa = pos; a.pb = pos; a.pos_dim = pos_dim;
        a.lambda_h = lambda_h; a.lambda_g = lambda_g; a.metric = metric; a.k = k;
        a.row_begin = row_begin; a.row_end = row_end; a.exclude_self = exclude_self;
        a.out_idx = out_idx; a.out_val = out_val;
        return launch_topk_fp32(a, st);
    }
    HG_REQUIRE(precision == 0, HG_ERR_BAD_ARG, "hg_knn_feature: unknown precision %d", precision);
    HG_REQUIRE(k <= 32, HG_ERR_UNSUPPORTED, "hg_knn_feature: the tensor-core path supports k <= 32 (got %d)", k);
    HG_REQUIRE(metric != HG_METRIC_GAUSSIAN || lambda_h > 0.f || pos == nullptr, HG_ERR_UNSUPPORTED,
               "hg_knn_feature: lambda_h must be > 0 when positions are given");

    const KnnWs w = plan_ws(n, d, k, row_begin, row_end, pos != nullptr);
    HG_REQUIRE(workspace && workspace_bytes >= w.bytes, HG_ERR_WORKSPACE,
               "hg_knn_feature: workspace too small (%zu < %zu)", workspace_bytes, w.bytes);
    HG_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 1023) == 0, HG_ERR_BAD_ARG,
               "hg_knn_feature: workspace must be 1024-byte aligned");
    char* ws = static_cast<char*>(workspace);
    op16_t* ops = reinterpret_cast<op16_t*>(ws + w.off_ops);
    float* norms = reinterpret_cast<float*>(ws + w.off_norms);
    float2* colp = reinterpret_cast<float2*>(ws + w.off_colp);
    float4* pos4 = pos ? reinterpret_cast<float4*>(ws + w.off_pos4) : nullptr;

    const bool sym = w.sym_stride > 0;
    float* thr = sym ? reinterpret_cast<float*>(ws + w.off_thr) : nullptr;
    int32_t* bkt_cnt = sym ? reinterpret_cast<int32_t*>(ws + w.off_cnt) : nullptr;
    int* ovf = reinterpret_cast<int*>(ws + w.off_ovf);
    uint2* bkt = sym ? reinterpret_cast<uint2*>(ws + w.off_bkt) : nullptr;

    float* norm_max = reinterpret_cast<float*>(ws + w.off_ovf) + 1;
    float* rho_max = reinterpret_cast<float*>(ws + w.off_ovf) + 4;      // largest rounding residual of an operand row (tc_eps)
    HG_CHECK_CUDA(cudaMemsetAsync(ovf, 0, 32, st));
    rc = launch_split(feats, n, d, ops, norms, st, false, rho_max);
    if (rc != HG_OK) return rc;
    rc = launch_colparams(norms, pos, pos_dim, n, w.n_pad, metric, colp, pos4, st, thr, bkt_cnt, norm_max);
    if (rc != HG_OK) return rc;

    TcParams tp{};
    tp.n = n; tp.row_begin = row_begin; tp.row_end = row_end; tp.d_pad = int(w.d_pad);
    tp.plan = w.plan; tp.colp = colp; tp.pos4 = pos4;
    tp.norms_a = norms; tp.prods = sym ? 3 : classic_prods(n);      // (sym: the guarded fallback pass)
    tp.rho_max = rho_max;
    tp.gscale = (pos && lambda_h > 0.f) ? lambda_g / lambda_h : 0.f;
    tp.exclude_self = exclude_self;
    tp.part_val = reinterpret_cast<float*>(ws + w.off_pval);
    tp.part_idx = reinterpret_cast<int32_t*>(ws + w.off_pidx);
    prof_begin(st);
    if (sym) {
        // (1) pre-pass over every stride-th column: thr[row] = an upper bound of the row's k-th best
        //     (the (k+1)-th best of the sample, self not excluded, plus error slack), (2) the half pass,
        //     (3) the classic full pass, which exits at once unless a bucket overflowed
        TcParams pre = tp;
        pre.plan = w.pre_plan; pre.col_stride = w.sym_stride; pre.exclude_self = 0;
        const int cosine = (metric == HG_METRIC_COSINE) ? 1 : 0;
        pre.thr_rank = k; pre.thr_margin_cos = cosine; pre.norm_max = norm_max;
        pre.prods = prepass_prods(n); pre.pre_main_prods = sym_prods(n);
        pre.thr_out = (w.pre_plan.splits == 1) ? thr : nullptr;     // several splits: lists, merged below
        rc = tc_launch(ops, w.n_pad, pre, st);
        if (rc != HG_OK) return rc;
        if (w.pre_plan.splits > 1) {
            rc = thr_merge_launch(pre.part_val, w.pre_plan.splits, n, w.pre_plan.kl, k, colp, norm_max, cosine,
                                  w.d_pad, pre.prods, pre.pre_main_prods, rho_max, thr, st);
            if (rc != HG_OK) return rc;
        }
        TcParams sp = tp;
        sp.prods = sym_prods(n);
        sp.plan = w.sym_plan; sp.thr = thr; sp.bkt_cnt = bkt_cnt; sp.bkt = bkt; sp.bkt_cap = bucket_cap(); sp.ovf = ovf;
        rc = tc_launch(ops, w.n_pad, sp, st);
        if (rc != HG_OK) return rc;
        tp.guard = ovf;
    }
    rc = tc_launch(ops, w.n_pad, tp, st);
    prof_end(st);
    if (rc != HG_OK) return rc;

    RerankParams rp{};
    rp.feats = feats; rp.n = n; rp.d = d; rp.pos = pos; rp.pos_dim = pos_dim;
    rp.sq_norms = norms; rp.norm_max = norm_max; rp.rho_max = rho_max;
    rp.metric = metric; rp.lambda_h = lambda_h; rp.lambda_g = pos ? lambda_g : 0.f;
    rp.part_val = tp.part_val; rp.part_idx = tp.part_idx;
    rp.splits = w.plan.splits; rp.kl = w.plan.kl; rp.n_rows = w.plan.n_rows; rp.row_begin = row_begin;
    rp.cols_per_split = int64_t(w.plan.ct_per_split) * TC_BN;
    rp.k = k; rp.out_idx = out_idx; rp.out_val = out_val; rp.out_cos = out_cos;
    rp.prods = tp.prods;
    if (sym) { rp.sym_plan = w.sym_plan; rp.sym_prods = sym_prods(n); rp.bkt_cnt = bkt_cnt; rp.bkt = bkt; rp.bkt_cap = bucket_cap(); rp.ovf = ovf; }
    rp.unc_cnt = ovf + 2;
    rp.unc_rows = reinterpret_cast<int32_t*>(ws + w.off_unc_rows);
    rp.more_cnt = ovf + 3; rp.more_rows = rp.unc_rows + w.plan.n_rows;
    rp.unc_part = reinterpret_cast<uint2*>(ws + w.off_unc_part);
    rp.exclude_self = exclude_self;
    rc = rerank_launch(rp, st);
    if (rc == HG_OK && sym && getenv("HG_B200_TC_DEBUG")) {      // tuning aid: bucket occupancy (synchronises)
        std::vector<int32_t> cnt(static_cast<size_t>(n), 0);
        int h_ovf = 0;
        HG_CHECK_CUDA(cudaStreamSynchronize(st));
        HG_CHECK_CUDA(cudaMemcpy(cnt.data(), bkt_cnt, size_t(n) * 4, cudaMemcpyDeviceToHost));
        HG_CHECK_CUDA(cudaMemcpy(&h_ovf, ovf, 4, cudaMemcpyDeviceToHost));
        int64_t sum = 0; int mx = 0;
        for (int32_t c : cnt) { sum += c; mx = c > mx ? c : mx; }
        fprintf(stderr, "hg_b200 sym: n=%lld stride=%d splits=%d cps=%d pre_ct=%d bucket mean=%.1f max=%d ovf=%d\n",
                (long long)n, w.sym_stride, w.sym_plan.splits, w.sym_plan.ct_per_split, w.pre_plan.num_ct,
                double(sum) / double(n), mx, h_ovf);
    }
    if (rc == HG_OK && getenv("HG_B200_TC_DEBUG")) {             // tuning aid: rows sent to the exact fallback (synchronises)
        int h_unc[2] = {0, 0};
        float h_rho = 0.f;
        HG_CHECK_CUDA(cudaStreamSynchronize(st));
        HG_CHECK_CUDA(cudaMemcpy(h_unc, ovf + 2, 8, cudaMemcpyDeviceToHost));
        HG_CHECK_CUDA(cudaMemcpy(&h_rho, rho_max, 4, cudaMemcpyDeviceToHost));
        fprintf(stderr, "hg_b200 rerank: %d of %lld rows uncertified -> exact fallback, %d rows multi-round, rho = %.3g (2^-11 = 4.88e-4)\n",
                h_unc[0], (long long)(row_end - row_begin), h_unc[1], double(h_rho));
    }
    return rc;
}

// ---- descriptor checks shared by the row-block entry points (symmetric pass over partitioned rows, below) --------
namespace hg {

static int part_check(const hg_knn_part* s) {
    HG_REQUIRE(s, HG_ERR_BAD_ARG, "hg_knn_part: null descriptor");
    int rc = check_knn_args(s->n, s->d, s->k, s->row_begin, s->row_end);
    if (rc != HG_OK) return rc;
    HG_REQUIRE(s->k <= 32, HG_ERR_UNSUPPORTED, "hg_knn_part: the tensor-core path supports k <= 32 (got %d)", s->k);
    HG_REQUIRE(s->metric == HG_METRIC_EUCLIDEAN || s->metric == HG_METRIC_GAUSSIAN || s->metric == HG_METRIC_COSINE,
               HG_ERR_BAD_ARG, "hg_knn_part: unknown metric %d", s->metric);
    HG_REQUIRE(s->pos == nullptr || (s->pos_dim >= 1 && s->pos_dim <= 3), HG_ERR_UNSUPPORTED, "hg_knn_part: pos_dim must be 1..3");
    HG_REQUIRE(s->metric != HG_METRIC_GAUSSIAN || s->lambda_h > 0.f || s->pos == nullptr, HG_ERR_UNSUPPORTED,
               "hg_knn_part: lambda_h must be > 0 when positions are given");
    return HG_OK;
}

static bool part_has_pos(const hg_knn_part* s) {
    return s->pos != nullptr && s->metric == HG_METRIC_GAUSSIAN && s->lambda_g != 0.f;
}

}  // namespace hg

// ---- grouped feature kNN: many small slides, a handful of launches ------------------------------------------
// A slide of 2-8k patches is 16-64 row blocks: launched on its own it leaves SMs idle (128 units on 148 SMs at
// N = 4,096), pays the ragged end of a persistent launch and the empty-list warm-up with no epilogue overlap
// behind its last tile -- 512 such slides spent 86 ms in 512 launches of the fused kernel.  Here the operands of
// a whole group of slides sit in ONE buffer (each slide at a 256-row aligned offset, so column tiles never
// straddle two slides) and one split launch, one column-parameter launch, ONE persistent launch of the fused
// kernel over a unit table spanning all the slides (longest units first) and one re-rank launch do the front
// half of every slide of the group.  Results are those of the per-slide calls bit for bit: a unit sweeps
// exactly the column tiles of its own slide, in the same order, with the same list length.
namespace hg {

// Pinned host staging for the per-group tables.  cudaMemcpyAsync from PAGEABLE memory is staged by a driver worker
// thread some time after the call returns, so the source must outlive the call; a small ring of pinned buffers,
// each guarded by an event recorded behind its copy, gives a source that does and a copy that is a plain DMA.
static std::mutex g_stage_mu;
static StageSlot g_stage[8];
static int g_stage_next = 0;

int stage_acquire(size_t bytes, StageSlot** out) {
    std::lock_guard<std::mutex> lock(g_stage_mu);
    StageSlot& s = g_stage[g_stage_next];
    g_stage_next = (g_stage_next + 1) % 8;
    if (s.done != nullptr) {                       // the copy that last read this slot
        HG_CHECK_CUDA(cudaEventSynchronize(s.done));
        HG_CHECK_CUDA(cudaEventDestroy(s.done));
        s.done = nullptr;
    }
    if (s.cap < bytes) {
        if (s.host != nullptr) HG_CHECK_CUDA(cudaFreeHost(s.host));
        s.host = nullptr; s.cap = 0;
        const size_t cap = (bytes + (size_t(1) << 20) - 1) & ~((size_t(1) << 20) - 1);
        HG_CHECK_CUDA(cudaHostAlloc(&s.host, cap, cudaHostAllocPortable));
        s.cap = cap;
    }
    *out = &s;
    return HG_OK;
}
int stage_release(StageSlot* s, cudaStream_t st) {
    HG_CHECK_CUDA(cudaEventCreateWithFlags(&s->done, cudaEventDisableTiming));
    HG_CHECK_CUDA(cudaEventRecord(s->done, st));
    return HG_OK;
}

struct GroupLayout {
    int64_t n_pad, d_pad;                  // operand rows of the whole group
    int kl;
    std::vector<int64_t> off;              // first operand row of every slide
    std::vector<int> splits;               // column splits (candidate lists per row) of every slide
    std::vector<int64_t> part_off;         // element offset of every slide's list block
    int64_t part_elems;
    int num_units;
    size_t o_ops, o_norms, o_colp, o_pos4, o_pval, o_pidx, o_flags, o_unc_rows, o_unc_part, o_units, o_slides, o_map, bytes;
};

// A unit sweeps at most this many column tiles: short units keep the tail of the persistent launch short and let
// slides of different sizes interleave; a slide of up to 4,096 rows keeps ONE list per row.
constexpr int kGroupTilesPerUnit = 16;

static GroupLayout group_layout(const KnnGroupSlide* slides, int num, int64_t d, int k, bool has_pos) {
    GroupLayout g{};
    g.d_pad = ops_cols(d);
    g.kl = tc_list_len(k);
    g.off.resize(num); g.splits.resize(num); g.part_off.resize(num);
    int64_t rows = 0, part = 0;
    int units = 0;
    for (int i = 0; i < num; ++i) {
        g.off[i] = rows;
        rows += ops_rows(slides[i].n);
        const int tiles = int(ceil_div(slides[i].n, TC_BN));
        int s = int(ceil_div(tiles, kGroupTilesPerUnit));
        if (s > TC_MAX_SPLITS) s = TC_MAX_SPLITS;
        g.splits[i] = s;
        g.part_off[i] = part;
        part += int64_t(s) * slides[i].n * g.kl;
        units += int(ceil_div(slides[i].n, TC_BM)) * s;
    }
    g.n_pad = rows; g.part_elems = part; g.num_units = units;
    size_t off = 0;
    auto take = [&](size_t b) { size_t o = off; off += (b + 1023) & ~size_t(1023); return o; };
    g.o_ops = take(size_t(g.n_pad) * size_t(ops_row_elems(g.d_pad)) * 2);
    g.o_norms = take(size_t(g.n_pad) * 4);
    g.o_colp = take(size_t(g.n_pad) * 8);
    g.o_pos4 = take(has_pos ? size_t(g.n_pad) * 16 : 0);
    g.o_pval = take(size_t(part) * 4);
    g.o_pidx = take(size_t(part) * 4);
    g.o_flags = take(32);
    g.o_unc_rows = take(size_t(g.n_pad) * 12);
    g.o_unc_part = take(rerank_unc_part_bytes(g.n_pad, k));
    g.o_units = take(size_t(units) * sizeof(TcUnit));
    g.o_slides = take(size_t(num) * sizeof(RrSlide));
    g.o_map = take(size_t(g.n_pad / kRowPad) * 4);
    g.bytes = off;
    return g;
}

// HG_B200_GROUP_DEBUG: print the slide table as the device sees it and check every list entry's range
__global__ void group_debug_kernel(const RrSlide* sl, const int32_t* map, int num, int64_t n_pad, const float* part_val,
                                   const int32_t* part_idx, int kl) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        for (int i = 0; i < num; ++i)
            printf("slide %d: feats %p pos %p off %lld n %lld part_off %lld splits %d out %p %p %p\n", i, (const void*)sl[i].feats,
                   (const void*)sl[i].pos, (long long)sl[i].off, (long long)sl[i].n, (long long)sl[i].part_off, sl[i].splits,
                   (void*)sl[i].out_idx, (void*)sl[i].out_val, (void*)sl[i].out_cos);
        for (int64_t b = 0; b < n_pad / 256; ++b) printf("map[%lld] = %d\n", (long long)b, map[b]);
    }
    for (int i = 0; i < num; ++i) {
        const int64_t cnt = int64_t(sl[i].splits) * sl[i].n * kl;
        for (int64_t e = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; e < cnt; e += int64_t(gridDim.x) * blockDim.x) {
            const int idx = part_idx[sl[i].part_off + e];
            if (idx < -1 || idx >= sl[i].n)
                printf("slide %d list entry %lld: index %d out of [-1, %lld) score %f\n", i, (long long)e, idx, (long long)sl[i].n,
                       part_val[sl[i].part_off + e]);
        }
    }
}

size_t knn_group_workspace_bytes(const KnnGroupSlide* slides, int num, int64_t d, int k, bool has_pos) {
    return group_layout(slides, num, d, k, has_pos).bytes;
}

// st_rerank: stream of the exact re-rank (null = st).  The re-rank streams ~(k + 1) feature rows per query out of
// HBM (2.5 M queries of a 512-slide batch: ~40 ms against 61 ms in the fused kernel); on a stream of its own it
// runs next to the NEXT group's fused kernel, which is tensor-bound and leaves the memory system idle.
int knn_group_enqueue(const KnnGroupSlide* slides, int num, int64_t d, int pos_dim, int metric, float lambda_h,
                      float lambda_g, int k, int exclude_self, void* workspace, size_t workspace_bytes, cudaStream_t st,
                      cudaStream_t st_rerank) {
    HG_REQUIRE(slides && num > 0 && d > 0 && k >= 1 && k <= 32, HG_ERR_BAD_ARG, "knn_group: bad argument");
    bool has_pos = metric == HG_METRIC_GAUSSIAN && lambda_g != 0.f;
    for (int i = 0; i < num; ++i) {
        HG_REQUIRE(slides[i].feats && slides[i].n > 1 && slides[i].out_idx && slides[i].out_val, HG_ERR_BAD_ARG,
                   "knn_group: slide %d: null buffer or fewer than 2 rows", i);
        has_pos = has_pos && slides[i].pos != nullptr;
    }
    HG_REQUIRE(!has_pos || (pos_dim >= 1 && pos_dim <= 3), HG_ERR_UNSUPPORTED, "knn_group: pos_dim must be 1..3");
    HG_REQUIRE(!has_pos || lambda_h > 0.f, HG_ERR_UNSUPPORTED, "knn_group: lambda_h must be > 0 when positions are given");
    const GroupLayout g = group_layout(slides, num, d, k, has_pos);
    HG_REQUIRE(g.n_pad < (int64_t(1) << 31) - 512, HG_ERR_UNSUPPORTED, "knn_group: group too large");
    HG_REQUIRE(workspace && workspace_bytes >= g.bytes, HG_ERR_WORKSPACE, "knn_group: workspace too small (%zu < %zu)",
               workspace_bytes, g.bytes);
    HG_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 1023) == 0, HG_ERR_BAD_ARG, "knn_group: workspace must be 1024-byte aligned");
    char* ws = static_cast<char*>(workspace);
    op16_t* ops = reinterpret_cast<op16_t*>(ws + g.o_ops);
    float* norms = reinterpret_cast<float*>(ws + g.o_norms);
    float2* colp = reinterpret_cast<float2*>(ws + g.o_colp);
    float4* pos4 = has_pos ? reinterpret_cast<float4*>(ws + g.o_pos4) : nullptr;
    int* flags = reinterpret_cast<int*>(ws + g.o_flags);
    TcUnit* d_units = reinterpret_cast<TcUnit*>(ws + g.o_units);
    RrSlide* d_slides = reinterpret_cast<RrSlide*>(ws + g.o_slides);
    int32_t* d_map = reinterpret_cast<int32_t*>(ws + g.o_map);

    // host tables (unit table, slide table, row-block -> slide map) in one pinned staging block laid out like the
    // device side, so a single DMA moves them
    StageSlot* stage = nullptr;
    const size_t tab_bytes = g.bytes - g.o_units;
    int rc = stage_acquire(tab_bytes, &stage);
    if (rc != HG_OK) return rc;
    char* hs = static_cast<char*>(stage->host);
    RrSlide* rs = reinterpret_cast<RrSlide*>(hs + (g.o_slides - g.o_units));
    int32_t* map = reinterpret_cast<int32_t*>(hs + (g.o_map - g.o_units));
    std::vector<TcUnit> units;
    units.reserve(size_t(g.num_units));
    bool vec = d % 4 == 0;
    int max_splits = 1;
    for (int i = 0; i < num; ++i) {
        const KnnGroupSlide& s = slides[i];
        vec = vec && (reinterpret_cast<uintptr_t>(s.feats) & 15) == 0;
        rs[i] = RrSlide{s.feats, has_pos ? s.pos : nullptr, g.off[i], s.n, g.part_off[i], g.splits[i], 0, s.out_idx, s.out_val, s.out_cos,
                        int64_t(ceil_div(ceil_div(s.n, TC_BN), g.splits[i])) * TC_BN};
        for (int64_t b = g.off[i] / kRowPad; b < (g.off[i] + ops_rows(s.n)) / kRowPad; ++b) map[size_t(b)] = i;
        const int tiles = int(ceil_div(s.n, TC_BN)), cps = int(ceil_div(tiles, g.splits[i]));
        const int ct_base = int(g.off[i] / TC_BN);
        if (g.splits[i] > max_splits) max_splits = g.splits[i];
        for (int sp = 0; sp < g.splits[i]; ++sp) {
            const int c0 = sp * cps, c1 = std::min(tiles, c0 + cps);
            for (int64_t rb = 0; rb < ceil_div(s.n, TC_BM); ++rb) {
                TcUnit u{};
                u.row0 = int32_t(g.off[i] + rb * TC_BM); u.row_end = int32_t(g.off[i] + s.n);
                u.ct0 = ct_base + c0; u.ct1 = ct_base + std::max(c0, c1);
                u.col_off = int32_t(g.off[i]);
                u.part_off = g.part_off[i] + int64_t(sp) * s.n * g.kl;         // the kernel adds (row - col_off) * kl
                units.push_back(u);
            }
        }
    }
    // longest units first (stable: units of a slide stay together, so CTAs running side by side share B tiles)
    std::stable_sort(units.begin(), units.end(), [](const TcUnit& a, const TcUnit& b) { return (a.ct1 - a.ct0) > (b.ct1 - b.ct0); });
    memcpy(hs, units.data(), units.size() * sizeof(TcUnit));
    HG_CHECK_CUDA(cudaMemcpyAsync(d_units, hs, tab_bytes, cudaMemcpyHostToDevice, st));
    rc = stage_release(stage, st);
    if (rc != HG_OK) return rc;
    HG_CHECK_CUDA(cudaMemsetAsync(flags, 0, 32, st));
    float* rho_max = reinterpret_cast<float*>(flags) + 4;           // measured rounding residual of the group's rows (tc_eps)

    const bool dbg = getenv("HG_B200_GROUP_DEBUG") != nullptr;      // synchronise after every phase and say which one failed
    auto phase = [&](const char* what) {
        if (!dbg) return HG_OK;
        const cudaError_t e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { set_error("knn_group: %s failed: %s", what, cudaGetErrorString(e)); return int(HG_ERR_CUDA); }
        return int(HG_OK);
    };
    rc = phase("table upload");
    if (rc != HG_OK) return rc;
    rc = launch_split_group(d_slides, d_map, g.n_pad, d, vec, ops, norms, st, rho_max);
    if (rc != HG_OK) return rc;
    if ((rc = phase("operand split")) != HG_OK) return rc;
    float* norm_max = reinterpret_cast<float*>(flags) + 1;
    rc = launch_colparams_group(norms, pos_dim, g.n_pad, metric, colp, pos4, norm_max, d_slides, d_map, st);
    if (rc != HG_OK) return rc;
    if ((rc = phase("column parameters")) != HG_OK) return rc;

    TcParams tp{};
    tp.d_pad = int(g.d_pad);
    tp.colp = colp; tp.pos4 = pos4;
    tp.norms_a = norms; tp.prods = classic_prods(GB_MAX_N); tp.rho_max = rho_max;      // every slide of a group is at most GB_MAX_N rows
    tp.gscale = (has_pos && lambda_h > 0.f) ? lambda_g / lambda_h : 0.f;
    tp.exclude_self = exclude_self;
    tp.part_val = reinterpret_cast<float*>(ws + g.o_pval);
    tp.part_idx = reinterpret_cast<int32_t*>(ws + g.o_pidx);
    prof_begin(st);
    rc = tc_launch_units(ops, g.n_pad, k, d_units, int(units.size()), tp, st);
    prof_end(st);
    if (rc != HG_OK) return rc;
    if ((rc = phase("fused kernel")) != HG_OK) return rc;
    if (dbg) {
        group_debug_kernel<<<32, 256, 0, st>>>(d_slides, d_map, num, g.n_pad, tp.part_val, tp.part_idx, g.kl);
        if ((rc = phase("list check")) != HG_OK) return rc;
    }

    RerankParams rp{};
    rp.d = d; rp.n = g.n_pad; rp.pos_dim = pos_dim;
    rp.sq_norms = norms; rp.norm_max = norm_max; rp.rho_max = rho_max;
    rp.metric = metric; rp.lambda_h = lambda_h; rp.lambda_g = has_pos ? lambda_g : 0.f;
    rp.part_val = tp.part_val; rp.part_idx = tp.part_idx;
    rp.splits = max_splits; rp.kl = g.kl; rp.n_rows = g.n_pad; rp.row_begin = 0;
    rp.k = k; rp.prods = tp.prods;
    const float* any_cos = nullptr;
    for (int i = 0; i < num; ++i) if (slides[i].out_cos != nullptr) any_cos = slides[i].out_cos;
    rp.out_cos = const_cast<float*>(any_cos);                       // only selects the kernel variant; per slide in the kernel
    rp.out_idx = slides[0].out_idx; rp.out_val = slides[0].out_val;
    rp.unc_cnt = flags + 2;
    rp.unc_rows = reinterpret_cast<int32_t*>(ws + g.o_unc_rows);
    rp.more_cnt = flags + 3; rp.more_rows = rp.unc_rows + g.n_pad;
    rp.unc_part = reinterpret_cast<uint2*>(ws + g.o_unc_part);
    rp.exclude_self = exclude_self;
    rp.g_slides = d_slides; rp.g_blk2slide = d_map;
    if (st_rerank != nullptr && st_rerank != st) {
        cudaEvent_t swept;
        HG_CHECK_CUDA(cudaEventCreateWithFlags(&swept, cudaEventDisableTiming));
        HG_CHECK_CUDA(cudaEventRecord(swept, st));
        HG_CHECK_CUDA(cudaStreamWaitEvent(st_rerank, swept, 0));
        HG_CHECK_CUDA(cudaEventDestroy(swept));                    // released once the wait has been satisfied
        rc = rerank_launch(rp, st_rerank);
    } else {
        rc = rerank_launch(rp, st);
    }
    if (rc == HG_OK && getenv("HG_B200_TC_DEBUG")) {             // tuning aid: rows sent to the exact fallback (synchronises)
        int h[4] = {0, 0, 0, 0};
        HG_CHECK_CUDA(cudaDeviceSynchronize());
        HG_CHECK_CUDA(cudaMemcpy(h, flags, 16, cudaMemcpyDeviceToHost));
        fprintf(stderr, "hg_b200 group of %d slides, %lld operand rows: %d rows uncertified -> exact fallback, %d rows multi-round\n",
                num, (long long)g.n_pad, h[2], h[3]);
    }
    return rc;
}

}  // namespace hg

// ---- symmetric pass over a slide whose ROWS are partitioned over several GPUs (C5) ---------------------------------
// hg_knn_feature on a row block runs the classic full pass: every rank computes all tiles of its rows although the
// tile (a, c) on the owner of a and the tile (c, a) on the owner of c hold the same products.  Here the checkerboard
// of the single-GPU symmetric mode (knn_tc.cu) is cut along the row blocks instead: the rank that owns row block a
// computes the tiles (a, c) of the checkerboard only (about half), keeps the row direction in its register lists and
// sends the transposed survivors -- (score, row i) for column j -- into bucket j ON THE GPU THAT OWNS ROW j, with a
// remote atomic + an 8-byte store over NVLink (peer memory mapped by the host side).  Per slide:
//   prepass  operand split of all n rows (every rank holds them after the all-gather), column parameters, own bucket
//            counters cleared, threshold pre-pass for the own rows -> thr[row_begin, row_end)
//   (host)   all-gather of thr: every rank needs the bound of every column; also the barrier behind which all
//            bucket counters are cleared
//   sweep    the symmetric kernel over the own rows; pushes into the peers' buckets
//   (host)   all-reduce(max) of the overflow flags: the barrier behind which all pushes have landed, and every rank
//            learns whether ANY bucket overflowed (then all of them take the classic pass, as on one GPU)
//   finish   classic pass guarded by the flag, merge of lists + own buckets, exact re-rank
// Row blocks are 256-aligned chunks of `chunk` rows (multigpu.row_range_aligned); results equal hg_knn_feature's.
namespace hg {

struct SymPartWs {
    size_t off_ops, off_norms, off_colp, off_pos4, off_pval, off_pidx, off_flags, off_unc_rows, off_unc_part, bytes;
    TcPlan plan, sym_plan, pre_plan;
    int stride;
    int64_t n_pad, d_pad, n_rows;
};

static int sympart_plan(const hg_knn_part* s, SymPartWs* w) {
    int rc = part_check(s);
    if (rc != HG_OK) return rc;
    HG_REQUIRE(s->row_begin % 256 == 0, HG_ERR_BAD_ARG, "hg_knn_sympart: row_begin must be a multiple of 256");
    w->n_pad = ops_rows(s->n);
    w->d_pad = ops_cols(s->d);
    w->n_rows = s->row_end - s->row_begin;
    w->stride = (s->n >= 32768) ? 16 : 8;
    HG_REQUIRE(s->n / w->stride >= 4 * (s->k + 8), HG_ERR_UNSUPPORTED, "hg_knn_sympart: slide too small for the symmetric pass");
    w->plan = tc_make_plan(s->n, w->n_rows, s->k, w->d_pad);
    w->sym_plan = tc_make_plan_sym_rows(s->n, w->n_rows, s->k, w->d_pad);
    w->pre_plan = tc_make_plan_prepass_rows(s->n, w->n_rows, s->k, w->d_pad, w->stride);
    const int max_splits = std::max(w->plan.splits, w->sym_plan.splits);
    size_t off = 0;
    auto take = [&](size_t b) { size_t o = off; off += (b + 1023) & ~size_t(1023); return o; };
    w->off_ops = take(hg_operand_bytes(s->n, s->d));
    w->off_norms = take(size_t(w->n_pad) * 4);
    w->off_colp = take(size_t(w->n_pad) * 8);
    w->off_pos4 = take(part_has_pos(s) ? size_t(w->n_pad) * 16 : 0);
    const size_t part = size_t(max_splits) * size_t(w->n_rows) * size_t(w->plan.kl);
    w->off_pval = take(part * 4);
    w->off_pidx = take(part * 4);
    w->off_flags = take(32);                    // [guard, largest squared norm, uncertified rows, multi-round rows, measured rho]
    w->off_unc_rows = take(size_t(w->n_rows) * 12);
    w->off_unc_part = take(rerank_unc_part_bytes(w->n_rows, s->k));
    w->bytes = off;
    return HG_OK;
}

static int sympart_ws(const hg_knn_part* s, void* workspace, size_t workspace_bytes, SymPartWs* w) {
    int rc = sympart_plan(s, w);
    if (rc != HG_OK) return rc;
    HG_REQUIRE(workspace && workspace_bytes >= w->bytes, HG_ERR_WORKSPACE, "hg_knn_sympart: workspace too small (%zu < %zu)",
               workspace_bytes, w->bytes);
    HG_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 1023) == 0, HG_ERR_BAD_ARG, "hg_knn_sympart: workspace must be 1024-byte aligned");
    return HG_OK;
}

static TcParams sympart_params(const hg_knn_part* s, const SymPartWs& w, char* ws) {
    const bool has_pos = part_has_pos(s);
    TcParams tp{};
    tp.n = s->n; tp.row_begin = s->row_begin; tp.row_end = s->row_end; tp.d_pad = int(w.d_pad);
    tp.colp = reinterpret_cast<float2*>(ws + w.off_colp);
    tp.pos4 = has_pos ? reinterpret_cast<float4*>(ws + w.off_pos4) : nullptr;
    tp.norms_a = reinterpret_cast<float*>(ws + w.off_norms); tp.prods = 3;
    tp.rho_max = reinterpret_cast<float*>(ws + w.off_flags) + 4;
    tp.gscale = (has_pos && s->lambda_h > 0.f) ? s->lambda_g / s->lambda_h : 0.f;
    tp.exclude_self = s->exclude_self;
    tp.part_val = reinterpret_cast<float*>(ws + w.off_pval);
    tp.part_idx = reinterpret_cast<int32_t*>(ws + w.off_pidx);
    return tp;
}
}  // namespace hg

extern "C" int hg_knn_sympart_workspace_bytes(const hg_knn_part* s, size_t* bytes) {
    HG_REQUIRE(bytes, HG_ERR_BAD_ARG, "hg_knn_sympart_workspace_bytes: null");
    SymPartWs w{};
    int rc = sympart_plan(s, &w);
    if (rc != HG_OK) return rc;
    *bytes = w.bytes;
    return HG_OK;
}

extern "C" int hg_knn_sympart_prepass(const hg_knn_part* s, const float* feats_all, float* thr_all, int32_t* own_bkt_cnt,
                                      int* own_flags, void* workspace, size_t workspace_bytes, void* stream) {
    SymPartWs w{};
    int rc = sympart_ws(s, workspace, workspace_bytes, &w);
    if (rc != HG_OK) return rc;
    HG_REQUIRE(feats_all && thr_all && own_bkt_cnt && own_flags, HG_ERR_BAD_ARG, "hg_knn_sympart_prepass: null buffer");
    char* ws = static_cast<char*>(workspace);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool has_pos = part_has_pos(s);
    op16_t* ops = reinterpret_cast<op16_t*>(ws + w.off_ops);
    float* norms = reinterpret_cast<float*>(ws + w.off_norms);
    float* norm_max = reinterpret_cast<float*>(ws + w.off_flags) + 1;
    HG_CHECK_CUDA(cudaMemsetAsync(ws + w.off_flags, 0, 32, st));
    HG_CHECK_CUDA(cudaMemsetAsync(own_flags, 0, 8, st));
    HG_CHECK_CUDA(cudaMemsetAsync(own_bkt_cnt, 0, size_t(w.n_rows) * 4, st));
    // every rank splits ALL rows itself, so the measured rounding residual (rho) is the same number on every rank
    rc = launch_split(feats_all, s->n, s->d, ops, norms, st, false, reinterpret_cast<float*>(ws + w.off_flags) + 4);
    if (rc != HG_OK) return rc;
    TcParams tp = sympart_params(s, w, ws);
    rc = launch_colparams(norms, has_pos ? s->pos : nullptr, s->pos_dim, s->n, w.n_pad, s->metric,
                          const_cast<float2*>(tp.colp), const_cast<float4*>(tp.pos4), st, nullptr, nullptr, norm_max);
    if (rc != HG_OK) return rc;
    TcParams pre = tp;
    pre.plan = w.pre_plan; pre.col_stride = w.stride; pre.exclude_self = 0;
    pre.thr_rank = s->k; pre.thr_margin_cos = (s->metric == HG_METRIC_COSINE) ? 1 : 0; pre.norm_max = norm_max;
    pre.thr_out = thr_all;                                    // indexed by the global row
    pre.prods = sympart_prods(); pre.pre_main_prods = sympart_prods();
    prof_begin(st);
    rc = tc_launch(ops, w.n_pad, pre, st);
    prof_end(st);
    return rc;
}

extern "C" int hg_knn_sympart_sweep(const hg_knn_part* s, const float* thr_all, int64_t chunk, int num_peers,
                                    void* const* peer_bkt_cnt, void* const* peer_bkt, int* own_flags,
                                    void* workspace, size_t workspace_bytes, void* stream) {
    SymPartWs w{};
    int rc = sympart_ws(s, workspace, workspace_bytes, &w);
    if (rc != HG_OK) return rc;
    HG_REQUIRE(thr_all && peer_bkt_cnt && peer_bkt && own_flags && num_peers >= 1 && num_peers <= 8 && chunk > 0 &&
               chunk % 256 == 0 && chunk * num_peers >= s->n && w.n_rows <= chunk, HG_ERR_BAD_ARG,
               "hg_knn_sympart_sweep: bad partition (1..8 peers, 256-aligned chunks covering n)");
    char* ws = static_cast<char*>(workspace);
    TcParams sp = sympart_params(s, w, ws);
    sp.plan = w.sym_plan; sp.prods = sympart_prods();
    sp.thr = thr_all; sp.bkt_cap = bucket_cap(); sp.ovf = own_flags;
    sp.sblk_base = int(s->row_begin / 256);
    sp.peer_chunk = chunk;
    for (int r = 0; r < num_peers; ++r) {
        HG_REQUIRE(peer_bkt_cnt[r] && peer_bkt[r], HG_ERR_BAD_ARG, "hg_knn_sympart_sweep: null peer buffer");
        sp.peer_cnt[r] = static_cast<int32_t*>(peer_bkt_cnt[r]);
        sp.peer_bkt[r] = static_cast<uint2*>(peer_bkt[r]);
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    prof_begin(st);
    rc = tc_launch(reinterpret_cast<op16_t*>(ws + w.off_ops), w.n_pad, sp, st);
    prof_end(st);
    return rc;
}

extern "C" int hg_knn_sympart_finish(const hg_knn_part* s, const float* feats_all, const int* any_overflow,
                                     const int32_t* own_bkt_cnt, const void* own_bkt,
                                     int32_t* out_idx, float* out_val, float* out_cos,
                                     void* workspace, size_t workspace_bytes, void* stream) {
    SymPartWs w{};
    int rc = sympart_ws(s, workspace, workspace_bytes, &w);
    if (rc != HG_OK) return rc;
    HG_REQUIRE(feats_all && any_overflow && own_bkt_cnt && own_bkt && out_idx && out_val, HG_ERR_BAD_ARG,
               "hg_knn_sympart_finish: null buffer");
    char* ws = static_cast<char*>(workspace);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool has_pos = part_has_pos(s);
    // the classic pass over the own rows, which exits at once unless a bucket overflowed somewhere
    TcParams tp = sympart_params(s, w, ws);
    tp.plan = w.plan;
    tp.guard = any_overflow;
    prof_begin(st);
    rc = tc_launch(reinterpret_cast<op16_t*>(ws + w.off_ops), w.n_pad, tp, st);
    prof_end(st);
    if (rc != HG_OK) return rc;
    RerankParams rp{};
    rp.feats = feats_all; rp.n = s->n; rp.d = s->d; rp.pos = has_pos ? s->pos : nullptr; rp.pos_dim = s->pos_dim;
    rp.sq_norms = reinterpret_cast<float*>(ws + w.off_norms);
    rp.norm_max = reinterpret_cast<float*>(ws + w.off_flags) + 1;
    rp.rho_max = reinterpret_cast<float*>(ws + w.off_flags) + 4;
    rp.metric = s->metric; rp.lambda_h = s->lambda_h; rp.lambda_g = has_pos ? s->lambda_g : 0.f;
    rp.part_val = tp.part_val; rp.part_idx = tp.part_idx;
    rp.splits = w.plan.splits; rp.kl = w.plan.kl; rp.n_rows = w.n_rows; rp.row_begin = s->row_begin;
    rp.cols_per_split = int64_t(w.plan.ct_per_split) * TC_BN;
    rp.k = s->k; rp.out_idx = out_idx; rp.out_val = out_val; rp.out_cos = out_cos;
    rp.prods = 3; rp.sym_prods = sympart_prods();
    rp.sym_plan = w.sym_plan; rp.bkt_cnt = own_bkt_cnt; rp.bkt = static_cast<const uint2*>(own_bkt); rp.bkt_cap = bucket_cap();
    rp.ovf = any_overflow; rp.bkt_row0 = s->row_begin;
    rp.unc_cnt = reinterpret_cast<int32_t*>(ws + w.off_flags) + 2;
    rp.unc_rows = reinterpret_cast<int32_t*>(ws + w.off_unc_rows);
    rp.more_cnt = rp.unc_cnt + 1; rp.more_rows = rp.unc_rows + w.n_rows;
    rp.unc_part = reinterpret_cast<uint2*>(ws + w.off_unc_part);
    rp.exclude_self = s->exclude_self;
    rc = rerank_launch(rp, st);
    if (rc == HG_OK && getenv("HG_B200_TC_DEBUG")) {             // tuning aid: rows sent to the exact fallback (synchronises)
        int h[2] = {0, 0};
        HG_CHECK_CUDA(cudaStreamSynchronize(st));
        HG_CHECK_CUDA(cudaMemcpy(h, rp.unc_cnt, 8, cudaMemcpyDeviceToHost));
        fprintf(stderr, "hg_b200 sympart rows [%lld, %lld): splits=%d, %d rows uncertified -> exact fallback, %d rows multi-round\n",
                (long long)s->row_begin, (long long)s->row_end, w.sym_plan.splits, h[0], h[1]);
    }
    return rc;
}
