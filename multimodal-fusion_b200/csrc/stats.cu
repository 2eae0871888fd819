// A9: mean / unbiased std / min / max / lower median of a float buffer.
//
// Reference: preprocess_hypergraph.py:190-196 (`K_wsi.mean/std/min/max/median`),
// :259-265 (WSI x TMA matrix) and similarity_kernel.py:183-186 (lower median of
// the off-diagonal).  torch.median returns the LOWER middle element for an even
// count; std is the unbiased estimator.
//
// One streaming pass for the moments (double accumulators), then a 4-pass MSB
// radix select on the order-preserving uint32 image of the floats for the median:
// 5 reads of the buffer in total, no sort, no copy.
#include "common.cuh"
#include "knn_tc.cuh"
#include <stdlib.h>

namespace hg {

struct StatsState {
    double sum, sumsq;
    unsigned int min_key, max_key;
    unsigned int prefix;             // selected high bytes so far
    unsigned int pad;
    unsigned long long rank;         // rank still to descend inside the selected bucket
    unsigned long long count;
    unsigned int hist[256];
};

__device__ __forceinline__ unsigned int ordered_key(float f) {
    const unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key_to_float(unsigned int k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__device__ __forceinline__ bool stats_live(int64_t i, int64_t diag_n) {
    return diag_n <= 0 || (i % (diag_n + 1)) != 0;
}

__global__ void stats_init_kernel(StatsState* s, unsigned long long count) {
    if (threadIdx.x == 0) {
        s->sum = 0.0; s->sumsq = 0.0; s->min_key = 0xffffffffu; s->max_key = 0u;
        s->prefix = 0u; s->rank = (count - 1) / 2; s->count = count;
    }
    s->hist[threadIdx.x] = 0u;
}

__global__ void __launch_bounds__(256)
stats_moments_kernel(const float* __restrict__ v, int64_t len, int64_t diag_n, StatsState* s) {
    double sum = 0.0, sq = 0.0;
    unsigned int mn = 0xffffffffu, mx = 0u;
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < len; i += int64_t(gridDim.x) * blockDim.x) {
        if (!stats_live(i, diag_n)) continue;
        const float x = v[i];
        sum += double(x); sq += double(x) * double(x);
        const unsigned int k = ordered_key(x);
        mn = min(mn, k); mx = max(mx, k);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        sq += __shfl_xor_sync(0xffffffffu, sq, o);
        mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&s->sum, sum); atomicAdd(&s->sumsq, sq);
        atomicMin(&s->min_key, mn); atomicMax(&s->max_key, mx);
    }
}

// histogram of byte `byte_idx` (3 = most significant) over keys matching the prefix above it
__global__ void __launch_bounds__(256)
stats_hist_kernel(const float* __restrict__ v, int64_t len, int64_t diag_n, int byte_idx, StatsState* s) {
    __shared__ unsigned int h[256];
    h[threadIdx.x] = 0u;
    __syncthreads();
    const unsigned int prefix = s->prefix;
    const int hi_shift = (byte_idx + 1) * 8;
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < len; i += int64_t(gridDim.x) * blockDim.x) {
        if (!stats_live(i, diag_n)) continue;
        const unsigned int k = ordered_key(v[i]);
        if (byte_idx == 3 || (k >> hi_shift) == (prefix >> hi_shift)) atomicAdd(&h[(k >> (byte_idx * 8)) & 255u], 1u);
    }
    __syncthreads();
    if (h[threadIdx.x]) atomicAdd(&s->hist[threadIdx.x], h[threadIdx.x]);
}

__global__ void stats_select_kernel(int byte_idx, StatsState* s) {
    __shared__ unsigned int h[256];
    h[threadIdx.x] = s->hist[threadIdx.x];
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long r = s->rank;
        int b = 0;
        for (; b < 255; ++b) {
            if (r < h[b]) break;
            r -= h[b];
        }
        s->rank = r;
        s->prefix |= (unsigned int)b << (byte_idx * 8);
    }
    __syncthreads();
    s->hist[threadIdx.x] = 0u;
}

__global__ void stats_final_kernel(const StatsState* s, double* out) {
    if (threadIdx.x != 0) return;
    const double n = double(s->count);
    const double mean = s->sum / n;
    double var = (s->sumsq - s->sum * s->sum / n) / (n - 1.0);
    if (var < 0.0) var = 0.0;
    out[0] = mean;
    out[1] = (s->count > 1) ? sqrt(var) : nan("");
    out[2] = double(key_to_float(s->min_key));
    out[3] = double(key_to_float(s->max_key));
    out[4] = double(key_to_float(s->prefix));
}

}  // namespace hg

using namespace hg;

extern "C" int hg_stats_workspace_bytes(int64_t count, size_t* bytes) {
    HG_REQUIRE(bytes && count > 0, HG_ERR_BAD_ARG, "hg_stats_workspace_bytes: bad argument");
    *bytes = (sizeof(StatsState) + 255) & ~size_t(255);
    return HG_OK;
}

extern "C" int hg_similarity_stats(const float* values, int64_t count, int64_t exclude_diag_n,
                                   double* stats_out, void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(values && count > 0 && stats_out, HG_ERR_BAD_ARG, "hg_similarity_stats: bad argument");
    HG_REQUIRE(exclude_diag_n <= 0 || exclude_diag_n * exclude_diag_n == count, HG_ERR_BAD_ARG,
               "hg_similarity_stats: exclude_diag_n^2 must equal count");
    HG_REQUIRE(workspace && workspace_bytes >= sizeof(StatsState), HG_ERR_WORKSPACE, "hg_similarity_stats: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    StatsState* s = static_cast<StatsState*>(workspace);
    const unsigned long long live = exclude_diag_n > 0 ? (unsigned long long)(count - exclude_diag_n) : (unsigned long long)count;
    HG_REQUIRE(live > 0, HG_ERR_BAD_ARG, "hg_similarity_stats: nothing left after excluding the diagonal");
    const int blocks = int(min64(ceil_div(count, 256 * 8), 148 * 8));
    stats_init_kernel<<<1, 256, 0, st>>>(s, live);
    stats_moments_kernel<<<blocks, 256, 0, st>>>(values, count, exclude_diag_n, s);
    for (int b = 3; b >= 0; --b) {
        stats_hist_kernel<<<blocks, 256, 0, st>>>(values, count, exclude_diag_n, b, s);
        stats_select_kernel<<<1, 256, 0, st>>>(b, s);
    }
    stats_final_kernel<<<1, 32, 0, st>>>(s, stats_out);
    HG_CHECK_LAUNCH();
    count_launches(11);
    return HG_OK;
}


// ---- A9 for slides whose N x N kernel does not fit: streaming statistics -----------------------------
// K = K_h * K_g is produced chunk by chunk (row ranges) by the tensor-core kernel in dense-output
// mode into a bounded scratch buffer and folded into the same moment / radix-select state, one
// sweep over the whole matrix per select pass (4 sweeps; the first also takes the moments).  The
// N x N matrix never exists in full; HBM traffic per sweep is one write + one read of the chunk.
namespace hg {
struct StreamWs { size_t off_ops, off_norms, off_colp, off_pos4, off_state, off_chunk, bytes; int64_t chunk_rows; };

static StreamWs plan_stream(int64_t n, int64_t d, bool has_pos) {
    StreamWs w{};
    const int64_t n_pad = ops_rows(n), d_pad = ops_cols(d);
    size_t off = 0;
    auto take = [&](size_t b) { size_t o = off; off += (b + 1023) & ~size_t(1023); return o; };
    w.off_ops = take(hg_operand_bytes(n, d));
    w.off_norms = take(size_t(n_pad) * 4);
    w.off_colp = take(size_t(n_pad) * 8);
    w.off_pos4 = take(has_pos ? size_t(n_pad) * 16 : 0);
    w.off_state = take(sizeof(StatsState));
    int64_t budget = int64_t(2) << 30;                                   // <= 2 GiB of K at a time
    if (const char* mb = getenv("HG_B200_STATS_CHUNK_MB")) budget = (int64_t(atoi(mb)) > 0 ? int64_t(atoi(mb)) : 1) << 20;
    int64_t rows = (budget / (n * 4)) / TC_BM * TC_BM;
    if (rows < TC_BM) rows = TC_BM;
    if (rows > round_up(n, TC_BM)) rows = round_up(n, TC_BM);
    w.chunk_rows = rows;
    w.off_chunk = take(size_t(rows) * size_t(n) * 4);
    w.bytes = off;
    return w;
}
}  // namespace hg

extern "C" int hg_similarity_stats_streaming_workspace_bytes(int64_t n, int64_t d, size_t* bytes) {
    HG_REQUIRE(bytes && n > 0 && d > 0, HG_ERR_BAD_ARG, "hg_similarity_stats_streaming_workspace_bytes: bad argument");
    *bytes = plan_stream(n, d, true).bytes;
    return HG_OK;
}

extern "C" int hg_similarity_stats_streaming(const float* feats, int64_t n, int64_t d, const float* pos, int pos_dim,
                                             float lambda_h, float lambda_g, double* stats_out,
                                             void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(feats && stats_out && n > 0 && d > 0, HG_ERR_BAD_ARG, "hg_similarity_stats_streaming: bad argument");
    HG_REQUIRE(lambda_h > 0.f, HG_ERR_UNSUPPORTED, "hg_similarity_stats_streaming: lambda_h must be > 0");
    if (lambda_g == 0.f) pos = nullptr;
    HG_REQUIRE(pos == nullptr || (pos_dim >= 1 && pos_dim <= 3), HG_ERR_UNSUPPORTED, "pos_dim must be 1..3");
    const StreamWs w = plan_stream(n, d, pos != nullptr);
    HG_REQUIRE(workspace && workspace_bytes >= w.bytes, HG_ERR_WORKSPACE,
               "hg_similarity_stats_streaming: workspace too small (%zu < %zu)", workspace_bytes, w.bytes);
    HG_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 1023) == 0, HG_ERR_BAD_ARG, "workspace must be 1024-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* ws = static_cast<char*>(workspace);
    op16_t* ops = reinterpret_cast<op16_t*>(ws + w.off_ops);
    float* norms = reinterpret_cast<float*>(ws + w.off_norms);
    float2* colp = reinterpret_cast<float2*>(ws + w.off_colp);
    float4* pos4 = pos ? reinterpret_cast<float4*>(ws + w.off_pos4) : nullptr;
    StatsState* state = reinterpret_cast<StatsState*>(ws + w.off_state);
    float* chunk = reinterpret_cast<float*>(ws + w.off_chunk);
    const int64_t n_pad = ops_rows(n), d_pad = ops_cols(d);

    int rc = launch_split(feats, n, d, ops, norms, st, true);      // dense-output operands: 8-bit hi halves
    if (rc != HG_OK) return rc;
    rc = launch_colparams(norms, pos, pos_dim, n, n_pad, HG_METRIC_GAUSSIAN, colp, pos4, st);
    if (rc != HG_OK) return rc;
    stats_init_kernel<<<1, 256, 0, st>>>(state, (unsigned long long)n * (unsigned long long)n);
    count_launches(1);
    for (int pass = 0; pass < 4; ++pass) {
        const int byte_idx = 3 - pass;
        for (int64_t r0 = 0; r0 < n; r0 += w.chunk_rows) {
            const int64_t r1 = min64(n, r0 + w.chunk_rows);
            TcParams tp{};
            tp.n = n; tp.row_begin = r0; tp.row_end = r1; tp.d_pad = int(d_pad);
            tp.plan = tc_make_plan(n, r1 - r0, 1, d_pad, false);
            tp.plan.kl = 0;                                        // dense-output mode
            tp.colp = colp; tp.pos4 = pos4; tp.norms_a = norms; tp.prods = 3;
            tp.gscale = pos ? lambda_g / lambda_h : 0.f;
            tp.exclude_self = 0;
            tp.dense_out = chunk; tp.dense_ld = n; tp.lambda_h = lambda_h; tp.dense_self = 1;
            rc = tc_launch(ops, n_pad, tp, st);
            if (rc != HG_OK) return rc;
            const int64_t cnt = (r1 - r0) * n;
            const int blocks = int(min64(ceil_div(cnt, 256 * 8), 148 * 8));
            if (pass == 0) stats_moments_kernel<<<blocks, 256, 0, st>>>(chunk, cnt, 0, state);
            stats_hist_kernel<<<blocks, 256, 0, st>>>(chunk, cnt, 0, byte_idx, state);
            count_launches(pass == 0 ? 2 : 1);
        }
        stats_select_kernel<<<1, 256, 0, st>>>(byte_idx, state);
        count_launches(1);
    }
    stats_final_kernel<<<1, 32, 0, st>>>(state, stats_out);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}


// ---- A1 / A2 dense output for large square inputs: the tensor-core kernel in dense-output mode ----------------
// K[i, j] = exp(-lambda_h (|x_i|^2 + |x_j|^2 - 2 x_i.x_j) - lambda_g |p_i - p_j|^2) for one matrix against itself
// (similarity_kernel.py:43-52, :122), the Gram product as three split-FP16 MMAs per k-step with FP32 accumulation.  13-20x the
// throughput of hg_similarity_dense from a few thousand rows up, but its entries are within ~1.5e-5 relative of
// float64, not 1e-5 (tests/test_gpu_parity.py::test_dense_tc_opt_in_accuracy): opt-in on the Python side.
extern "C" int hg_similarity_dense_tc_workspace_bytes(int64_t n, int64_t d, size_t* bytes) {
    HG_REQUIRE(bytes && n > 0 && d > 0, HG_ERR_BAD_ARG, "hg_similarity_dense_tc_workspace_bytes: bad argument");
    const int64_t n_pad = ops_rows(n);
    auto al = [](size_t b) { return (b + 1023) & ~size_t(1023); };
    *bytes = al(hg_operand_bytes(n, d)) + al(size_t(n_pad) * 4) + al(size_t(n_pad) * 8) + al(size_t(n_pad) * 16);
    return HG_OK;
}

extern "C" int hg_similarity_dense_tc(const float* feats, int64_t n, int64_t d, const float* pos, int pos_dim,
                                      float lambda_h, float lambda_g, float* out, int64_t ld_out,
                                      void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(feats && out && n > 0 && d > 0 && ld_out >= n, HG_ERR_BAD_ARG, "hg_similarity_dense_tc: bad argument");
    HG_REQUIRE(lambda_h > 0.f, HG_ERR_UNSUPPORTED, "hg_similarity_dense_tc: lambda_h must be > 0");
    if (lambda_g == 0.f) pos = nullptr;
    HG_REQUIRE(pos == nullptr || (pos_dim >= 1 && pos_dim <= 3), HG_ERR_UNSUPPORTED, "pos_dim must be 1..3");
    size_t need = 0;
    hg_similarity_dense_tc_workspace_bytes(n, d, &need);
    HG_REQUIRE(workspace && workspace_bytes >= need, HG_ERR_WORKSPACE, "hg_similarity_dense_tc: workspace too small");
    HG_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 1023) == 0, HG_ERR_BAD_ARG, "workspace must be 1024-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t n_pad = ops_rows(n), d_pad = ops_cols(d);
    char* ws = static_cast<char*>(workspace);
    auto al = [](size_t b) { return (b + 1023) & ~size_t(1023); };
    op16_t* ops = reinterpret_cast<op16_t*>(ws);
    ws += al(hg_operand_bytes(n, d));
    float* norms = reinterpret_cast<float*>(ws);
    ws += al(size_t(n_pad) * 4);
    float2* colp = reinterpret_cast<float2*>(ws);
    ws += al(size_t(n_pad) * 8);
    float4* pos4 = pos ? reinterpret_cast<float4*>(ws) : nullptr;
    int rc = launch_split(feats, n, d, ops, norms, st, true);      // dense-output operands: 8-bit hi halves
    if (rc != HG_OK) return rc;
    rc = launch_colparams(norms, pos, pos_dim, n, n_pad, HG_METRIC_GAUSSIAN, colp, pos4, st);
    if (rc != HG_OK) return rc;
    TcParams tp{};
    tp.n = n; tp.row_begin = 0; tp.row_end = n; tp.d_pad = int(d_pad);
    tp.plan = tc_make_plan(n, n, 1, d_pad, false);
    tp.plan.kl = 0;                                            // dense-output mode
    tp.colp = colp; tp.pos4 = pos4; tp.norms_a = norms; tp.prods = 3;
    tp.gscale = pos ? lambda_g / lambda_h : 0.f;
    tp.exclude_self = 0;
    tp.dense_out = out; tp.dense_ld = ld_out; tp.lambda_h = lambda_h; tp.dense_self = 1;
    return tc_launch(ops, n_pad, tp, st);
}
