// A3: spatial-coordinate kNN over patch (x, y[, z]) positions.
//
// Semantics pinned by the reference's spatial kernel K_g = exp(-lambda_g |g_i-g_j|^2)
// (similarity_kernel.py:57-86): the top-k of a K_g row is the k smallest squared
// distances.  Distances are taken in direct-difference FP32 (exact for integer
// grid coordinates), self is excluded by index, ties resolve by (d2, index)
// ascending -- on a patch grid 4 neighbours tie at d=1 and 4 at sqrt(2), so the
// rule matters and the oracle uses the same one.
//
// One query per thread, candidate points staged through shared memory as packed
// float2 / SoA tiles read with broadcast LDS.128; the candidate range is split
// across blockIdx.y so small slides still fill 148 SMs, and a second kernel
// merges the per-split lists.
#include "common.cuh"
#include "topk_list.cuh"

namespace hg {

constexpr int SP_THREADS = 128;
constexpr int SP_TILE = 2048;      // candidate points per shared-memory tile

template <int PD, int KL>
__global__ void __launch_bounds__(SP_THREADS)
spatial_knn_kernel(const float* __restrict__ pos, int64_t n, int64_t row_begin, int64_t row_end,
                   int64_t cols_per_split, float* __restrict__ part_val, int32_t* __restrict__ part_idx) {
    __shared__ __align__(16) float tile[PD][SP_TILE];
    const int64_t row = row_begin + int64_t(blockIdx.x) * SP_THREADS + threadIdx.x;
    const bool live = row < row_end;
    float q[PD];
#pragma unroll
    for (int c = 0; c < PD; ++c) q[c] = live ? __ldg(pos + row * PD + c) : 0.f;
    float lv[KL];
    int li[KL];
#pragma unroll
    for (int t = 0; t < KL; ++t) { lv[t] = INFINITY; li[t] = -1; }

    const int64_t c_begin = int64_t(blockIdx.y) * cols_per_split;
    const int64_t c_end = min(n, c_begin + cols_per_split);
    for (int64_t c0 = c_begin; c0 < c_end; c0 += SP_TILE) {
        const int cnt = int(min64(SP_TILE, c_end - c0));
        __syncthreads();
        for (int t = threadIdx.x; t < cnt * PD; t += SP_THREADS) {
            const float v = __ldg(pos + c0 * PD + t);        // coalesced AoS read
            tile[t % PD][t / PD] = v;                         // SoA in shared memory
        }
        for (int t = cnt + threadIdx.x; t < round_up(cnt, 4); t += SP_THREADS)
#pragma unroll
            for (int c = 0; c < PD; ++c) tile[c][t] = 3.0e18f;   // far away: never selected
        __syncthreads();
        if (!live) continue;
        const int self = (row >= c0 && row < c0 + cnt) ? int(row - c0) : -1;
        for (int t = 0; t < cnt; t += 4) {
            float4 cx[PD];
#pragma unroll
            for (int c = 0; c < PD; ++c) cx[c] = *reinterpret_cast<const float4*>(&tile[c][t]);
            float d2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int c = 0; c < PD; ++c) {
                const float e0 = q[c] - cx[c].x, e1 = q[c] - cx[c].y, e2 = q[c] - cx[c].z, e3 = q[c] - cx[c].w;
                d2[0] = fmaf(e0, e0, d2[0]); d2[1] = fmaf(e1, e1, d2[1]);
                d2[2] = fmaf(e2, e2, d2[2]); d2[3] = fmaf(e3, e3, d2[3]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (d2[u] < lv[KL - 1] && t + u != self) list_insert<KL>(lv, li, d2[u], int(c0 + t + u));
            }
        }
    }
    if (live) {
        const int64_t o = (int64_t(blockIdx.y) * (row_end - row_begin) + (row - row_begin)) * KL;
#pragma unroll
        for (int t = 0; t < KL; ++t) { part_val[o + t] = lv[t]; part_idx[o + t] = li[t]; }
    }
}

// Merge `splits` ascending (val, idx) lists of length KL per row into the best k.
// Splits cover increasing index ranges, so on equal values the earlier split wins.
__global__ void merge_lists_kernel(const float* __restrict__ part_val, const int32_t* __restrict__ part_idx,
                                   int64_t n_rows, int splits, int kl, int k,
                                   float* __restrict__ out_val, int32_t* __restrict__ out_idx) {
    const int64_t r = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (r >= n_rows) return;
    int head[32];
    for (int s = 0; s < splits; ++s) head[s] = 0;
    for (int t = 0; t < k; ++t) {
        float best = INFINITY;
        int bs = -1;
        for (int s = 0; s < splits; ++s) {
            if (head[s] >= kl) continue;
            const int64_t o = (int64_t(s) * n_rows + r) * kl + head[s];
            const float v = part_val[o];
            if (part_idx[o] >= 0 && v < best) { best = v; bs = s; }
        }
        if (bs < 0) { out_val[r * k + t] = NAN; out_idx[r * k + t] = -1; continue; }
        const int64_t o = (int64_t(bs) * n_rows + r) * kl + head[bs];
        out_val[r * k + t] = best;
        out_idx[r * k + t] = part_idx[o];
        ++head[bs];
    }
}

template <int PD>
static int launch_spatial(const float* pos, int64_t n, int k, int64_t r0, int64_t r1, int splits,
                          int64_t cols_per_split, float* pv, int32_t* pi, int kl, cudaStream_t st) {
    dim3 grid((unsigned)ceil_div(r1 - r0, SP_THREADS), (unsigned)splits);
    if (kl == 8)       spatial_knn_kernel<PD, 8><<<grid, SP_THREADS, 0, st>>>(pos, n, r0, r1, cols_per_split, pv, pi);
    else if (kl == 16) spatial_knn_kernel<PD, 16><<<grid, SP_THREADS, 0, st>>>(pos, n, r0, r1, cols_per_split, pv, pi);
    else               spatial_knn_kernel<PD, 32><<<grid, SP_THREADS, 0, st>>>(pos, n, r0, r1, cols_per_split, pv, pi);
    HG_CHECK_LAUNCH();
    return HG_OK;
}

static int spatial_plan(int64_t n, int64_t n_rows, int k, int* kl, int* splits, int64_t* cols_per_split) {
    *kl = k <= 8 ? 8 : (k <= 16 ? 16 : 32);
    const int64_t row_blocks = ceil_div(n_rows, SP_THREADS);
    int s = int(ceil_div(4 * kNumSMs, row_blocks));          // aim at >= 4 CTAs per SM
    s = max(1, min(s, 32));
    s = int(min64(s, ceil_div(n, SP_TILE)));
    *cols_per_split = round_up(ceil_div(n, s), 4);
    *splits = int(ceil_div(n, *cols_per_split));
    return HG_OK;
}

}  // namespace hg

extern "C" int hg_knn_spatial_workspace_bytes(int64_t n, int k, int64_t row_begin, int64_t row_end, size_t* bytes) {
    HG_REQUIRE(bytes && n > 0 && k >= 1 && k <= 32 && row_begin >= 0 && row_end <= n && row_begin < row_end,
               HG_ERR_BAD_ARG, "hg_knn_spatial_workspace_bytes: bad argument (k must be 1..32)");
    int kl, splits; int64_t cps;
    hg::spatial_plan(n, row_end - row_begin, k, &kl, &splits, &cps);
    const size_t per = size_t(splits) * size_t(row_end - row_begin) * size_t(kl);
    *bytes = 2 * ((per * 4 + 255) & ~size_t(255));
    return HG_OK;
}

extern "C" int hg_knn_spatial(const float* pos, int64_t n, int pos_dim, int k,
                              int64_t row_begin, int64_t row_end,
                              int32_t* out_idx, float* out_d2,
                              void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(pos && out_idx && out_d2 && n > 0, HG_ERR_BAD_ARG, "hg_knn_spatial: null argument");
    HG_REQUIRE(pos_dim == 2 || pos_dim == 3, HG_ERR_UNSUPPORTED, "hg_knn_spatial: pos_dim must be 2 or 3");
    size_t need = 0;
    int rc = hg_knn_spatial_workspace_bytes(n, k, row_begin, row_end, &need);
    if (rc != HG_OK) return rc;
    HG_REQUIRE(workspace && workspace_bytes >= need, HG_ERR_WORKSPACE, "hg_knn_spatial: workspace too small (%zu < %zu)",
               workspace_bytes, need);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int kl, splits; int64_t cps;
    const int64_t n_rows = row_end - row_begin;
    hg::spatial_plan(n, n_rows, k, &kl, &splits, &cps);
    hg::Carver cv(workspace);
    float* pv = cv.take<float>(size_t(splits) * n_rows * kl);
    int32_t* pi = cv.take<int32_t>(size_t(splits) * n_rows * kl);
    rc = (pos_dim == 2) ? hg::launch_spatial<2>(pos, n, k, row_begin, row_end, splits, cps, pv, pi, kl, st)
                        : hg::launch_spatial<3>(pos, n, k, row_begin, row_end, splits, cps, pv, pi, kl, st);
    if (rc != HG_OK) return rc;
    hg::merge_lists_kernel<<<(unsigned)hg::ceil_div(n_rows, 128), 128, 0, st>>>(pv, pi, n_rows, splits, kl, k,
                                                                                out_d2, out_idx);
    HG_CHECK_LAUNCH();
    hg::count_launches(2);
    return HG_OK;
}
