// Super-patch aggregation on the device: segmented means of features and positions by
// cluster label, and the average intra-cluster similarity, without the per-cluster host loop.
//
// Reference semantics (preprocess_hypergraph.py):
//   :157-170   super_feature[c] = wsi_features[labels == c].mean(0), same for positions;
//              an empty cluster raises ValueError
//   :175-184   for each cluster with m > 1 members: mean of K_wsi[idx][:, idx] off the diagonal;
//              'avg_intra_cluster_similarity' = mean of those per-cluster means (0.0 if none)
//
// Plan: one stable radix sort of (label, row) puts the members of every cluster next to each other in
// ascending row order, so every sum below runs in a fixed order (deterministic, no float atomics).
//   agg_keys -> radix sort -> agg_bounds (cluster starts) -> agg_mean (row-sliced partial sums) ->
//   agg_rowsum_{dense|tiles} (per-row sum of same-cluster kernel values) -> agg_finish.
// The intra-cluster values come from the caller's dense N x N kernel when there is one, else they are
// recomputed in FP32 64 x 64 tiles over the block diagonal of the sorted order only (N^2 / S pairs):
// same expansion form and accumulation order as the dense kernel of sim_simt.cu.
#include "common.cuh"
#include "sort_scan.cuh"

namespace hg {

constexpr int AG_THREADS = 256;
constexpr int AG_MAX_SPLIT = 8;

__global__ void agg_keys_kernel(const int64_t* __restrict__ labels, int64_t n, int num_clusters,
                                uint64_t* __restrict__ keys, uint32_t* __restrict__ rows) {
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        const int64_t l = labels[i];
        keys[i] = (l >= 0 && l < num_clusters) ? uint64_t(l) : uint64_t(num_clusters);   // strays sort last
        rows[i] = uint32_t(i);
    }
}

// start[c] = first sorted position whose label is >= c, for c in [0, S]; start[S] = members of all clusters
__global__ void agg_bounds_kernel(const uint64_t* __restrict__ sorted_keys, int64_t n, int num_clusters,
                                  uint32_t* __restrict__ start) {
    for (int64_t p = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; p <= n; p += int64_t(gridDim.x) * blockDim.x) {
        const int64_t lab = (p < n) ? int64_t(sorted_keys[p]) : int64_t(num_clusters);
        const int64_t prev = (p > 0) ? int64_t(sorted_keys[p - 1]) : -1;
        for (int64_t c = prev + 1; c <= lab && c <= num_clusters; ++c) start[c] = uint32_t(p);
    }
}

int launch_cluster_bounds(const uint64_t* sorted_keys, int64_t n, int num_clusters, uint32_t* start, cudaStream_t st) {
    agg_bounds_kernel<<<int(min64(ceil_div(n + 1, 256), int64_t(kNumSMs) * 8)), 256, 0, st>>>(sorted_keys, n, num_clusters, start);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

// partial[z][c][col] = sum of feats[row][col] over slice z of cluster c's members (ascending row order).
// grid (S, ceil(D / 256), split); the column block is one coalesced 1 KB read per member row.
__global__ void __launch_bounds__(AG_THREADS)
agg_mean_kernel(const float* __restrict__ feats, int64_t d, const uint32_t* __restrict__ order,
                const uint32_t* __restrict__ start, int num_clusters, float* __restrict__ partial) {
    const int c = blockIdx.x, z = blockIdx.z, split = gridDim.z;
    const int64_t col = int64_t(blockIdx.y) * AG_THREADS + threadIdx.x;
    const uint32_t s = start[c], e = start[c + 1];
    const uint32_t per = (e - s + split - 1) / split;
    const uint32_t lo = min(e, s + z * per), hi = min(e, lo + per);
    float acc = 0.f;
    if (col < d) {
        uint32_t q = lo;
        for (; q + 8 <= hi; q += 8) {
            float v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) v[u] = __ldg(feats + int64_t(order[q + u]) * d + col);
#pragma unroll
            for (int u = 0; u < 8; ++u) acc += v[u];
        }
        for (; q < hi; ++q) acc += __ldg(feats + int64_t(order[q]) * d + col);
        partial[(int64_t(z) * num_clusters + c) * d + col] = acc;
    }
}

// out[c][col] = (sum over slices) / count; positions are summed here directly (2-3 columns)
__global__ void agg_mean_finish_kernel(const float* __restrict__ partial, int split, int64_t d, int num_clusters,
                                       const uint32_t* __restrict__ start, float* __restrict__ out) {
    const int c = blockIdx.x;
    const float m = float(start[c + 1] - start[c]);
    for (int64_t col = threadIdx.x; col < d; col += blockDim.x) {
        float acc = 0.f;
        for (int z = 0; z < split; ++z) acc += partial[(int64_t(z) * num_clusters + c) * d + col];
        out[int64_t(c) * d + col] = acc / m;                      // 0/0 = NaN marks an empty cluster
    }
}

__global__ void __launch_bounds__(AG_THREADS)
agg_pos_mean_kernel(const float* __restrict__ pos, int pos_dim, const uint32_t* __restrict__ order,
                    const uint32_t* __restrict__ start, float* __restrict__ out) {
    __shared__ float red[3][AG_THREADS];
    const int c = blockIdx.x;
    const uint32_t s = start[c], e = start[c + 1];
    float acc[3] = {0.f, 0.f, 0.f};
    for (uint32_t q = s + threadIdx.x; q < e; q += AG_THREADS)
        for (int a = 0; a < pos_dim; ++a) acc[a] += __ldg(pos + int64_t(order[q]) * pos_dim + a);
    for (int a = 0; a < 3; ++a) red[a][threadIdx.x] = acc[a];
    __syncthreads();
    for (int w = AG_THREADS / 2; w > 0; w >>= 1) {
        if (threadIdx.x < w)
            for (int a = 0; a < 3; ++a) red[a][threadIdx.x] += red[a][threadIdx.x + w];
        __syncthreads();
    }
    if (threadIdx.x < pos_dim) out[int64_t(c) * pos_dim + threadIdx.x] = red[threadIdx.x][0] / float(e - s);
}

// ---- per-row sums of same-cluster kernel values ----------------------------------------------------
// dense K given: one warp per sorted position, lanes stride over the members of its cluster
__global__ void __launch_bounds__(AG_THREADS)
agg_rowsum_dense_kernel(const float* __restrict__ K, int64_t ld, const uint64_t* __restrict__ sorted_keys,
                        const uint32_t* __restrict__ order, const uint32_t* __restrict__ start, int num_clusters,
                        double* __restrict__ rowsum) {
    const int lane = threadIdx.x & 31;
    const int64_t members = start[num_clusters];
    for (int64_t p = int64_t(blockIdx.x) * (AG_THREADS / 32) + (threadIdx.x >> 5); p < members;
         p += int64_t(gridDim.x) * (AG_THREADS / 32)) {
        const int c = int(sorted_keys[p]);
        const uint32_t s = start[c], e = start[c + 1];
        const float* krow = K + int64_t(order[p]) * ld;
        double acc = 0.0;
        for (uint32_t q = s + lane; q < e; q += 32)
            if (q != p) acc += double(__ldg(krow + order[q]));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) rowsum[p] = acc;
    }
}

// no dense K: FP32 tiles over the block diagonal of the sorted order.  CTA = 64 sorted rows; its column
// range is the union of the clusters those rows belong to.  4 x 4 micro-tile per thread, BK = 16.
constexpr int TB = 64, TK = 16, TPAD = 4;

template <bool VEC>
__device__ __forceinline__ void gather4(const float* __restrict__ base, int64_t d, int64_t row, int64_t c, float (&v)[4]) {
    if (VEC) {
        if (row >= 0 && c + 3 < d) {
            const float4 t = __ldg(reinterpret_cast<const float4*>(base + row * d + c));
            v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
            return;
        }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) v[q] = (row >= 0 && c + q < d) ? __ldg(base + row * d + c + q) : 0.f;
}

template <bool VEC>
__global__ void __launch_bounds__(256)
agg_rowsum_tiles_kernel(const float* __restrict__ feats, int64_t d, const float* __restrict__ pos, int pos_dim,
                        float lambda_h, float lambda_g, const uint64_t* __restrict__ sorted_keys,
                        const uint32_t* __restrict__ order, const uint32_t* __restrict__ start, int num_clusters,
                        double* __restrict__ rowsum) {
    __shared__ __align__(16) float As[TK][TB + TPAD];
    __shared__ __align__(16) float Bs[TK][TB + TPAD];
    __shared__ int lab_a[TB], lab_b[TB];
    __shared__ float pos_a[TB][3], pos_b[TB][3];
    const int64_t members = start[num_clusters];
    const int64_t row0 = int64_t(blockIdx.x) * TB;
    if (row0 >= members) return;
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int lr = tid >> 2, lk = (tid & 3) * 4;
    const int64_t last = min64(row0 + TB, members) - 1;
    const int64_t col_begin = start[int(sorted_keys[row0])];
    const int64_t col_end = start[int(sorted_keys[last]) + 1];

    const int64_t my_a = (row0 + lr < members) ? int64_t(order[row0 + lr]) : -1;
    if (tid < TB) {
        const int64_t p = row0 + tid;
        lab_a[tid] = p < members ? int(sorted_keys[p]) : -1;
        for (int a = 0; a < 3; ++a) pos_a[tid][a] = (p < members && pos != nullptr && a < pos_dim) ? __ldg(pos + int64_t(order[p]) * pos_dim + a) : 0.f;
    }
    float rsum[4] = {0.f, 0.f, 0.f, 0.f};
    for (int64_t col0 = col_begin; col0 < col_end; col0 += TB) {
        const int64_t my_b = (col0 + lr < members) ? int64_t(order[col0 + lr]) : -1;
        __syncthreads();                                          // previous tile's lab_b / pos_b readers are done
        if (tid < TB) {
            const int64_t p = col0 + tid;
            lab_b[tid] = p < members ? int(sorted_keys[p]) : -2;
            for (int a = 0; a < 3; ++a) pos_b[tid][a] = (p < members && pos != nullptr && a < pos_dim) ? __ldg(pos + int64_t(order[p]) * pos_dim + a) : 0.f;
        }
        float acc[4][4] = {};
        float na[4] = {}, nb[4] = {};
        for (int64_t k0 = 0; k0 < d; k0 += TK) {
            float va[4], vb[4];
            gather4<VEC>(feats, d, my_a, k0 + lk, va);
            gather4<VEC>(feats, d, my_b, k0 + lk, vb);
            __syncthreads();
#pragma unroll
            for (int q = 0; q < 4; ++q) { As[lk + q][lr] = va[q]; Bs[lk + q][lr] = vb[q]; }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < TK; ++kk) {
                const float4 a4 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
                const float4 b4 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
                const float av[4] = {a4.x, a4.y, a4.z, a4.w};
                const float bv[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    na[i] = fmaf(av[i], av[i], na[i]);
                    nb[i] = fmaf(bv[i], bv[i], nb[i]);
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int ra = ty * 4 + i;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int cb = tx * 4 + j;
                if (lab_a[ra] == lab_b[cb] && row0 + ra != col0 + cb) {
                    const float d2 = (na[i] + nb[j]) - 2.f * acc[i][j];
                    float d2g = 0.f;
                    for (int a = 0; a < pos_dim; ++a) {
                        const float df = pos_a[ra][a] - pos_b[cb][a];
                        d2g = fmaf(df, df, d2g);
                    }
                    rsum[i] += expf(-(lambda_h * d2 + lambda_g * d2g));
                }
            }
        }
    }
    // the 16 threads that share a row are one half-warp
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float v = rsum[i];
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        const int64_t p = row0 + ty * 4 + i;
        if (tx == 0 && p < members) rowsum[p] = double(v);
    }
}

// intra[c] = sum(rowsum over members) / (m (m - 1)) for m > 1 (else NaN), counts, first empty cluster,
// and the mean of the defined intra values.  One block; clusters are few.
__global__ void __launch_bounds__(AG_THREADS)
agg_finish_kernel(const double* __restrict__ rowsum, const uint32_t* __restrict__ start, int num_clusters,
                  int64_t* __restrict__ counts, float* __restrict__ intra, float* __restrict__ summary) {
    __shared__ double sm_sum[AG_THREADS / 32];
    __shared__ int sm_cnt[AG_THREADS / 32];
    __shared__ int sm_empty[AG_THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double tot = 0.0;
    int defined = 0, first_empty = num_clusters;
    for (int c = warp; c < num_clusters; c += AG_THREADS / 32) {
        const uint32_t s = start[c], e = start[c + 1];
        double acc = 0.0;
        if (rowsum != nullptr)
            for (uint32_t q = s + lane; q < e; q += 32) acc += rowsum[q];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        const double m = double(e - s);
        if (lane == 0) {
            counts[c] = int64_t(e - s);
            const float v = (e - s > 1) ? float(acc / (m * (m - 1.0))) : NAN;
            intra[c] = v;
            if (e - s > 1) { tot += double(v); ++defined; }
            if (e == s && c < first_empty) first_empty = c;
        }
    }
    if (lane == 0) { sm_sum[warp] = tot; sm_cnt[warp] = defined; sm_empty[warp] = first_empty; }
    __syncthreads();
    if (threadIdx.x == 0) {
        // per-warp totals interleave clusters; re-sum in cluster order so the result does not depend on AG_THREADS
        double t = 0.0;
        int dn = 0, fe = num_clusters;
        for (int w = 0; w < AG_THREADS / 32; ++w) { dn += sm_cnt[w]; fe = min(fe, sm_empty[w]); }
        for (int c = 0; c < num_clusters; ++c)
            if (start[c + 1] - start[c] > 1) t += double(intra[c]);
        summary[0] = dn ? float(t / double(dn)) : 0.f;            // reference :190
        summary[1] = float(dn);
        summary[2] = float(fe == num_clusters ? -1 : fe);         // first empty cluster (reference raises)
        summary[3] = float(int64_t(start[num_clusters]));         // rows with a label in range
    }
}

struct AggWs {
    uint64_t *keys, *keys_tmp;
    uint32_t *rows, *rows_tmp, *table, *start;
    float* partial;
    double* rowsum;
};

static int agg_split(int num_clusters, int64_t d) {
    const int64_t ctas = int64_t(num_clusters) * ceil_div(d, AG_THREADS);
    int s = int(ceil_div(int64_t(kNumSMs) * 4, ctas));
    return s < 1 ? 1 : (s > AG_MAX_SPLIT ? AG_MAX_SPLIT : s);
}

static size_t agg_carve(void* base, int64_t n, int64_t d, int num_clusters, AggWs* w) {
    Carver c(base);
    AggWs t;
    t.keys = c.take<uint64_t>(n);
    t.keys_tmp = c.take<uint64_t>(n);
    t.rows = c.take<uint32_t>(n);
    t.rows_tmp = c.take<uint32_t>(n);
    t.table = c.take<uint32_t>(sort_table_entries(n));
    t.start = c.take<uint32_t>(size_t(num_clusters) + 2);
    t.partial = c.take<float>(size_t(AG_MAX_SPLIT) * num_clusters * d);
    t.rowsum = c.take<double>(n);
    if (w) *w = t;
    return c.off;
}

}  // namespace hg

extern "C" int hg_aggregate_workspace_bytes(int64_t n, int64_t d, int num_clusters, size_t* bytes) {
    HG_REQUIRE(bytes && n > 0 && d > 0 && num_clusters > 0, HG_ERR_BAD_ARG, "hg_aggregate_workspace_bytes: bad argument");
    *bytes = hg::agg_carve(nullptr, n, d, num_clusters, nullptr);
    return HG_OK;
}

extern "C" int hg_aggregate_super_patches(const float* feats, int64_t n, int64_t d, const float* pos, int pos_dim,
                                          const int64_t* labels, int num_clusters, float lambda_h, float lambda_g,
                                          const float* dense_k, int64_t ld_k, int want_intra,
                                          float* super_feats, float* super_pos, int64_t* counts, float* intra,
                                          float* summary, int32_t* order_out,
                                          void* workspace, size_t workspace_bytes, void* stream) {
    using namespace hg;
    HG_REQUIRE(feats && labels && super_feats && counts && intra && summary && workspace, HG_ERR_BAD_ARG,
               "hg_aggregate_super_patches: null argument");
    HG_REQUIRE(n > 0 && n < (int64_t(1) << 30) && d > 0 && num_clusters > 0 && num_clusters <= (1 << 24), HG_ERR_BAD_ARG,
               "hg_aggregate_super_patches: bad n / d / num_clusters");
    HG_REQUIRE(pos == nullptr || (pos_dim >= 1 && pos_dim <= 3 && super_pos), HG_ERR_BAD_ARG, "pos_dim must be 1..3");
    HG_REQUIRE(dense_k == nullptr || ld_k >= n, HG_ERR_BAD_ARG, "hg_aggregate_super_patches: ld_k < n");
    AggWs w;
    HG_REQUIRE(agg_carve(workspace, n, d, num_clusters, &w) <= workspace_bytes, HG_ERR_WORKSPACE,
               "hg_aggregate_super_patches: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int blocks = int(min64(ceil_div(n + 1, 256), int64_t(kNumSMs) * 8));

    agg_keys_kernel<<<blocks, 256, 0, st>>>(labels, n, num_clusters, w.keys, w.rows);
    HG_CHECK_LAUNCH();
    count_launches(1);
    uint64_t* sk; uint32_t* order;
    int rc = radix_sort_u64(w.keys, w.keys_tmp, w.rows, w.rows_tmp, n, bits_for(uint64_t(num_clusters)), w.table, &sk,
                            &order, st);
    if (rc != HG_OK) return rc;
    agg_bounds_kernel<<<blocks, 256, 0, st>>>(sk, n, num_clusters, w.start);

    const int split = agg_split(num_clusters, d);
    agg_mean_kernel<<<dim3(num_clusters, (unsigned)ceil_div(d, AG_THREADS), split), AG_THREADS, 0, st>>>(
        feats, d, order, w.start, num_clusters, w.partial);
    agg_mean_finish_kernel<<<num_clusters, 256, 0, st>>>(w.partial, split, d, num_clusters, w.start, super_feats);
    HG_CHECK_LAUNCH();
    count_launches(3);
    if (pos != nullptr) {
        agg_pos_mean_kernel<<<num_clusters, AG_THREADS, 0, st>>>(pos, pos_dim, order, w.start, super_pos);
        count_launches(1);
    }
    if (want_intra) {
        if (dense_k != nullptr) {
            agg_rowsum_dense_kernel<<<int(min64(ceil_div(n, AG_THREADS / 32), int64_t(kNumSMs) * 16)), AG_THREADS, 0, st>>>(
                dense_k, ld_k, sk, order, w.start, num_clusters, w.rowsum);
        } else {
            const bool vec = d % 4 == 0 && (reinterpret_cast<uintptr_t>(feats) & 15) == 0;
            const unsigned tiles = (unsigned)ceil_div(n, TB);
            if (vec) agg_rowsum_tiles_kernel<true><<<tiles, 256, 0, st>>>(feats, d, pos, pos_dim, lambda_h, lambda_g, sk, order,
                                                                           w.start, num_clusters, w.rowsum);
            else     agg_rowsum_tiles_kernel<false><<<tiles, 256, 0, st>>>(feats, d, pos, pos_dim, lambda_h, lambda_g, sk, order,
                                                                            w.start, num_clusters, w.rowsum);
        }
        count_launches(1);
    }
    agg_finish_kernel<<<1, AG_THREADS, 0, st>>>(want_intra ? w.rowsum : nullptr, w.start, num_clusters, counts, intra, summary);
    HG_CHECK_LAUNCH();
    count_launches(1);
    if (order_out != nullptr)
        HG_CHECK_CUDA(cudaMemcpyAsync(order_out, order, size_t(n) * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
    return HG_OK;
}
