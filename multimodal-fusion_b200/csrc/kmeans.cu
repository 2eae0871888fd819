// Row (f)1 of the scope table ("next"): the KMeans stages of the pipeline on the GPU.
//
// Reference: sklearn.cluster.KMeans(n_clusters, random_state=42, n_init=10) at
// preprocess_hypergraph.py:150-151 (super-patches), :299-300 (grouping), :391-392 (hyperedges).
// Lloyd iterations are the same distance contraction as the kNN path with k = 1:
//   assign   : argmin_c |x - mu_c|^2         -> the FP32 SIMT tile kernel in top-1 mode
//   update   : mu_c = mean of its members    -> shared-memory privatised segmented sum
//   kmeans++ : d^2 of every row to a handful of candidate rows (greedy seeding)
// Parity with scikit-learn can only be statistical (different RNG stream, different
// summation order): tests compare the inertia, not the labels.
#include "sim_simt.cuh"

namespace hg {

constexpr int KM_ROWS = 2048;     // rows per CTA in the update kernel
constexpr int KM_DS = 32;         // feature columns per CTA (one per lane)

// sums[c][d] += x[i][d] for label[i] == c.  Each CTA owns KM_ROWS rows x 32 columns, accumulates
// per-cluster partials in shared memory (float, conflict-free across lanes), then flushes the
// non-empty ones with double atomics so the cross-CTA order cannot move a centroid visibly.
__global__ void __launch_bounds__(256)
kmeans_accumulate_kernel(const float* __restrict__ x, int64_t n, int64_t d, const int32_t* __restrict__ labels,
                         int c, double* __restrict__ sums, unsigned int* __restrict__ counts) {
    extern __shared__ float part[];                      // [c][32]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int t = threadIdx.x; t < c * KM_DS; t += blockDim.x) part[t] = 0.f;
    __syncthreads();
    const int64_t r0 = int64_t(blockIdx.x) * KM_ROWS;
    const int64_t r1 = min64(n, r0 + KM_ROWS);
    const int64_t col = int64_t(blockIdx.y) * KM_DS + lane;
    for (int64_t r = r0 + warp; r < r1; r += 8) {
        const int lb = labels[r];
        if (lb < 0 || lb >= c) continue;
        if (col < d) atomicAdd(&part[lb * KM_DS + lane], __ldg(x + r * d + col));
        if (blockIdx.y == 0 && lane == 0) atomicAdd(&counts[lb], 1u);
    }
    __syncthreads();
    for (int t = threadIdx.x; t < c * KM_DS; t += blockDim.x) {
        const int cc = t / KM_DS, l = t % KM_DS;
        const int64_t dc = int64_t(blockIdx.y) * KM_DS + l;
        const float v = part[t];
        if (v != 0.f && dc < d) atomicAdd(&sums[int64_t(cc) * d + dc], double(v));
    }
}

// centroid <- sums / count (empty clusters keep their centre); shift2 += |new - old|^2
__global__ void kmeans_finalize_kernel(const double* __restrict__ sums, const unsigned int* __restrict__ counts,
                                       int c, int64_t d, float* __restrict__ centroids, double* __restrict__ shift2) {
    const int cc = blockIdx.x;
    const unsigned int cnt = counts[cc];
    double local = 0.0;
    if (cnt > 0) {
        for (int64_t j = threadIdx.x; j < d; j += blockDim.x) {
            const float nv = float(sums[int64_t(cc) * d + j] / double(cnt));
            const float ov = centroids[int64_t(cc) * d + j];
            local += double(nv - ov) * double(nv - ov);
            centroids[int64_t(cc) * d + j] = nv;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0 && local != 0.0) atomicAdd(shift2, local);
}

__global__ void kmeans_inertia_kernel(const float* __restrict__ d2, int64_t n, double* __restrict__ out) {
    double s = 0.0;
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x)
        s += double(d2[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, s);
}

// out[t][i] = |x_i - x_{cand[t]}|^2 for up to 16 candidate rows, one warp per row, direct difference
__global__ void __launch_bounds__(256)
kmeans_candidate_d2_kernel(const float* __restrict__ x, int64_t n, int64_t d, const int64_t* __restrict__ cand,
                           int t, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    if (r >= n) return;
    float acc[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) acc[q] = 0.f;
    for (int64_t j = lane; j < d; j += 32) {
        const float v = __ldg(x + r * d + j);
#pragma unroll
        for (int q = 0; q < 16; ++q) {
            if (q < t) {
                const float df = v - __ldg(x + cand[q] * d + j);
                acc[q] = fmaf(df, df, acc[q]);
            }
        }
    }
#pragma unroll
    for (int q = 0; q < 16; ++q) {
        if (q < t) {
            float a = acc[q];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
            if (lane == 0) out[int64_t(q) * n + r] = a;
        }
    }
}

}  // namespace hg

using namespace hg;

// labels[i] = argmin_c |x_i - centroids_c|^2 (ties -> lowest c), d2[i] = that squared distance
extern "C" int hg_kmeans_assign(const float* x, int64_t n, int64_t d, const float* centroids, int c,
                                int32_t* labels, float* d2, void* stream) {
    HG_REQUIRE(x && centroids && labels && d2 && n > 0 && d > 0 && c > 0, HG_ERR_BAD_ARG, "hg_kmeans_assign: bad argument");
    SimArgs a{};
    a.a = x; a.b = centroids; a.m = n; a.n = c; a.d = d; a.metric = HG_METRIC_EUCLIDEAN; a.k = 1;
    a.row_begin = 0; a.row_end = n; a.exclude_self = 0; a.out_idx = labels; a.out_val = d2;
    return launch_topk_fp32(a, static_cast<cudaStream_t>(stream));
}

extern "C" int hg_kmeans_workspace_bytes(int64_t d, int c, size_t* bytes) {
    HG_REQUIRE(bytes && d > 0 && c > 0, HG_ERR_BAD_ARG, "hg_kmeans_workspace_bytes: bad argument");
    *bytes = ((size_t(c) * size_t(d) * 8 + 255) & ~size_t(255)) + ((size_t(c) * 4 + 255) & ~size_t(255)) + 256;
    return HG_OK;
}

// One Lloyd update: centroids <- per-cluster means of x under `labels`; *shift2 (device double)
// <- sum of squared centre movements; counts_out (device, c entries, may be NULL) <- cluster sizes.
extern "C" int hg_kmeans_update(const float* x, int64_t n, int64_t d, const int32_t* labels, int c,
                                float* centroids, double* shift2, uint32_t* counts_out,
                                void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(x && labels && centroids && shift2 && n > 0 && d > 0 && c > 0, HG_ERR_BAD_ARG, "hg_kmeans_update: bad argument");
    HG_REQUIRE(size_t(c) * KM_DS * 4 <= 200 * 1024, HG_ERR_UNSUPPORTED, "hg_kmeans_update: at most 1600 clusters");
    size_t need = 0;
    hg_kmeans_workspace_bytes(d, c, &need);
    HG_REQUIRE(workspace && workspace_bytes >= need, HG_ERR_WORKSPACE, "hg_kmeans_update: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    Carver cv(workspace);
    double* sums = cv.take<double>(size_t(c) * d);
    unsigned int* counts = cv.take<unsigned int>(c);
    HG_CHECK_CUDA(cudaMemsetAsync(sums, 0, size_t(c) * d * 8, st));
    HG_CHECK_CUDA(cudaMemsetAsync(counts, 0, size_t(c) * 4, st));
    HG_CHECK_CUDA(cudaMemsetAsync(shift2, 0, 8, st));
    const size_t smem = size_t(c) * KM_DS * 4;
    if (smem > 48 * 1024)
        HG_CHECK_CUDA(cudaFuncSetAttribute(kmeans_accumulate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    dim3 grid((unsigned)ceil_div(n, KM_ROWS), (unsigned)ceil_div(d, KM_DS));
    kmeans_accumulate_kernel<<<grid, 256, smem, st>>>(x, n, d, labels, c, sums, counts);
    kmeans_finalize_kernel<<<c, 128, 0, st>>>(sums, counts, c, d, centroids, shift2);
    HG_CHECK_LAUNCH();
    count_launches(2);
    if (counts_out) HG_CHECK_CUDA(cudaMemcpyAsync(counts_out, counts, size_t(c) * 4, cudaMemcpyDeviceToDevice, st));
    return HG_OK;
}

extern "C" int hg_kmeans_inertia(const float* d2, int64_t n, double* inertia, void* stream) {
    HG_REQUIRE(d2 && inertia && n > 0, HG_ERR_BAD_ARG, "hg_kmeans_inertia: bad argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    HG_CHECK_CUDA(cudaMemsetAsync(inertia, 0, 8, st));
    kmeans_inertia_kernel<<<int(min64(ceil_div(n, 1024), 148 * 4)), 256, 0, st>>>(d2, n, inertia);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

// kmeans++ helper: out[t][i] = |x_i - x_{cand[t]}|^2, t <= 16 candidate row ids (device int64)
extern "C" int hg_kmeans_candidate_d2(const float* x, int64_t n, int64_t d, const int64_t* cand, int t,
                                      float* out, void* stream) {
    HG_REQUIRE(x && cand && out && n > 0 && d > 0 && t >= 1 && t <= 16, HG_ERR_BAD_ARG,
               "hg_kmeans_candidate_d2: bad argument (1 <= t <= 16)");
    kmeans_candidate_d2_kernel<<<(unsigned)ceil_div(n * 32, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        x, n, d, cand, t, out);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}
