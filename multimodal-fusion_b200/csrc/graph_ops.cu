// A6 cosine edge weights, CSR incidence, A8 dense-threshold edge builder.
#include "common.cuh"
#include "rowdot.cuh"
#include "sort_scan.cuh"

namespace hg {

// ---- A6: w_e = max(0, cos(f_i, f_j)) ---------------------------------------------------
// preprocess_hypergraph.py:413-422, F.cosine_similarity eps = 1e-8 (each norm clamped).
// One warp per edge, 128-bit loads; edges arrive sorted by (min, max) so the row
// of `min` is re-read from L1/L2 by consecutive warps.  Algorithmic bytes per edge:
// 16 (indices) + 4 (weight) + the two gathered rows (2 * 4 D, L2-resident when the
// slide fits the 126 MB L2).
__device__ __forceinline__ float warp_cosine(const float* __restrict__ feats, int64_t d, int64_t ia, int64_t ib,
                                             int lane, bool vec) {
    float dot, na, nb;
    warp_dot_norms(feats + ia * d, feats + ib * d, d, lane, vec, dot, na, nb);
    return fmaxf(0.f, cosine_from(dot, na, nb));
}

__global__ void __launch_bounds__(256)
edge_cosine_kernel(const float* __restrict__ feats, int64_t d, const int64_t* __restrict__ edge_index,
                   int64_t num_edges, float* __restrict__ w) {
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
    const bool vec = rows_vectorisable(feats, d);
    for (int64_t e = warp0; e < num_edges; e += nwarps) {
        const float c = warp_cosine(feats, d, edge_index[e], edge_index[num_edges + e], lane, vec);
        if (lane == 0) w[e] = c;
    }
}

// The same weights when the feature-kNN stage already produced the cosine of every list entry
// (rerank.cu, out_cos -- bit-identical arithmetic): an edge (a, b) is looked up in a's list, then
// in b's; only edges that come from elsewhere (spatial lists, cliques) gather their two rows.
// Per edge: 16 B of indices + up to 2 k list entries (L1/L2-resident, edges are sorted by a)
// instead of two feature rows.
__global__ void __launch_bounds__(256)
edge_weights_from_knn_kernel(const float* __restrict__ feats, int64_t d, const int32_t* __restrict__ knn_idx,
                             const float* __restrict__ knn_cos, int k, const int64_t* __restrict__ edge_index,
                             int64_t num_edges, const int64_t* __restrict__ num_edges_dev, float* __restrict__ w) {
    if (num_edges_dev != nullptr) num_edges = *num_edges_dev;          // stream-ordered build: count still on the device
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
    const bool vec = rows_vectorisable(feats, d);
    for (int64_t base = warp0 * 32; base < num_edges; base += nwarps * 32) {
        const int64_t e = base + lane;
        int64_t a = 0, b = 0;
        bool found = true;
        if (e < num_edges) {
            a = edge_index[e]; b = edge_index[num_edges + e];
            found = false;
            float c = 0.f;
            for (int s = 0; s < k; ++s)
                if (__ldg(knn_idx + a * k + s) == int32_t(b)) { c = __ldg(knn_cos + a * k + s); found = true; break; }
            if (!found)
                for (int s = 0; s < k; ++s)
                    if (__ldg(knn_idx + b * k + s) == int32_t(a)) { c = __ldg(knn_cos + b * k + s); found = true; break; }
            if (found) w[e] = fmaxf(0.f, c);
        }
        uint32_t pending = __ballot_sync(0xffffffffu, !found);
        while (pending) {
            const int l = __ffs(pending) - 1;
            pending &= pending - 1;
            const int64_t ia = __shfl_sync(0xffffffffu, a, l), ib = __shfl_sync(0xffffffffu, b, l);
            const float c = warp_cosine(feats, d, ia, ib, lane, vec);
            if (lane == 0) w[base + l] = c;
        }
    }
}

// ---- CSR ---------------------------------------------------------------------------------
// num_edges_dev != null: the count is on the device; slots [E, capacity) get the sentinel key n*n,
// which sorts behind every real key
__global__ void csr_emit_kernel(const int64_t* __restrict__ edge_index, int64_t num_edges,
                                const int64_t* __restrict__ num_edges_dev, int64_t capacity, uint64_t n,
                                uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    if (num_edges_dev != nullptr) num_edges = *num_edges_dev;
    for (int64_t e = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; e < capacity; e += int64_t(gridDim.x) * blockDim.x) {
        if (e < num_edges) {
            const uint64_t u = uint64_t(edge_index[e]), v = uint64_t(edge_index[num_edges + e]);
            keys[2 * e] = u * n + v;     vals[2 * e] = uint32_t(e);
            keys[2 * e + 1] = v * n + u; vals[2 * e + 1] = uint32_t(e);
        } else {
            keys[2 * e] = n * n;     vals[2 * e] = 0u;
            keys[2 * e + 1] = n * n; vals[2 * e + 1] = 0u;
        }
    }
}

// sorted (node*n + nbr, eid) -> col, eid, rowptr (first position of every node; gaps = empty rows)
__global__ void csr_finish_kernel(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals, int64_t nnz,
                                  const int64_t* __restrict__ num_edges_dev, uint64_t n, int64_t* __restrict__ rowptr,
                                  int64_t* __restrict__ col, int64_t* __restrict__ eid) {
    if (num_edges_dev != nullptr) nnz = 2 * *num_edges_dev;
    for (int64_t p = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; p <= nnz; p += int64_t(gridDim.x) * blockDim.x) {
        const int64_t node = (p < nnz) ? int64_t(keys[p] / n) : int64_t(n);        // virtual end marker
        const int64_t prev = (p > 0) ? int64_t(keys[p - 1] / n) : -1;
        for (int64_t r = prev + 1; r <= node; ++r) rowptr[r] = p;
        if (p < nnz) {
            col[p] = int64_t(keys[p] - uint64_t(node) * n);
            eid[p] = int64_t(vals[p]);
        }
    }
}

// ---- A8: threshold compaction ----------------------------------------------------------
__global__ void __launch_bounds__(SC_THREADS)
thr_count_kernel(const float* __restrict__ K, int64_t count, float thr, uint32_t* __restrict__ tile_sums) {
    __shared__ uint32_t sm[33];
    const int64_t base = int64_t(blockIdx.x) * SC_TILE + int64_t(threadIdx.x) * SC_ITEMS;
    uint32_t c = 0;
#pragma unroll
    for (int q = 0; q < SC_ITEMS; ++q)
        if (base + q < count) c += !(K[base + q] < thr) ? 1u : 0u;     // similarity_kernel.py:198
    uint32_t total;
    block_exclusive_scan(c, &total, sm);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(SC_THREADS)
thr_scatter_kernel(const float* __restrict__ K, int64_t n, int64_t count, float thr,
                   const uint32_t* __restrict__ tile_offsets, int64_t num_edges,
                   int64_t* __restrict__ edge_index, float* __restrict__ weights) {
    __shared__ uint32_t sm[33];
    const int64_t base = int64_t(blockIdx.x) * SC_TILE + int64_t(threadIdx.x) * SC_ITEMS;
    float v[SC_ITEMS];
    uint32_t c = 0;
#pragma unroll
    for (int q = 0; q < SC_ITEMS; ++q) {
        v[q] = (base + q < count) ? K[base + q] : -INFINITY;
        c += (base + q < count && !(v[q] < thr)) ? 1u : 0u;
    }
    uint32_t total;
    int64_t pos = int64_t(block_exclusive_scan(c, &total, sm)) + tile_offsets[blockIdx.x];
#pragma unroll
    for (int q = 0; q < SC_ITEMS; ++q) {
        if (base + q < count && !(v[q] < thr) && pos < num_edges) {
            const int64_t i = (base + q) / n;
            edge_index[pos] = i;
            edge_index[num_edges + pos] = (base + q) - i * n;
            weights[pos] = v[q];
            ++pos;
        }
    }
}

}  // namespace hg

using namespace hg;

extern "C" int hg_edge_weights_cosine(const float* feats, int64_t n, int64_t d,
                                      const int64_t* edge_index, int64_t num_edges,
                                      float* weights, void* stream) {
    HG_REQUIRE(feats && n > 0 && d > 0 && num_edges >= 0, HG_ERR_BAD_ARG, "hg_edge_weights_cosine: bad argument");
    if (num_edges == 0) return HG_OK;
    HG_REQUIRE(edge_index && weights, HG_ERR_BAD_ARG, "hg_edge_weights_cosine: null buffer");
    const int blocks = int(min64(ceil_div(num_edges, 8), 148 * 8));
    edge_cosine_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(feats, d, edge_index, num_edges, weights);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

extern "C" int hg_edge_weights_from_knn(const float* feats, int64_t n, int64_t d,
                                        const int32_t* knn_idx, const float* knn_cos, int k,
                                        const int64_t* edge_index, int64_t num_edges,
                                        float* weights, void* stream) {
    HG_REQUIRE(feats && n > 0 && d > 0 && num_edges >= 0 && k >= 1, HG_ERR_BAD_ARG, "hg_edge_weights_from_knn: bad argument");
    if (num_edges == 0) return HG_OK;
    HG_REQUIRE(edge_index && weights && knn_idx && knn_cos, HG_ERR_BAD_ARG, "hg_edge_weights_from_knn: null buffer");
    const int blocks = int(min64(ceil_div(num_edges, 256), 148 * 8));
    edge_weights_from_knn_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        feats, d, knn_idx, knn_cos, k, edge_index, num_edges, nullptr, weights);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

int hg::edge_weights_from_knn_enqueue(const float* feats, int64_t n, int64_t d, const int32_t* knn_idx,
                                      const float* knn_cos, int k, const int64_t* edge_index,
                                      const int64_t* num_edges_dev, int64_t capacity, float* weights, cudaStream_t st) {
    if (capacity <= 0) return HG_OK;
    const int blocks = int(min64(ceil_div(capacity, 256), 148 * 8));
    edge_weights_from_knn_kernel<<<blocks, 256, 0, st>>>(feats, d, knn_idx, knn_cos, k, edge_index, 0, num_edges_dev,
                                                         weights);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

extern "C" int hg_csr_workspace_bytes(int64_t n, int64_t num_edges, size_t* bytes) {
    HG_REQUIRE(bytes && n > 0 && num_edges >= 0, HG_ERR_BAD_ARG, "hg_csr_workspace_bytes: bad argument");
    const size_t nnz = size_t(2 * num_edges);
    auto al = [](size_t b) { return (b + 255) & ~size_t(255); };
    *bytes = 2 * al(nnz * 8) + 2 * al(nnz * 4) + al(sort_table_entries(int64_t(nnz)) * 4) + 256;
    return HG_OK;
}

extern "C" int hg_csr_from_coo(const int64_t* edge_index, int64_t num_edges, int64_t n,
                               int64_t* rowptr, int64_t* col, int64_t* eid,
                               void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(rowptr && n > 0 && num_edges >= 0, HG_ERR_BAD_ARG, "hg_csr_from_coo: bad argument");
    HG_REQUIRE(n < (int64_t(1) << 31) && num_edges < (int64_t(1) << 31), HG_ERR_UNSUPPORTED, "hg_csr_from_coo: too large");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    size_t need = 0;
    hg_csr_workspace_bytes(n, num_edges, &need);
    HG_REQUIRE(num_edges == 0 || (workspace && workspace_bytes >= need), HG_ERR_WORKSPACE,
               "hg_csr_from_coo: workspace too small (%zu < %zu)", workspace_bytes, need);
    const int64_t nnz = 2 * num_edges;
    if (nnz == 0) {
        HG_CHECK_CUDA(cudaMemsetAsync(rowptr, 0, size_t(n + 1) * 8, st));
        return HG_OK;
    }
    HG_REQUIRE(edge_index && col && eid, HG_ERR_BAD_ARG, "hg_csr_from_coo: null buffer");
    Carver cv(workspace);
    uint64_t* keys = cv.take<uint64_t>(nnz);
    uint64_t* tmp = cv.take<uint64_t>(nnz);
    uint32_t* vals = cv.take<uint32_t>(nnz);
    uint32_t* vtmp = cv.take<uint32_t>(nnz);
    uint32_t* table = cv.take<uint32_t>(sort_table_entries(nnz));
    const int blocks = int(min64(ceil_div(num_edges, 256), 148 * 16));
    csr_emit_kernel<<<blocks, 256, 0, st>>>(edge_index, num_edges, nullptr, num_edges, uint64_t(n), keys, vals);
    HG_CHECK_LAUNCH();
    count_launches(1);
    uint64_t* sk; uint32_t* sv;
    int rc = radix_sort_u64(keys, tmp, vals, vtmp, nnz, bits_for(uint64_t(n) * uint64_t(n)), table, &sk, &sv, st);
    if (rc != HG_OK) return rc;
    csr_finish_kernel<<<int(min64(ceil_div(nnz + 1, 256), 148 * 16)), 256, 0, st>>>(sk, sv, nnz, nullptr, uint64_t(n),
                                                                                         rowptr, col, eid);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

// hg_csr_from_coo with the edge count still on the device: the directed-key sort runs over 2 * capacity
// slots (the tail holds sentinels), the finish kernel stops at 2 E.
int hg::csr_from_coo_enqueue(const int64_t* edge_index, const int64_t* num_edges_dev, int64_t capacity, int64_t n,
                             int64_t* rowptr, int64_t* col, int64_t* eid, void* workspace, size_t workspace_bytes,
                             cudaStream_t st) {
    HG_REQUIRE(rowptr && n > 0 && capacity >= 0 && num_edges_dev, HG_ERR_BAD_ARG, "csr_from_coo_enqueue: bad argument");
    HG_REQUIRE(n < (int64_t(1) << 31) && capacity < (int64_t(1) << 29), HG_ERR_UNSUPPORTED, "csr_from_coo_enqueue: too large");
    size_t need = 0;
    hg_csr_workspace_bytes(n, capacity, &need);
    HG_REQUIRE(capacity == 0 || (workspace && workspace_bytes >= need), HG_ERR_WORKSPACE,
               "csr_from_coo_enqueue: workspace too small (%zu < %zu)", workspace_bytes, need);
    const int64_t nnz = 2 * capacity;
    if (nnz == 0) {
        HG_CHECK_CUDA(cudaMemsetAsync(rowptr, 0, size_t(n + 1) * 8, st));
        return HG_OK;
    }
    HG_REQUIRE(edge_index && col && eid, HG_ERR_BAD_ARG, "csr_from_coo_enqueue: null buffer");
    Carver cv(workspace);
    uint64_t* keys = cv.take<uint64_t>(nnz);
    uint64_t* tmp = cv.take<uint64_t>(nnz);
    uint32_t* vals = cv.take<uint32_t>(nnz);
    uint32_t* vtmp = cv.take<uint32_t>(nnz);
    uint32_t* table = cv.take<uint32_t>(sort_table_entries(nnz));
    const int blocks = int(min64(ceil_div(capacity, 256), 148 * 16));
    csr_emit_kernel<<<blocks, 256, 0, st>>>(edge_index, 0, num_edges_dev, capacity, uint64_t(n), keys, vals);
    HG_CHECK_LAUNCH();
    count_launches(1);
    uint64_t* sk; uint32_t* sv;
    int rc = radix_sort_u64(keys, tmp, vals, vtmp, nnz, bits_for(uint64_t(n) * uint64_t(n)), table, &sk, &sv, st);
    if (rc != HG_OK) return rc;
    csr_finish_kernel<<<int(min64(ceil_div(nnz + 1, 256), 148 * 16)), 256, 0, st>>>(sk, sv, 0, num_edges_dev, uint64_t(n),
                                                                                         rowptr, col, eid);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

extern "C" int hg_threshold_workspace_bytes(int64_t n, size_t* bytes) {
    HG_REQUIRE(bytes && n > 0, HG_ERR_BAD_ARG, "hg_threshold_workspace_bytes: bad argument");
    *bytes = ((unique_tiles(n * n) * 4 + 255) & ~size_t(255)) + 256;
    return HG_OK;
}

extern "C" int hg_threshold_edges_count(const float* K, int64_t n, float threshold, int64_t* num_edges_out_host,
                                        void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(K && n > 0 && num_edges_out_host, HG_ERR_BAD_ARG, "hg_threshold_edges_count: bad argument");
    size_t need = 0;
    hg_threshold_workspace_bytes(n, &need);
    HG_REQUIRE(workspace && workspace_bytes >= need, HG_ERR_WORKSPACE, "hg_threshold_edges_count: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    Carver cv(workspace);
    unsigned long long* d_count = cv.take<unsigned long long>(1);
    uint32_t* tiles = cv.take<uint32_t>(unique_tiles(n * n));
    const int nt = int(unique_tiles(n * n));
    thr_count_kernel<<<nt, SC_THREADS, 0, st>>>(K, n * n, threshold, tiles);
    HG_CHECK_LAUNCH();
    count_launches(1);
    int rc = scan_tile_sums(tiles, nt, d_count, st);
    if (rc != HG_OK) return rc;
    unsigned long long h = 0;
    HG_CHECK_CUDA(cudaMemcpyAsync(&h, d_count, 8, cudaMemcpyDeviceToHost, st));
    HG_CHECK_CUDA(cudaStreamSynchronize(st));
    *num_edges_out_host = int64_t(h);
    return HG_OK;
}

extern "C" int hg_threshold_edges_fill(const float* K, int64_t n, float threshold, int64_t num_edges,
                                       int64_t* edge_index, float* edge_weights,
                                       const void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(K && n > 0 && num_edges >= 0 && workspace, HG_ERR_BAD_ARG, "hg_threshold_edges_fill: bad argument");
    if (num_edges == 0) return HG_OK;
    HG_REQUIRE(edge_index && edge_weights, HG_ERR_BAD_ARG, "hg_threshold_edges_fill: null buffer");
    Carver cv(const_cast<void*>(workspace));
    cv.take<unsigned long long>(1);
    const uint32_t* tiles = cv.take<uint32_t>(unique_tiles(n * n));
    const int nt = int(unique_tiles(n * n));
    thr_scatter_kernel<<<nt, SC_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(K, n, n * n, threshold, tiles, num_edges,
                                                                               edge_index, edge_weights);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}
