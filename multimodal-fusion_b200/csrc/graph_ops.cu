// A6 cosine edge weights, CSR incidence, A8 dense-threshold edge builder.
#include "common.cuh"
#include "rowdot.cuh"
#include <mutex>
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
// The edge list arrives sorted by (u, v) with u < v.  Row x of the incidence holds, by ascending
// neighbour: first the edges (u', x) in which x is the LARGER endpoint (ascending u' = ascending
// edge id), then the edges (x, v) in which it is the smaller one (a contiguous run of the edge
// list).  Only the first part needs data movement: ONE stable radix sort of the E pairs
// (key = larger endpoint, log2 n bits; payload = edge id) transposes the list.  Row starts follow
// without a scan: with first_hi[x] / first_lo[x] = number of edges whose smaller / larger endpoint
// is < x (both found by boundary detection on a sorted run), rowptr[x] = first_hi[x] + first_lo[x].
// num_edges_dev != null: the count is still on the device and slots [E, capacity) carry the
// sentinel key n, which sorts behind every node.
__global__ void csr_keys_kernel(const int64_t* __restrict__ edge_index, int64_t num_edges,
                                const int64_t* __restrict__ num_edges_dev, int64_t capacity, int64_t n,
                                uint64_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t* __restrict__ first_hi) {
    if (num_edges_dev != nullptr) num_edges = *num_edges_dev;
    const int64_t span = capacity > num_edges ? capacity : num_edges + 1;      // e = num_edges closes first_hi
    for (int64_t e = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; e < span; e += int64_t(gridDim.x) * blockDim.x) {
        if (e < capacity) {
            keys[e] = (e < num_edges) ? uint64_t(edge_index[num_edges + e]) : uint64_t(n);
            vals[e] = uint32_t(e);
        }
        if (e <= num_edges) {
            const int64_t node = (e < num_edges) ? edge_index[e] : n;
            const int64_t prev = (e > 0) ? edge_index[e - 1] : -1;
            for (int64_t r = prev + 1; r <= node; ++r) first_hi[r] = uint32_t(e);
        }
    }
}

// boundaries of the run sorted by larger endpoint -> first_lo, and the row starts
__global__ void csr_bounds_kernel(const uint64_t* __restrict__ sorted_keys, int64_t num_edges,
                                  const int64_t* __restrict__ num_edges_dev, int64_t n,
                                  const uint32_t* __restrict__ first_hi, uint32_t* __restrict__ first_lo,
                                  int64_t* __restrict__ rowptr) {
    if (num_edges_dev != nullptr) num_edges = *num_edges_dev;
    for (int64_t p = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; p <= num_edges; p += int64_t(gridDim.x) * blockDim.x) {
        const int64_t node = (p < num_edges) ? int64_t(sorted_keys[p]) : n;
        const int64_t prev = (p > 0) ? int64_t(sorted_keys[p - 1]) : -1;
        for (int64_t r = prev + 1; r <= node; ++r) {
            first_lo[r] = uint32_t(p);
            rowptr[r] = int64_t(first_hi[r]) + p;
        }
    }
}

__global__ void csr_fill_kernel(const int64_t* __restrict__ edge_index, const uint64_t* __restrict__ sorted_keys,
                                const uint32_t* __restrict__ sorted_eids, int64_t num_edges,
                                const int64_t* __restrict__ num_edges_dev, const uint32_t* __restrict__ first_hi,
                                const uint32_t* __restrict__ first_lo, const int64_t* __restrict__ rowptr,
                                int64_t* __restrict__ col, int64_t* __restrict__ eid) {
    if (num_edges_dev != nullptr) num_edges = *num_edges_dev;
    for (int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; t < num_edges; t += int64_t(gridDim.x) * blockDim.x) {
        {   // position t of the transposed run: edge (u, x) seen from its larger endpoint x
            const int64_t x = int64_t(sorted_keys[t]);
            const int64_t e = int64_t(sorted_eids[t]);
            const int64_t dst = rowptr[x] + (t - int64_t(first_lo[x]));
            col[dst] = edge_index[e];
            eid[dst] = e;
        }
        {   // edge t = (u, v) seen from its smaller endpoint u, after u's first part
            const int64_t u = edge_index[t];
            const int64_t dst = rowptr[u] + int64_t(first_lo[u + 1] - first_lo[u]) + (t - int64_t(first_hi[u]));
            col[dst] = edge_index[num_edges + t];
            eid[dst] = t;
        }
    }
}

// ---- (f)3: consumer-side propagation over the incidence ------------------------------------------
// The reference's consumer hands edge_index [2, E] to torch_geometric's HypergraphConv
// (downstream_survival/models/cust_omics.py:58-63, 100-102), which reads row 0 as NODE ids and row 1
// as HYPEREDGE ids and computes  out = D^-1 H B^-1 H^T (X Theta)  (no attention, unit hyperedge
// weights): m_j = (1/B_j) sum_{(i, j)} x_i over the pairs whose row 1 is j, then
// out_v = (1/D_v) sum_{(v, j)} m_j over the pairs whose row 0 is v; B_j, D_v = pair counts, 1/0 -> 0.
// With the (min, max) pair list this library emits, row 0 < row 1, so in the CSR incidence of node
// x (neighbours ascending) the members of hyperedge x are exactly the neighbours < x and the
// hyperedges incident to node x the neighbours > x: both passes are segmented gather-sums over CSR
// rows, one warp per row, 128-bit loads, sums in ascending neighbour order (deterministic).
// HBM/L2-bound: per pass E gathered rows of F floats + N written.
template <bool UPPER>
__global__ void __launch_bounds__(256)
hconv_gather_kernel(const float* __restrict__ src, int64_t n, int64_t f, const int64_t* __restrict__ rowptr,
                    const int64_t* __restrict__ col, float* __restrict__ dst) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    if (row >= n) return;
    const int64_t b = rowptr[row], e = rowptr[row + 1];
    // neighbours ascend: [b, mid) are < row, [mid, e) are > row
    int64_t lo = b, hi = e;
    while (lo < hi) {
        const int64_t m = (lo + hi) >> 1;
        if (col[m] < row) lo = m + 1; else hi = m;
    }
    const int64_t s0 = UPPER ? lo : b, s1 = UPPER ? e : lo;
    const float inv = (s1 > s0) ? 1.f / float(s1 - s0) : 0.f;
    const bool vec = (f & 3) == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
    if (vec) {
        for (int64_t c0 = int64_t(lane) * 4; c0 < f; c0 += 128) {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int64_t p = s0; p < s1; ++p) {
                const float4 v = __ldg(reinterpret_cast<const float4*>(src + col[p] * f + c0));
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
            *reinterpret_cast<float4*>(dst + row * f + c0) = make_float4(acc.x * inv, acc.y * inv, acc.z * inv, acc.w * inv);
        }
    } else {
        for (int64_t c = lane; c < f; c += 32) {
            float acc = 0.f;
            for (int64_t p = s0; p < s1; ++p) acc += __ldg(src + col[p] * f + c);
            dst[row * f + c] = acc * inv;
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
    const size_t cap = size_t(num_edges);
    auto al = [](size_t b) { return (b + 255) & ~size_t(255); };
    *bytes = 2 * al(cap * 8) + 2 * al(cap * 4) + al(sort_table_entries(int64_t(cap)) * 4) + 2 * al(size_t(n + 1) * 4) + 256;
    return HG_OK;
}

// shared by the host-count and device-count entry points (exactly one of num_edges / num_edges_dev is used)
static int csr_build(const int64_t* edge_index, int64_t num_edges, const int64_t* num_edges_dev, int64_t capacity,
                     int64_t n, int64_t* rowptr, int64_t* col, int64_t* eid, void* workspace, size_t workspace_bytes,
                     cudaStream_t st) {
    HG_REQUIRE(rowptr && n > 0 && capacity >= 0, HG_ERR_BAD_ARG, "hg_csr_from_coo: bad argument");
    HG_REQUIRE(n < (int64_t(1) << 31) - 1 && capacity < (int64_t(1) << 30), HG_ERR_UNSUPPORTED, "hg_csr_from_coo: too large");
    if (capacity == 0) {
        HG_CHECK_CUDA(cudaMemsetAsync(rowptr, 0, size_t(n + 1) * 8, st));
        return HG_OK;
    }
    size_t need = 0;
    hg_csr_workspace_bytes(n, capacity, &need);
    HG_REQUIRE(workspace && workspace_bytes >= need, HG_ERR_WORKSPACE, "hg_csr_from_coo: workspace too small (%zu < %zu)",
               workspace_bytes, need);
    HG_REQUIRE(edge_index && col && eid, HG_ERR_BAD_ARG, "hg_csr_from_coo: null buffer");
    Carver cv(workspace);
    uint64_t* keys = cv.take<uint64_t>(capacity);
    uint64_t* tmp = cv.take<uint64_t>(capacity);
    uint32_t* vals = cv.take<uint32_t>(capacity);
    uint32_t* vtmp = cv.take<uint32_t>(capacity);
    uint32_t* table = cv.take<uint32_t>(sort_table_entries(capacity));
    uint32_t* first_hi = cv.take<uint32_t>(n + 1);
    uint32_t* first_lo = cv.take<uint32_t>(n + 1);
    const int blocks = int(min64(ceil_div(capacity + 1, 256), 148 * 16));
    csr_keys_kernel<<<blocks, 256, 0, st>>>(edge_index, num_edges, num_edges_dev, capacity, n, keys, vals, first_hi);
    HG_CHECK_LAUNCH();
    count_launches(1);
    uint64_t* sk = keys; uint32_t* sv = vals;
    int rc = radix_sort_u64(keys, tmp, vals, vtmp, capacity, bits_for(uint64_t(n)), table, &sk, &sv, st);
    if (rc != HG_OK) return rc;
    csr_bounds_kernel<<<blocks, 256, 0, st>>>(sk, num_edges, num_edges_dev, n, first_hi, first_lo, rowptr);
    csr_fill_kernel<<<blocks, 256, 0, st>>>(edge_index, sk, sv, num_edges, num_edges_dev, first_hi, first_lo, rowptr, col, eid);
    HG_CHECK_LAUNCH();
    count_launches(2);
    return HG_OK;
}

extern "C" int hg_csr_from_coo(const int64_t* edge_index, int64_t num_edges, int64_t n,
                               int64_t* rowptr, int64_t* col, int64_t* eid,
                               void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(num_edges >= 0, HG_ERR_BAD_ARG, "hg_csr_from_coo: bad argument");
    return csr_build(edge_index, num_edges, nullptr, num_edges, n, rowptr, col, eid, workspace, workspace_bytes,
                     static_cast<cudaStream_t>(stream));
}

// ---- hg_csr_from_coo without any allocation on the caller's side ------------------------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it): the scratch of the CSR build --
// sort keys and payloads, digit tables, first_hi / first_lo -- lives in a file-scope device array instead of a
// caller-provided workspace.  Every GPU that loads the library holds HG_CSR_STATIC_SCRATCH_BYTES (18 MiB) until
// the process exits.  The array is shared by every call on a device and keeps the last call's contents: the build
// initialises everything it reads (the same kernels run on an uninitialised caller workspace in hg_csr_from_coo),
// but two calls that overlap on the GPU (different streams, or threads) race on it -- serialise them.
// In host code the array's name is only a stand-in for the symbol, not a device address: the address is looked up
// once per device with cudaGetSymbolAddress (that device current) and passed to the kernels.
static __device__ __align__(256) unsigned char g_csr_static_scratch[HG_CSR_STATIC_SCRATCH_BYTES];

static std::mutex g_csr_static_mu;
static void* g_csr_static_addr[64] = {};          // per device

extern "C" int hg_csr_from_coo_static(const int64_t* edge_index, int64_t num_edges, int64_t n,
                                      int64_t* rowptr, int64_t* col, int64_t* eid, void* stream) {
    HG_REQUIRE(num_edges >= 0 && n > 0, HG_ERR_BAD_ARG, "hg_csr_from_coo_static: bad argument");
    HG_REQUIRE(n <= HG_CSR_STATIC_MAX_N && num_edges <= HG_CSR_STATIC_MAX_EDGES, HG_ERR_UNSUPPORTED,
               "hg_csr_from_coo_static: n = %lld / num_edges = %lld beyond the static scratch (n <= %lld, num_edges <= %lld); "
               "use hg_csr_from_coo with a caller workspace", (long long)n, (long long)num_edges,
               (long long)HG_CSR_STATIC_MAX_N, (long long)HG_CSR_STATIC_MAX_EDGES);
    size_t need = 0;
    int rc = hg_csr_workspace_bytes(n, num_edges, &need);
    if (rc != HG_OK) return rc;
    HG_REQUIRE(need <= size_t(HG_CSR_STATIC_SCRATCH_BYTES), HG_ERR_UNSUPPORTED,
               "hg_csr_from_coo_static: %zu bytes of scratch needed, %zu available", need, size_t(HG_CSR_STATIC_SCRATCH_BYTES));
    int dev = 0;
    HG_CHECK_CUDA(cudaGetDevice(&dev));
    HG_REQUIRE(dev >= 0 && dev < 64, HG_ERR_UNSUPPORTED, "hg_csr_from_coo_static: device ordinal %d", dev);
    void* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_csr_static_mu);
        if (g_csr_static_addr[dev] == nullptr) {
            void* ptr = nullptr;
            HG_CHECK_CUDA(cudaGetSymbolAddress(&ptr, g_csr_static_scratch));       // this device's copy of the array
            HG_REQUIRE(ptr != nullptr && (reinterpret_cast<uintptr_t>(ptr) & 255) == 0, HG_ERR_CUDA,
                       "hg_csr_from_coo_static: bad symbol address");
            g_csr_static_addr[dev] = ptr;
        }
        scratch = g_csr_static_addr[dev];
    }
    return csr_build(edge_index, num_edges, nullptr, num_edges, n, rowptr, col, eid, scratch,
                     size_t(HG_CSR_STATIC_SCRATCH_BYTES), static_cast<cudaStream_t>(stream));
}

// hg_csr_from_coo with the edge count still on the device: the sort runs over `capacity` slots (the
// tail holds sentinels), every kernel stops at E.
int hg::csr_from_coo_enqueue(const int64_t* edge_index, const int64_t* num_edges_dev, int64_t capacity, int64_t n,
                             int64_t* rowptr, int64_t* col, int64_t* eid, void* workspace, size_t workspace_bytes,
                             cudaStream_t st) {
    HG_REQUIRE(num_edges_dev, HG_ERR_BAD_ARG, "csr_from_coo_enqueue: bad argument");
    return csr_build(edge_index, 0, num_edges_dev, capacity, n, rowptr, col, eid, workspace, workspace_bytes, st);
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

extern "C" int hg_hypergraph_propagate(const float* x, int64_t n, int64_t f, const int64_t* csr_rowptr,
                                       const int64_t* csr_col, float* hyperedge_feat, float* out, void* stream) {
    HG_REQUIRE(x && csr_rowptr && hyperedge_feat && out && n > 0 && f > 0, HG_ERR_BAD_ARG, "hg_hypergraph_propagate: bad argument");
    HG_REQUIRE(hyperedge_feat != out && hyperedge_feat != x, HG_ERR_BAD_ARG, "hg_hypergraph_propagate: buffers must not alias");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned blocks = unsigned(ceil_div(n * 32, 256));
    hconv_gather_kernel<false><<<blocks, 256, 0, st>>>(x, n, f, csr_rowptr, csr_col, hyperedge_feat);
    hconv_gather_kernel<true><<<blocks, 256, 0, st>>>(hyperedge_feat, n, f, csr_rowptr, csr_col, out);
    HG_CHECK_LAUNCH();
    count_launches(2);
    return HG_OK;
}

extern "C" int hg_threshold_edges_build(const float* K, int64_t n, float threshold, int64_t num_edges,
                                        int64_t* edge_index, float* edge_weights,
                                        const void* workspace, size_t workspace_bytes, void* stream) {
    return hg_threshold_edges_fill(K, n, threshold, num_edges, edge_index, edge_weights, workspace, workspace_bytes, stream);
}
