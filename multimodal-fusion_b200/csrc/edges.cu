// A5 / A6 / CSR: hyperedge assembly from neighbour lists and cluster labels.
//
// Reference (preprocess_hypergraph.py):
//   :385-388  directed kNN pairs (i, nbr)
//   :394-400  every ordered pair inside each KMeans cluster
//   :403      canonicalise to (min, max) and set-deduplicate
//   :410      edge_index int64 [2, E]
//   :413-422  w_e = max(0, cos(f_i, f_j))
// The canonical order produced here is lexicographic by (min, max) -- the
// reference order is CPython set-iteration order and carries no meaning.
//
// Pipeline: emit 64-bit keys min*n+max -> LSD radix sort on the needed bits ->
// adjacent-unique compaction -> decode.  All stages are HBM-bound streaming
// kernels (8 B per key per pass).
#include "common.cuh"
#include "sort_scan.cuh"

namespace hg {

struct EdgesHeader {            // lives at the start of the caller's workspace (device)
    unsigned long long num_edges;
    unsigned long long unique_off;   // byte offset of the unique key array
    unsigned long long n;
    unsigned long long magic;
};
constexpr unsigned long long kEdgesMagic = 0x4847423230304544ull;
constexpr int kMaxLists = 4;

struct KnnLists {
    const int32_t* idx[kMaxLists];
    int k[kMaxLists];
    int64_t slot_begin[kMaxLists + 1];   // prefix of n*k over lists
    int num;
};

__global__ void emit_knn_keys_kernel(KnnLists L, int64_t n, uint64_t sentinel, uint64_t* __restrict__ keys) {
    const int64_t total = L.slot_begin[L.num];
    for (int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += int64_t(gridDim.x) * blockDim.x) {
        int l = 0;
        while (l + 1 < L.num && t >= L.slot_begin[l + 1]) ++l;
        const int64_t local = t - L.slot_begin[l];
        const int64_t i = local / L.k[l];
        const int64_t j = L.idx[l][local];
        uint64_t key = sentinel;
        if (j >= 0 && j < n && j != i) {
            const uint64_t lo = uint64_t(min(i, j)), hi = uint64_t(max(i, j));
            key = lo * uint64_t(n) + hi;
        }
        keys[t] = key;
    }
}

// ---- cliques ----------------------------------------------------------------------
__global__ void label_keys_kernel(const int32_t* __restrict__ labels, int64_t n, int num_h,
                                  uint64_t* __restrict__ lkeys, uint32_t* __restrict__ lvals,
                                  unsigned int* __restrict__ counts) {
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int32_t lb = labels[i];
    const int c = (lb >= 0 && lb < num_h) ? lb : num_h;       // out-of-range labels form no clique
    lkeys[i] = uint64_t(c);
    lvals[i] = uint32_t(i);
    atomicAdd(&counts[c], 1u);
}

// one thread: exclusive prefixes of cluster sizes and of pair counts s(s-1)/2
__global__ void clique_offsets_kernel(const unsigned int* __restrict__ counts, int num_h,
                                      unsigned long long* __restrict__ member_off,
                                      unsigned long long* __restrict__ pair_off) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    unsigned long long m = 0, p = 0;
    for (int c = 0; c < num_h; ++c) {
        member_off[c] = m; pair_off[c] = p;
        const unsigned long long s = counts[c];
        m += s; p += s * (s - (s ? 1 : 0)) / 2;
    }
    member_off[num_h] = m; pair_off[num_h] = p;
}

__global__ void expand_cliques_kernel(const uint32_t* __restrict__ members, const unsigned int* __restrict__ counts,
                                      const unsigned long long* __restrict__ member_off,
                                      const unsigned long long* __restrict__ pair_off, int num_h, int64_t n,
                                      int64_t pairs_expected, uint64_t sentinel, uint64_t* __restrict__ keys) {
    const unsigned long long total = pair_off[num_h];
    for (int64_t p = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; p < pairs_expected;
         p += int64_t(gridDim.x) * blockDim.x) {
        if (uint64_t(p) >= total) { keys[p] = sentinel; continue; }
        int lo = 0, hi = num_h;                       // largest c with pair_off[c] <= p
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (pair_off[mid] <= uint64_t(p)) lo = mid; else hi = mid;
        }
        const int c = lo;
        const long long q = p - (long long)pair_off[c];
        const long long s = counts[c];
        // row a of the strict upper triangle holds s-1-a pairs; rows before a hold a(2s-a-1)/2
        long long a = (long long)floor(((2.0 * s - 1.0) - sqrt((2.0 * s - 1.0) * (2.0 * s - 1.0) - 8.0 * double(q))) * 0.5);
        a = max(0ll, min(a, s - 2));
        while (a * (2 * s - a - 1) / 2 > q) --a;
        while ((a + 1) * (2 * s - a - 2) / 2 <= q) ++a;
        const long long b = q - a * (2 * s - a - 1) / 2 + a + 1;
        const uint64_t ma = members[member_off[c] + a], mb = members[member_off[c] + b];
        keys[p] = ma * uint64_t(n) + mb;              // members ascend inside a cluster: ma < mb
    }
}

__global__ void write_header_kernel(EdgesHeader* h, const unsigned long long* count, unsigned long long unique_off,
                                    unsigned long long n) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        h->num_edges = *count; h->unique_off = unique_off; h->n = n; h->magic = kEdgesMagic;
    }
}

// num_edges < 0: the count is taken from the header (stream-ordered build, no host round trip) and
// also published to *count_out
__global__ void decode_edges_kernel(const EdgesHeader* __restrict__ h, const char* __restrict__ ws,
                                    int64_t num_edges, int64_t* __restrict__ edge_index,
                                    int64_t* __restrict__ count_out) {
    const uint64_t* uniq = reinterpret_cast<const uint64_t*>(ws + h->unique_off);
    const uint64_t n = h->n;
    if (num_edges < 0) num_edges = int64_t(h->num_edges);
    if (count_out != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *count_out = num_edges;
    for (int64_t e = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; e < num_edges; e += int64_t(gridDim.x) * blockDim.x) {
        const uint64_t key = uniq[e];
        const uint64_t u = key / n;
        edge_index[e] = int64_t(u);
        edge_index[num_edges + e] = int64_t(key - u * n);
    }
}

struct EdgesPlan {
    int64_t knn_slots, total_keys;
    size_t off_header, off_keys, off_tmp, off_table, off_tiles, off_count;
    size_t off_lkeys, off_ltmp, off_lvals, off_lvtmp, off_counts, off_moff, off_poff, off_ltable;
    size_t bytes;
};

static EdgesPlan plan_edges(int64_t n, int64_t knn_slots, int num_h, int64_t clique_pairs) {
    EdgesPlan p{};
    p.knn_slots = knn_slots;
    p.total_keys = knn_slots + clique_pairs;
    size_t off = 0;
    auto take = [&](size_t b) { size_t o = off; off += (b + 255) & ~size_t(255); return o; };
    p.off_header = take(sizeof(EdgesHeader));
    p.off_keys = take(size_t(p.total_keys) * 8);
    p.off_tmp = take(size_t(p.total_keys) * 8);
    p.off_table = take(sort_table_entries(p.total_keys) * 4);
    p.off_tiles = take(unique_tiles(p.total_keys) * 4 + 4);
    p.off_count = take(8);
    if (num_h > 0) {
        p.off_lkeys = take(size_t(n) * 8);
        p.off_ltmp = take(size_t(n) * 8);
        p.off_lvals = take(size_t(n) * 4);
        p.off_lvtmp = take(size_t(n) * 4);
        p.off_counts = take(size_t(num_h + 1) * 4);
        p.off_moff = take(size_t(num_h + 1) * 8);
        p.off_poff = take(size_t(num_h + 1) * 8);
        p.off_ltable = take(sort_table_entries(n) * 4);
    }
    p.bytes = off;
    return p;
}

}  // namespace hg

using namespace hg;

extern "C" int hg_edges_workspace_bytes(int64_t n, int64_t knn_slots, int num_hyperedges, int64_t clique_pairs,
                                        size_t* bytes) {
    HG_REQUIRE(bytes && n > 0 && knn_slots >= 0 && clique_pairs >= 0 && num_hyperedges >= 0, HG_ERR_BAD_ARG,
               "hg_edges_workspace_bytes: bad argument");
    *bytes = plan_edges(n, knn_slots, num_hyperedges, clique_pairs).bytes;
    return HG_OK;
}

extern "C" int hg_clique_pairs(const int32_t* labels, int64_t n, int num_hyperedges, int64_t* pairs_out_host,
                               void* stream) {
    HG_REQUIRE(labels && pairs_out_host && n > 0 && num_hyperedges > 0 && num_hyperedges <= (1 << 20), HG_ERR_BAD_ARG,
               "hg_clique_pairs: bad argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // host-side tally: labels are N int32 (<= 1 MB), the count gates an allocation anyway
    int32_t* host = static_cast<int32_t*>(malloc(size_t(n) * 4));
    HG_REQUIRE(host, HG_ERR_BAD_ARG, "hg_clique_pairs: out of host memory");
    cudaError_t e = cudaMemcpyAsync(host, labels, size_t(n) * 4, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { free(host); set_error("hg_clique_pairs: %s", cudaGetErrorString(e)); return HG_ERR_CUDA; }
    int64_t* cnt = static_cast<int64_t*>(calloc(size_t(num_hyperedges), 8));
    for (int64_t i = 0; i < n; ++i)
        if (host[i] >= 0 && host[i] < num_hyperedges) ++cnt[host[i]];
    int64_t pairs = 0;
    for (int c = 0; c < num_hyperedges; ++c) pairs += cnt[c] * (cnt[c] - 1) / 2;
    free(cnt); free(host);
    *pairs_out_host = pairs;
    return HG_OK;
}

// Everything of hg_edges_count up to (not including) the host read-back of the count: the unique keys
// and the header are left in the workspace, *d_count_out = device address of the 64-bit count.
int hg::edges_enqueue(const int32_t* const* knn_lists, const int* ks, int num_lists, int64_t n,
                      const int32_t* labels, int num_hyperedges, int64_t clique_pairs,
                      void* workspace, size_t workspace_bytes, cudaStream_t st,
                      const unsigned long long** d_count_out) {
    HG_REQUIRE(n > 0 && workspace, HG_ERR_BAD_ARG, "hg_edges_count: bad argument");
    HG_REQUIRE(num_lists >= 0 && num_lists <= kMaxLists, HG_ERR_UNSUPPORTED, "hg_edges_count: at most %d lists", kMaxLists);
    HG_REQUIRE(n < (int64_t(1) << 31), HG_ERR_UNSUPPORTED, "hg_edges_count: n must be < 2^31");
    if (labels == nullptr) { num_hyperedges = 0; clique_pairs = 0; }
    KnnLists L{};
    L.num = num_lists;
    L.slot_begin[0] = 0;
    for (int l = 0; l < num_lists; ++l) {
        HG_REQUIRE(knn_lists && knn_lists[l] && ks && ks[l] > 0, HG_ERR_BAD_ARG, "hg_edges_count: bad list %d", l);
        L.idx[l] = knn_lists[l]; L.k[l] = ks[l];
        L.slot_begin[l + 1] = L.slot_begin[l] + n * ks[l];
    }
    const EdgesPlan p = plan_edges(n, L.slot_begin[num_lists], num_hyperedges, clique_pairs);
    HG_REQUIRE(workspace_bytes >= p.bytes, HG_ERR_WORKSPACE, "hg_edges_count: workspace too small (%zu < %zu)",
               workspace_bytes, p.bytes);
    char* ws = static_cast<char*>(workspace);
    uint64_t* keys = reinterpret_cast<uint64_t*>(ws + p.off_keys);
    uint64_t* tmp = reinterpret_cast<uint64_t*>(ws + p.off_tmp);
    const uint64_t sentinel = uint64_t(n) * uint64_t(n);
    const int bits = bits_for(sentinel);

    if (p.knn_slots > 0) {
        const int blocks = int(min64(ceil_div(p.knn_slots, 256), 148 * 16));
        emit_knn_keys_kernel<<<blocks, 256, 0, st>>>(L, n, sentinel, keys);
        HG_CHECK_LAUNCH();
        count_launches(1);
    }
    if (num_hyperedges > 0 && clique_pairs > 0) {
        uint64_t* lkeys = reinterpret_cast<uint64_t*>(ws + p.off_lkeys);
        uint64_t* ltmp = reinterpret_cast<uint64_t*>(ws + p.off_ltmp);
        uint32_t* lvals = reinterpret_cast<uint32_t*>(ws + p.off_lvals);
        uint32_t* lvtmp = reinterpret_cast<uint32_t*>(ws + p.off_lvtmp);
        unsigned int* counts = reinterpret_cast<unsigned int*>(ws + p.off_counts);
        unsigned long long* moff = reinterpret_cast<unsigned long long*>(ws + p.off_moff);
        unsigned long long* poff = reinterpret_cast<unsigned long long*>(ws + p.off_poff);
        HG_CHECK_CUDA(cudaMemsetAsync(counts, 0, size_t(num_hyperedges + 1) * 4, st));
        label_keys_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, st>>>(labels, n, num_hyperedges, lkeys, lvals, counts);
        HG_CHECK_LAUNCH();
        uint64_t* sk; uint32_t* sv;
        int rc = radix_sort_u64(lkeys, ltmp, lvals, lvtmp, n, bits_for(uint64_t(num_hyperedges)),
                                reinterpret_cast<uint32_t*>(ws + p.off_ltable), &sk, &sv, st);
        if (rc != HG_OK) return rc;
        clique_offsets_kernel<<<1, 32, 0, st>>>(counts, num_hyperedges, moff, poff);
        const int blocks = int(min64(ceil_div(clique_pairs, 256), 148 * 32));
        expand_cliques_kernel<<<blocks, 256, 0, st>>>(sv, counts, moff, poff, num_hyperedges, n, clique_pairs,
                                                      sentinel, keys + p.knn_slots);
        HG_CHECK_LAUNCH();
        count_launches(3);
    }
    unsigned long long* d_count = reinterpret_cast<unsigned long long*>(ws + p.off_count);
    uint64_t* sorted = keys;
    int rc = radix_sort_u64(keys, tmp, nullptr, nullptr, p.total_keys, bits,
                            reinterpret_cast<uint32_t*>(ws + p.off_table), &sorted, nullptr, st);
    if (rc != HG_OK) return rc;
    uint64_t* uniq = (sorted == keys) ? tmp : keys;
    rc = unique_sorted_u64(sorted, p.total_keys, sentinel, uniq, reinterpret_cast<uint32_t*>(ws + p.off_tiles),
                           d_count, st);
    if (rc != HG_OK) return rc;
    write_header_kernel<<<1, 32, 0, st>>>(reinterpret_cast<EdgesHeader*>(ws + p.off_header), d_count,
                                          (unsigned long long)(reinterpret_cast<char*>(uniq) - ws), (unsigned long long)n);
    HG_CHECK_LAUNCH();
    count_launches(1);
    if (d_count_out) *d_count_out = d_count;
    return HG_OK;
}

extern "C" int hg_edges_count(const int32_t* const* knn_lists, const int* ks, int num_lists, int64_t n,
                              const int32_t* labels, int num_hyperedges, int64_t clique_pairs,
                              int64_t* num_edges_out_host,
                              void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(num_edges_out_host, HG_ERR_BAD_ARG, "hg_edges_count: bad argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned long long* d_count = nullptr;
    int rc = edges_enqueue(knn_lists, ks, num_lists, n, labels, num_hyperedges, clique_pairs, workspace, workspace_bytes,
                           st, &d_count);
    if (rc != HG_OK) return rc;
    unsigned long long h_count = 0;
    HG_CHECK_CUDA(cudaMemcpyAsync(&h_count, d_count, 8, cudaMemcpyDeviceToHost, st));
    HG_CHECK_CUDA(cudaStreamSynchronize(st));
    *num_edges_out_host = int64_t(h_count);
    return HG_OK;
}

// hg_edges_fill for a count that is still on the device: decodes h->num_edges (<= capacity) edges
// as edge_index [2, E] (row 1 at offset E) and publishes E to *count_out_dev.
int hg::edges_fill_enqueue(int64_t capacity, int64_t* edge_index, const void* workspace, int64_t* count_out_dev,
                           cudaStream_t st) {
    if (capacity <= 0) {
        if (count_out_dev) HG_CHECK_CUDA(cudaMemsetAsync(count_out_dev, 0, 8, st));
        return HG_OK;
    }
    const int blocks = int(min64(ceil_div(capacity, 256), 148 * 16));
    decode_edges_kernel<<<blocks, 256, 0, st>>>(static_cast<const EdgesHeader*>(workspace),
                                                static_cast<const char*>(workspace), -1, edge_index, count_out_dev);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

extern "C" int hg_edges_fill(int64_t num_edges, int64_t* edge_index,
                             const void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(workspace && workspace_bytes >= sizeof(EdgesHeader) && num_edges >= 0, HG_ERR_BAD_ARG,
               "hg_edges_fill: bad argument");
    if (num_edges == 0) return HG_OK;
    HG_REQUIRE(edge_index, HG_ERR_BAD_ARG, "hg_edges_fill: null edge_index");
    const int blocks = int(min64(ceil_div(num_edges, 256), 148 * 16));
    decode_edges_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const EdgesHeader*>(workspace), static_cast<const char*>(workspace), num_edges, edge_index, nullptr);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

extern "C" int hg_edges_build(int64_t num_edges, int64_t* edge_index,
                              const void* workspace, size_t workspace_bytes, void* stream) {
    return hg_edges_fill(num_edges, edge_index, workspace, workspace_bytes, stream);
}
