// Declarations of the sort / scan / compaction primitives (radix_sort.cu).
#pragma once
#include "common.cuh"

namespace hg {

size_t sort_table_entries(int64_t count);     // uint32 entries needed for `table`
int radix_sort_u64(uint64_t* keys, uint64_t* keys_tmp, uint32_t* vals, uint32_t* vals_tmp, int64_t count,
                   int bits, uint32_t* table, uint64_t** sorted_keys, uint32_t** sorted_vals, cudaStream_t st);

// start[c] = first position of a run sorted by cluster label whose label is >= c, c in [0, num_clusters] (aggregate.cu)
int launch_cluster_bounds(const uint64_t* sorted_keys, int64_t n, int num_clusters, uint32_t* start, cudaStream_t st);

size_t unique_tiles(int64_t count);           // uint32 entries needed for `tile_sums` (SC_TILE items per tile)
// distinct keys of a sorted run, skipping `sentinel`; *count_out (device) <- how many
int unique_sorted_u64(const uint64_t* keys, int64_t count, uint64_t sentinel, uint64_t* out, uint32_t* tile_sums,
                      unsigned long long* count_out, cudaStream_t st);
// in-place exclusive scan of per-tile counts; grand total -> *total_out (device)
int scan_tile_sums(uint32_t* sums, int64_t len, unsigned long long* total_out, cudaStream_t st);

// ---- stream-ordered halves of the count -> fill pairs (build_api.cu: no host round trip) ----------
int edges_enqueue(const int32_t* const* knn_lists, const int* ks, int num_lists, int64_t n,
                  const int32_t* labels, int num_hyperedges, int64_t clique_pairs,
                  void* workspace, size_t workspace_bytes, cudaStream_t st, const unsigned long long** d_count_out);
int edges_fill_enqueue(int64_t capacity, int64_t* edge_index, const void* workspace, int64_t* count_out_dev,
                       cudaStream_t st);
// A6 / CSR with the edge count read from *num_edges_dev; grids and the CSR sort are sized for `capacity`
int edge_weights_from_knn_enqueue(const float* feats, int64_t n, int64_t d, const int32_t* knn_idx, const float* knn_cos,
                                  int k, const int64_t* edge_index, const int64_t* num_edges_dev, int64_t capacity,
                                  float* weights, cudaStream_t st);
int csr_from_coo_enqueue(const int64_t* edge_index, const int64_t* num_edges_dev, int64_t capacity, int64_t n,
                         int64_t* rowptr, int64_t* col, int64_t* eid, void* workspace, size_t workspace_bytes,
                         cudaStream_t st);

// ---- device-wide exclusive scan of uint32 -> uint64 totals --------------------------
constexpr int SC_THREADS = 256;
constexpr int SC_ITEMS = 8;
constexpr int SC_TILE = SC_THREADS * SC_ITEMS;

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t x, uint32_t* total, uint32_t* smem33) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = x;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += y;
    }
    if (lane == 31) smem33[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = (lane < SC_THREADS / 32) ? smem33[lane] : 0u;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += y;
        }
        smem33[lane] = w;
    }
    __syncthreads();
    const uint32_t pre = warp ? smem33[warp - 1] : 0u;
    *total = smem33[SC_THREADS / 32 - 1];
    __syncthreads();
    return pre + inc - x;
}


inline int bits_for(uint64_t max_value) {     // number of bits needed to hold max_value
    int b = 1;
    while (b < 64 && (max_value >> b) != 0) ++b;
    return b;
}

}  // namespace hg
