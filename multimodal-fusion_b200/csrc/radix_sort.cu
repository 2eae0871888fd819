// Device-wide LSD radix sort of 64-bit keys (optional 32-bit payload), exclusive
// scan, and adjacent-unique compaction -- the building blocks of edge assembly
// (reference preprocess_hypergraph.py:403 `set(tuple(sorted(edge)))`).
//
// Per 8-bit pass: (1) per-tile digit histogram -> table[digit][tile],
// (2) exclusive scan of the table, (3) stable scatter.  Stability inside a tile
// comes from warp-synchronous ranking: each warp owns a contiguous run of the
// tile, ranks its keys digit by digit with __match_any_sync + popc, and the
// per-warp digit totals are prefix-summed across warps.
// HBM-bound: each pass reads the keys twice and writes them once.
#include "common.cuh"
#include "sort_scan.cuh"

namespace hg {

constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;   // 4096 keys per CTA
constexpr int RS_WARPS = RS_THREADS / 32;

__global__ void __launch_bounds__(RS_THREADS)
rs_hist_kernel(const uint64_t* __restrict__ keys, int64_t count, int shift, uint32_t* __restrict__ table,
               int num_tiles) {
    __shared__ uint32_t hist[256];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = int64_t(blockIdx.x) * RS_TILE;
#pragma unroll 4
    for (int it = 0; it < RS_ITEMS; ++it) {
        const int64_t i = base + it * RS_THREADS + threadIdx.x;
        if (i < count) atomicAdd(&hist[(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    table[size_t(threadIdx.x) * num_tiles + blockIdx.x] = hist[threadIdx.x];
}

// Exclusive scan of the digit-major table in two levels: CTA `d` scans row d (one entry per
// tile) in place and publishes the row total; the scatter kernel turns the 256 row totals into
// digit bases itself.  256 CTAs instead of one long single-CTA sweep.
__global__ void __launch_bounds__(256)
rs_scan_rows_kernel(uint32_t* __restrict__ table, int num_tiles, uint32_t* __restrict__ digit_totals) {
    __shared__ uint32_t warp_tot[8];
    __shared__ uint32_t carry_s;
    uint32_t* row = table + size_t(blockIdx.x) * num_tiles;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < num_tiles; base += 256) {
        const int i = base + threadIdx.x;
        const uint32_t v = (i < num_tiles) ? row[i] : 0u;
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        uint32_t pre = carry_s;
        for (int w = 0; w < warp; ++w) pre += warp_tot[w];
        if (i < num_tiles) row[i] = pre + x - v;
        __syncthreads();
        if (threadIdx.x == 255) carry_s = pre + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) digit_totals[blockIdx.x] = carry_s;
}

template <bool HAS_VALS>
__global__ void __launch_bounds__(RS_THREADS)
rs_scatter_kernel(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals, int64_t count, int shift,
                  const uint32_t* __restrict__ table, const uint32_t* __restrict__ digit_totals, int num_tiles,
                  uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out) {
    __shared__ uint32_t cnt[RS_WARPS][256];
    __shared__ uint32_t gbase[256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int w = 0; w < RS_WARPS; ++w) cnt[w][threadIdx.x] = 0;
    __syncthreads();

    // warp `warp` owns keys [base + warp*512, +512) in (it, lane) order
    const int64_t wbase = int64_t(blockIdx.x) * RS_TILE + int64_t(warp) * (32 * RS_ITEMS);
    uint64_t key[RS_ITEMS];
    uint32_t rank[RS_ITEMS];
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int it = 0; it < RS_ITEMS; ++it) {
        const int64_t i = wbase + it * 32 + lane;
        const bool ok = i < count;
        key[it] = ok ? keys[i] : ~0ull;
        const uint32_t dgt = ok ? uint32_t((key[it] >> shift) & 255u) : 256u;   // 256: inactive lanes group apart
        const uint32_t peers = __match_any_sync(0xffffffffu, dgt);
        const int leader = __ffs(peers) - 1;
        uint32_t pre = 0;
        if (ok && lane == leader) {
            pre = cnt[warp][dgt];
            cnt[warp][dgt] = pre + __popc(peers);
        }
        pre = __shfl_sync(0xffffffffu, pre, leader);
        rank[it] = pre + __popc(peers & lt);
        __syncwarp();
    }
    __syncthreads();
    {   // thread t = digit t: exclusive prefix over warps + global base of (digit, tile)
        const int dgt = threadIdx.x;
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) {
            const uint32_t c = cnt[w][dgt];
            cnt[w][dgt] = run;
            run += c;
        }
        // digit base = exclusive prefix of the 256 digit totals (8 warp scans + 8 partials)
        const uint32_t tot = digit_totals[dgt];
        uint32_t inc = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += y;
        }
        if (lane == 31) gbase[warp] = inc;            // gbase doubles as scratch for the warp totals
        __syncthreads();
        uint32_t pre = 0;
        for (int w = 0; w < warp; ++w) pre += gbase[w];
        __syncthreads();
        gbase[dgt] = pre + inc - tot + table[size_t(dgt) * num_tiles + blockIdx.x];
    }
    __syncthreads();
#pragma unroll
    for (int it = 0; it < RS_ITEMS; ++it) {
        const int64_t i = wbase + it * 32 + lane;
        if (i < count) {
            const uint32_t dgt = uint32_t((key[it] >> shift) & 255u);
            const uint32_t pos = gbase[dgt] + cnt[warp][dgt] + rank[it];
            keys_out[pos] = key[it];
            if (HAS_VALS) vals_out[pos] = vals[i];
        }
    }
}

size_t sort_table_entries(int64_t count) { return size_t(256) * size_t(ceil_div(count, RS_TILE)) + 256; }

// Sorts ascending on bits [0, bits).  Returns (through *sorted_keys / *sorted_vals)
// which of the two ping-pong buffers holds the result.
int radix_sort_u64(uint64_t* keys, uint64_t* keys_tmp, uint32_t* vals, uint32_t* vals_tmp, int64_t count,
                   int bits, uint32_t* table, uint64_t** sorted_keys, uint32_t** sorted_vals, cudaStream_t st) {
    *sorted_keys = keys;
    if (sorted_vals) *sorted_vals = vals;
    if (count <= 1) return HG_OK;
    HG_REQUIRE(count < (int64_t(1) << 32), HG_ERR_UNSUPPORTED, "radix sort: more than 2^32 keys");
    const int num_tiles = int(ceil_div(count, RS_TILE));
    uint64_t* src = keys; uint64_t* dst = keys_tmp;
    uint32_t* vsrc = vals; uint32_t* vdst = vals_tmp;
    uint32_t* digit_totals = table + size_t(256) * num_tiles;
    for (int shift = 0; shift < bits; shift += 8) {
        rs_hist_kernel<<<num_tiles, RS_THREADS, 0, st>>>(src, count, shift, table, num_tiles);
        rs_scan_rows_kernel<<<256, 256, 0, st>>>(table, num_tiles, digit_totals);
        if (vals)
            rs_scatter_kernel<true><<<num_tiles, RS_THREADS, 0, st>>>(src, vsrc, count, shift, table, digit_totals, num_tiles, dst, vdst);
        else
            rs_scatter_kernel<false><<<num_tiles, RS_THREADS, 0, st>>>(src, nullptr, count, shift, table, digit_totals, num_tiles, dst, nullptr);
        HG_CHECK_LAUNCH();
        count_launches(3);
        uint64_t* t = src; src = dst; dst = t;
        uint32_t* tv = vsrc; vsrc = vdst; vdst = tv;
    }
    *sorted_keys = src;
    if (sorted_vals) *sorted_vals = vsrc;
    return HG_OK;
}

// ---- device-wide exclusive scan of uint32 -> uint64 totals --------------------------
// flags[i] = 1 where keys[i] starts a new run; per-tile totals -> tile_sums
__global__ void __launch_bounds__(SC_THREADS)
unique_count_kernel(const uint64_t* __restrict__ keys, int64_t count, uint64_t sentinel,
                    uint32_t* __restrict__ tile_sums) {
    __shared__ uint32_t sm[33];
    const int64_t base = int64_t(blockIdx.x) * SC_TILE + int64_t(threadIdx.x) * SC_ITEMS;
    uint32_t c = 0;
#pragma unroll
    for (int q = 0; q < SC_ITEMS; ++q) {
        const int64_t i = base + q;
        if (i < count) c += (keys[i] != sentinel && (i == 0 || keys[i] != keys[i - 1])) ? 1u : 0u;
    }
    uint32_t total;
    block_exclusive_scan(c, &total, sm);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(SC_THREADS)
unique_scatter_kernel(const uint64_t* __restrict__ keys, int64_t count, uint64_t sentinel,
                      const uint32_t* __restrict__ tile_offsets, uint64_t* __restrict__ out) {
    __shared__ uint32_t sm[33];
    const int64_t base = int64_t(blockIdx.x) * SC_TILE + int64_t(threadIdx.x) * SC_ITEMS;
    uint32_t flag[SC_ITEMS];
    uint64_t kv[SC_ITEMS];
    uint32_t c = 0;
#pragma unroll
    for (int q = 0; q < SC_ITEMS; ++q) {
        const int64_t i = base + q;
        flag[q] = 0;
        if (i < count) {
            kv[q] = keys[i];
            flag[q] = (kv[q] != sentinel && (i == 0 || kv[q] != keys[i - 1])) ? 1u : 0u;
        }
        c += flag[q];
    }
    uint32_t total;
    uint32_t pos = block_exclusive_scan(c, &total, sm) + tile_offsets[blockIdx.x];
#pragma unroll
    for (int q = 0; q < SC_ITEMS; ++q)
        if (flag[q]) out[pos++] = kv[q];
}

// exclusive scan of tile sums (in place); grand total written to *total_out (uint64)
__global__ void __launch_bounds__(1024)
scan_sums_kernel(uint32_t* __restrict__ sums, int64_t len, unsigned long long* __restrict__ total_out) {
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t carry_s;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t base = 0; base < len; base += 1024) {
        const int64_t i = base + threadIdx.x;
        const uint32_t v = (i < len) ? sums[i] : 0u;
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = warp_tot[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t y = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += y;
            }
            warp_tot[lane] = w;
        }
        __syncthreads();
        const uint32_t carry = carry_s;
        if (i < len) sums[i] = (warp ? warp_tot[warp - 1] : 0u) + carry + x - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = carry + warp_tot[31];
        __syncthreads();
    }
    if (threadIdx.x == 0 && total_out) *total_out = carry_s;
}

size_t unique_tiles(int64_t count) { return size_t(ceil_div(count, SC_TILE)); }

// out <- distinct keys of the sorted run `keys`; *count_out (device) <- how many
int scan_tile_sums(uint32_t* sums, int64_t len, unsigned long long* total_out, cudaStream_t st) {
    scan_sums_kernel<<<1, 1024, 0, st>>>(sums, len, total_out);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
}

int unique_sorted_u64(const uint64_t* keys, int64_t count, uint64_t sentinel, uint64_t* out, uint32_t* tile_sums,
                      unsigned long long* count_out, cudaStream_t st) {
    if (count == 0) {
        HG_CHECK_CUDA(cudaMemsetAsync(count_out, 0, sizeof(unsigned long long), st));
        return HG_OK;
    }
    const int tiles = int(unique_tiles(count));
    unique_count_kernel<<<tiles, SC_THREADS, 0, st>>>(keys, count, sentinel, tile_sums);
    scan_sums_kernel<<<1, 1024, 0, st>>>(tile_sums, tiles, count_out);
    unique_scatter_kernel<<<tiles, SC_THREADS, 0, st>>>(keys, count, sentinel, tile_sums, out);
    HG_CHECK_LAUNCH();
    count_launches(3);
    return HG_OK;
}

}  // namespace hg
