// Device-wide LSD radix sort of 64-bit keys (optional 32-bit payload), exclusive
// scan, and adjacent-unique compaction -- the building blocks of edge assembly
// (reference preprocess_hypergraph.py:403 `set(tuple(sorted(edge)))`).
//
// "Onesweep" organisation: one kernel histograms every digit position up front
// (digit totals do not depend on the order of the keys), then ONE kernel per 8-bit
// pass ranks and scatters.  A tile learns how many keys of each digit precede it
// from its predecessors through a chained look-back on per-(tile, digit) status
// words (aggregate -> inclusive prefix), so there is no separate scan kernel and
// the keys are read once per pass.  Stability inside a tile comes from
// warp-synchronous ranking: each warp owns a contiguous run of the tile, ranks its
// keys with __match_any_sync + popc, and per-warp digit totals are prefix-summed.
// HBM-bound: (1 + passes) reads and `passes` writes of the key array.
#include "common.cuh"
#include "sort_scan.cuh"

namespace hg {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_MAX_PASSES = 8;
constexpr int RS_SMALL_ITEMS = 4;      // 1024-key tiles: more CTAs when a slide has few edges
constexpr int RS_LARGE_ITEMS = 16;     // 4096-key tiles
constexpr int64_t RS_SMALL_COUNT = int64_t(1) << 19;
constexpr uint32_t RS_FLAG_AGG = 1u << 30, RS_FLAG_PREFIX = 2u << 30, RS_VALUE_MASK = (1u << 30) - 1u;

static int rs_items_for(int64_t count) { return count <= RS_SMALL_COUNT ? RS_SMALL_ITEMS : RS_LARGE_ITEMS; }
static int rs_tiles_for(int64_t count) { return int(ceil_div(count, int64_t(RS_THREADS) * rs_items_for(count))); }

// hist[pass][digit] += occurrences, for every pass at once
__global__ void __launch_bounds__(RS_THREADS)
rs_global_hist_kernel(const uint64_t* __restrict__ keys, int64_t count, int passes, uint32_t* __restrict__ hist) {
    __shared__ uint32_t h[RS_MAX_PASSES][256];
    for (int p = 0; p < passes; ++p) h[p][threadIdx.x] = 0;
    __syncthreads();
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < count; i += int64_t(gridDim.x) * blockDim.x) {
        const uint64_t key = keys[i];
        for (int p = 0; p < passes; ++p) atomicAdd(&h[p][(key >> (8 * p)) & 255u], 1u);
    }
    __syncthreads();
    for (int p = 0; p < passes; ++p)
        if (h[p][threadIdx.x]) atomicAdd(&hist[p * 256 + threadIdx.x], h[p][threadIdx.x]);
}

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_volatile_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <bool HAS_VALS, int ITEMS>
__global__ void __launch_bounds__(RS_THREADS)
rs_onesweep_kernel(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals, int64_t count, int shift,
                   const uint32_t* __restrict__ digit_hist,   // [256] totals of this pass
                   uint32_t* __restrict__ status,             // [tiles][256], zeroed; status[tiles*256] = tile ticket
                   int num_tiles, uint64_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out) {
    __shared__ uint32_t cnt[RS_WARPS][256];
    __shared__ uint32_t gbase[256];
    __shared__ uint32_t tile_s;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // tickets are handed out in launch order, so every predecessor tile is already running
    if (threadIdx.x == 0) tile_s = atomicAdd(status + size_t(num_tiles) * 256, 1u);
    for (int w = 0; w < RS_WARPS; ++w) cnt[w][threadIdx.x] = 0;
    __syncthreads();
    const uint32_t tile = tile_s;

    // warp `warp` owns keys [tile*TILE + warp*32*ITEMS, +32*ITEMS) in (it, lane) order
    const int64_t wbase = int64_t(tile) * (RS_THREADS * ITEMS) + int64_t(warp) * (32 * ITEMS);
    uint64_t key[ITEMS];
    uint32_t rank[ITEMS];
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int it = 0; it < ITEMS; ++it) {
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
    {   // thread t = digit t
        const int dgt = threadIdx.x;
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) {
            const uint32_t c = cnt[w][dgt];
            cnt[w][dgt] = run;                        // exclusive prefix over the warps of this tile
            run += c;
        }
        uint32_t* my = status + size_t(tile) * 256 + dgt;
        uint32_t excl = 0;
        if (tile == 0) {
            st_volatile_u32(my, RS_FLAG_PREFIX | run);
        } else {
            st_volatile_u32(my, RS_FLAG_AGG | run);
            for (int t = int(tile) - 1; t >= 0; --t) {            // chained look-back
                const uint32_t* q = status + size_t(t) * 256 + dgt;
                uint32_t s = ld_volatile_u32(q);
                while ((s & ~RS_VALUE_MASK) == 0u) { __nanosleep(20); s = ld_volatile_u32(q); }
                excl += s & RS_VALUE_MASK;
                if (s & RS_FLAG_PREFIX) break;
            }
            st_volatile_u32(my, RS_FLAG_PREFIX | (excl + run));
        }
        // digit base = exclusive prefix of the 256 digit totals (8 warp scans + 8 partials)
        const uint32_t tot = digit_hist[dgt];
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
        gbase[dgt] = pre + inc - tot + excl;
    }
    __syncthreads();
#pragma unroll
    for (int it = 0; it < ITEMS; ++it) {
        const int64_t i = wbase + it * 32 + lane;
        if (i < count) {
            const uint32_t dgt = uint32_t((key[it] >> shift) & 255u);
            const uint32_t pos = gbase[dgt] + cnt[warp][dgt] + rank[it];
            keys_out[pos] = key[it];
            if (HAS_VALS) vals_out[pos] = vals[i];
        }
    }
}

// uint32 scratch entries: per-pass digit totals + one status array (tiles x 256 + the ticket) per
// pass, so that a single memset prepares the whole sort
static size_t rs_status_entries(int64_t count) { return size_t(rs_tiles_for(count)) * 256 + 64; }
size_t sort_table_entries(int64_t count) {
    return size_t(RS_MAX_PASSES) * 256 + size_t(RS_MAX_PASSES) * rs_status_entries(count);
}

template <bool HAS_VALS>
static void launch_onesweep(int items, int tiles, const uint64_t* src, const uint32_t* vsrc, int64_t count, int shift,
                            const uint32_t* hist, uint32_t* status, uint64_t* dst, uint32_t* vdst, cudaStream_t st) {
    if (items == RS_SMALL_ITEMS)
        rs_onesweep_kernel<HAS_VALS, RS_SMALL_ITEMS><<<tiles, RS_THREADS, 0, st>>>(src, vsrc, count, shift, hist, status,
                                                                                   tiles, dst, vdst);
    else
        rs_onesweep_kernel<HAS_VALS, RS_LARGE_ITEMS><<<tiles, RS_THREADS, 0, st>>>(src, vsrc, count, shift, hist, status,
                                                                                   tiles, dst, vdst);
}

// Sorts ascending on bits [0, bits).  Returns (through *sorted_keys / *sorted_vals)
// which of the two ping-pong buffers holds the result.
int radix_sort_u64(uint64_t* keys, uint64_t* keys_tmp, uint32_t* vals, uint32_t* vals_tmp, int64_t count,
                   int bits, uint32_t* table, uint64_t** sorted_keys, uint32_t** sorted_vals, cudaStream_t st) {
    *sorted_keys = keys;
    if (sorted_vals) *sorted_vals = vals;
    if (count <= 1) return HG_OK;
    HG_REQUIRE(count < (int64_t(1) << 30), HG_ERR_UNSUPPORTED, "radix sort: more than 2^30 keys");
    const int passes = (bits + 7) / 8;
    HG_REQUIRE(passes <= RS_MAX_PASSES, HG_ERR_BAD_ARG, "radix sort: bad bit count %d", bits);
    const int items = rs_items_for(count), tiles = rs_tiles_for(count);
    uint32_t* hist = table;
    uint32_t* status0 = table + size_t(RS_MAX_PASSES) * 256;
    const size_t status_stride = rs_status_entries(count);
    HG_CHECK_CUDA(cudaMemsetAsync(table, 0, (size_t(RS_MAX_PASSES) * 256 + size_t(passes) * status_stride) * sizeof(uint32_t),
                                  st));
    const int hist_blocks = int(min64(ceil_div(count, 256 * 8), 148 * 4));
    rs_global_hist_kernel<<<hist_blocks, RS_THREADS, 0, st>>>(keys, count, passes, hist);
    count_launches(1);
    uint64_t* src = keys; uint64_t* dst = keys_tmp;
    uint32_t* vsrc = vals; uint32_t* vdst = vals_tmp;
    for (int p = 0; p < passes; ++p) {
        uint32_t* status = status0 + size_t(p) * status_stride;
        if (vals) launch_onesweep<true>(items, tiles, src, vsrc, count, 8 * p, hist + p * 256, status, dst, vdst, st);
        else      launch_onesweep<false>(items, tiles, src, nullptr, count, 8 * p, hist + p * 256, status, dst, nullptr, st);
        HG_CHECK_LAUNCH();
        count_launches(1);
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
