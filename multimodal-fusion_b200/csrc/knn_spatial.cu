// A3: spatial-coordinate kNN over patch (x, y[, z]) positions.
//
// Semantics pinned by the reference's spatial kernel K_g = exp(-lambda_g |g_i-g_j|^2)
// (similarity_kernel.py:57-86): the top-k of a K_g row is the k smallest squared
// distances.  Distances are taken in direct-difference FP32 (exact for integer
// grid coordinates), self is excluded by index, ties resolve by (d2, index)
// ascending -- on a patch grid 4 neighbours tie at d=1 and 4 at sqrt(2), so the
// rule matters and the oracle uses the same one.
//
// Two paths, same results bit for bit:
//  * uniform-grid path (2-D positions, n >= SP_GRID_MIN_N): the points are binned into square
//    cells of about SP_CELL_TARGET points (bounding box -> cell histogram -> exclusive scan ->
//    counting-sort scatter of packed (x, y, index) float4 records), then one thread per point, in
//    cell order, searches growing square rings of cells and stops as soon as its k-th best is
//    strictly closer than anything outside the searched square can be.  O(N) work, every
//    array is streamed once with 128-bit loads: HBM/L2-bandwidth-bound (N * (8 + 8k) compulsory
//    bytes) instead of N^2 distance evaluations;
//  * brute force (3-D positions, small slides, and the fallback when a cell holds more than
//    SP_CELL_MAX points, e.g. many duplicate coordinates): one query per thread, candidate
//    points staged through shared memory as SoA tiles read with broadcast LDS.128; the candidate
//    range is split across blockIdx.y so small slides still fill 148 SMs, and a second kernel
//    merges the per-split lists.
#include "common.cuh"
#include "sort_scan.cuh"
#include "topk_list.cuh"
#include <stdlib.h>

namespace hg {

constexpr int SP_THREADS = 128;
constexpr int SP_TILE = 2048;      // candidate points per shared-memory tile

template <int PD, int KL>
__global__ void __launch_bounds__(SP_THREADS)
spatial_knn_kernel(const float* __restrict__ pos, int64_t n, int64_t row_begin, int64_t row_end,
                   int64_t cols_per_split, float* __restrict__ part_val, int32_t* __restrict__ part_idx,
                   const int* __restrict__ guard) {
    __shared__ __align__(16) float tile[PD][SP_TILE];
    if (guard != nullptr && *guard == 0) return;      // fallback launch that is not needed (grid-uniform)
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
            for (int c = 0; c < PD; ++c) tile[c][t] = INFINITY;  // d2 = inf: never below the list's last entry
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
                                   float* __restrict__ out_val, int32_t* __restrict__ out_idx,
                                   const int* __restrict__ guard) {
    const int64_t r = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (r >= n_rows || (guard != nullptr && *guard == 0)) return;
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
                          int64_t cols_per_split, float* pv, int32_t* pi, int kl, const int* guard, cudaStream_t st) {
    dim3 grid((unsigned)ceil_div(r1 - r0, SP_THREADS), (unsigned)splits);
    if (kl == 8)       spatial_knn_kernel<PD, 8><<<grid, SP_THREADS, 0, st>>>(pos, n, r0, r1, cols_per_split, pv, pi, guard);
    else if (kl == 16) spatial_knn_kernel<PD, 16><<<grid, SP_THREADS, 0, st>>>(pos, n, r0, r1, cols_per_split, pv, pi, guard);
    else               spatial_knn_kernel<PD, 32><<<grid, SP_THREADS, 0, st>>>(pos, n, r0, r1, cols_per_split, pv, pi, guard);
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


// ================================ uniform-grid path (2-D) ================================
constexpr int64_t SP_GRID_MIN_N = 8192;    // below this the brute-force pass costs less than the binning launches
constexpr float SP_CELL_TARGET = 2.0f;     // aimed average number of points per cell
constexpr int SP_CELL_MAX = 1024;          // a fuller cell (duplicates, extreme clustering) -> brute-force fallback
constexpr int SP_GRID_DIM_MAX = 32000;     // cells per axis: keeps the FP32 rounding of a cell coordinate < 2^-7
constexpr int SP_SCAN_TILE = 2048;         // cells per block of the start-offset scan

// Device-resident header: the bounding box as order-preserving uint keys (mins stored complemented,
// so all four are atomicMax from 0) and the fallback flag.
struct SpGridHdr { uint32_t nxmin, xmax, nymin, ymax; int fallback; int pad[3]; };

__device__ __forceinline__ uint32_t f2key(float f) {           // monotone float -> uint
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key2f(uint32_t k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

struct SpGeom { float x0, y0, inv_s; int gx, gy; };

// Cell geometry from the bounding box; every thread of every kernel derives the same values.
// gx * gy <= cell_cap = 3 n / target + 2 by construction (s >= sqrt(w h target / n) and
// s >= max(w, h) target / n), enforced by the loop against rounding.
__device__ __forceinline__ SpGeom sp_geom(const SpGridHdr* h, int64_t n, float target, int64_t cell_cap) {
    SpGeom g;
    const float xmin = key2f(~h->nxmin), xmax = key2f(h->xmax), ymin = key2f(~h->nymin), ymax = key2f(h->ymax);
    g.x0 = xmin; g.y0 = ymin; g.inv_s = 0.f; g.gx = 1; g.gy = 1;
    const float w = xmax - xmin, ht = ymax - ymin;
    if (!(w >= 0.f) || !(ht >= 0.f) || w > 1e30f || ht > 1e30f) return g;      // non-finite box: one cell
    const float per = target / float(n);
    float s = fmaxf(sqrtf(w * per) * sqrtf(ht), fmaxf(w, ht) * per);
    s = fmaxf(s, fmaxf(w, ht) * (1.f / float(SP_GRID_DIM_MAX)));
    if (!(s > 0.f)) return g;                                                   // all points coincide
    for (int it = 0; it < 8; ++it) {
        const float inv = 1.f / s;
        const int gx = int(w * inv) + 1, gy = int(ht * inv) + 1;
        if (int64_t(gx) * gy <= cell_cap) { g.inv_s = inv; g.gx = gx; g.gy = gy; return g; }
        s *= 1.25f;
    }
    return g;                                                                   // unreachable in practice: one cell
}
__device__ __forceinline__ int sp_cell(float u, int gdim) { return min(max(int(u), 0), gdim - 1); }   // int(NaN) = 0

__global__ void sp_bbox_kernel(const float* __restrict__ pos, int64_t n, SpGridHdr* __restrict__ hdr) {
    uint32_t a = 0, b = 0, c = 0, d = 0;
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        const float2 p = __ldg(reinterpret_cast<const float2*>(pos) + i);
        if (p.x == p.x) { a = max(a, ~f2key(p.x)); b = max(b, f2key(p.x)); }   // NaN coordinates do not shape the box
        if (p.y == p.y) { c = max(c, ~f2key(p.y)); d = max(d, f2key(p.y)); }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a = max(a, __shfl_xor_sync(0xffffffffu, a, o)); b = max(b, __shfl_xor_sync(0xffffffffu, b, o));
        c = max(c, __shfl_xor_sync(0xffffffffu, c, o)); d = max(d, __shfl_xor_sync(0xffffffffu, d, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMax(&hdr->nxmin, a); atomicMax(&hdr->xmax, b); atomicMax(&hdr->nymin, c); atomicMax(&hdr->ymax, d);
    }
}

// cell of every point + its arrival rank inside the cell (the counting sort needs no second atomic)
__global__ void sp_hist_kernel(const float* __restrict__ pos, int64_t n, SpGridHdr* __restrict__ hdr, float target,
                               int64_t cell_cap, uint32_t* __restrict__ cnt, int32_t* __restrict__ cell_of,
                               uint32_t* __restrict__ rank_of) {
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const SpGeom g = sp_geom(hdr, n, target, cell_cap);
    const float2 p = __ldg(reinterpret_cast<const float2*>(pos) + i);
    const int c = sp_cell((p.y - g.y0) * g.inv_s, g.gy) * g.gx + sp_cell((p.x - g.x0) * g.inv_s, g.gx);
    const uint32_t r = atomicAdd(cnt + c, 1u);
    cell_of[i] = c;
    rank_of[i] = r;
    if (r >= uint32_t(SP_CELL_MAX)) hdr->fallback = 1;
}

// exclusive scan of the cell counts inside tiles of SP_SCAN_TILE cells (in place) + tile totals
__global__ void __launch_bounds__(SC_THREADS)
sp_scan_tiles_kernel(uint32_t* __restrict__ cnt, int64_t len, uint32_t* __restrict__ tile_sums) {
    __shared__ uint32_t sm[33];
    const int64_t base = int64_t(blockIdx.x) * SP_SCAN_TILE + int64_t(threadIdx.x) * SC_ITEMS;
    uint32_t v[SC_ITEMS], c = 0;
#pragma unroll
    for (int q = 0; q < SC_ITEMS; ++q) { v[q] = (base + q < len) ? cnt[base + q] : 0u; c += v[q]; }
    uint32_t total;
    uint32_t run = block_exclusive_scan(c, &total, sm);
#pragma unroll
    for (int q = 0; q < SC_ITEMS; ++q) {
        if (base + q < len) cnt[base + q] = run;
        run += v[q];
    }
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}
static_assert(SP_SCAN_TILE == SC_THREADS * SC_ITEMS, "one scan tile per block");

__device__ __forceinline__ uint32_t sp_start(const uint32_t* __restrict__ cell_start, const uint32_t* __restrict__ tile_off,
                                             int c) {
    return __ldg(cell_start + c) + __ldg(tile_off + (c / SP_SCAN_TILE));
}

// counting-sort scatter: records (x, y, original index) in cell order, 16 bytes each
__global__ void sp_scatter_kernel(const float* __restrict__ pos, int64_t n, const int32_t* __restrict__ cell_of,
                                  const uint32_t* __restrict__ rank_of, const uint32_t* __restrict__ cell_start,
                                  const uint32_t* __restrict__ tile_off, const SpGridHdr* __restrict__ hdr,
                                  float4* __restrict__ sorted) {
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n || hdr->fallback) return;
    const float2 p = __ldg(reinterpret_cast<const float2*>(pos) + i);
    sorted[sp_start(cell_start, tile_off, cell_of[i]) + rank_of[i]] = make_float4(p.x, p.y, __int_as_float(int(i)), 0.f);
}

__device__ __forceinline__ bool lex_less(float va, int ia, float vb, int ib) {
    return va < vb || (va == vb && ia < ib);
}
// (value, index)-ordered insertion: arrival order carries no meaning here.  Precondition: (s, idx) < last entry.
template <int KL>
__device__ __forceinline__ void list_insert_lex(float (&v)[KL], int (&ix)[KL], float s, int idx) {
#pragma unroll
    for (int q = KL - 1; q > 0; --q) {
        const bool from_prev = lex_less(s, idx, v[q - 1], ix[q - 1]);
        const bool here = lex_less(s, idx, v[q], ix[q]);
        const int ni = from_prev ? ix[q - 1] : (here ? idx : ix[q]);
        const float nv = from_prev ? v[q - 1] : (here ? s : v[q]);
        ix[q] = ni; v[q] = nv;
    }
    if (lex_less(s, idx, v[0], ix[0])) { v[0] = s; ix[0] = idx; }
}

// One thread per point in cell order (neighbouring threads read the same cells: L1 hits).  Ring r = the
// cells at Chebyshev distance r of the query's cell: its top and bottom rows are one contiguous record
// range each (cells are row-major), the rows between contribute their two end cells.
template <int KL>
__global__ void __launch_bounds__(128)
sp_query_kernel(const float4* __restrict__ sorted, const uint32_t* __restrict__ cell_start,
                const uint32_t* __restrict__ tile_off, const SpGridHdr* __restrict__ hdr, int64_t n, float target,
                int64_t cell_cap, int64_t row_begin, int64_t row_end, int k,
                int32_t* __restrict__ out_idx, float* __restrict__ out_d2) {
    const int64_t j = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (j >= n || hdr->fallback) return;
    const float4 me = __ldg(sorted + j);
    const int self = __float_as_int(me.z);
    if (self < row_begin || self >= row_end) return;
    const SpGeom g = sp_geom(hdr, n, target, cell_cap);
    const float ux = (me.x - g.x0) * g.inv_s, uy = (me.y - g.y0) * g.inv_s;
    const int cx = sp_cell(ux, g.gx), cy = sp_cell(uy, g.gy);
    float lv[KL];
    int li[KL];
#pragma unroll
    for (int t = 0; t < KL; ++t) { lv[t] = INFINITY; li[t] = 0x7fffffff; }

    for (int r = 0;; ++r) {
        const int xa = max(cx - r, 0), xb = min(cx + r, g.gx - 1);
        for (int row = max(cy - r, 0); row <= min(cy + r, g.gy - 1); ++row) {
            const bool full = (row == cy - r) || (row == cy + r);    // top / bottom row of the ring: all its cells
            for (int part = 0; part < (full ? 1 : 2); ++part) {
                int ca = xa, cb = xb;
                if (!full) {
                    ca = cb = part ? cx + r : cx - r;
                    if (ca < 0 || ca >= g.gx) continue;
                }
                const int c0 = row * g.gx;
                const uint32_t b = sp_start(cell_start, tile_off, c0 + ca), e = sp_start(cell_start, tile_off, c0 + cb + 1);
                for (uint32_t t = b; t < e; ++t) {
                    const float4 p = __ldg(sorted + t);
                    const float dx = me.x - p.x, dy = me.y - p.y;
                    const float d2 = fmaf(dy, dy, dx * dx);          // the brute-force kernel's order of operations
                    const int pi = __float_as_int(p.z);
                    if (pi != self && lex_less(d2, pi, lv[KL - 1], li[KL - 1])) list_insert_lex<KL>(lv, li, d2, pi);
                }
            }
        }
        const bool l_open = cx - r > 0, r_open = cx + r < g.gx - 1, t_open = cy - r > 0, b_open = cy + r < g.gy - 1;
        if (!(l_open || r_open || t_open || b_open)) break;          // the whole grid has been searched
        // Every unsearched point lies beyond an open side of the searched square; in cell units it is at
        // least `edge` away (FP32 rounding of a cell coordinate < 2^-7, taken off), so its computed d2
        // is >= bd^2 (1 - 2^-22).  Stop when the k-th best is strictly below that.
        float edge = INFINITY;
        if (l_open) edge = fminf(edge, ux - float(cx - r));
        if (r_open) edge = fminf(edge, float(cx + r + 1) - ux);
        if (t_open) edge = fminf(edge, uy - float(cy - r));
        if (b_open) edge = fminf(edge, float(cy + r + 1) - uy);
        float kth = INFINITY;
#pragma unroll
        for (int t = 0; t < KL; ++t) kth = (t == k - 1) ? lv[t] : kth;
        const float bd = (edge - 0.015625f) / g.inv_s;
        if (bd > 0.f && kth < bd * bd * 0.99999f) break;
    }
    const int64_t o = (int64_t(self) - row_begin) * k;
#pragma unroll
    for (int t = 0; t < KL; ++t) {
        if (t < k) {
            const bool have = li[t] != 0x7fffffff;
            out_idx[o + t] = have ? li[t] : -1;
            out_d2[o + t] = have ? lv[t] : NAN;
        }
    }
}

struct SpGridWs { size_t off_hdr, off_cnt, off_tiles, off_cell, off_rank, off_sorted, bytes; int64_t cell_cap, tiles; };

static float sp_cell_target() {            // tuning override: HG_B200_SP_CELL=<points per cell>
    if (const char* e = getenv("HG_B200_SP_CELL")) {
        const float v = float(atof(e));
        if (v >= 0.25f && v <= 64.f) return v;
    }
    return SP_CELL_TARGET;
}

static SpGridWs sp_grid_ws(int64_t n) {
    SpGridWs w{};
    w.cell_cap = int64_t(3.0 * double(n) / double(sp_cell_target())) + 2;
    w.tiles = ceil_div(w.cell_cap + 1, SP_SCAN_TILE);
    size_t off = 0;
    auto take = [&](size_t b) { size_t o = off; off += (b + 255) & ~size_t(255); return o; };
    w.off_hdr = take(sizeof(SpGridHdr));
    w.off_cnt = take(size_t(w.cell_cap + 1) * 4);            // hdr + cnt are zeroed by one memset
    w.off_tiles = take(size_t(w.tiles + 1) * 4);
    w.off_cell = take(size_t(n) * 4);
    w.off_rank = take(size_t(n) * 4);
    w.off_sorted = take(size_t(n) * 16);
    w.bytes = off;
    return w;
}

// HG_B200_SP_GRID=0 forces brute force, =1 forces the grid for any n (tests)
static bool sp_use_grid(int64_t n, int pos_dim) {
    if (pos_dim != 2 || n >= (int64_t(1) << 31) - 1) return false;
    if (const char* e = getenv("HG_B200_SP_GRID")) return atoi(e) != 0;
    return n >= SP_GRID_MIN_N;
}

static int launch_grid(const float* pos, int64_t n, int k, int kl, int64_t r0, int64_t r1, int32_t* out_idx,
                       float* out_d2, char* ws, const SpGridWs& w, cudaStream_t st) {
    SpGridHdr* hdr = reinterpret_cast<SpGridHdr*>(ws + w.off_hdr);
    uint32_t* cnt = reinterpret_cast<uint32_t*>(ws + w.off_cnt);
    uint32_t* tiles = reinterpret_cast<uint32_t*>(ws + w.off_tiles);
    int32_t* cell_of = reinterpret_cast<int32_t*>(ws + w.off_cell);
    uint32_t* rank_of = reinterpret_cast<uint32_t*>(ws + w.off_rank);
    float4* sorted = reinterpret_cast<float4*>(ws + w.off_sorted);
    const float target = sp_cell_target();
    HG_CHECK_CUDA(cudaMemsetAsync(ws + w.off_hdr, 0, w.off_tiles - w.off_hdr, st));
    const unsigned pb = (unsigned)ceil_div(n, 256);
    sp_bbox_kernel<<<(unsigned)min64(pb, 2 * kNumSMs), 256, 0, st>>>(pos, n, hdr);
    sp_hist_kernel<<<pb, 256, 0, st>>>(pos, n, hdr, target, w.cell_cap, cnt, cell_of, rank_of);
    sp_scan_tiles_kernel<<<(unsigned)w.tiles, SC_THREADS, 0, st>>>(cnt, w.cell_cap + 1, tiles);
    HG_CHECK_LAUNCH();
    count_launches(3);
    int rc = scan_tile_sums(tiles, w.tiles, nullptr, st);
    if (rc != HG_OK) return rc;
    sp_scatter_kernel<<<pb, 256, 0, st>>>(pos, n, cell_of, rank_of, cnt, tiles, hdr, sorted);
    const unsigned qb = (unsigned)ceil_div(n, 128);
    if (kl == 8)       sp_query_kernel<8><<<qb, 128, 0, st>>>(sorted, cnt, tiles, hdr, n, target, w.cell_cap, r0, r1, k, out_idx, out_d2);
    else if (kl == 16) sp_query_kernel<16><<<qb, 128, 0, st>>>(sorted, cnt, tiles, hdr, n, target, w.cell_cap, r0, r1, k, out_idx, out_d2);
    else               sp_query_kernel<32><<<qb, 128, 0, st>>>(sorted, cnt, tiles, hdr, n, target, w.cell_cap, r0, r1, k, out_idx, out_d2);
    HG_CHECK_LAUNCH();
    count_launches(2);
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
    if (hg::sp_use_grid(n, 2)) *bytes += hg::sp_grid_ws(n).bytes;      // the grid's arrays follow the brute-force lists
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
    HG_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 15) == 0, HG_ERR_BAD_ARG,
               "hg_knn_spatial: workspace must be 16-byte aligned");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int kl, splits; int64_t cps;
    const int64_t n_rows = row_end - row_begin;
    hg::spatial_plan(n, n_rows, k, &kl, &splits, &cps);
    hg::Carver cv(workspace);
    float* pv = cv.take<float>(size_t(splits) * n_rows * kl);
    int32_t* pi = cv.take<int32_t>(size_t(splits) * n_rows * kl);
    const int* guard = nullptr;
    if (hg::sp_use_grid(n, pos_dim) && (reinterpret_cast<uintptr_t>(pos) & 7) == 0) {     // float2 loads
        // grid path; the brute-force launches below then exit at once unless a cell overflowed
        const hg::SpGridWs w = hg::sp_grid_ws(n);
        char* gws = static_cast<char*>(workspace) + cv.off;
        rc = hg::launch_grid(pos, n, k, kl, row_begin, row_end, out_idx, out_d2, gws, w, st);
        if (rc != HG_OK) return rc;
        guard = &reinterpret_cast<const hg::SpGridHdr*>(gws + w.off_hdr)->fallback;
    }
    rc = (pos_dim == 2) ? hg::launch_spatial<2>(pos, n, k, row_begin, row_end, splits, cps, pv, pi, kl, guard, st)
                        : hg::launch_spatial<3>(pos, n, k, row_begin, row_end, splits, cps, pv, pi, kl, guard, st);
    if (rc != HG_OK) return rc;
    hg::merge_lists_kernel<<<(unsigned)hg::ceil_div(n_rows, 128), 128, 0, st>>>(pv, pi, n_rows, splits, kl, k,
                                                                                out_d2, out_idx, guard);
    HG_CHECK_LAUNCH();
    hg::count_launches(2);
    return HG_OK;
}
