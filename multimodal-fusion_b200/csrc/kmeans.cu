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
#include "knn_tc.cuh"
#include "rowdot.cuh"
#include "sort_scan.cuh"
#include <algorithm>
#include <stdlib.h>

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


// The same distances for the one-call fit: the (few) candidate rows are staged in shared memory once per
// CTA, every row of x is read once with 128-bit loads, packed FP32 pairs.  HBM-bound: 4 d bytes per row.
template <int TMAX>
__global__ void __launch_bounds__(256, TMAX <= 8 ? 3 : 1)
km_cand_d2_smem_kernel(const float* __restrict__ x, int64_t n, int64_t d, const int64_t* __restrict__ cand, int t,
                       int64_t rows_per_block, float* __restrict__ out) {
    extern __shared__ __align__(16) unsigned char km_smem_raw[];
    ulonglong2* cs = reinterpret_cast<ulonglong2*>(km_smem_raw);          // [t][d / 4]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t nv = d >> 2;
    for (int64_t i = threadIdx.x; i < int64_t(TMAX) * nv; i += blockDim.x) {       // slots t .. TMAX-1: zeros
        const int q = int(i / nv);
        cs[i] = (q < t) ? __ldg(reinterpret_cast<const ulonglong2*>(x + cand[q] * d) + (i - q * nv)) : make_ulonglong2(0ull, 0ull);
    }
    __syncthreads();
    const int64_t r0 = int64_t(blockIdx.x) * rows_per_block, r1 = min64(n, r0 + rows_per_block);
    // a warp takes two rows per pass: every candidate value read from shared memory serves both
    for (int64_t r = r0 + 2 * warp; r < r1; r += 16) {
        const bool two = r + 1 < r1;
        uint64_t acc0[TMAX], acc1[TMAX];
#pragma unroll
        for (int q = 0; q < TMAX; ++q) { acc0[q] = 0ull; acc1[q] = 0ull; }
        const ulonglong2* xr0 = reinterpret_cast<const ulonglong2*>(x + r * d);
        const ulonglong2* xr1 = reinterpret_cast<const ulonglong2*>(x + (two ? r + 1 : r) * d);
        for (int64_t v0 = lane; v0 < nv; v0 += 64) {                // two 16-byte loads of each row in flight per lane
            ulonglong2 a0[2], a1[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const bool ok = v0 + 32 * u < nv;
                a0[u] = ok ? __ldg(xr0 + v0 + 32 * u) : make_ulonglong2(0ull, 0ull);
                a1[u] = ok ? __ldg(xr1 + v0 + 32 * u) : make_ulonglong2(0ull, 0ull);
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                if (v0 + 32 * u >= nv) break;
#pragma unroll
                for (int q = 0; q < TMAX; ++q) {
                    const ulonglong2 b = cs[q * nv + v0 + 32 * u];
                    uint64_t e0 = pk_sub(a0[u].x, b.x), e1 = pk_sub(a0[u].y, b.y);
                    acc0[q] = pk_fma(e0, e0, acc0[q]); acc0[q] = pk_fma(e1, e1, acc0[q]);
                    e0 = pk_sub(a1[u].x, b.x); e1 = pk_sub(a1[u].y, b.y);
                    acc1[q] = pk_fma(e0, e0, acc1[q]); acc1[q] = pk_fma(e1, e1, acc1[q]);
                }
            }
        }
#pragma unroll
        for (int q = 0; q < TMAX; ++q) {
            if (q < t) {
                const float s0 = warp_butterfly(pk_hsum(acc0[q]));
                const float s1 = warp_butterfly(pk_hsum(acc1[q]));
                if (lane == 0) {
                    out[int64_t(q) * n + r] = s0;
                    if (two) out[int64_t(q) * n + r + 1] = s1;
                }
            }
        }
    }
}

// =================================================================================================
// One-call fit (hg_kmeans_fit): every restart's k-means++ seeding and Lloyd loop stays on the device.
// The host only polls a "converged" flag every KM_POLL iterations.  All sums run in a fixed order
// (per-warp slabs, per-row-block partials, trees of fixed shape), so a fit is reproducible bit for
// bit for a given seed -- what `random_state=42` buys in the reference.
// =================================================================================================
constexpr int KM_TILE = 1024;     // elements per block of the O(n) seeding kernels
constexpr int KM_POLL = 4;        // Lloyd iterations enqueued between two reads of the flag

struct KmState {                  // device scalars of a fit (one 256-byte slot of the workspace)
    double tol_abs, best_inertia;
    unsigned long long moved;
    int done, n_iter, best_t, take;
};

__device__ __forceinline__ uint64_t km_mix(uint64_t z) {        // splitmix64 finaliser
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ uint64_t km_rand(uint64_t seed, int run, int step, int t) {
    return km_mix(km_mix(seed ^ (uint64_t(uint32_t(run)) << 40)) ^ (uint64_t(uint32_t(step)) << 8) ^ uint64_t(uint32_t(t)));
}

// block-wide sum of doubles in a fixed tree (deterministic); result valid in thread 0
__device__ __forceinline__ double km_block_sum(double v, double* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[warp] = v;
    __syncthreads();
    double s = 0.0;
    if (threadIdx.x == 0) for (int w = 0; w < nw; ++w) s += sh[w];
    return s;
}

// column sums / sums of squares of x, per row block: part[rb][0][j] = sum, part[rb][1][j] = sum of squares
__global__ void __launch_bounds__(256)
km_colstats_kernel(const float* __restrict__ x, int64_t n, int64_t d, int rows_per, double* __restrict__ part) {
    __shared__ double sh[8][2][KM_DS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t r0 = int64_t(blockIdx.x) * rows_per, r1 = min64(n, r0 + rows_per);
    const int64_t col = int64_t(blockIdx.y) * KM_DS + lane;
    double s = 0.0, q = 0.0;
    if (col < d)
        for (int64_t r = r0 + warp; r < r1; r += 8) { const double v = __ldg(x + r * d + col); s += v; q += v * v; }
    sh[warp][0][lane] = s; sh[warp][1][lane] = q;
    __syncthreads();
    if (warp < 2 && col < d) {
        double a = 0.0;
        for (int w = 0; w < 8; ++w) a += sh[w][warp][lane];
        part[(int64_t(blockIdx.x) * 2 + warp) * d + col] = a;
    }
}

// tol_abs = tol * mean_j var(x[:, j])   (scikit-learn's _tolerance)
__global__ void km_tol_kernel(const double* __restrict__ part, int nrb, int64_t n, int64_t d, double tol,
                              KmState* __restrict__ st, float* __restrict__ mean_out) {
    __shared__ double sh[32];
    double acc = 0.0;
    for (int64_t j = threadIdx.x; j < d; j += blockDim.x) {
        double s = 0.0, q = 0.0;
        for (int rb = 0; rb < nrb; ++rb) { s += part[(int64_t(rb) * 2) * d + j]; q += part[(int64_t(rb) * 2 + 1) * d + j]; }
        const double mean = s / double(n);
        acc += fmax(q / double(n) - mean * mean, 0.0);
        mean_out[j] = float(mean);
    }
    const double tot = km_block_sum(acc, sh);
    if (threadIdx.x == 0) { st->tol_abs = tol * tot / double(d); st->best_inertia = 0.0; }
}

// xc = x - column means: distances do not change, but the FP32 expansion |x|^2 + |c|^2 - 2 x.c and the FP16
// split stop cancelling on data that sits far from the origin (scikit-learn's fit centres X for the same reason;
// without it nearly constant data, e.g. a similarity matrix with entries around 0.98, never stops moving labels)
__global__ void __launch_bounds__(256)
km_center_kernel(const float* __restrict__ x, int64_t n, int64_t d, const float* __restrict__ mean, int rows_per,
                 float* __restrict__ xc) {
    const int64_t r0 = int64_t(blockIdx.x) * rows_per, r1 = min64(n, r0 + rows_per);
    for (int64_t j = threadIdx.x; j < d; j += blockDim.x) {
        const float m = mean[j];
        for (int64_t r = r0; r < r1; ++r) xc[r * d + j] = x[r * d + j] - m;
    }
}

// start of a restart: flags cleared, first centre drawn uniformly
__global__ void km_seed_first_kernel(KmState* __restrict__ st, int64_t* __restrict__ cand, int64_t* __restrict__ center_idx,
                                     uint64_t seed, int run, int64_t n) {
    st->done = 0; st->n_iter = 0; st->moved = 0ull; st->best_t = 0; st->take = 0;
    const int64_t first = int64_t(km_rand(seed, run, 0, 0) % uint64_t(n));
    cand[0] = first; center_idx[0] = first;
}

// closest <- cand_d2[best_t] (first) or min(closest, cand_d2[best_t]); tile_sums[b] = sum of the tile
__global__ void __launch_bounds__(256)
km_pp_apply_kernel(float* __restrict__ closest, const float* __restrict__ cand_d2, int64_t n,
                   const KmState* __restrict__ st, int first, double* __restrict__ tile_sums) {
    __shared__ double sh[8];
    const float* src = cand_d2 + int64_t(st->best_t) * n;
    const int64_t base = int64_t(blockIdx.x) * KM_TILE;
    double acc = 0.0;
#pragma unroll
    for (int q = 0; q < KM_TILE / 256; ++q) {
        const int64_t i = base + q * 256 + threadIdx.x;
        if (i < n) {
            float v = src[i];
            if (!first) v = fminf(v, closest[i]);
            closest[i] = v;
            acc += double(v);
        }
    }
    const double tot = km_block_sum(acc, sh);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

// D^2 sampling: warp t draws candidate t with probability closest[i] / sum(closest) -- the first i whose
// running sum reaches r = u * total (searchsorted on the cumulative sum, as scikit-learn's _kmeans_plusplus)
__global__ void km_pp_sample_kernel(const double* __restrict__ tile_sums, int64_t ntiles, const float* __restrict__ closest,
                                    int64_t n, int trials, uint64_t seed, int run, int step, int64_t* __restrict__ cand) {
    const int lane = threadIdx.x & 31, t = threadIdx.x >> 5;
    if (t >= trials) return;
    double total = 0.0;
    for (int64_t b0 = 0; b0 < ntiles; b0 += 32) {
        double v = (b0 + lane < ntiles) ? tile_sums[b0 + lane] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        total += v;
    }
    const double u = double(km_rand(seed, run, step, t) >> 11) * (1.0 / 9007199254740992.0);
    const double r = u * total;
    // tile whose inclusive running sum first reaches r
    double before = 0.0;
    int64_t tile = ntiles - 1;
    bool found = false;
    for (int64_t b0 = 0; b0 < ntiles && !found; b0 += 32) {
        const double v = (b0 + lane < ntiles) ? tile_sums[b0 + lane] : 0.0;
        double inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const double y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
        const unsigned hit = __ballot_sync(0xffffffffu, (b0 + lane < ntiles) && before + inc >= r);
        if (hit) {
            const int l = __ffs(hit) - 1;
            tile = b0 + l;
            before += __shfl_sync(0xffffffffu, inc - v, l);
            found = true;
        } else {
            before += __shfl_sync(0xffffffffu, inc, 31);
        }
    }
    int64_t pick = n - 1;                            // rounding left r above the last running sum: clip, as scikit-learn does
    if (found) {
        const int64_t base = tile * KM_TILE;
        pick = min64(n, base + KM_TILE) - 1;
        float vals[KM_TILE / 32];                    // all loads in flight at once: the scan below waits once, not 32 times
#pragma unroll
        for (int k = 0; k < KM_TILE / 32; ++k) {
            const int64_t i = base + 32 * k + lane;
            vals[k] = (i < n) ? closest[i] : 0.f;
        }
        bool got = false;
#pragma unroll
        for (int k = 0; k < KM_TILE / 32; ++k) {
            if (got) continue;                       // warp-uniform
            const int64_t i = base + 32 * k + lane;
            double inc = double(vals[k]);
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const double y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
            const unsigned hit = __ballot_sync(0xffffffffu, i < n && before + inc >= r);
            if (hit) { pick = base + 32 * k + (__ffs(hit) - 1); got = true; }
            else before += __shfl_sync(0xffffffffu, inc, 31);
        }
    }
    if (lane == 0) cand[t] = pick;
}

// pot_tile[t][b] = sum over tile b of min(closest, cand_d2[t])
__global__ void __launch_bounds__(256)
km_pp_potential_kernel(const float* __restrict__ closest, const float* __restrict__ cand_d2, int64_t n, int trials,
                       int64_t ntiles, double* __restrict__ pot_tile) {
    __shared__ double sh[8];
    const int64_t base = int64_t(blockIdx.x) * KM_TILE;
    float cl[KM_TILE / 256];
#pragma unroll
    for (int q = 0; q < KM_TILE / 256; ++q) {
        const int64_t i = base + q * 256 + threadIdx.x;
        cl[q] = (i < n) ? closest[i] : 0.f;
    }
    for (int t = 0; t < trials; ++t) {
        double acc = 0.0;
#pragma unroll
        for (int q = 0; q < KM_TILE / 256; ++q) {
            const int64_t i = base + q * 256 + threadIdx.x;
            if (i < n) acc += double(fminf(cl[q], cand_d2[int64_t(t) * n + i]));
        }
        const double tot = km_block_sum(acc, sh);
        if (threadIdx.x == 0) pot_tile[int64_t(t) * ntiles + blockIdx.x] = tot;
    }
}

// the candidate with the smallest potential becomes centre `step` (first minimum, as numpy's argmin)
__global__ void km_pp_select_kernel(const double* __restrict__ pot_tile, int64_t ntiles, int trials,
                                    const int64_t* __restrict__ cand, int step, int64_t* __restrict__ center_idx,
                                    KmState* __restrict__ st) {
    __shared__ double pot[16];
    const int lane = threadIdx.x & 31, t = threadIdx.x >> 5;
    if (t < trials) {
        double acc = 0.0;
        for (int64_t b0 = 0; b0 < ntiles; b0 += 32) {
            double v = (b0 + lane < ntiles) ? pot_tile[int64_t(t) * ntiles + b0 + lane] : 0.0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            acc += v;
        }
        if (lane == 0) pot[t] = acc;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int best = 0;
        for (int q = 1; q < trials; ++q) if (pot[q] < pot[best]) best = q;
        st->best_t = best;
        center_idx[step] = cand[best];
    }
}

__global__ void km_gather_centers_kernel(const float* __restrict__ x, int64_t d, const int64_t* __restrict__ center_idx,
                                         float* __restrict__ centers) {
    const float* src = x + center_idx[blockIdx.x] * d;
    for (int64_t j = threadIdx.x; j < d; j += blockDim.x) centers[int64_t(blockIdx.x) * d + j] = src[j];
}

// Lloyd update, deterministic.  One stable radix sort of (label, row) makes every cluster a contiguous run in
// ascending row order (the plan of aggregate.cu); a CTA then sums one slice of one cluster, whole rows at a time
// (256 threads x float4 = 4 KB per member row, 8 rows in flight), in registers -- no atomics, fixed order.
__global__ void __launch_bounds__(256)
km_keys_kernel(const int32_t* __restrict__ labels, int32_t* __restrict__ prev, int64_t n, int c,
               uint64_t* __restrict__ keys, uint32_t* __restrict__ rows, KmState* __restrict__ st) {
    const bool live = st->done == 0;
    unsigned int moved = 0;
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        const int l = labels[i];
        keys[i] = (l >= 0 && l < c) ? uint64_t(l) : uint64_t(c);         // strays sort last
        rows[i] = uint32_t(i);
        if (live) { moved += (prev[i] != l) ? 1u : 0u; prev[i] = l; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) moved += __shfl_xor_sync(0xffffffffu, moved, o);
    if ((threadIdx.x & 31) == 0 && moved) atomicAdd(&st->moved, (unsigned long long)moved);
}

// partial[z][cc][:] = sum of the rows in slice z of cluster cc (grid (c, split))
template <bool VEC>
__global__ void __launch_bounds__(256)
km_segsum_kernel(const float* __restrict__ x, int64_t d, const uint32_t* __restrict__ order,
                 const uint32_t* __restrict__ start, int c, float* __restrict__ partial, const KmState* __restrict__ st) {
    if (st->done) return;
    const int cc = blockIdx.x, z = blockIdx.y, split = gridDim.y;
    const uint32_t s = start[cc], e = start[cc + 1];
    const uint32_t per = (e - s + split - 1) / split;
    const uint32_t lo = min(e, s + z * per), hi = min(e, lo + per);
    float* out = partial + (int64_t(z) * c + cc) * d;
    if (VEC) {
        const int64_t nv = d >> 2;
        for (int64_t v = threadIdx.x; v < nv; v += blockDim.x) {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            uint32_t q = lo;
            for (; q + 8 <= hi; q += 8) {
                float4 t[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) t[u] = __ldg(reinterpret_cast<const float4*>(x + int64_t(order[q + u]) * d) + v);
#pragma unroll
                for (int u = 0; u < 8; ++u) { acc.x += t[u].x; acc.y += t[u].y; acc.z += t[u].z; acc.w += t[u].w; }
            }
            for (; q < hi; ++q) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(x + int64_t(order[q]) * d) + v);
                acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
            }
            reinterpret_cast<float4*>(out)[v] = acc;
        }
    } else {
        for (int64_t col = threadIdx.x; col < d; col += blockDim.x) {
            float acc = 0.f;
            for (uint32_t q = lo; q < hi; ++q) acc += __ldg(x + int64_t(order[q]) * d + col);
            out[col] = acc;
        }
    }
}

// centre cc <- mean of its members (empty clusters keep their centre here; km_relocate_kernel moves them);
// shift_part[cc] = |new - old|^2
__global__ void __launch_bounds__(128)
km_finalize_det_kernel(const float* __restrict__ partial, const uint32_t* __restrict__ start, int split, int c,
                       int64_t d, float* __restrict__ centers, double* __restrict__ shift_part,
                       const KmState* __restrict__ st) {
    if (st->done) return;
    __shared__ double sh[4];
    const int cc = blockIdx.x;
    const unsigned int cnt = start[cc + 1] - start[cc];
    double local = 0.0;
    if (cnt > 0) {
        for (int64_t j = threadIdx.x; j < d; j += blockDim.x) {
            double s = 0.0;
            for (int z = 0; z < split; ++z) s += double(partial[(int64_t(z) * c + cc) * d + j]);
            const float nv = float(s / double(cnt));
            const float ov = centers[int64_t(cc) * d + j];
            local += double(nv - ov) * double(nv - ov);
            centers[int64_t(cc) * d + j] = nv;
        }
    }
    const double tot = km_block_sum(local, sh);
    if (threadIdx.x == 0) shift_part[cc] = tot;
}

// scikit-learn's _relocate_empty_clusters_dense (cluster/_k_means_common.pyx): the e-th empty cluster takes the point
// that lies e-th farthest from its own centre (distances of this iteration's assignment), and that point leaves
// the mean of the cluster it was counted in.  One block; a no-op (one pass over `start`) unless a cluster is
// empty, which is rare -- but without it a restart that loses a cluster keeps it lost, and the reference pipeline
// (preprocess_hypergraph.py:164) raises "Cluster i is empty" on slides scikit-learn handles.
constexpr int KM_RELOC_MAX_C = 4096;
__global__ void __launch_bounds__(1024)
km_relocate_kernel(const float* __restrict__ x, int64_t n, int64_t d, const int32_t* __restrict__ labels,
                   float* __restrict__ d2, const uint32_t* __restrict__ start, int c, float* __restrict__ centers,
                   double* __restrict__ shift_part, KmState* __restrict__ st) {
    if (st->done) return;
    __shared__ int s_any, s_far;
    __shared__ float s_bv[32];
    __shared__ int s_bi[32];
    __shared__ double s_sum[32];
    __shared__ unsigned short s_removed[KM_RELOC_MAX_C];
    if (threadIdx.x == 0) s_any = 0;
    for (int q = threadIdx.x; q < KM_RELOC_MAX_C; q += blockDim.x) s_removed[q] = 0;
    __syncthreads();
    for (int q = threadIdx.x; q < c; q += blockDim.x)
        if (start[q + 1] == start[q]) s_any = 1;
    __syncthreads();
    if (!s_any) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int e = 0; e < c; ++e) {
        if (start[e + 1] != start[e]) continue;                      // block-uniform
        float bv = -1.f; int bi = 0x7fffffff;                       // largest distance, lowest index among equals
        for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
            const float v = d2[i];
            if (v > bv) { bv = v; bi = int(i); }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { s_bv[warp] = bv; s_bi[warp] = bi; }
        __syncthreads();
        if (warp == 0) {
            bv = s_bv[lane]; bi = s_bi[lane];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
            }
            if (lane == 0) s_far = (bv >= 0.f) ? bi : -1;
        }
        __syncthreads();
        const int far = s_far;
        if (far < 0) break;                                           // every point already taken
        const int old = labels[far];
        int cnt_old = int(start[old + 1] - start[old]);
        if (old < KM_RELOC_MAX_C) cnt_old -= int(s_removed[old]);
        double local = 0.0;
        for (int64_t j = threadIdx.x; j < d; j += blockDim.x) {
            const float nv = x[int64_t(far) * d + j];
            const float ov = centers[int64_t(e) * d + j];
            local += double(nv - ov) * double(nv - ov);
            centers[int64_t(e) * d + j] = nv;
            if (cnt_old > 1) {
                const double m = double(centers[int64_t(old) * d + j]);
                centers[int64_t(old) * d + j] = float((m * double(cnt_old) - double(nv)) / double(cnt_old - 1));
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
        if (lane == 0) s_sum[warp] = local;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int w = 0; w < 32; ++w) t += s_sum[w];
            shift_part[e] = t;
            d2[far] = -1.f;                                           // taken
            if (old < KM_RELOC_MAX_C) s_removed[old] += 1;
            st->moved = 1ull;                                         // the labels will change: not converged
        }
        __syncthreads();
    }
}

__global__ void km_test_far_centre_kernel(float* __restrict__ centre, int64_t d) {
    for (int64_t j = threadIdx.x; j < d; j += blockDim.x) centre[j] = 1.0e6f;
}

// scikit-learn's two stopping tests: no label moved (strict convergence) or centre shift <= tol
__global__ void km_check_kernel(const double* __restrict__ shift_part, int c, KmState* __restrict__ st) {
    if (st->done) return;
    double s = 0.0;
    for (int q = 0; q < c; ++q) s += shift_part[q];
    st->n_iter += 1;
    if (s <= st->tol_abs || st->moved == 0ull) st->done = 1;
    st->moved = 0ull;
}

__global__ void __launch_bounds__(256)
km_tile_sums_kernel(const float* __restrict__ v, int64_t n, double* __restrict__ tile_sums) {
    __shared__ double sh[8];
    const int64_t base = int64_t(blockIdx.x) * KM_TILE;
    double acc = 0.0;
#pragma unroll
    for (int q = 0; q < KM_TILE / 256; ++q) {
        const int64_t i = base + q * 256 + threadIdx.x;
        if (i < n) acc += double(v[i]);
    }
    const double tot = km_block_sum(acc, sh);
    if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

// inertia of the finished restart; it becomes the result if it is the first or strictly better
__global__ void km_best_kernel(const double* __restrict__ tile_sums, int64_t ntiles, int run, KmState* __restrict__ st,
                               double* __restrict__ inertia_out, int32_t* __restrict__ n_iter_out) {
    double s = 0.0;
    for (int64_t b = 0; b < ntiles; ++b) s += tile_sums[b];
    const int take = (run == 0 || s < st->best_inertia) ? 1 : 0;
    st->take = take;
    if (take) {
        st->best_inertia = s;
        *inertia_out = s;
        if (n_iter_out) *n_iter_out = st->n_iter;
    }
}

__global__ void km_take_kernel(const KmState* __restrict__ st, const int32_t* __restrict__ labels, int64_t n,
                               const float* __restrict__ centers, int64_t cd, int64_t d, const float* __restrict__ mean,
                               int32_t* __restrict__ labels_out, float* __restrict__ centers_out) {
    if (!st->take) return;
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < n) labels_out[i] = labels[i];
    if (i < cd) centers_out[i] = centers[i] + mean[i % d];        // centres of the centred data, moved back
}

// Tensor-core assignment, second half: the fused kernel (knn_tc.cu, rectangular launch) left every row the
// KL centres with the smallest split-FP16 (3-product) scores |c|^2 - 2 x.c.  Every candidate whose score lies within the
// guaranteed error slack of the best one is recomputed as sum (x - c)^2 in FP32 (almost always one), and the
// smallest exact distance wins, ties to the lower centre.  One warp per row.
__global__ void __launch_bounds__(256, 4)
km_pick_kernel(const float* __restrict__ x, const float* __restrict__ centers, int64_t n, int64_t d,
               const float* __restrict__ part_val, const int32_t* __restrict__ part_idx, int kl,
               const float* __restrict__ norms_x, const float* __restrict__ norm_max,
               int32_t* __restrict__ labels, float* __restrict__ d2_out) {
    const int lane = threadIdx.x & 31;
    const int64_t r = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    if (r >= n) return;
    int idx = -1;
    float sc = INFINITY;
    if (lane < kl) {
        idx = part_idx[r * kl + lane];
        if (idx >= 0) sc = part_val[r * kl + lane];
    }
    float best = sc;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = fminf(best, __shfl_xor_sync(0xffffffffu, best, o));
    const float slack = tc_slack(round_up(d, kColPad), 3, 3) * (norms_x[r] + __ldg(norm_max));
    unsigned todo = __ballot_sync(0xffffffffu, idx >= 0 && sc <= best + slack);
    const float* xi = x + r * d;
    const bool vec = rows_vectorisable(x, d) && rows_vectorisable(centers, d);
    float win_d2 = INFINITY;
    int win = -1;
    while (todo) {
        const int l = __ffs(todo) - 1;
        todo &= todo - 1;
        const int j = __shfl_sync(0xffffffffu, idx, l);
        const float* cj = centers + int64_t(j) * d;
        float acc;
        if (vec) {
            uint64_t ad = 0;
            const int64_t nv = d >> 2;
            for (int64_t v0 = lane; v0 < nv; v0 += 128) {           // four loads of each row in flight, sums in index order
                ulonglong2 a[4], b[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const bool ok = v0 + 32 * u < nv;
                    a[u] = ok ? __ldg(reinterpret_cast<const ulonglong2*>(xi) + v0 + 32 * u) : make_ulonglong2(0ull, 0ull);
                    b[u] = ok ? *(reinterpret_cast<const ulonglong2*>(cj) + v0 + 32 * u) : make_ulonglong2(0ull, 0ull);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const uint64_t e0 = pk_sub(a[u].x, b[u].x), e1 = pk_sub(a[u].y, b[u].y);
                    ad = pk_fma(e0, e0, ad); ad = pk_fma(e1, e1, ad);
                }
            }
            acc = pk_hsum(ad);
        } else {
            acc = 0.f;
            for (int64_t c = lane; c < d; c += 32) { const float df = __ldg(xi + c) - cj[c]; acc = fmaf(df, df, acc); }
        }
        acc = warp_butterfly(acc);
        if (acc < win_d2 || (acc == win_d2 && j < win)) { win_d2 = acc; win = j; }
    }
    if (lane == 0) { labels[r] = win; d2_out[r] = (win >= 0) ? win_d2 : 0.f; }
}

struct KmFitPlan {
    int rows_per, nrb, split, trials;
    int64_t ntiles;
    size_t bytes;
    // assignment on the tensor cores (rectangular launch of the fused kNN kernel) from this much work up
    bool tc;
    int64_t n_pad_x, c_pad, d_pad;
    TcPlan tcplan;
    size_t tc_bytes;
};

// HG_B200_KMEANS_TC=0 / 1 forces the FP32 SIMT / the tensor-core assignment (tests, tuning)
static bool km_use_tc(int64_t n, int64_t d, int c) {
    if (const char* e = getenv("HG_B200_KMEANS_TC")) return atoi(e) != 0;
    return double(n) * double(d) * double(c) >= double(int64_t(1) << 27);
}

static KmFitPlan km_fit_plan(int64_t n, int64_t d, int c) {
    KmFitPlan pl{};
    pl.rows_per = (n >= 32768) ? 2048 : 512;
    pl.nrb = int(ceil_div(n, pl.rows_per));
    pl.split = int(std::max<int64_t>(1, std::min<int64_t>(32, ceil_div(int64_t(kNumSMs) * 4, c))));   // slices per cluster
    pl.trials = std::min(16, 2 + int(log(double(c))));   // scikit-learn: 2 + int(ln k)
    pl.ntiles = ceil_div(n, KM_TILE);
    auto al = [](size_t b) { return (b + 1023) & ~size_t(1023); };
    const size_t part = std::max(size_t(pl.split) * c * d * 4, size_t(pl.nrb) * 2 * d * 8);
    pl.bytes = al(sizeof(KmState)) + al(part) + 2 * al(size_t(n) * 8) + 2 * al(size_t(n) * 4) /* sort keys, rows */ +
               al(sort_table_entries(n) * 4) + al((size_t(c) + 2) * 4) + al(size_t(c) * 8) /* shift_part */ +
               3 * al(size_t(n) * 4) /* cur, prev, d2 */ + al(size_t(n) * 4) /* closest */ +
               al(size_t(16) * n * 4) /* cand_d2 */ + al(size_t(pl.ntiles) * 8) + al(size_t(16) * pl.ntiles * 8) +
               al(16 * 8) + al(size_t(c) * 8) + al(size_t(c) * d * 4) + al(size_t(d) * 4) + al(size_t(n) * d * 4) /* mean, xc */ + 1024;
    pl.tc = km_use_tc(n, d, c);
    if (pl.tc) {
        pl.n_pad_x = ops_rows(n); pl.c_pad = ops_rows(c); pl.d_pad = ops_cols(d);
        pl.tcplan = tc_make_plan_rect(c, n, 1, pl.d_pad);
        pl.tc_bytes = al(hg_operand_bytes(n, d)) + al(hg_operand_bytes(c, d)) + al(size_t(pl.n_pad_x) * 4) +
                      al(size_t(pl.c_pad) * 4) + al(size_t(pl.c_pad) * 8) + al(16) +
                      2 * al(size_t(n) * pl.tcplan.kl * 4);
        pl.bytes += pl.tc_bytes;
    }
    return pl;
}

// workspace carving with the 1024-byte alignment the TMA operand buffers want
struct KmCarver {
    char* base; size_t off;
    explicit KmCarver(void* p) : base(static_cast<char*>(p)), off(0) {}
    template <typename T> T* take(size_t count) {
        T* p = reinterpret_cast<T*>(base + off);
        off += (count * sizeof(T) + 1023) & ~size_t(1023);
        return p;
    }
};

struct KmAssign {                 // everything one assignment pass needs
    SimArgs sim;                  // FP32 SIMT path (and the shapes / outputs of both paths)
    bool tc;
    op16_t *ops_x, *ops_c;
    float *norms_x, *norms_c, *norm_max;
    float2* colp;
    float* part_val; int32_t* part_idx;
    int64_t n_pad_x, c_pad, d_pad;
    TcPlan plan;
};

// labels[i] = nearest centre, d2[i] = squared distance to it
static int km_assign(const KmAssign& a, cudaStream_t st) {
    if (!a.tc) return launch_topk_fp32(a.sim, st);
    const int64_t n = a.sim.m, d = a.sim.d;
    const int c = int(a.sim.n);
    int rc = launch_split(a.sim.b, c, d, a.ops_c, a.norms_c, st);
    if (rc != HG_OK) return rc;
    HG_CHECK_CUDA(cudaMemsetAsync(a.norm_max, 0, 4, st));
    rc = launch_colparams(a.norms_c, nullptr, 0, c, a.c_pad, HG_METRIC_EUCLIDEAN, a.colp, nullptr, st, nullptr, nullptr,
                          a.norm_max);
    if (rc != HG_OK) return rc;
    TcParams tp{};
    tp.n = c; tp.row_begin = 0; tp.row_end = n; tp.d_pad = int(a.d_pad);
    tp.plan = a.plan; tp.colp = a.colp; tp.part_val = a.part_val; tp.part_idx = a.part_idx;
    tp.norms_a = a.norms_x; tp.norms_b = a.norms_c; tp.prods = 3;
    rc = tc_launch_rect(a.ops_x, a.n_pad_x, a.ops_c, a.c_pad, tp, st);
    if (rc != HG_OK) return rc;
    km_pick_kernel<<<(unsigned)ceil_div(n * 32, 256), 256, 0, st>>>(a.sim.a, a.sim.b, n, d, a.part_val, a.part_idx, a.plan.kl,
                                                                    a.norms_x, a.norm_max, a.sim.out_idx, a.sim.out_val);
    HG_CHECK_LAUNCH();
    count_launches(1);
    return HG_OK;
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

// candidate distances of the seeding: shared-memory kernel when the rows are 16-byte aligned float4 runs and the
// candidates fit 96 KB, else the plain kernel
static int km_launch_cand_d2(const float* x, int64_t n, int64_t d, const int64_t* cand, int t, float* out, cudaStream_t st) {
    const size_t smem = size_t(t <= 8 ? 8 : 16) * d * 4;
    const bool vec = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
    if (vec && smem <= 96 * 1024) {
        const int64_t grid = std::max<int64_t>(1, std::min<int64_t>(ceil_div(n, 16), int64_t(kNumSMs) * 6));
        const int64_t rows_per_block = ceil_div(n, grid);
        if (t <= 8) {
            if (smem > 48 * 1024)
                HG_CHECK_CUDA(cudaFuncSetAttribute(km_cand_d2_smem_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
            km_cand_d2_smem_kernel<8><<<(unsigned)grid, 256, smem, st>>>(x, n, d, cand, t, rows_per_block, out);
        } else {
            if (smem > 48 * 1024)
                HG_CHECK_CUDA(cudaFuncSetAttribute(km_cand_d2_smem_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
            km_cand_d2_smem_kernel<16><<<(unsigned)grid, 256, smem, st>>>(x, n, d, cand, t, rows_per_block, out);
        }
    } else {
        kmeans_candidate_d2_kernel<<<(unsigned)ceil_div(n * 32, 256), 256, 0, st>>>(x, n, d, cand, t, out);
    }
    HG_CHECK_LAUNCH();
    return HG_OK;
}

extern "C" int hg_kmeans_fit_workspace_bytes(int64_t n, int64_t d, int c, size_t* bytes) {
    HG_REQUIRE(bytes && n > 0 && d > 0 && c > 0, HG_ERR_BAD_ARG, "hg_kmeans_fit_workspace_bytes: bad argument");
    *bytes = km_fit_plan(n, d, c).bytes;
    return HG_OK;
}

// KMeans(n_clusters=c, n_init, max_iter, tol, random_state=seed).fit(x): greedy k-means++ seeding,
// Lloyd iterations until no label moves or the squared centre shift is <= tol * mean(var(x)), best
// restart by inertia.  labels [n] int32, centroids [c, d], *inertia (double) and *n_iter (int32, may be
// NULL) are device buffers.  The call blocks on `stream` while it polls the convergence flag.
extern "C" int hg_kmeans_fit(const float* x, int64_t n, int64_t d, int c, int n_init, int max_iter, double tol,
                             uint64_t seed, int32_t* labels, float* centroids, double* inertia, int32_t* n_iter,
                             void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(x && labels && centroids && inertia && n > 0 && d > 0 && c > 0 && n_init > 0 && max_iter > 0 && tol >= 0.0,
               HG_ERR_BAD_ARG, "hg_kmeans_fit: bad argument");
    HG_REQUIRE(n >= c, HG_ERR_BAD_ARG, "n_samples=%lld should be >= n_clusters=%d.", (long long)n, c);
    HG_REQUIRE(c <= 1536, HG_ERR_UNSUPPORTED, "hg_kmeans_fit: at most 1536 clusters");
    const KmFitPlan pl = km_fit_plan(n, d, c);
    HG_REQUIRE(workspace && workspace_bytes >= pl.bytes, HG_ERR_WORKSPACE, "hg_kmeans_fit: workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    HG_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 1023) == 0, HG_ERR_BAD_ARG, "hg_kmeans_fit: workspace must be 1024-byte aligned");
    KmCarver cv(workspace);
    KmState* state = cv.take<KmState>(1);
    const size_t part_bytes = std::max(size_t(pl.split) * c * d * 4, size_t(pl.nrb) * 2 * d * 8);
    float* part = reinterpret_cast<float*>(cv.take<char>(part_bytes));
    uint64_t* keys = cv.take<uint64_t>(n);
    uint64_t* keys_tmp = cv.take<uint64_t>(n);
    uint32_t* rows = cv.take<uint32_t>(n);
    uint32_t* rows_tmp = cv.take<uint32_t>(n);
    uint32_t* table = cv.take<uint32_t>(sort_table_entries(n));
    uint32_t* start = cv.take<uint32_t>(size_t(c) + 2);
    double* shift_part = cv.take<double>(c);
    int32_t* cur = cv.take<int32_t>(n);
    int32_t* prev = cv.take<int32_t>(n);
    float* d2 = cv.take<float>(n);
    float* closest = cv.take<float>(n);
    float* cand_d2 = cv.take<float>(size_t(16) * n);
    double* tile_sums = cv.take<double>(pl.ntiles);
    double* pot_tile = cv.take<double>(size_t(16) * pl.ntiles);
    int64_t* cand = cv.take<int64_t>(16);
    int64_t* center_idx = cv.take<int64_t>(c);
    float* centers = cv.take<float>(size_t(c) * d);
    float* mean = cv.take<float>(d);
    float* xc = cv.take<float>(size_t(n) * d);
    KmAssign ka{};
    ka.tc = pl.tc;
    if (pl.tc) {
        ka.ops_x = reinterpret_cast<op16_t*>(cv.take<char>(hg_operand_bytes(n, d)));
        ka.ops_c = reinterpret_cast<op16_t*>(cv.take<char>(hg_operand_bytes(c, d)));
        ka.norms_x = cv.take<float>(pl.n_pad_x);
        ka.norms_c = cv.take<float>(pl.c_pad);
        ka.colp = cv.take<float2>(pl.c_pad);
        ka.norm_max = cv.take<float>(4);
        ka.part_val = cv.take<float>(size_t(n) * pl.tcplan.kl);
        ka.part_idx = cv.take<int32_t>(size_t(n) * pl.tcplan.kl);
        ka.n_pad_x = pl.n_pad_x; ka.c_pad = pl.c_pad; ka.d_pad = pl.d_pad; ka.plan = pl.tcplan;
    }

    const bool vec = (d % 4 == 0);                      // the segmented sums read the centred copy, which is aligned
    const int label_bits = bits_for(uint64_t(c));
    const int key_blocks = int(min64(ceil_div(n, 256), int64_t(kNumSMs) * 8));
    const dim3 grid_rc((unsigned)pl.nrb, (unsigned)ceil_div(d, KM_DS));
    const unsigned nt = (unsigned)pl.ntiles;

    km_colstats_kernel<<<grid_rc, 256, 0, st>>>(x, n, d, pl.rows_per, reinterpret_cast<double*>(part));
    km_tol_kernel<<<1, 256, 0, st>>>(reinterpret_cast<const double*>(part), pl.nrb, n, d, tol, state, mean);
    km_center_kernel<<<(unsigned)ceil_div(n, 64), 256, 0, st>>>(x, n, d, mean, 64, xc);
    HG_CHECK_LAUNCH();
    count_launches(3);
    x = xc;                                                                 // everything below works on the centred rows
    if (pl.tc) {
        int rc = launch_split(x, n, d, ka.ops_x, ka.norms_x, st);          // the rows are split once per fit
        if (rc != HG_OK) return rc;
    }

    SimArgs& a = ka.sim;
    a.a = x; a.b = centers; a.m = n; a.n = c; a.d = d; a.metric = HG_METRIC_EUCLIDEAN; a.k = 1;
    a.row_begin = 0; a.row_end = n; a.exclude_self = 0; a.out_idx = cur; a.out_val = d2;

    // HG_B200_KMEANS_RELOCATE=0: empty clusters keep their centre (tests compare both behaviours)
    const char* reloc_env = getenv("HG_B200_KMEANS_RELOCATE");
    const bool relocate = !(reloc_env && atoi(reloc_env) == 0);
    const bool test_far_seed = getenv("HG_B200_KMEANS_TEST_FAR_SEED") != nullptr;           // tests/test_gpu_kmeans.py only
    for (int run = 0; run < n_init; ++run) {
        // ---- greedy k-means++ ----
        km_seed_first_kernel<<<1, 1, 0, st>>>(state, cand, center_idx, seed, run, n);
        { int rc = km_launch_cand_d2(x, n, d, cand, 1, cand_d2, st); if (rc != HG_OK) return rc; }
        km_pp_apply_kernel<<<nt, 256, 0, st>>>(closest, cand_d2, n, state, 1, tile_sums);
        count_launches(3);
        for (int step = 1; step < c; ++step) {
            km_pp_sample_kernel<<<1, 32 * pl.trials, 0, st>>>(tile_sums, pl.ntiles, closest, n, pl.trials, seed, run, step, cand);
            { int rc = km_launch_cand_d2(x, n, d, cand, pl.trials, cand_d2, st); if (rc != HG_OK) return rc; }
            km_pp_potential_kernel<<<nt, 256, 0, st>>>(closest, cand_d2, n, pl.trials, pl.ntiles, pot_tile);
            km_pp_select_kernel<<<1, 32 * pl.trials, 0, st>>>(pot_tile, pl.ntiles, pl.trials, cand, step, center_idx, state);
            km_pp_apply_kernel<<<nt, 256, 0, st>>>(closest, cand_d2, n, state, 0, tile_sums);
            count_launches(5);
        }
        km_gather_centers_kernel<<<c, 128, 0, st>>>(x, d, center_idx, centers);
        if (test_far_seed && c > 1)        // test hook: the last centre starts far from every point, its cluster is empty
            km_test_far_centre_kernel<<<1, 128, 0, st>>>(centers + int64_t(c - 1) * d, d);
        HG_CHECK_CUDA(cudaMemsetAsync(prev, 0xff, size_t(n) * 4, st));
        HG_CHECK_LAUNCH();
        count_launches(1);
        // ---- Lloyd ----
        for (int it = 0; it < max_iter;) {
            const int chunk = std::min(KM_POLL, max_iter - it);
            for (int q = 0; q < chunk; ++q) {
                int rc = km_assign(ka, st);
                if (rc != HG_OK) return rc;
                km_keys_kernel<<<key_blocks, 256, 0, st>>>(cur, prev, n, c, keys, rows, state);
                uint64_t* sk; uint32_t* order;
                rc = radix_sort_u64(keys, keys_tmp, rows, rows_tmp, n, label_bits, table, &sk, &order, st);
                if (rc != HG_OK) return rc;
                rc = launch_cluster_bounds(sk, n, c, start, st);
                if (rc != HG_OK) return rc;
                if (vec) km_segsum_kernel<true><<<dim3(c, pl.split), 256, 0, st>>>(x, d, order, start, c, part, state);
                else     km_segsum_kernel<false><<<dim3(c, pl.split), 256, 0, st>>>(x, d, order, start, c, part, state);
                km_finalize_det_kernel<<<c, 128, 0, st>>>(part, start, pl.split, c, d, centers, shift_part, state);
                if (relocate) km_relocate_kernel<<<1, 1024, 0, st>>>(x, n, d, cur, d2, start, c, centers, shift_part, state);
                km_check_kernel<<<1, 1, 0, st>>>(shift_part, c, state);
                count_launches(relocate ? 5 : 4);
            }
            HG_CHECK_LAUNCH();
            it += chunk;
            int done = 0;
            HG_CHECK_CUDA(cudaMemcpyAsync(&done, &state->done, sizeof(int), cudaMemcpyDeviceToHost, st));
            HG_CHECK_CUDA(cudaStreamSynchronize(st));
            if (done) break;
        }
        // ---- labels of the final centres, inertia, keep the best restart ----
        int rc = km_assign(ka, st);
        if (rc != HG_OK) return rc;
        km_tile_sums_kernel<<<nt, 256, 0, st>>>(d2, n, tile_sums);
        km_best_kernel<<<1, 1, 0, st>>>(tile_sums, pl.ntiles, run, state, inertia, n_iter);
        const int64_t cd = int64_t(c) * d;
        km_take_kernel<<<(unsigned)ceil_div(std::max<int64_t>(n, cd), 256), 256, 0, st>>>(state, cur, n, centers, cd, d, mean, labels, centroids);
        HG_CHECK_LAUNCH();
        count_launches(3);
    }
    return HG_OK;
}
