// Grouped back half: edge_index, cosine weights and CSR incidence of EVERY slide of a group of small slides from
// their neighbour lists, in ~22 launches per group instead of ~18 per slide.
//
// Reference: the per-slide body of build_hypergraph_knn_kmeans, preprocess_hypergraph.py:385-422 (pair loop, set
// de-duplication, cosine weights), run once per slide by the dataset loop :651-670.  A slide of a few thousand
// patches gives each of those stages a few microseconds of GPU work; launched slide by slide, a batch of 512
// slides was bound by the ~9,000 kernel launches of its back halves (97 ms of host enqueue against 61 ms in the
// fused tensor-core kernel).  Here the stages run ONCE over the whole group in group coordinates:
//   * slide s owns the rows [off_s, off_s + n_s) of the group's row space (off_s a multiple of 256, NP rows in
//     all; the layout of the grouped feature kNN, knn_api.cu), blk2slide[row / 256] names the slide of a row;
//   * pair key = (off_s + min(i, j)) << 14 | max(i, j)  (n_s <= 9,216 < 2^14): ONE radix sort orders all slides'
//     pairs slide by slide and by (min, max) inside a slide, ONE adjacent-unique pass de-duplicates them, the
//     slides' edge ranges are found by boundary detection, and the decode kernel writes every slide's compact
//     [2, E_s] edge_index, its weights and its edge count (on the device);
//   * CSR: one stable sort of (larger endpoint in group rows, edge id) over the group, row starts in group
//     coordinates, shifted by the slide's first edge when written out.
// Outputs equal the per-slide entry points' bit for bit (tests/test_gpu_enqueue.py).
#include "knn_tc.cuh"
#include "rowdot.cuh"
#include "sort_scan.cuh"
#include <string.h>
#include <vector>

namespace hg {

constexpr int GB_BSHIFT = 14;                 // bits of the in-slide column of a pair key
static_assert(GB_MAX_N <= (int64_t(1) << GB_BSHIFT), "in-slide columns must fit GB_BSHIFT bits");

struct GbDev {                                // one slide as the kernels see it (device table)
    int64_t n, off;
    const int32_t* knn_idx; const float* knn_cos;
    const int32_t* sidx; int64_t sk;          // spatial lists [n, sk] or null
    const float* feats;
    int64_t* edge_index; float* edge_weights; int64_t* num_edges;
    int64_t* rowptr; int64_t* col; int64_t* eid;
};

// ---- pairs ----------------------------------------------------------------------------------------------------
// thread t = (group row g, list slot q): the canonical pair key of the entry, or the sentinel
__global__ void gb_emit_kernel(const GbDev* __restrict__ tab, const int32_t* __restrict__ map, int64_t np, int k, int kt,
                               uint64_t sentinel, uint64_t* __restrict__ keys) {
    const int64_t total = np * kt;
    for (int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += int64_t(gridDim.x) * blockDim.x) {
        const int64_t g = t / kt;
        const int q = int(t - g * kt);
        const GbDev* sl = tab + map[g >> 8];
        const int64_t off = sl->off, n = sl->n;
        const int64_t i = g - off;
        uint64_t key = sentinel;
        if (i < n) {
            int64_t j = -1;
            if (q < k) {
                j = sl->knn_idx[i * k + q];
            } else {
                const int64_t sk = sl->sk;
                const int32_t* sidx = sl->sidx;
                if (sidx != nullptr && q - k < sk) j = sidx[i * sk + (q - k)];
            }
            if (j >= 0 && j < n && j != i)
                key = (uint64_t(off + min(i, j)) << GB_BSHIFT) | uint64_t(max(i, j));
        }
        keys[t] = key;
    }
}

__device__ __forceinline__ int gb_slide_of(const int32_t* __restrict__ map, uint64_t key) {
    return map[(key >> GB_BSHIFT) >> 8];
}

// first_edge[s] = position of the first unique pair of slide s (or of the next slide that has one), s in [0, S]
__global__ void gb_starts_kernel(const uint64_t* __restrict__ uniq, const unsigned long long* __restrict__ count,
                                 const int32_t* __restrict__ map, int num_slides, uint32_t* __restrict__ first_edge) {
    const int64_t eg = int64_t(*count);
    for (int64_t p = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; p <= eg; p += int64_t(gridDim.x) * blockDim.x) {
        const int s = (p < eg) ? gb_slide_of(map, uniq[p]) : num_slides;
        const int prev = (p > 0) ? gb_slide_of(map, uniq[p - 1]) : -1;
        for (int q = prev + 1; q <= s; ++q) first_edge[q] = uint32_t(p);
    }
}

__global__ void gb_counts_kernel(const GbDev* __restrict__ tab, int num_slides, const uint32_t* __restrict__ first_edge) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < num_slides) *tab[s].num_edges = int64_t(first_edge[s + 1] - first_edge[s]);
}

// ---- decode + weights -----------------------------------------------------------------------------------------
// Position p of the unique list -> slide s, local edge e = p - first_edge[s]: edge_index_s[e] / [E_s + e], and the
// weight max(0, cos) looked up in the lists of the smaller, then of the larger endpoint (the re-rank's by-product,
// rerank.cu); pairs that only the spatial lists hold are computed from the two feature rows by the whole warp, in
// the canonical order of rowdot.cuh -- the arithmetic of hg_edge_weights_from_knn (graph_ops.cu).
__global__ void __launch_bounds__(256)
gb_decode_kernel(const GbDev* __restrict__ tab, const int32_t* __restrict__ map, const uint64_t* __restrict__ uniq,
                 const unsigned long long* __restrict__ count, const uint32_t* __restrict__ first_edge, int k, int64_t d) {
    const int64_t eg = int64_t(*count);
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = (int64_t(gridDim.x) * blockDim.x) >> 5;
    for (int64_t base = warp0 * 32; base < eg; base += nwarps * 32) {
        const int64_t p = base + lane;
        bool found = true;
        int s = 0;
        int64_t a = 0, b = 0, e = 0;
        if (p < eg) {
            const uint64_t key = uniq[p];
            s = gb_slide_of(map, key);
            const GbDev* sl = tab + s;
            a = int64_t(key >> GB_BSHIFT) - sl->off;
            b = int64_t(key & ((uint64_t(1) << GB_BSHIFT) - 1));
            const int64_t fe = first_edge[s], es = int64_t(first_edge[s + 1]) - fe;
            e = p - fe;
            int64_t* ei = sl->edge_index;
            ei[e] = a;
            ei[es + e] = b;
            const int32_t* idx = sl->knn_idx;
            const float* cs = sl->knn_cos;
            found = false;
            float c = 0.f;
            for (int q = 0; q < k; ++q)
                if (__ldg(idx + a * k + q) == int32_t(b)) { c = __ldg(cs + a * k + q); found = true; break; }
            if (!found)
                for (int q = 0; q < k; ++q)
                    if (__ldg(idx + b * k + q) == int32_t(a)) { c = __ldg(cs + b * k + q); found = true; break; }
            if (found) sl->edge_weights[e] = fmaxf(0.f, c);
        }
        uint32_t pending = __ballot_sync(0xffffffffu, !found);
        while (pending) {
            const int l = __ffs(pending) - 1;
            pending &= pending - 1;
            const int ss = __shfl_sync(0xffffffffu, s, l);
            const int64_t ia = __shfl_sync(0xffffffffu, a, l), ib = __shfl_sync(0xffffffffu, b, l);
            const int64_t ee = __shfl_sync(0xffffffffu, e, l);
            const GbDev* sl = tab + ss;
            const float* feats = sl->feats;
            float dot, na, nb;
            warp_dot_norms(feats + ia * d, feats + ib * d, d, lane, rows_vectorisable(feats, d), dot, na, nb);
            if (lane == 0) sl->edge_weights[ee] = fmaxf(0.f, cosine_from(dot, na, nb));
        }
    }
}

// ---- CSR in group coordinates ------------------------------------------------------------------------------------
// keys2[p] = larger endpoint (group row) of pair p, payload p; slots [E, cap) carry the sentinel row NP.
// first_hi[x] = number of pairs whose smaller endpoint (group row) is < x, x in [0, NP].
__global__ void gb_csr_keys_kernel(const GbDev* __restrict__ tab, const int32_t* __restrict__ map,
                                   const uint64_t* __restrict__ uniq, const unsigned long long* __restrict__ count,
                                   int64_t cap, int64_t np, uint64_t* __restrict__ keys2, uint32_t* __restrict__ vals2,
                                   uint32_t* __restrict__ first_hi) {
    const int64_t eg = int64_t(*count);
    const int64_t span = cap > eg ? cap : eg + 1;
    for (int64_t p = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; p < span; p += int64_t(gridDim.x) * blockDim.x) {
        if (p < cap) {
            uint64_t kk = uint64_t(np);
            if (p < eg) {
                const uint64_t key = uniq[p];
                kk = uint64_t(tab[gb_slide_of(map, key)].off) + (key & ((uint64_t(1) << GB_BSHIFT) - 1));
            }
            keys2[p] = kk;
            vals2[p] = uint32_t(p);
        }
        if (p <= eg) {
            const int64_t node = (p < eg) ? int64_t(uniq[p] >> GB_BSHIFT) : np;
            const int64_t prev = (p > 0) ? int64_t(uniq[p - 1] >> GB_BSHIFT) : -1;
            for (int64_t r = prev + 1; r <= node; ++r) first_hi[r] = uint32_t(p);
        }
    }
}

// first_lo[x] = number of pairs whose larger endpoint is < x; rowg[x] = first_hi[x] + first_lo[x], x in [0, NP]
__global__ void gb_csr_bounds_kernel(const uint64_t* __restrict__ sorted_keys, const unsigned long long* __restrict__ count,
                                     int64_t np, const uint32_t* __restrict__ first_hi, uint32_t* __restrict__ first_lo,
                                     uint32_t* __restrict__ rowg) {
    const int64_t eg = int64_t(*count);
    for (int64_t p = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; p <= eg; p += int64_t(gridDim.x) * blockDim.x) {
        const int64_t node = (p < eg) ? int64_t(sorted_keys[p]) : np;
        const int64_t prev = (p > 0) ? int64_t(sorted_keys[p - 1]) : -1;
        for (int64_t r = prev + 1; r <= node; ++r) {
            first_lo[r] = uint32_t(p);
            rowg[r] = first_hi[r] + uint32_t(p);
        }
    }
}

// every slide's own rowptr: the group row starts minus twice the slide's first edge
__global__ void gb_rowptr_kernel(const GbDev* __restrict__ tab, const int32_t* __restrict__ map, int64_t np,
                                 const uint32_t* __restrict__ rowg, const uint32_t* __restrict__ first_edge) {
    for (int64_t g = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; g < np; g += int64_t(gridDim.x) * blockDim.x) {
        const int s = map[g >> 8];
        const GbDev* sl = tab + s;
        const int64_t off = sl->off, n = sl->n, r = g - off;
        int64_t* rowptr = sl->rowptr;
        if (rowptr == nullptr || r >= n) continue;
        const int64_t base = 2 * int64_t(first_edge[s]);
        rowptr[r] = int64_t(rowg[g]) - base;
        if (r == 0) rowptr[n] = int64_t(rowg[off + n]) - base;
    }
}

__global__ void gb_csr_fill_kernel(const GbDev* __restrict__ tab, const int32_t* __restrict__ map,
                                   const uint64_t* __restrict__ uniq, const unsigned long long* __restrict__ count,
                                   const uint64_t* __restrict__ sorted_keys, const uint32_t* __restrict__ sorted_ids,
                                   const uint32_t* __restrict__ first_hi, const uint32_t* __restrict__ first_lo,
                                   const uint32_t* __restrict__ rowg, const uint32_t* __restrict__ first_edge) {
    const int64_t eg = int64_t(*count);
    constexpr uint64_t BMASK = (uint64_t(1) << GB_BSHIFT) - 1;
    for (int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; t < eg; t += int64_t(gridDim.x) * blockDim.x) {
        {   // position t of the transposed run: pair id seen from its larger endpoint x
            const int64_t x = int64_t(sorted_keys[t]);
            const int64_t id = int64_t(sorted_ids[t]);
            const int s = map[x >> 8];
            const GbDev* sl = tab + s;
            const int64_t fe = first_edge[s];
            const int64_t dst = int64_t(rowg[x]) + (t - int64_t(first_lo[x])) - 2 * fe;
            sl->col[dst] = int64_t(uniq[id] >> GB_BSHIFT) - sl->off;
            sl->eid[dst] = id - fe;
        }
        {   // pair t = (u, v) seen from its smaller endpoint u, after u's first part
            const uint64_t key = uniq[t];
            const int64_t u = int64_t(key >> GB_BSHIFT);
            const int s = map[u >> 8];
            const GbDev* sl = tab + s;
            const int64_t fe = first_edge[s];
            const int64_t dst = int64_t(rowg[u]) + int64_t(first_lo[u + 1] - first_lo[u]) + (t - int64_t(first_hi[u])) - 2 * fe;
            sl->col[dst] = int64_t(key & BMASK);
            sl->eid[dst] = t - fe;
        }
    }
}

// ---- host side ---------------------------------------------------------------------------------------------------
struct GbLayout {
    int64_t np, m;                 // group rows, pair slots (np * kt)
    int kt;
    std::vector<int64_t> off;
    size_t o_keys, o_tmp, o_uniq, o_tiles, o_count, o_table, o_vals, o_vtmp, o_fe, o_fhi, o_flo, o_rowg, o_tab, o_map, bytes;
};

static GbLayout gb_layout(const GbSlide* slides, int num, int k, int sk) {
    GbLayout g{};
    g.kt = k + sk;
    g.off.resize(size_t(num));
    int64_t rows = 0;
    for (int i = 0; i < num; ++i) { g.off[size_t(i)] = rows; rows += ops_rows(slides[i].n); }
    g.np = rows;
    g.m = rows * g.kt;
    size_t off = 0;
    auto take = [&](size_t b) { size_t o = off; off += (b + 255) & ~size_t(255); return o; };
    g.o_keys = take(size_t(g.m) * 8);          // pair keys, later the CSR sort keys
    g.o_tmp = take(size_t(g.m) * 8);
    g.o_uniq = take(size_t(g.m) * 8);
    g.o_tiles = take(unique_tiles(g.m) * 4);
    g.o_count = take(8);
    g.o_table = take(sort_table_entries(g.m) * 4);
    g.o_vals = take(size_t(g.m) * 4);
    g.o_vtmp = take(size_t(g.m) * 4);
    g.o_fe = take(size_t(num + 1) * 4);
    g.o_fhi = take(size_t(g.np + 2) * 4);
    g.o_flo = take(size_t(g.np + 2) * 4);
    g.o_rowg = take(size_t(g.np + 2) * 4);
    g.o_tab = take(size_t(num) * sizeof(GbDev));
    g.o_map = take(size_t(g.np / kRowPad) * 4);
    g.bytes = off;
    return g;
}

size_t group_back_workspace_bytes(const GbSlide* slides, int num, int k, int spatial_k) {
    return gb_layout(slides, num, k, spatial_k).bytes;
}

int group_back_enqueue(const GbSlide* slides, int num, int64_t d, int k, int spatial_k, void* workspace,
                       size_t workspace_bytes, cudaStream_t st) {
    HG_REQUIRE(slides && num > 0 && k >= 1 && spatial_k >= 0 && d > 0, HG_ERR_BAD_ARG, "group_back: bad argument");
    bool want_csr = slides[0].rowptr != nullptr;
    for (int i = 0; i < num; ++i) {
        const GbSlide& s = slides[i];
        HG_REQUIRE(s.n > 1 && s.n <= GB_MAX_N && s.feats && s.knn_idx && s.knn_cos && s.edge_index && s.edge_weights &&
                   s.num_edges, HG_ERR_BAD_ARG, "group_back: slide %d: null buffer or unsupported size", i);
        HG_REQUIRE((s.rowptr != nullptr) == want_csr && (!want_csr || (s.col && s.eid)), HG_ERR_BAD_ARG,
                   "group_back: CSR buffers for all slides of a group or for none");
    }
    const GbLayout g = gb_layout(slides, num, k, spatial_k);
    HG_REQUIRE(g.m < (int64_t(1) << 30), HG_ERR_UNSUPPORTED, "group_back: group too large");
    HG_REQUIRE(workspace && workspace_bytes >= g.bytes, HG_ERR_WORKSPACE, "group_back: workspace too small (%zu < %zu)",
               workspace_bytes, g.bytes);
    HG_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, HG_ERR_BAD_ARG, "group_back: workspace must be 256-byte aligned");
    char* ws = static_cast<char*>(workspace);
    uint64_t* keys = reinterpret_cast<uint64_t*>(ws + g.o_keys);
    uint64_t* tmp = reinterpret_cast<uint64_t*>(ws + g.o_tmp);
    uint64_t* uniq = reinterpret_cast<uint64_t*>(ws + g.o_uniq);
    uint32_t* tiles = reinterpret_cast<uint32_t*>(ws + g.o_tiles);
    unsigned long long* count = reinterpret_cast<unsigned long long*>(ws + g.o_count);
    uint32_t* table = reinterpret_cast<uint32_t*>(ws + g.o_table);
    uint32_t* vals = reinterpret_cast<uint32_t*>(ws + g.o_vals);
    uint32_t* vtmp = reinterpret_cast<uint32_t*>(ws + g.o_vtmp);
    uint32_t* first_edge = reinterpret_cast<uint32_t*>(ws + g.o_fe);
    uint32_t* first_hi = reinterpret_cast<uint32_t*>(ws + g.o_fhi);
    uint32_t* first_lo = reinterpret_cast<uint32_t*>(ws + g.o_flo);
    uint32_t* rowg = reinterpret_cast<uint32_t*>(ws + g.o_rowg);
    GbDev* d_tab = reinterpret_cast<GbDev*>(ws + g.o_tab);
    int32_t* d_map = reinterpret_cast<int32_t*>(ws + g.o_map);

    // slide table + row-block map: one pinned staging block laid out like the device side, one DMA
    StageSlot* stage = nullptr;
    const size_t tab_bytes = g.bytes - g.o_tab;
    int rc = stage_acquire(tab_bytes, &stage);
    if (rc != HG_OK) return rc;
    char* hs = static_cast<char*>(stage->host);
    GbDev* h_tab = reinterpret_cast<GbDev*>(hs);
    int32_t* h_map = reinterpret_cast<int32_t*>(hs + (g.o_map - g.o_tab));
    for (int i = 0; i < num; ++i) {
        const GbSlide& s = slides[i];
        const bool sp = s.spatial_idx != nullptr && spatial_k > 0;
        h_tab[i] = GbDev{s.n, g.off[size_t(i)], s.knn_idx, s.knn_cos, sp ? s.spatial_idx : nullptr, sp ? int64_t(spatial_k) : 0,
                         s.feats, s.edge_index, s.edge_weights, s.num_edges, s.rowptr, s.col, s.eid};
        for (int64_t b = g.off[size_t(i)] / kRowPad; b < (g.off[size_t(i)] + ops_rows(s.n)) / kRowPad; ++b) h_map[b] = i;
    }
    HG_CHECK_CUDA(cudaMemcpyAsync(d_tab, hs, tab_bytes, cudaMemcpyHostToDevice, st));
    rc = stage_release(stage, st);
    if (rc != HG_OK) return rc;

    const uint64_t sentinel = uint64_t(g.np) << GB_BSHIFT;
    const int blocks_m = int(min64(ceil_div(g.m, 256), 148 * 16));
    gb_emit_kernel<<<blocks_m, 256, 0, st>>>(d_tab, d_map, g.np, k, g.kt, sentinel, keys);
    HG_CHECK_LAUNCH();
    count_launches(1);
    uint64_t* sk = keys;
    rc = radix_sort_u64(keys, tmp, nullptr, nullptr, g.m, bits_for(sentinel), table, &sk, nullptr, st);
    if (rc != HG_OK) return rc;
    rc = unique_sorted_u64(sk, g.m, sentinel, uniq, tiles, count, st);
    if (rc != HG_OK) return rc;
    gb_starts_kernel<<<blocks_m, 256, 0, st>>>(uniq, count, d_map, num, first_edge);
    gb_counts_kernel<<<unsigned(ceil_div(num, 256)), 256, 0, st>>>(d_tab, num, first_edge);
    gb_decode_kernel<<<blocks_m, 256, 0, st>>>(d_tab, d_map, uniq, count, first_edge, k, d);
    HG_CHECK_LAUNCH();
    count_launches(3);
    if (!want_csr) return HG_OK;

    // the pair-key buffers are free again: they carry the CSR sort
    gb_csr_keys_kernel<<<blocks_m, 256, 0, st>>>(d_tab, d_map, uniq, count, g.m, g.np, keys, vals, first_hi);
    HG_CHECK_LAUNCH();
    count_launches(1);
    uint64_t* sk2 = keys; uint32_t* sv2 = vals;
    rc = radix_sort_u64(keys, tmp, vals, vtmp, g.m, bits_for(uint64_t(g.np)), table, &sk2, &sv2, st);
    if (rc != HG_OK) return rc;
    gb_csr_bounds_kernel<<<blocks_m, 256, 0, st>>>(sk2, count, g.np, first_hi, first_lo, rowg);
    gb_rowptr_kernel<<<unsigned(ceil_div(g.np, 256)), 256, 0, st>>>(d_tab, d_map, g.np, rowg, first_edge);
    gb_csr_fill_kernel<<<blocks_m, 256, 0, st>>>(d_tab, d_map, uniq, count, sk2, sv2, first_hi, first_lo, rowg, first_edge);
    HG_CHECK_LAUNCH();
    count_launches(3);
    return HG_OK;
}

}  // namespace hg
