// One-call per-slide build: feature kNN (+ spatial kNN) -> edge assembly -> cosine weights ->
// CSR.  It only composes the stage entry points of this library on one stream, so a slide costs
// one host call and one stream synchronisation (the edge count) instead of a dozen round trips
// through the host language -- what matters for batches of small slides (BASELINE config C4).
// Mirrors the body of the reference's build_hypergraph_knn_kmeans (preprocess_hypergraph.py:374-433).
#include "knn_tc.cuh"
#include "sort_scan.cuh"
#include <stdlib.h>
#include <string>
#include <thread>
#include <vector>

namespace hg {
static size_t al1k(size_t b) { return (b + 1023) & ~size_t(1023); }

struct BuildWs { size_t off_knn, off_spatial, off_edges, off_csr, off_cos, bytes; size_t knn, spatial, edges, csr; };

static int plan_build(int64_t n, int64_t d, int k, int spatial_k, int num_hyperedges, int64_t clique_pairs,
                      int64_t max_edges, bool want_csr, BuildWs* w) {
    int rc = hg_knn_feature_workspace_bytes(n, d, k, 0, n, &w->knn);
    if (rc != HG_OK) return rc;
    w->spatial = 0;
    if (spatial_k > 0) {
        rc = hg_knn_spatial_workspace_bytes(n, spatial_k, 0, n, &w->spatial);
        if (rc != HG_OK) return rc;
    }
    rc = hg_edges_workspace_bytes(n, n * int64_t(k) + n * int64_t(spatial_k), num_hyperedges, clique_pairs, &w->edges);
    if (rc != HG_OK) return rc;
    w->csr = 0;
    if (want_csr) {
        rc = hg_csr_workspace_bytes(n, max_edges, &w->csr);
        if (rc != HG_OK) return rc;
    }
    // the kNN scratch (operand planes, candidate lists) is dead once the lists exist: the edge
    // and CSR stages reuse the same bytes
    w->off_knn = 0;
    w->off_spatial = al1k(w->knn);
    const size_t stage1 = w->off_spatial + al1k(w->spatial);
    w->off_edges = 0;
    w->off_csr = al1k(w->edges);
    const size_t stage2 = w->off_csr + al1k(w->csr);
    // cos(f_i, f_j) of every feature-kNN entry lives through both stages (edge weights read it)
    w->off_cos = stage1 > stage2 ? stage1 : stage2;
    w->bytes = w->off_cos + al1k(size_t(n) * size_t(k) * 4);
    return HG_OK;
}
}  // namespace hg

using namespace hg;

extern "C" int hg_build_workspace_bytes(int64_t n, int64_t d, int k, int spatial_k, int num_hyperedges,
                                        int64_t clique_pairs, int64_t max_edges, int want_csr, size_t* bytes) {
    HG_REQUIRE(bytes && max_edges >= 0, HG_ERR_BAD_ARG, "hg_build_workspace_bytes: bad argument");
    BuildWs w{};
    int rc = plan_build(n, d, k, spatial_k, num_hyperedges, clique_pairs, max_edges, want_csr != 0, &w);
    if (rc != HG_OK) return rc;
    *bytes = w.bytes;
    return HG_OK;
}

static int build_back_half(const float* feats, int64_t n, int64_t d, const float* pos, int pos_dim, int k, int spatial_k,
                           const int32_t* labels, int num_hyperedges, int64_t clique_pairs,
                           const int32_t* knn_idx, const float* knn_cos, int32_t* spatial_idx, float* spatial_d2,
                           int64_t max_edges, int64_t* edge_index, float* edge_weights,
                           int64_t* csr_rowptr, int64_t* csr_col, int64_t* csr_eid,
                           int64_t* num_edges_out_host, int64_t* num_edges_out_dev,
                           char* ws, const BuildWs& w, void* stream);

// num_edges_out_host != null: synchronising form (compact outputs sized by the host-known count);
// else stream-ordered form: the count stays on the device (*num_edges_out_dev), nothing blocks.
static int build_impl(const float* feats, int64_t n, int64_t d, const float* pos, int pos_dim,
                      int metric, float lambda_h, float lambda_g, int k, int spatial_k,
                      const int32_t* labels, int num_hyperedges, int64_t clique_pairs,
                      int32_t* knn_idx, float* knn_val, int32_t* spatial_idx, float* spatial_d2,
                      int64_t max_edges, int64_t* edge_index, float* edge_weights,
                      int64_t* csr_rowptr, int64_t* csr_col, int64_t* csr_eid,
                      int64_t* num_edges_out_host, int64_t* num_edges_out_dev,
                      void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(feats && knn_idx && knn_val && edge_index && edge_weights && (num_edges_out_host || num_edges_out_dev),
               HG_ERR_BAD_ARG, "hg_build_knn_hypergraph: null buffer");
    HG_REQUIRE(n > 1, HG_ERR_BAD_ARG, "number of nodes must be greater than 1, got N=%lld", (long long)n);
    HG_REQUIRE(spatial_k == 0 || (pos && spatial_idx && spatial_d2), HG_ERR_BAD_ARG,
               "hg_build_knn_hypergraph: spatial_k > 0 needs positions and spatial output buffers");
    const bool want_csr = csr_rowptr != nullptr;
    HG_REQUIRE(!want_csr || (csr_col && csr_eid), HG_ERR_BAD_ARG, "hg_build_knn_hypergraph: incomplete CSR buffers");
    if (labels == nullptr) { num_hyperedges = 0; clique_pairs = 0; }
    const int64_t bound = n * int64_t(k) + n * int64_t(spatial_k) + clique_pairs;
    HG_REQUIRE(max_edges >= bound, HG_ERR_BAD_ARG,
               "hg_build_knn_hypergraph: max_edges %lld below the bound n*(k+spatial_k)+clique_pairs = %lld",
               (long long)max_edges, (long long)bound);
    BuildWs w{};
    int rc = plan_build(n, d, k, spatial_k, num_hyperedges, clique_pairs, max_edges, want_csr, &w);
    if (rc != HG_OK) return rc;
    HG_REQUIRE(workspace && workspace_bytes >= w.bytes, HG_ERR_WORKSPACE,
               "hg_build_knn_hypergraph: workspace too small (%zu < %zu)", workspace_bytes, w.bytes);
    char* ws = static_cast<char*>(workspace);
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    float* knn_cos = reinterpret_cast<float*>(ws + w.off_cos);
    rc = hg_knn_feature_cos(feats, n, d, pos, pos_dim, metric, lambda_h, lambda_g, k, 0, n, 1, 0, knn_idx, knn_val,
                            knn_cos, ws + w.off_knn, w.knn, stream);
    if (rc != HG_OK) return rc;
    return build_back_half(feats, n, d, pos, pos_dim, k, spatial_k, labels, num_hyperedges, clique_pairs, knn_idx, knn_cos,
                           spatial_idx, spatial_d2, max_edges, edge_index, edge_weights, csr_rowptr, csr_col, csr_eid,
                           num_edges_out_host, num_edges_out_dev, ws, w, stream);
}

// everything behind the feature kNN: spatial kNN, edge assembly, weights, CSR (knn_cos: cos of every list entry)
static int build_back_half(const float* feats, int64_t n, int64_t d, const float* pos, int pos_dim, int k, int spatial_k,
                           const int32_t* labels, int num_hyperedges, int64_t clique_pairs,
                           const int32_t* knn_idx, const float* knn_cos, int32_t* spatial_idx, float* spatial_d2,
                           int64_t max_edges, int64_t* edge_index, float* edge_weights,
                           int64_t* csr_rowptr, int64_t* csr_col, int64_t* csr_eid,
                           int64_t* num_edges_out_host, int64_t* num_edges_out_dev,
                           char* ws, const BuildWs& w, void* stream) {
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool want_csr = csr_rowptr != nullptr;
    const int64_t bound = n * int64_t(k) + n * int64_t(spatial_k) + clique_pairs;
    int rc = HG_OK;
    const int32_t* lists[2] = {knn_idx, nullptr};
    int ks[2] = {k, 0};
    int num_lists = 1;
    if (spatial_k > 0) {
        rc = hg_knn_spatial(pos, n, pos_dim, spatial_k, 0, n, spatial_idx, spatial_d2, ws + w.off_spatial, w.spatial,
                            stream);
        if (rc != HG_OK) return rc;
        lists[1] = spatial_idx; ks[1] = spatial_k; num_lists = 2;
    }
    if (num_edges_out_host != nullptr) {
        rc = hg_edges_count(lists, ks, num_lists, n, labels, num_hyperedges, clique_pairs, num_edges_out_host,
                            ws + w.off_edges, w.edges, stream);
        if (rc != HG_OK) return rc;
        const int64_t e = *num_edges_out_host;
        HG_REQUIRE(e <= max_edges, HG_ERR_BAD_ARG, "hg_build_knn_hypergraph: %lld edges exceed max_edges", (long long)e);
        rc = hg_edges_fill(e, edge_index, ws + w.off_edges, w.edges, stream);
        if (rc != HG_OK) return rc;
        rc = hg_edge_weights_from_knn(feats, n, d, knn_idx, knn_cos, k, edge_index, e, edge_weights, stream);
        if (rc != HG_OK) return rc;
        if (want_csr) rc = hg_csr_from_coo(edge_index, e, n, csr_rowptr, csr_col, csr_eid, ws + w.off_csr, w.csr, stream);
        return rc;
    }
    // stream-ordered: E <= bound <= max_edges by construction; every later stage reads it from the device
    rc = edges_enqueue(lists, ks, num_lists, n, labels, num_hyperedges, clique_pairs, ws + w.off_edges, w.edges, st, nullptr);
    if (rc != HG_OK) return rc;
    rc = edges_fill_enqueue(bound, edge_index, ws + w.off_edges, num_edges_out_dev, st);
    if (rc != HG_OK) return rc;
    rc = edge_weights_from_knn_enqueue(feats, n, d, knn_idx, knn_cos, k, edge_index, num_edges_out_dev, bound, edge_weights, st);
    if (rc != HG_OK) return rc;
    if (want_csr)
        rc = csr_from_coo_enqueue(edge_index, num_edges_out_dev, bound, n, csr_rowptr, csr_col, csr_eid, ws + w.off_csr,
                                  w.csr, st);
    return rc;
}

extern "C" int hg_build_knn_hypergraph(const float* feats, int64_t n, int64_t d, const float* pos, int pos_dim,
                                       int metric, float lambda_h, float lambda_g, int k, int spatial_k,
                                       const int32_t* labels, int num_hyperedges, int64_t clique_pairs,
                                       int32_t* knn_idx, float* knn_val, int32_t* spatial_idx, float* spatial_d2,
                                       int64_t max_edges, int64_t* edge_index, float* edge_weights,
                                       int64_t* csr_rowptr, int64_t* csr_col, int64_t* csr_eid,
                                       int64_t* num_edges_out_host,
                                       void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(num_edges_out_host, HG_ERR_BAD_ARG, "hg_build_knn_hypergraph: null buffer");
    return build_impl(feats, n, d, pos, pos_dim, metric, lambda_h, lambda_g, k, spatial_k, labels, num_hyperedges,
                      clique_pairs, knn_idx, knn_val, spatial_idx, spatial_d2, max_edges, edge_index, edge_weights,
                      csr_rowptr, csr_col, csr_eid, num_edges_out_host, nullptr, workspace, workspace_bytes, stream);
}

extern "C" int hg_build_knn_hypergraph_enqueue(const float* feats, int64_t n, int64_t d, const float* pos, int pos_dim,
                                               int metric, float lambda_h, float lambda_g, int k, int spatial_k,
                                               const int32_t* labels, int num_hyperedges, int64_t clique_pairs,
                                               int32_t* knn_idx, float* knn_val, int32_t* spatial_idx, float* spatial_d2,
                                               int64_t max_edges, int64_t* edge_index, float* edge_weights,
                                               int64_t* csr_rowptr, int64_t* csr_col, int64_t* csr_eid,
                                               int64_t* num_edges_out_device,
                                               void* workspace, size_t workspace_bytes, void* stream) {
    HG_REQUIRE(num_edges_out_device, HG_ERR_BAD_ARG, "hg_build_knn_hypergraph_enqueue: null buffer");
    return build_impl(feats, n, d, pos, pos_dim, metric, lambda_h, lambda_g, k, spatial_k, labels, num_hyperedges,
                      clique_pairs, knn_idx, knn_val, spatial_idx, spatial_d2, max_edges, edge_index, edge_weights,
                      csr_rowptr, csr_col, csr_eid, nullptr, num_edges_out_device, workspace, workspace_bytes, stream);
}

// ---- batches of slides ---------------------------------------------------------------------------
// The dataset loop of the reference (preprocess_hypergraph.py:651-670) visits slides one by one; a
// slide of a few thousand patches is a few hundred microseconds of GPU work spread over ~25 kernels,
// so the host has to stay ahead: one C call enqueues the whole batch, slide i on lane i % num_lanes,
// optionally from several host threads (thread t drives the lanes l with l % num_threads == t).
extern "C" int hg_build_batch_workspace_bytes(const hg_slide* slides, int num_slides, int64_t d, int k, int spatial_k,
                                              int want_csr, size_t* bytes_per_lane) {
    HG_REQUIRE(slides && num_slides > 0 && bytes_per_lane, HG_ERR_BAD_ARG, "hg_build_batch_workspace_bytes: bad argument");
    size_t need = 0;
    for (int i = 0; i < num_slides; ++i) {
        BuildWs w{};
        int rc = plan_build(slides[i].n, d, k, spatial_k, 0, 0, slides[i].max_edges, want_csr != 0, &w);
        if (rc != HG_OK) return rc;
        if (w.bytes > need) need = w.bytes;
    }
    *bytes_per_lane = al1k(need);
    return HG_OK;
}

extern "C" int hg_build_batch_enqueue(const hg_slide* slides, int num_slides, int64_t d, int pos_dim,
                                      int metric, float lambda_h, float lambda_g, int k, int spatial_k,
                                      void* const* lane_workspaces, size_t bytes_per_lane,
                                      void* const* lane_streams, int num_lanes, int num_threads) {
    HG_REQUIRE(slides && num_slides > 0 && lane_workspaces && lane_streams && num_lanes > 0, HG_ERR_BAD_ARG,
               "hg_build_batch_enqueue: bad argument");
    if (num_threads < 1) num_threads = 1;
    if (num_threads > num_lanes) num_threads = num_lanes;
    int device = 0;
    HG_CHECK_CUDA(cudaGetDevice(&device));
    std::vector<int> rcs(size_t(num_threads), HG_OK);
    std::vector<std::string> msgs(static_cast<size_t>(num_threads));
    auto drive = [&](int tid) {
        if (tid != 0 && cudaSetDevice(device) != cudaSuccess) { rcs[tid] = HG_ERR_CUDA; msgs[tid] = "cudaSetDevice failed"; return; }
        for (int i = 0; i < num_slides; ++i) {
            const int lane = i % num_lanes;
            if (lane % num_threads != tid) continue;
            const hg_slide& s = slides[i];
            const int rc = build_impl(s.feats, s.n, d, s.pos, pos_dim, metric, lambda_h, lambda_g, k,
                                      s.pos ? spatial_k : 0, nullptr, 0, 0, s.knn_idx, s.knn_val, s.spatial_idx, s.spatial_d2,
                                      s.max_edges, s.edge_index, s.edge_weights, s.csr_rowptr, s.csr_col, s.csr_eid,
                                      nullptr, s.num_edges_out_device, lane_workspaces[lane], bytes_per_lane,
                                      lane_streams[lane]);
            if (rc != HG_OK) {
                rcs[tid] = rc;
                msgs[tid] = std::string("slide ") + std::to_string(i) + ": " + get_error();
                return;
            }
        }
    };
    if (num_threads == 1) {
        drive(0);
    } else {
        std::vector<std::thread> pool;
        for (int t = 1; t < num_threads; ++t) pool.emplace_back(drive, t);
        drive(0);
        for (auto& th : pool) th.join();
    }
    for (int t = 0; t < num_threads; ++t)
        if (rcs[t] != HG_OK) { set_error("hg_build_batch_enqueue: %s", msgs[t].c_str()); return rcs[t]; }
    return HG_OK;
}

// ---- batches of SMALL slides: grouped front half ----------------------------------------------------------------
// hg_build_batch_enqueue runs ~25 launches per slide, one of them the persistent fused kernel, which a slide of a
// few thousand rows cannot fill.  Here consecutive slides are packed into groups of up to max_group_rows operand
// rows; a group's feature kNN is FOUR launches on the front stream (knn_group_enqueue, knn_api.cu) and the slides'
// back halves (spatial kNN, edge assembly, weights, CSR) then run on the lanes, next to the following group's
// fused kernel.  Two front workspaces alternate, so group g + 2 waits for the back halves of group g.  Slides of
// more than kGroupMaxN rows (where the symmetric single-slide kernel wins) take the per-slide path on a lane.
namespace hg {
constexpr int64_t kGroupMaxN = 9215;              // tc_sym_stride: symmetric mode from 9,216 rows up

struct BatchGroup { int first, count; int64_t rows; };

static std::vector<BatchGroup> make_groups(const hg_slide* slides, int num, int64_t max_group_rows) {
    std::vector<BatchGroup> groups;
    BatchGroup cur{0, 0, 0};
    for (int i = 0; i < num; ++i) {
        const bool big = slides[i].n > kGroupMaxN;
        const int64_t rows = ops_rows(slides[i].n);
        if (cur.count > 0 && (big || cur.rows + rows > max_group_rows)) { groups.push_back(cur); cur = BatchGroup{i, 0, 0}; }
        if (big) { groups.push_back(BatchGroup{i, 1, -1}); cur = BatchGroup{i + 1, 0, 0}; continue; }     // rows = -1: per-slide path
        if (cur.count == 0) cur.first = i;
        cur.count++; cur.rows += rows;
    }
    if (cur.count > 0) groups.push_back(cur);
    return groups;
}

static int plan_back(int64_t n, int k, int spatial_k, int64_t max_edges, bool want_csr, BuildWs* w) {
    *w = BuildWs{};
    int rc = HG_OK;
    if (spatial_k > 0) {
        rc = hg_knn_spatial_workspace_bytes(n, spatial_k, 0, n, &w->spatial);
        if (rc != HG_OK) return rc;
    }
    rc = hg_edges_workspace_bytes(n, n * int64_t(k) + n * int64_t(spatial_k), 0, 0, &w->edges);
    if (rc != HG_OK) return rc;
    if (want_csr) {
        rc = hg_csr_workspace_bytes(n, max_edges, &w->csr);
        if (rc != HG_OK) return rc;
    }
    // the spatial search is over before the edge stage starts: both at offset 0
    w->off_spatial = 0; w->off_edges = 0; w->off_csr = al1k(w->edges);
    const size_t a = al1k(w->spatial), b = w->off_csr + al1k(w->csr);
    w->bytes = a > b ? a : b;
    return HG_OK;
}

static std::vector<KnnGroupSlide> group_slides(const hg_slide* slides, const BatchGroup& g, int k, char* cos_base,
                                               size_t* cos_bytes) {
    std::vector<KnnGroupSlide> gs(static_cast<size_t>(g.count));
    size_t off = 0;
    for (int j = 0; j < g.count; ++j) {
        const hg_slide& s = slides[g.first + j];
        gs[size_t(j)] = KnnGroupSlide{s.feats, s.pos, s.n, s.knn_idx, s.knn_val,
                                      cos_base ? reinterpret_cast<float*>(cos_base + off) : nullptr};
        off += al1k(size_t(s.n) * size_t(k) * 4);
    }
    if (cos_bytes) *cos_bytes = off;
    return gs;
}
// the group's slides as the grouped back half (group_back.cu) wants them; cos_base as in group_slides
static std::vector<GbSlide> back_slides(const hg_slide* slides, const BatchGroup& g, int k, int spatial_k, char* cos_base) {
    std::vector<GbSlide> gb(static_cast<size_t>(g.count));
    size_t off = 0;
    for (int j = 0; j < g.count; ++j) {
        const hg_slide& s = slides[g.first + j];
        const bool sp = spatial_k > 0 && s.pos != nullptr;
        gb[size_t(j)] = GbSlide{s.n, s.feats, s.knn_idx, reinterpret_cast<const float*>(cos_base + off), sp ? s.spatial_idx : nullptr,
                                s.edge_index, s.edge_weights, s.num_edges_out_device, s.csr_rowptr, s.csr_col, s.csr_eid};
        off += al1k(size_t(s.n) * size_t(k) * 4);
    }
    return gb;
}

// HG_B200_GROUP_BACK=0: per-slide back halves on the lanes (the first version; kept for A/B runs)
static bool grouped_back_half() {
    const char* e = getenv("HG_B200_GROUP_BACK");
    return !(e && atoi(e) == 0);
}

// scratch of the largest spatial search of a group (the searches of a group run one after the other on its lane)
static int group_spatial_bytes(const hg_slide* slides, const BatchGroup& g, int spatial_k, size_t* bytes) {
    *bytes = 0;
    if (spatial_k <= 0) return HG_OK;
    for (int j = 0; j < g.count; ++j) {
        const hg_slide& s = slides[g.first + j];
        if (s.pos == nullptr) continue;
        size_t b = 0;
        int rc = hg_knn_spatial_workspace_bytes(s.n, spatial_k, 0, s.n, &b);
        if (rc != HG_OK) return rc;
        if (b > *bytes) *bytes = b;
    }
    *bytes = al1k(*bytes);
    return HG_OK;
}
}  // namespace hg

extern "C" int hg_build_batch_grouped_workspace_bytes(const hg_slide* slides, int num_slides, int64_t d, int k, int spatial_k,
                                                      int want_csr, int64_t max_group_rows, size_t* front_bytes,
                                                      size_t* bytes_per_lane) {
    HG_REQUIRE(slides && num_slides > 0 && front_bytes && bytes_per_lane && max_group_rows >= 256, HG_ERR_BAD_ARG,
               "hg_build_batch_grouped_workspace_bytes: bad argument");
    size_t front = 0, lane = 0;
    for (const BatchGroup& g : make_groups(slides, num_slides, max_group_rows)) {
        if (g.rows < 0) {                                            // a big slide: whole build on a lane
            BuildWs w{};
            const hg_slide& s = slides[g.first];
            int rc = plan_build(s.n, d, k, s.pos ? spatial_k : 0, 0, 0, s.max_edges, want_csr != 0, &w);
            if (rc != HG_OK) return rc;
            if (w.bytes > lane) lane = w.bytes;
            continue;
        }
        size_t cos_bytes = 0;
        const std::vector<KnnGroupSlide> gs = group_slides(slides, g, k, nullptr, &cos_bytes);
        bool has_pos = true;
        for (const KnnGroupSlide& s : gs) has_pos = has_pos && s.pos != nullptr;
        const size_t need = al1k(knn_group_workspace_bytes(gs.data(), g.count, d, k, has_pos)) + cos_bytes;
        if (need > front) front = need;
        for (int j = 0; j < g.count; ++j) {
            BuildWs w{};
            const hg_slide& s = slides[g.first + j];
            int rc = plan_back(s.n, k, s.pos ? spatial_k : 0, s.max_edges, want_csr != 0, &w);
            if (rc != HG_OK) return rc;
            if (w.bytes > lane) lane = w.bytes;
        }
        // grouped back half: the spatial searches' scratch, then the group's sort / CSR scratch
        size_t sp_bytes = 0;
        int rc = group_spatial_bytes(slides, g, spatial_k, &sp_bytes);
        if (rc != HG_OK) return rc;
        const std::vector<GbSlide> gb = back_slides(slides, g, k, spatial_k, nullptr);
        const size_t back = sp_bytes + al1k(group_back_workspace_bytes(gb.data(), g.count, k, spatial_k));
        if (back > lane) lane = back;
    }
    *front_bytes = al1k(front);
    *bytes_per_lane = al1k(lane);
    return HG_OK;
}

extern "C" int hg_build_batch_grouped_enqueue(const hg_slide* slides, int num_slides, int64_t d, int pos_dim,
                                              int metric, float lambda_h, float lambda_g, int k, int spatial_k,
                                              int64_t max_group_rows,
                                              void* const* front_workspaces, size_t front_bytes, void* front_stream,
                                              void* const* lane_workspaces, size_t bytes_per_lane,
                                              void* const* lane_streams, int num_lanes) {
    HG_REQUIRE(slides && num_slides > 0 && front_workspaces && front_workspaces[0] && front_workspaces[1] &&
               lane_workspaces && lane_streams && num_lanes > 0 && max_group_rows >= 256, HG_ERR_BAD_ARG,
               "hg_build_batch_grouped_enqueue: bad argument");
    cudaStream_t fst = static_cast<cudaStream_t>(front_stream);
    const std::vector<BatchGroup> groups = make_groups(slides, num_slides, max_group_rows);
    std::vector<cudaEvent_t> events;                                  // destroyed at the end: CUDA defers the release
    auto new_event = [&](cudaEvent_t* ev) {
        cudaError_t e = cudaEventCreateWithFlags(ev, cudaEventDisableTiming);
        if (e == cudaSuccess) events.push_back(*ev);
        return e;
    };
    // back-half events of the last group that used each front workspace
    std::vector<cudaEvent_t> busy[2];
    int rc = HG_OK, slot = 0, lane_rr = 0;
    for (size_t gi = 0; gi < groups.size() && rc == HG_OK; ++gi) {
        const BatchGroup& g = groups[gi];
        if (g.rows < 0) {                                            // big slide: the per-slide build on the next lane
            const hg_slide& s = slides[g.first];
            const int lane = lane_rr++ % num_lanes;
            rc = build_impl(s.feats, s.n, d, s.pos, pos_dim, metric, lambda_h, lambda_g, k, s.pos ? spatial_k : 0, nullptr, 0, 0,
                            s.knn_idx, s.knn_val, s.spatial_idx, s.spatial_d2, s.max_edges, s.edge_index, s.edge_weights,
                            s.csr_rowptr, s.csr_col, s.csr_eid, nullptr, s.num_edges_out_device, lane_workspaces[lane],
                            bytes_per_lane, lane_streams[lane]);
            continue;
        }
        char* fws = static_cast<char*>(front_workspaces[slot]);
        for (cudaEvent_t ev : busy[slot]) HG_CHECK_CUDA(cudaStreamWaitEvent(fst, ev, 0));
        busy[slot].clear();
        size_t cos_bytes = 0;
        group_slides(slides, g, k, nullptr, &cos_bytes);
        bool has_pos = true;
        for (int j = 0; j < g.count; ++j) has_pos = has_pos && slides[g.first + j].pos != nullptr;
        HG_REQUIRE(front_bytes >= cos_bytes, HG_ERR_WORKSPACE, "hg_build_batch_grouped_enqueue: front workspace too small");
        const std::vector<KnnGroupSlide> gs = group_slides(slides, g, k, fws, nullptr);
        const size_t knn_ws = front_bytes - cos_bytes;
        // the fused kernel on the front stream, the exact re-rank on the group's lane (next to the next group's sweep)
        const int lane0 = lane_rr % num_lanes;
        cudaStream_t lst0 = static_cast<cudaStream_t>(lane_streams[lane0]);
        rc = knn_group_enqueue(gs.data(), g.count, d, pos_dim, metric, lambda_h, lambda_g, k, 1, fws + cos_bytes, knn_ws, fst, lst0);
        if (rc != HG_OK) break;
        cudaEvent_t front_done;
        HG_CHECK_CUDA(new_event(&front_done));
        HG_CHECK_CUDA(cudaEventRecord(front_done, lst0));
        std::vector<char> lane_used(size_t(num_lanes), 0);
        if (grouped_back_half()) {
            // the whole group's back half on ONE lane: spatial searches slide by slide, then group_back.cu
            const int lane = lane_rr++ % num_lanes;
            cudaStream_t lst = static_cast<cudaStream_t>(lane_streams[lane]);
            char* lws = static_cast<char*>(lane_workspaces[lane]);
            HG_CHECK_CUDA(cudaStreamWaitEvent(lst, front_done, 0));
            lane_used[size_t(lane)] = 1;
            size_t sp_bytes = 0;
            rc = group_spatial_bytes(slides, g, spatial_k, &sp_bytes);
            if (rc != HG_OK) break;
            for (int j = 0; j < g.count && rc == HG_OK; ++j) {
                const hg_slide& s = slides[g.first + j];
                HG_REQUIRE(s.edge_index && s.edge_weights && s.num_edges_out_device &&
                           s.max_edges >= s.n * int64_t(k + (s.pos ? spatial_k : 0)), HG_ERR_BAD_ARG,
                           "hg_build_batch_grouped_enqueue: slide %d: bad output buffers", g.first + j);
                if (spatial_k > 0 && s.pos != nullptr) {
                    HG_REQUIRE(s.spatial_idx && s.spatial_d2, HG_ERR_BAD_ARG,
                               "hg_build_batch_grouped_enqueue: slide %d: spatial output buffers missing", g.first + j);
                    rc = hg_knn_spatial(s.pos, s.n, pos_dim, spatial_k, 0, s.n, s.spatial_idx, s.spatial_d2, lws, sp_bytes, lst);
                }
            }
            if (rc != HG_OK) break;
            const std::vector<GbSlide> gb = back_slides(slides, g, k, spatial_k, fws);
            HG_REQUIRE(bytes_per_lane >= sp_bytes, HG_ERR_WORKSPACE, "hg_build_batch_grouped_enqueue: lane workspace too small");
            rc = group_back_enqueue(gb.data(), g.count, d, k, spatial_k, lws + sp_bytes, bytes_per_lane - sp_bytes, lst);
            if (rc != HG_OK) break;
        } else {
        for (int j = 0; j < g.count && rc == HG_OK; ++j) {
            const hg_slide& s = slides[g.first + j];
            const int lane = lane_rr++ % num_lanes;
            cudaStream_t lst = static_cast<cudaStream_t>(lane_streams[lane]);
            if (!lane_used[size_t(lane)]) { HG_CHECK_CUDA(cudaStreamWaitEvent(lst, front_done, 0)); lane_used[size_t(lane)] = 1; }
            BuildWs w{};
            const int sk = s.pos ? spatial_k : 0;
            rc = plan_back(s.n, k, sk, s.max_edges, s.csr_rowptr != nullptr, &w);
            if (rc != HG_OK) break;
            HG_REQUIRE(bytes_per_lane >= w.bytes, HG_ERR_WORKSPACE, "hg_build_batch_grouped_enqueue: lane workspace too small");
            HG_REQUIRE(s.edge_index && s.edge_weights && s.num_edges_out_device && s.max_edges >= s.n * int64_t(k + sk),
                       HG_ERR_BAD_ARG, "hg_build_batch_grouped_enqueue: slide %d: bad output buffers", g.first + j);
            rc = build_back_half(s.feats, s.n, d, s.pos, pos_dim, k, sk, nullptr, 0, 0, s.knn_idx, gs[size_t(j)].out_cos,
                                 s.spatial_idx, s.spatial_d2, s.max_edges, s.edge_index, s.edge_weights, s.csr_rowptr,
                                 s.csr_col, s.csr_eid, nullptr, s.num_edges_out_device,
                                 static_cast<char*>(lane_workspaces[lane]), w, lane_streams[lane]);
        }
        }
        if (rc != HG_OK) break;
        for (int l = 0; l < num_lanes; ++l) {
            if (!lane_used[size_t(l)]) continue;
            cudaEvent_t ev;
            HG_CHECK_CUDA(new_event(&ev));
            HG_CHECK_CUDA(cudaEventRecord(ev, static_cast<cudaStream_t>(lane_streams[l])));
            busy[slot].push_back(ev);
        }
        slot ^= 1;
    }
    for (cudaEvent_t ev : events) cudaEventDestroy(ev);
    return rc;
}
