/*
 * hg_b200.h -- C ABI of the B200-native hypergraph-construction hot path.
 *
 * Drop-in boundary for zz9tf/multimodal-fusion `build_hypergraph/` (reference
 * paths below are relative to the reference root).  The reference is pure
 * Python; its "FFI" for this path is the set of torch / scikit-learn calls the
 * functions below replace.  Every entry point takes plain device pointers and
 * sizes plus a `cudaStream_t` passed as `void*` -- no torch types.
 *
 * Conventions:
 *  - every function returns int: 0 = HG_OK, negative = HG_ERR_*;
 *    hg_last_error() returns a thread-local message for the last failure;
 *  - the CALLER allocates every buffer (with torch, cudaMalloc, ...); the
 *    library never allocates or frees user-visible memory; scratch comes from
 *    a caller workspace whose size is asked first (`*_workspace_bytes`);
 *  - calls are asynchronous on `stream` unless documented (the count half of a
 *    count -> fill pair synchronises the stream to return the count);
 *  - all matrices are row-major, float32, densely packed unless a leading
 *    dimension is given; indices out are int32 (kNN lists) or int64 (COO/CSR,
 *    the reference's `edge_index` dtype, preprocess_hypergraph.py:410).
 */
#ifndef HG_B200_H
#define HG_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HG_OK               0
#define HG_ERR_BAD_ARG     (-1)
#define HG_ERR_UNSUPPORTED (-2)
#define HG_ERR_CUDA        (-3)
#define HG_ERR_WORKSPACE   (-4)

/* ranking metric of the fused similarity + top-k kernel */
#define HG_METRIC_EUCLIDEAN 0 /* value out = squared distance (clamped >= 0)            */
#define HG_METRIC_GAUSSIAN  1 /* value out = exp(-lambda_h d2_h - lambda_g d2_g)         */
#define HG_METRIC_COSINE    2 /* value out = cos(f_i, f_j), norms clamped at 1e-8        */

/* form of the squared distance in hg_similarity_dense */
#define HG_FORM_EXPANSION 0   /* |a|^2 + |b|^2 - 2 a.b   similarity_kernel.py:43-49     */
#define HG_FORM_DIRECT    1   /* sum (a-b)^2             preprocess_hypergraph.py:254-255 */

int hg_version(void);
const char* hg_last_error(void);
/* kernels launched by this library since load (monotonic; bench.py's gpu_launches) */
int64_t hg_launch_count(void);
/* in-stream CUDA-event timer around the fused similarity+top-k kernel: enable, run,
 * then collect the summed duration (ms) and the number of launches it covers */
int hg_profile_enable(int on);
int hg_profile_collect(float* ms_total, int* launches);

/* ---- operand preparation -------------------------------------------------
 * Splits FP32 rows, scaled by a per-row power of two, into FP16 hi/lo halves (s x ~= hi + lo, 22 significant bits)
 * laid out [n_pad][d_pad / 32][2][32] for TMA (every 128 bytes = 32 hi | 32 lo values
 * of one 32-wide k-block), zero padded to n_pad = ceil256(n), d_pad = ceil64(d), and
 * writes per-row squared norms (FP32, from the original data).  Replaces `torch.sum(features**2)` (similarity_kernel.py:43) and is the
 * input of the tensor-core Gram product that replaces `torch.mm` (:46). */
size_t hg_operand_bytes(int64_t n, int64_t d);
int hg_split_operands(const float* feats, int64_t n, int64_t d,
                      void* ops, float* sq_norms, void* stream);

/* ---- A1/A2/A4: fused similarity + running top-k (N x N never materialised) --
 * For rows [row_begin, row_end) of `feats` [n, d]: the k best columns among all
 * n (self excluded by index when exclude_self != 0), ordered best first with
 * ties broken by ascending column index.
 *   EUCLIDEAN: sklearn NearestNeighbors(k+1,'euclidean') minus self,
 *              preprocess_hypergraph.py:379-388;
 *   GAUSSIAN : per-row top-k of K = K_h * K_g, similarity_kernel.py:52,84,122
 *              (pos may be NULL -> K_h only);
 *   COSINE   : sklearn metric='cosine' (preprocess_hypergraph.py:311) /
 *              F.cosine_similarity (:419).
 * out_idx [rows, k] int32 (-1 when fewer than k candidates exist),
 * out_val [rows, k] float32.  `precision`: 0 = split-FP16 tensor-core candidates +
 * exact FP32 re-rank (default), 1 = plain FP32 SIMT kernel (any shape).
 * Over the full row range of a large slide (n >= 9216) the tensor-core pass exploits the symmetry
 * of K: about half of the tiles are computed and each serves both (i, j) and (j, i); results are
 * identical to the full pass (HG_B200_TC_SYM=0 in the environment forces the full pass). */
int hg_knn_feature_workspace_bytes(int64_t n, int64_t d, int k,
                                   int64_t row_begin, int64_t row_end, size_t* bytes);
/* What hg_knn_feature will run for this shape: symmetric_stride > 0 means the symmetric
 * (half-of-the-tiles) pass with a threshold pre-pass over every stride-th column; mma_flop is the
 * FP16 tensor-core work actually issued (hg_knn_feature_products MMAs per product, padded tiles),
 * which is about 0.56 x 2 x 2 N^2 D in symmetric mode. */
int hg_knn_feature_plan(int64_t n, int64_t d, int k, int64_t row_begin, int64_t row_end,
                        int* symmetric_stride, double* mma_flop);
/* Tensor-core MMAs issued per logical product x_i . x_j by the main pass and by the threshold
 * pre-pass (0 when there is none) of hg_knn_feature for this shape: 3 = hi*hi + hi*lo + lo*hi
 * of the split-FP16 operands (classic pass over more than 16,384 columns, dense launches),
 * 2 = hi*hi + lo*hi (symmetric passes, classic and grouped launches over small slides;
 * HG_B200_TC_SYM_PRODS=3 / HG_B200_TC_CLASSIC_PRODS=3 in the environment keep three).  The roofline of the
 * fused kernel is the 16-bit tensor-core peak divided by this number. */
int hg_knn_feature_products(int64_t n, int64_t d, int k, int64_t row_begin, int64_t row_end,
                            int* main_pass, int* pre_pass);
int hg_knn_feature(const float* feats, int64_t n, int64_t d,
                   const float* pos, int pos_dim,
                   int metric, float lambda_h, float lambda_g,
                   int k, int64_t row_begin, int64_t row_end, int exclude_self,
                   int precision,
                   int32_t* out_idx, float* out_val,
                   void* workspace, size_t workspace_bytes, void* stream);

/* hg_knn_feature that also writes out_cos [rows, k] (may be NULL): cos(f_row, f_idx) of every list
 * entry, a by-product of the exact re-rank, bit-identical to what hg_edge_weights_cosine computes
 * for the pair -- the input of hg_edge_weights_from_knn.  Tensor-core path only. */
int hg_knn_feature_cos(const float* feats, int64_t n, int64_t d,
                       const float* pos, int pos_dim,
                       int metric, float lambda_h, float lambda_g,
                       int k, int64_t row_begin, int64_t row_end, int exclude_self,
                       int precision,
                       int32_t* out_idx, float* out_val, float* out_cos,
                       void* workspace, size_t workspace_bytes, void* stream);

/* ---- A1/A2/A7: dense (small-N / rectangular) Gaussian kernel -----------------
 * out[i, j] = exp(-lambda_h d2(a_i, b_j) - lambda_g d2(pa_i, pb_j)), un-clamped.
 * b == NULL means b = a (square).  pos_* may be NULL.
 * similarity_kernel.py:17-124, preprocess_hypergraph.py:248-257. */
int hg_similarity_dense(const float* a, int64_t m, const float* b, int64_t n, int64_t d,
                        const float* pos_a, const float* pos_b, int pos_dim,
                        float lambda_h, float lambda_g, int form,
                        float* out, int64_t ld_out, void* stream);

/* ---- A9: mean / unbiased std / min / max / lower median of `count` floats -----
 * preprocess_hypergraph.py:190-196, 259-265.  stats_out: 5 doubles on the
 * DEVICE.  If exclude_diag_n > 0 the buffer is an n x n matrix whose diagonal is
 * skipped (similarity_kernel.py:183-186). */
int hg_stats_workspace_bytes(int64_t count, size_t* bytes);
int hg_similarity_stats(const float* values, int64_t count, int64_t exclude_diag_n,
                        double* stats_out, void* workspace, size_t workspace_bytes, void* stream);

/* A9 when N x N does not fit: the same five statistics of K = K_h * K_g over all N^2 entries,
 * streamed chunk by chunk through the tensor-core kernel (K from split-FP16 dot products, relative
 * error ~1e-6) -- the matrix is never stored in full.  preprocess_hypergraph.py:137,190-196. */
/* A1 / A2 for large square inputs: K of one matrix against itself through the tensor-core kernel's
 * dense-output mode (split-FP16 Gram product, 3 MMAs per k-step, FP32 accumulation; same expansion form, un-clamped).  13-20x faster
 * than hg_similarity_dense, entries within ~1.5e-5 relative of float64 (measured) instead of 1e-5: opt-in.
 * ld_out >= n, lambda_h > 0, workspace 1024-byte aligned. */
int hg_similarity_dense_tc_workspace_bytes(int64_t n, int64_t d, size_t* bytes);
int hg_similarity_dense_tc(const float* feats, int64_t n, int64_t d, const float* pos, int pos_dim,
                           float lambda_h, float lambda_g, float* out, int64_t ld_out,
                           void* workspace, size_t workspace_bytes, void* stream);
int hg_similarity_stats_streaming_workspace_bytes(int64_t n, int64_t d, size_t* bytes);
int hg_similarity_stats_streaming(const float* feats, int64_t n, int64_t d, const float* pos, int pos_dim,
                                  float lambda_h, float lambda_g, double* stats_out,
                                  void* workspace, size_t workspace_bytes, void* stream);

/* ---- A3: spatial-coordinate kNN ------------------------------------------------
 * k smallest |g_i - g_j|^2 (direct-difference FP32), self excluded, ties by
 * (d2, index).  Semantics pinned by similarity_kernel.py:57-86 (top-k of a K_g
 * row).  pos [n, pos_dim], pos_dim in {2, 3}.  2-D inputs with n >= 8192 are binned
 * into a uniform cell grid (O(n) work); smaller or 3-D inputs, and inputs that put
 * more than 1024 points into one cell, take the brute-force pass.  Same results
 * bit for bit either way (HG_B200_SP_GRID=0/1 forces one path). */
int hg_knn_spatial_workspace_bytes(int64_t n, int k, int64_t row_begin, int64_t row_end, size_t* bytes);
int hg_knn_spatial(const float* pos, int64_t n, int pos_dim, int k,
                   int64_t row_begin, int64_t row_end,
                   int32_t* out_idx, float* out_d2,
                   void* workspace, size_t workspace_bytes, void* stream);

/* ---- A5: edge assembly ------------------------------------------------------------
 * Directed kNN pairs (i, knn_idx[i, :]) of up to 4 neighbour lists (feature
 * kNN, spatial kNN, ...; knn_lists is a HOST array of device pointers, entries
 * < 0 are skipped) plus all pairs inside every cluster
 * label in [0, num_hyperedges) (labels may be NULL), canonicalised to (min, max),
 * radix-sorted and de-duplicated.  preprocess_hypergraph.py:385-410.
 * hg_edges_count runs the pipeline into the workspace, synchronises `stream`
 * and returns E; hg_edges_fill then copies edge_index [2, E] int64 (row 0 < row
 * 1, lexicographic order) out of the same workspace. */
int hg_edges_workspace_bytes(int64_t n, int64_t knn_slots /* sum over lists of n*k */,
                             int num_hyperedges, int64_t clique_pairs, size_t* bytes);
/* number of intra-cluster pairs sum s_c (s_c - 1) / 2 (synchronises `stream`) */
int hg_clique_pairs(const int32_t* labels, int64_t n, int num_hyperedges,
                    int64_t* pairs_out_host, void* stream);
int hg_edges_count(const int32_t* const* knn_lists, const int* ks, int num_lists, int64_t n,
                   const int32_t* labels, int num_hyperedges, int64_t clique_pairs,
                   int64_t* num_edges_out_host,
                   void* workspace, size_t workspace_bytes, void* stream);
int hg_edges_fill(int64_t num_edges, int64_t* edge_index,
                  const void* workspace, size_t workspace_bytes, void* stream);
/* the name SURVEY.md 8(b) lists for the second half of the count -> allocate -> fill pair; same call */
int hg_edges_build(int64_t num_edges, int64_t* edge_index,
                   const void* workspace, size_t workspace_bytes, void* stream);

/* ---- A6: w_e = max(0, cos(f_i, f_j)), eps = 1e-8 -- preprocess_hypergraph.py:413-422 */
int hg_edge_weights_cosine(const float* feats, int64_t n, int64_t d,
                           const int64_t* edge_index, int64_t num_edges,
                           float* weights, void* stream);

/* A6 without re-reading feature rows: w_e looked up in the feature-kNN lists (knn_idx / knn_cos
 * [n, k] of a full-range hg_knn_feature_cos call); edges found in neither endpoint's list
 * (spatial kNN, cliques) are computed from the rows as above.  Same values bit for bit. */
int hg_edge_weights_from_knn(const float* feats, int64_t n, int64_t d,
                             const int32_t* knn_idx, const float* knn_cos, int k,
                             const int64_t* edge_index, int64_t num_edges,
                             float* weights, void* stream);

/* ---- CSR incidence of the undirected pair list ---------------------------------------
 * rowptr [n+1], col [2E] (neighbour), eid [2E] (edge id), rows sorted by neighbour.
 * Requires edge_index sorted as produced by hg_edges_fill. */
int hg_csr_workspace_bytes(int64_t n, int64_t num_edges, size_t* bytes);
int hg_csr_from_coo(const int64_t* edge_index, int64_t num_edges, int64_t n,
                    int64_t* rowptr, int64_t* col, int64_t* eid,
                    void* workspace, size_t workspace_bytes, void* stream);

/* hg_csr_from_coo for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO counterpart in the reference.  Same inputs,
 * outputs and result as hg_csr_from_coo, but the scratch lives in a file-scope device array of
 * HG_CSR_STATIC_SCRATCH_BYTES (18 MiB) that every GPU loading the library holds until exit; the call allocates nothing
 * (no cudaMalloc / cudaFree / allocating library routine).  Limits: n <= HG_CSR_STATIC_MAX_N and
 * num_edges <= HG_CSR_STATIC_MAX_EDGES, beyond which it returns HG_ERR_UNSUPPORTED.  All calls on a device share
 * the array (it keeps the previous call's contents; the build initialises what it reads): calls that overlap on
 * the GPU race -- keep them on one stream or order them with events.  The device must be current. */
#define HG_CSR_STATIC_MAX_N           131072
#define HG_CSR_STATIC_MAX_EDGES       524288
#define HG_CSR_STATIC_SCRATCH_BYTES   (18 * 1024 * 1024)
int hg_csr_from_coo_static(const int64_t* edge_index, int64_t num_edges, int64_t n,
                           int64_t* rowptr, int64_t* col, int64_t* eid, void* stream);


/* ---- A8: dense-threshold builder -------------------------------------------------------
 * Keep every ordered (i, j), diagonal included, with K[i, j] >= threshold, row
 * major.  similarity_kernel.py:193-210.  K is a dense n x n matrix (small N). */
int hg_threshold_edges_count(const float* K, int64_t n, float threshold,
                             int64_t* num_edges_out_host,
                             void* workspace, size_t workspace_bytes, void* stream);
int hg_threshold_edges_fill(const float* K, int64_t n, float threshold, int64_t num_edges,
                            int64_t* edge_index, float* edge_weights,
                            const void* workspace, size_t workspace_bytes, void* stream);
int hg_threshold_edges_build(const float* K, int64_t n, float threshold, int64_t num_edges,
                             int64_t* edge_index, float* edge_weights,
                             const void* workspace, size_t workspace_bytes, void* stream);   /* = _fill */
int hg_threshold_workspace_bytes(int64_t n, size_t* bytes);

/* ---- row (f)1, "next": KMeans building blocks (Lloyd assign / update, kmeans++ distances) -------
 * Replace sklearn.cluster.KMeans at preprocess_hypergraph.py:150-151, 299-300, 391-392; the
 * estimator itself (n_init restarts, kmeans++ seeding, tolerance) is hg_kmeans_fit below.  Parity
 * with scikit-learn is statistical (inertia), not label-exact. */
int hg_kmeans_assign(const float* x, int64_t n, int64_t d, const float* centroids, int c,
                     int32_t* labels, float* d2, void* stream);
int hg_kmeans_workspace_bytes(int64_t d, int c, size_t* bytes);
int hg_kmeans_update(const float* x, int64_t n, int64_t d, const int32_t* labels, int c,
                     float* centroids, double* shift2, uint32_t* counts_out,
                     void* workspace, size_t workspace_bytes, void* stream);
int hg_kmeans_inertia(const float* d2, int64_t n, double* inertia, void* stream);
int hg_kmeans_candidate_d2(const float* x, int64_t n, int64_t d, const int64_t* cand, int t,
                           float* out, void* stream);
/* The whole estimator in one call: KMeans(n_clusters=c, n_init, max_iter, tol, random_state=seed)
 * .fit(x) as the reference constructs it (scikit-learn defaults: greedy k-means++ with 2 + ln(c) local
 * trials, Lloyd until no label moves or the squared centre shift is <= tol * mean(var(x)), best restart
 * by inertia).  Seeding, iterations and the choice of the best restart stay on the device; the host
 * polls one flag every few iterations, so the call blocks on `stream`.  Every sum runs in a fixed
 * order: the same seed gives the same labels bit for bit.  labels [n] int32, centroids [c, d] float,
 * inertia (double) and n_iter (int32, may be NULL; iterations of the winning restart) are device
 * buffers.  HG_ERR_BAD_ARG when n < c (scikit-learn's ValueError text in hg_last_error). */
int hg_kmeans_fit_workspace_bytes(int64_t n, int64_t d, int c, size_t* bytes);
int hg_kmeans_fit(const float* x, int64_t n, int64_t d, int c, int n_init, int max_iter, double tol,
                  uint64_t seed, int32_t* labels, float* centroids, double* inertia, int32_t* n_iter,
                  void* workspace, size_t workspace_bytes, void* stream);

/* ---- one call per slide: A4 (+A3) -> A5 -> A6 -> CSR ---------------------------------------
 * The body of build_hypergraph_knn_kmeans (preprocess_hypergraph.py:374-433) on one stream with
 * a single synchronisation (the edge count).  Variable-size outputs are caller buffers sized for
 * `max_edges` >= n*(k+spatial_k)+clique_pairs; they are filled compactly: edge_index as [2, E]
 * (row 1 starts at offset E), edge_weights [E], csr_col/csr_eid [2E].  csr_* may be NULL (no
 * CSR), spatial_* may be NULL when spatial_k == 0, labels may be NULL (no cliques). */
int hg_build_workspace_bytes(int64_t n, int64_t d, int k, int spatial_k, int num_hyperedges,
                             int64_t clique_pairs, int64_t max_edges, int want_csr, size_t* bytes);
int hg_build_knn_hypergraph(const float* feats, int64_t n, int64_t d, const float* pos, int pos_dim,
                            int metric, float lambda_h, float lambda_g, int k, int spatial_k,
                            const int32_t* labels, int num_hyperedges, int64_t clique_pairs,
                            int32_t* knn_idx, float* knn_val, int32_t* spatial_idx, float* spatial_d2,
                            int64_t max_edges, int64_t* edge_index, float* edge_weights,
                            int64_t* csr_rowptr, int64_t* csr_col, int64_t* csr_eid,
                            int64_t* num_edges_out_host,
                            void* workspace, size_t workspace_bytes, void* stream);

/* The same build, fully stream-ordered: no host synchronisation at all.  The edge count E is
 * written to *num_edges_out_device (8 bytes of DEVICE memory) and every stage after the de-dup
 * reads it from there, so many slides can be in flight on several streams (datasets of small
 * slides, BASELINE config C4: the dataset loop preprocess_hypergraph.py:651-670).  Outputs have
 * the compact layout above relative to E (edge_index row 1 at offset E, csr_col / csr_eid 2E
 * entries); the caller reads E back when it collects the slide.  Same buffer capacities. */
int hg_build_knn_hypergraph_enqueue(const float* feats, int64_t n, int64_t d, const float* pos, int pos_dim,
                                    int metric, float lambda_h, float lambda_g, int k, int spatial_k,
                                    const int32_t* labels, int num_hyperedges, int64_t clique_pairs,
                                    int32_t* knn_idx, float* knn_val, int32_t* spatial_idx, float* spatial_d2,
                                    int64_t max_edges, int64_t* edge_index, float* edge_weights,
                                    int64_t* csr_rowptr, int64_t* csr_col, int64_t* csr_eid,
                                    int64_t* num_edges_out_device,
                                    void* workspace, size_t workspace_bytes, void* stream);

/* ---- descriptor of the feature kNN of a ROW BLOCK of a slide (plain data, the same for all calls of a slide) ---- */
typedef struct hg_knn_part {
    int64_t n, d;                 /* rows of the whole slide, feature width */
    int64_t row_begin, row_end;   /* query rows of this GPU */
    int64_t reserved0;
    const float* pos;             /* [n, pos_dim] positions of all rows (Gaussian metric with lambda_g) or NULL */
    int32_t pos_dim, metric, k, exclude_self;
    int32_t reserved1, reserved2;
    float lambda_h, lambda_g;
} hg_knn_part;

/* ---- row-partitioned slides in SYMMETRIC mode ---------------------------------------------
 * hg_knn_feature on a row block computes every tile of its rows (classic pass).  Here the GPUs share the
 * checkerboard of the single-GPU symmetric mode: the owner of a row block computes about half of its tiles, keeps
 * the row direction locally and pushes the transposed survivors into per-row buckets ON THE GPU THAT OWNS THE ROW
 * (remote atomics + stores over NVLink; the caller maps the peers' bucket arrays, e.g. with CUDA IPC).  Row blocks
 * are chunks of `chunk` rows, chunk a multiple of 256 (the last one may be shorter).  Per slide, stream-ordered:
 *   prepass  (all n rows in feats_all) operand split, column parameters, own bucket counters + flags cleared,
 *            threshold pre-pass of the own rows -> thr_all[row_begin, row_end)   (thr_all: [ceil256(n)] floats)
 *   caller   all-gather of thr_all (entries beyond n = NaN); it is also the barrier behind which every GPU has
 *            cleared its buckets
 *   sweep    symmetric kernel over the own rows; peer_bkt_cnt[r] / peer_bkt[r] = the arrays of GPU r
 *            ([chunk] int32, [chunk * 512] pairs of uint32; r = own rank: the own arrays)
 *   caller   all-reduce(max) of own_flags[0] over the GPUs: the barrier behind which all pushes have landed
 *   finish   classic pass if any bucket overflowed (any_overflow[0] != 0, then on every GPU alike), merge, exact
 *            re-rank -> out_idx / out_val / out_cos [row_end - row_begin, k], equal to hg_knn_feature_cos's.
 * Replaces the reference's sklearn kNN call,
 * preprocess_hypergraph.py:379-382, for slides too large for one GPU's time budget. */
/* peer access from the current device to `peer_device` (kernels of hg_knn_sympart_sweep store into mapped peer memory) */
int hg_enable_peer_access(int peer_device);
/* map an allocation another process exported (64-byte cudaIpcMemHandle_t) for kernels of the CURRENT device; the
 * library maps, it does not allocate.  One open per exported allocation and process. */
int hg_ipc_open(const void* handle64, void** base_out);
/* a cudaMalloc'ed exchange buffer on the current device plus its 64-byte IPC handle, for peers to hg_ipc_open (the
 * ONE allocation this library makes, on explicit request: framework allocators may return memory whose handles are
 * not cudaIpcMemHandle_t); hg_peer_free releases it */
int hg_peer_alloc(size_t bytes, void** ptr_out, void* handle64_out);
int hg_peer_free(void* ptr);
int hg_ipc_close(void* base);
int hg_knn_sympart_workspace_bytes(const hg_knn_part* s, size_t* bytes);
int hg_knn_sympart_prepass(const hg_knn_part* s, const float* feats_all, float* thr_all, int32_t* own_bkt_cnt,
                           int* own_flags, void* workspace, size_t workspace_bytes, void* stream);
int hg_knn_sympart_sweep(const hg_knn_part* s, const float* thr_all, int64_t chunk, int num_peers,
                         void* const* peer_bkt_cnt, void* const* peer_bkt, int* own_flags,
                         void* workspace, size_t workspace_bytes, void* stream);
int hg_knn_sympart_finish(const hg_knn_part* s, const float* feats_all, const int* any_overflow,
                          const int32_t* own_bkt_cnt, const void* own_bkt,
                          int32_t* out_idx, float* out_val, float* out_cos,
                          void* workspace, size_t workspace_bytes, void* stream);

/* ---- row (f)3, "next": what the consumer does with the incidence ---------------------------------
 * One HypergraphConv propagation  out = D^-1 H B^-1 H^T x  (torch_geometric semantics without
 * attention, unit hyperedge weights) the way downstream_survival/models/cust_omics.py:58-63,100-102
 * applies it to the stored pair list: row 0 of edge_index = node ids, row 1 = hyperedge ids.  Takes
 * the CSR incidence of hg_csr_from_coo (for a (min, max) pair list the members of hyperedge j are
 * row j's neighbours < j, the hyperedges of node v its neighbours > v).  x, hyperedge_feat (scratch /
 * by-product: B^-1 H^T x) and out are [n, f] float32; the dense Theta / bias stay with the caller. */
int hg_hypergraph_propagate(const float* x, int64_t n, int64_t f, const int64_t* csr_rowptr,
                            const int64_t* csr_col, float* hyperedge_feat, float* out, void* stream);

/* ---- row (f)4, "next": super-patch aggregation without the per-cluster host loop -----------------
 * Replaces preprocess_hypergraph.py:157-170 (mean feature / mean position of the rows of each
 * cluster label) and :175-184 (for each cluster with m > 1 members, the mean of the off-diagonal
 * entries of K_wsi[idx][:, idx]; 'avg_intra_cluster_similarity' = mean of those means, 0.0 if none).
 * labels [n] int64 on the device (KMeans output); rows whose label is outside [0, num_clusters) are
 * left out.  The intra-cluster kernel values are read from dense_k ([n, ld_k], the caller's N x N
 * kernel) when it is given, else recomputed in FP32 tiles over same-cluster pairs only (expansion
 * form x Gaussian of positions, as hg_similarity_dense).  want_intra = 0 skips that stage.
 * Outputs (device): super_feats [S, d], super_pos [S, pos_dim] (pos may be NULL), counts [S] int64,
 * intra [S] float32 (NaN for m <= 1), summary [4] float32 = {avg_intra, clusters with m > 1,
 * first empty cluster or -1 (the reference raises ValueError there; its means are NaN here),
 * rows with a label in range}, order_out [n] int32 or NULL = rows sorted by (label, row).
 * Stream-ordered, deterministic (no float atomics). */
int hg_aggregate_workspace_bytes(int64_t n, int64_t d, int num_clusters, size_t* bytes);
int hg_aggregate_super_patches(const float* feats, int64_t n, int64_t d, const float* pos, int pos_dim,
                               const int64_t* labels, int num_clusters, float lambda_h, float lambda_g,
                               const float* dense_k, int64_t ld_k, int want_intra,
                               float* super_feats, float* super_pos, int64_t* counts, float* intra,
                               float* summary, int32_t* order_out,
                               void* workspace, size_t workspace_bytes, void* stream);

/* ---- batches of slides: ONE host call enqueues every build ----------------------------------
 * The reference's dataset loop (preprocess_hypergraph.py:651-670) handles slides one at a time.
 * A slide of a few thousand patches is only a few hundred microseconds of GPU work, so here the
 * whole batch is enqueued stream-ordered (hg_build_knn_hypergraph_enqueue semantics, no cliques):
 * slide i runs on lane i % num_lanes = (lane_streams[l], lane_workspaces[l]); with num_threads > 1
 * the lanes are driven from that many host threads.  All fields are 8 bytes wide, so an array of
 * hg_slide is also an int64 matrix [num_slides][14].  Pointers are device pointers; pos,
 * spatial_*, csr_* may be NULL as in hg_build_knn_hypergraph. */
typedef struct hg_slide {
    const float* feats;            /* [n, d] */
    int64_t n;
    const float* pos;              /* [n, pos_dim] or NULL */
    int32_t* knn_idx;              /* [n, k] */
    float* knn_val;                /* [n, k] */
    int32_t* spatial_idx;          /* [n, spatial_k] or NULL */
    float* spatial_d2;
    int64_t max_edges;             /* >= n * (k + spatial_k) */
    int64_t* edge_index;           /* 2 * max_edges */
    float* edge_weights;           /* max_edges */
    int64_t* csr_rowptr;           /* n + 1, or NULL: no CSR */
    int64_t* csr_col;              /* 2 * max_edges */
    int64_t* csr_eid;              /* 2 * max_edges */
    int64_t* num_edges_out_device; /* 1 */
} hg_slide;

int hg_build_batch_workspace_bytes(const hg_slide* slides, int num_slides, int64_t d, int k, int spatial_k,
                                   int want_csr, size_t* bytes_per_lane);
int hg_build_batch_enqueue(const hg_slide* slides, int num_slides, int64_t d, int pos_dim,
                           int metric, float lambda_h, float lambda_g, int k, int spatial_k,
                           void* const* lane_workspaces, size_t bytes_per_lane,
                           void* const* lane_streams, int num_lanes, int num_threads);

/* The same batch with a GROUPED front half, for slides of a few thousand patches (BASELINE config C4): consecutive
 * slides are packed into groups of up to max_group_rows operand rows (every slide padded to a multiple of 256); a
 * group's feature kNN is four launches on front_stream -- operand split, column parameters, ONE persistent launch
 * of the fused tcgen05 kernel over a unit table spanning all the slides of the group, exact re-rank -- and the
 * slides' back halves (spatial kNN, edge assembly, weights, CSR) run on the lanes behind it, next to the following
 * group's fused kernel.  front_workspaces[0..1] (front_bytes each) alternate between groups.  Slides of 9,216
 * rows or more take the per-slide path on a lane.  Results equal hg_build_batch_enqueue's bit for bit.
 * Replaces the reference's one-slide-at-a-time loop, preprocess_hypergraph.py:651-670. */
int hg_build_batch_grouped_workspace_bytes(const hg_slide* slides, int num_slides, int64_t d, int k, int spatial_k,
                                           int want_csr, int64_t max_group_rows, size_t* front_bytes,
                                           size_t* bytes_per_lane);
int hg_build_batch_grouped_enqueue(const hg_slide* slides, int num_slides, int64_t d, int pos_dim,
                                   int metric, float lambda_h, float lambda_g, int k, int spatial_k,
                                   int64_t max_group_rows,
                                   void* const* front_workspaces, size_t front_bytes, void* front_stream,
                                   void* const* lane_workspaces, size_t bytes_per_lane,
                                   void* const* lane_streams, int num_lanes);

#ifdef __cplusplus
}
#endif
#endif /* HG_B200_H */
