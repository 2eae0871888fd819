// Fused pairwise-similarity + running top-k on the sm_100a tensor cores.
//
// Replaces, without ever writing the N x N matrix to HBM:
//   reference similarity_kernel.py:46   torch.mm(features, features.t())
//             similarity_kernel.py:49-52 n_i + n_j - 2 dot -> exp(-lambda .)
//             similarity_kernel.py:79-84,122  the spatial factor K_g
//             preprocess_hypergraph.py:379-382  sklearn brute-force kNN
//
// Arithmetic: FP32 rows are scaled by a power of two and pre-split into FP16 hi/lo halves (split.cu, knn_tc.cuh),
// interleaved per 32-wide k-block so that one 128-byte operand row carries 32 hi | 32 lo values.  The dot
// product is accumulated in FP32 in TMEM from tcgen05 MMAs (kind::f16, FP16 operands) per k-step:
//   3 products  hi*lo^T + lo*hi^T + hi*hi^T   (~2^-21 relative: classic, grouped and dense launches)
//   2 products  lo*hi^T + hi*hi^T             (the B rows rounded to 11 bits, 2^-11 relative: symmetric passes,
//                                              whose many short lists + buckets absorb the wider slack; their B
//                                              tiles are loaded without the lo halves)
// The epilogue turns the accumulator into a ranking score  a_j + b_j * dot
// (+ g * |pos_i - pos_j|^2)  and keeps a per-row sorted list of the KL best
// (score, column) pairs in registers.  The candidates are re-ranked exactly in
// FP32 by rerank.cu, so the tensor-core pass only has to find the right set.
//
// Structure (one CTA per SM, persistent over work units):
//   warp 0      TMA producer: A tile [128 rows][32 hi | 32 lo] + B tile [256 rows][32 hi | 32 lo] per stage
//               (48 KB; four stages in flight: with the earlier 96 KB stages only two fitted and the
//               reload of a stage had one stage of MMAs, 0.8 us, to land)
//   warp 1      TMEM allocator + single-thread tcgen05.mma issuer
//   warps 2..5  epilogue: tcgen05.ld 32 lanes x 32 columns, score, register top-k
// Two 256-column FP32 accumulators (all 512 TMEM columns) let the epilogue of tile
// t overlap the MMAs of tile t+1.
//
// Work unit = (row block of 128 queries, one of S column ranges).  Units are
// numbered split-major so the CTAs running concurrently sweep the SAME column
// range in the same order: a B tile is fetched from HBM once and then served to
// all 148 SMs out of L2, while each CTA's A block (128 rows) stays L2-resident.
//
// Symmetric mode (SYM, full row range): K is symmetric, so only about half of the tiles are
// computed -- half the MMAs.  With A = the 256-row super-block a row block belongs to and C = a
// column tile, tile (A, C) is computed iff C == A, or C > A with A + C even, or C < A with A + C
// odd (a checkerboard: of the two tiles (A, C) and (C, A) exactly one is computed, and every row
// block sweeps about half of the column tiles).  A computed off-diagonal tile serves two directions:
//   * row direction, as above: column j is a candidate of row i;
//   * transposed direction: row i is a candidate OF column j.  Its score a_i + b_i * dot is
//     compared with thr[j], an upper bound of j's final k-th best that a strided pre-pass (same
//     kernel, 1/stride of the columns) computed, and the rare survivors go through a per-warp
//     shared-memory queue into bucket j in global memory.
// Every ordered pair (i, j), i != j, is covered exactly once; every row receives about half of
// its candidates through each direction, so bucket occupancy is uniform.
#include "knn_tc.cuh"
#include "ptx.cuh"
#include "topk_list.cuh"
#include <cudaTypedefs.h>
#include <stdlib.h>
#include <algorithm>

namespace hg {

// CG = 1: one CTA per tile (M = 128).  CG = 2: a CTA pair (cluster of 2, tcgen05 cta_group::2) shares
// one 256 x 256 tile: each CTA stages its own 128 A rows and HALF of the B tile, the leader issues
// M = 256 MMAs that read both halves, so operand traffic per CTA drops from 96 to 64 KB per stage
// and a third pipeline stage fits.
// MC (multicast pair, CG = 1): two CTAs of a cluster work on the two 128-row halves of one 256-row block
// against the same column tiles.  Each issues its own M = 128 MMAs from its own shared memory, but stages
// only HALF of every B tile and multicasts it into both CTAs (TMA .multicast::cluster), so a B tile
// crosses the L2 -> SM fabric once per pair and the two CTAs advance in lock-step by construction.
template <int CG>
struct TcCfg {
    static constexpr int STAGES = CG == 1 ? 4 : 6;
    static constexpr int B_ROWS = TC_BN / CG;                       // B rows staged by this CTA
    static constexpr int A_TILE = TC_BM * 128;                      // 16 KB: 128-byte rows of 32 hi | 32 lo
    static constexpr int B_TILE = B_ROWS * 128;                     // 32 / 16 KB
    static constexpr int STAGE_BYTES = A_TILE + B_TILE;             // 48 / 32 KB
};
constexpr int TC_MAX_STAGES = 8;
constexpr int TC_THREADS = 192;
constexpr int TC_EPI_THREADS = 128;
// dense-output mode: eight epilogue warps (two per TMEM lane quarter, 128 of the 256 tile columns each).  Its
// single tile in flight (see the MMA issuer) serialises MMAs and epilogue, and the epilogue -- 256 exp + a 1 KB
// row segment stored per thread -- was 12 us of a 25 us tile with four warps.
constexpr int TC_THREADS_DENSE = 320;
constexpr int TC_EPI_THREADS_DENSE = 256;
constexpr int TC_TMEM_COLS = 512;

constexpr int TC_WQ_CAP = 256;        // transposed-candidate queue entries per epilogue warp

struct TcSmem {                       // lives after the (1024-byte aligned) stage buffers
    float2 colp[2][TC_BN];
    float4 cpos[2][TC_BN];
    float thr[2][TC_BN];              // SYM: admission threshold of each column of the tile
    float cinv[2][TC_BN];             // SYM: 1 / operand scale of each column (the transposed score needs it apart)
    uint2 wq[4][TC_WQ_CAP];           // SYM: (score bits, column-in-tile << 8 | row-in-block)
    int wq_cnt[4];
    uint64_t full[TC_MAX_STAGES], empty[TC_MAX_STAGES], tmem_full[2], tmem_empty[2];
    uint32_t tmem_base;
};
template <int CG>
constexpr size_t tc_smem_bytes() { return 1024 + size_t(TcCfg<CG>::STAGES) * TcCfg<CG>::STAGE_BYTES + sizeof(TcSmem); }

// Unit order (see the header comment): row blocks are taken in chunks of rc; inside a chunk the
// units run split-group-major, then split, then row block, so any 148 consecutive units cover
// about rc row blocks x cc column splits.
__device__ __forceinline__ void decode_unit(const TcPlan& pl, int u, int& rb, int& split) {
    const int per_chunk = pl.rc * pl.splits;
    const int g = u / per_chunk;
    const int rows_here = min(pl.rc, pl.num_u - g * pl.rc);
    int rem = u - g * per_chunk;
    const int h = rem / (rows_here * pl.cc);
    rem -= h * rows_here * pl.cc;
    const int c = rem / rows_here;
    rb = g * pl.rc + (rem - c * rows_here);
    split = h * pl.cc + c;
}

// What one work unit covers: from the plan (row blocks x column splits of one slide) or, in a grouped launch, from the
// unit table.
struct UnitInfo {
    int64_t row0, row_end;      // the CTA's 128 query rows start at row0; rows >= row_end are padding
    int ct0, ct1;               // column tiles swept
    int rb;                     // row block (symmetric tile ownership; plan mode only)
    int col_off;                // subtracted from a column to get the index stored in the lists
    int64_t list_off;           // element offset of the list of query row `row`: list_off + (row - list_row0) * kl
    int64_t list_row0;
};
template <int CL>
__device__ __forceinline__ UnitInfo unit_info(const TcParams& p, int u, uint32_t crank, int kl) {
    UnitInfo ui;
    if (p.unit_tab != nullptr) {
        const TcUnit t = p.unit_tab[u];
        ui.row0 = t.row0; ui.row_end = t.row_end; ui.ct0 = t.ct0; ui.ct1 = t.ct1; ui.rb = 0; ui.col_off = t.col_off;
        ui.list_off = t.part_off; ui.list_row0 = t.col_off;
        return ui;
    }
    int split, rb;
    decode_unit(p.plan, u, rb, split);
    ui.rb = rb;
    ui.row0 = p.row_begin + int64_t(rb) * (TC_BM * CL) + int64_t(crank) * TC_BM;
    ui.row_end = p.row_end;
    ui.ct0 = split * p.plan.ct_per_split;
    ui.ct1 = min(ui.ct0 + p.plan.ct_per_split, p.plan.num_ct);
    ui.col_off = 0;
    ui.list_off = int64_t(split) * p.plan.n_rows * kl; ui.list_row0 = p.row_begin;
    return ui;
}

// Does the unit of row block rb (classic: every tile; symmetric: the checkerboard) own column tile ct?
template <bool SYM>
__device__ __forceinline__ bool owns_tile(int a, int ct) {          // a = 256-row super-block of the row block
    if (!SYM) return true;
    return ct == a || (((a + ct) & 1) == (ct < a ? 1 : 0));
}

// r[j] for a run-time j without local memory: a binary tree of selects on the bits of j (31 SELs,
// depth 5 -- a linear chain of 31 dependent selects stalls the single epilogue warp of a scheduler
// for its whole length; ncu, N = 4096: 116 -> 109 us).
__device__ __forceinline__ uint32_t select32(const uint32_t (&r)[32], int j) {
    const bool b0 = (j & 1) != 0, b1 = (j & 2) != 0, b2 = (j & 4) != 0, b3 = (j & 8) != 0, b4 = (j & 16) != 0;
    uint32_t a[16], b[8], c[4];
#pragma unroll
    for (int q = 0; q < 16; ++q) a[q] = b0 ? r[2 * q + 1] : r[2 * q];
#pragma unroll
    for (int q = 0; q < 8; ++q) b[q] = b1 ? a[2 * q + 1] : a[2 * q];
#pragma unroll
    for (int q = 0; q < 4; ++q) c[q] = b2 ? b[2 * q + 1] : b[2 * q];
    const uint32_t d0 = b3 ? c[1] : c[0], d1 = b3 ? c[3] : c[2];
    return b4 ? d1 : d0;
}

__device__ __forceinline__ void bucket_push(const TcParams& p, int col, int row, uint32_t score_bits) {
    if (p.peer_chunk > 0) {
        // partitioned slide: column col is a row of GPU `owner`; the slot counter is a remote atomic over NVLink,
        // the entry a remote 8-byte store (peer memory mapped by the host side, or the own arrays)
        const int owner = int(col / p.peer_chunk);
        const int64_t local = col - owner * p.peer_chunk;
        const int slot = atomicAdd_system(p.peer_cnt[owner] + local, 1);
        if (slot < p.bkt_cap) p.peer_bkt[owner][local * p.bkt_cap + slot] = make_uint2(score_bits, uint32_t(row));
        else *p.ovf = 1;
        return;
    }
    const int slot = atomicAdd(p.bkt_cnt + col, 1);
    if (slot < p.bkt_cap) p.bkt[int64_t(col) * p.bkt_cap + slot] = make_uint2(score_bits, uint32_t(row));
    else *p.ovf = 1;
}

template <int KL, bool HAS_POS, int CG, bool SYM, bool MC>
__global__ void __launch_bounds__(KL == 0 ? TC_THREADS_DENSE : TC_THREADS, 1)
knn_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
              const TcParams p) {
    using Cfg = TcCfg<CG>;
    constexpr int TC_STAGES = Cfg::STAGES;
    constexpr int TC_A_TILE = Cfg::A_TILE, TC_STAGE_BYTES = Cfg::STAGE_BYTES;
    static_assert(!SYM || KL > 0, "symmetric mode keeps candidate lists");
    static_assert(!MC || CG == 1, "multicast pairs issue single-CTA MMAs");
    constexpr int CL = (CG == 2 || MC) ? 2 : 1;        // CTAs per cluster = 128-row blocks per work unit
    constexpr bool DENSE = KL == 0;                    // dense-output mode: K written out, no candidate lists
    if (p.guard != nullptr && *p.guard == 0) return;   // fallback launch that is not needed (grid-uniform)
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t* smem_al = smem_raw + (smem_base - ptx::smem_u32(smem_raw));
    TcSmem* sm = reinterpret_cast<TcSmem*>(smem_al + size_t(TC_STAGES) * TC_STAGE_BYTES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_kb = p.d_pad / TC_BK;
    const int units = p.unit_tab != nullptr ? p.num_tab_units : p.plan.num_u * p.plan.splits;
    const uint32_t crank = (CL == 2) ? ptx::cluster_ctarank() : 0u;     // 0 = leader of the CTA pair
    const int wid = blockIdx.x / CL, nworkers = gridDim.x / CL;          // a worker = one CTA or one pair

    if (threadIdx.x == 0) {
        for (int s = 0; s < TC_MAX_STAGES; ++s) {
            ptx::mbar_init(ptx::smem_u32(&sm->full[s]), 1);
            ptx::mbar_init(ptx::smem_u32(&sm->empty[s]), MC ? 2 : 1);      // MC: both CTAs' MMAs release a stage
        }
        for (int b = 0; b < 2; ++b) {
            ptx::mbar_init(ptx::smem_u32(&sm->tmem_full[b]), 1);
            ptx::mbar_init(ptx::smem_u32(&sm->tmem_empty[b]), (DENSE ? TC_EPI_THREADS_DENSE : TC_EPI_THREADS) * CG);
        }
        for (int q = 0; q < 4; ++q) sm->wq_cnt[q] = 0;
        ptx::fence_barrier_init();
        ptx::fence_proxy_async();
    }
    if (warp == 0 && lane == 0) {
        ptx::prefetch_tensormap(&tmap_a);
        ptx::prefetch_tensormap(&tmap_b);
    }
    if (CL == 2) ptx::cluster_sync();                  // the peer's barriers exist before anything signals them
    if (warp == 1) {
        ptx::tmem_alloc<CG>(ptx::smem_u32(&sm->tmem_base), TC_TMEM_COLS);
        ptx::tmem_relinquish<CG>();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = sm->tmem_base;

    if (warp == 0) {
        // ============================ TMA producer ============================
        // elect.sync instead of lane == 0: the compiler then knows that ONE thread runs the block and emits the TMA /
        // MMA / commit instructions bare; behind a plain lane test it wraps each of them in an ELECT + BRA.U.ANY loop
        // over the active threads -- ~150 instructions per pipeline stage in the MMA issuer, which made the issuing
        // thread, not the tensor pipe, the limit once a stage held two products instead of three (ncu source view)
        if (ptx::elect_one()) {
            // B tile: 128-byte rows (32 hi | 32 lo) or, when no product reads the lo halves of B, their first 64 bytes;
            // the smaller stages then come six or eight to the buffer instead of four or six
            const uint32_t b_bytes = p.b_hi_only ? Cfg::B_TILE / 2 : Cfg::B_TILE;
            const uint32_t stage_bytes = TC_A_TILE + b_bytes, stage_tx = stage_bytes;
            const uint32_t nstages = uint32_t(p.stages);
            uint32_t stage = 0, ph = 0;
            for (int u = wid; u < units; u += nworkers) {
                const UnitInfo ui = unit_info<CL>(p, u, crank, KL);
                const int row0 = int(ui.row0), rb = ui.rb;
                for (int ct = ui.ct0; ct < ui.ct1; ++ct) {
                    if (!owns_tile<SYM>(p.sblk_base + (CL == 2 ? rb : rb >> 1), ct)) continue;
                    for (int kb = 0; kb < num_kb; ++kb) {
                        if (!p.pair_stages) {
                            ptx::mbar_wait(ptx::smem_u32(&sm->empty[stage]), ph ^ 1u);
                        } else if ((stage & 1u) == 0) {    // stages are refilled in pairs (the stage count and num_kb are even)
                            ptx::mbar_wait(ptx::smem_u32(&sm->empty[stage]), ph ^ 1u);
                            ptx::mbar_wait(ptx::smem_u32(&sm->empty[stage + 1]), ph ^ 1u);
                        }
                        const uint32_t full = ptx::smem_u32(&sm->full[stage]);
                        const uint32_t dst = smem_base + stage * stage_bytes;
                        if (MC) {
                            // own A rows + the own half of the B tile, which also lands in the peer; the peer's
                            // half arrives here the same way (its bytes may precede this expect_tx: same phase)
                            const int HALF = int(b_bytes / 2);
                            ptx::mbar_expect_tx(full, stage_tx);
                            ptx::tma_load_3d(dst, &tmap_a, full, kb * (2 * TC_BK), row0, 0, p.hint_a);
                            ptx::tma_load_3d_mc(dst + TC_A_TILE + int(crank) * HALF, &tmap_b, full, kb * (2 * TC_BK),
                                                ct * TC_BN + int(crank) * (TC_BN / 2), 0, uint16_t(3), p.hint_b);
                        } else if (CG == 1) {
                            ptx::mbar_expect_tx(full, stage_tx);
                            ptx::tma_load_3d(dst, &tmap_a, full, kb * (2 * TC_BK), row0, 0, p.hint_a);
                            ptx::tma_load_3d(dst + TC_A_TILE, &tmap_b, full, kb * (2 * TC_BK), ct * TC_BN, 0, p.hint_b);
                        } else {
                            // both CTAs' bytes are counted on the LEADER's barrier (peer bit cleared)
                            const uint32_t lead_full = full & 0xFEFFFFFFu;
                            if (crank == 0) ptx::mbar_expect_tx(full, 2 * stage_tx);
                            ptx::tma_load_3d_2sm(dst, &tmap_a, lead_full, kb * (2 * TC_BK), row0, 0, p.hint_a);
                            ptx::tma_load_3d_2sm(dst + TC_A_TILE, &tmap_b, lead_full, kb * (2 * TC_BK),
                                                 ct * TC_BN + int(crank) * Cfg::B_ROWS, 0, p.hint_b);
                        }
                        if (++stage == nstages) { stage = 0; ph ^= 1u; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ============================ MMA issuer ==============================
        if ((crank == 0 || MC) && ptx::elect_one()) {
            constexpr uint32_t idesc = ptx::make_idesc_f16(TC_BM * CG, TC_BN);
            const int prods = p.prods;
            const uint32_t stage_bytes = TC_A_TILE + (p.b_hi_only ? Cfg::B_TILE / 2 : Cfg::B_TILE);
            const uint32_t nstages = uint32_t(p.stages);
            uint32_t stage = 0, ph = 0, tile = 0;
            for (int u = wid; u < units; u += nworkers) {
                const UnitInfo ui = unit_info<CL>(p, u, crank, KL);
                const int rb = ui.rb;
                for (int ct = ui.ct0; ct < ui.ct1; ++ct) {
                    if (!owns_tile<SYM>(p.sblk_base + (CL == 2 ? rb : rb >> 1), ct)) continue;
                    // dense-output mode: ONE tile in flight, its two accumulators fill the 512 columns (see below)
                    const uint32_t buf = DENSE ? 0u : (tile & 1u), tph = DENSE ? (tile & 1u) : ((tile >> 1) & 1u);
                    ptx::mbar_wait(ptx::smem_u32(&sm->tmem_empty[buf]), tph ^ 1u);
                    ptx::tc_fence_after();
                    const uint32_t d_tmem = tmem_base + buf * TC_BN;
                    // The tensor core adds every MMA into the FP32 accumulator with a truncating add, so a long chain
                    // drifts by ~1 ulp of the running sum per step (1.4e-5 relative on exp(-d2) after 192 steps at
                    // D = 1024).  List mode does not care (exact FP32 re-rank).  Dense mode sums the small products
                    // (2^-8 of the magnitude) in an accumulator of their own and adds the two in FP32 in the epilogue:
                    // the large accumulator sees D / 16 steps instead of 4 D / 16.
                    const uint32_t d_cross = DENSE ? d_tmem + TC_BN : d_tmem;
                    for (int kb = 0; kb < num_kb; ++kb) {
                        ptx::mbar_wait(ptx::smem_u32(&sm->full[stage]), ph);
                        ptx::tc_fence_after();
                        const uint32_t a_addr = smem_base + stage * stage_bytes;
                        const uint32_t b_addr = a_addr + TC_A_TILE;
                        // inside a 128-byte swizzled row: hi values at bytes [0, 64), lo at [64, 128)
                        const uint64_t a_hi = ptx::make_smem_desc_sw128(a_addr);
                        const uint64_t a_lo = a_hi + 4;                     // + 64 bytes, in 16-byte units
                        const uint64_t b_hi = p.b_hi_only ? ptx::make_smem_desc_sw64(b_addr) : ptx::make_smem_desc_sw128(b_addr);
                        const uint64_t b_lo = b_hi + 4;                     // (128-byte rows only)
#pragma unroll
                        for (int ks = 0; ks < TC_BK / 16; ++ks) {
                            const uint64_t off = uint64_t(ks * 2);        // 32 bytes per k-step, >> 4
                            const uint32_t acc0 = uint32_t((kb | ks) != 0);
                            if (prods >= 2) {
                                // lo*hi first: the small terms enter the (empty) accumulator before the large one;
                                // lo*lo (2^-22 of the magnitude) is left out.  Dense mode: operands with 8-bit hi
                                // halves (split.cu), so lo*lo is 2^-16 and counts -- on the diagonal and for
                                // near-duplicate rows it is a sum of squares, no cancellation -- and the small
                                // products get an accumulator of their own.
                                // Two products: lo*hi + hi*hi -- the B tiles then carry no lo halves at all.
                                if (DENSE) ptx::mma_f16<CG>(d_cross, a_lo + off, b_lo + off, idesc, acc0);
                                ptx::mma_f16<CG>(d_cross, a_lo + off, b_hi + off, idesc, DENSE ? 1u : acc0);
                                if (prods >= 3) ptx::mma_f16<CG>(d_cross, a_hi + off, b_lo + off, idesc, 1u);
                                ptx::mma_f16<CG>(d_tmem, a_hi + off, b_hi + off, idesc, DENSE ? acc0 : 1u);
                            } else {                          // coarse pre-pass: one product
                                ptx::mma_f16<CG>(d_tmem, a_hi + off, b_hi + off, idesc, acc0);
                            }
                        }
                        // frees the stage (in both CTAs of a pair) when the MMAs retire
                        if (MC) ptx::mma_commit_mc(ptx::smem_u32(&sm->empty[stage]), 3);
                        else if (CG == 1) ptx::mma_commit(ptx::smem_u32(&sm->empty[stage]));
                        else ptx::mma_commit_2sm(ptx::smem_u32(&sm->empty[stage]), 3);
                        if (++stage == nstages) { stage = 0; ph ^= 1u; }
                    }
                    if (CG == 1) ptx::mma_commit(ptx::smem_u32(&sm->tmem_full[buf]));
                    else ptx::mma_commit_2sm(ptx::smem_u32(&sm->tmem_full[buf]), 3);     // (buf = 0 in dense mode)
                    ++tile;
                }
            }
        }
    } else {
        // ============================ epilogue ================================
        constexpr int EPI = DENSE ? TC_EPI_THREADS_DENSE : TC_EPI_THREADS;
        const int et = threadIdx.x - 64;                 // 0..EPI-1
        const int quarter = warp & 3;                    // TMEM lane quarter this warp may read
        const int chalf = DENSE ? (warp - 2) >> 2 : 0;   // dense mode: which 128 columns of the tile this warp writes
        const uint32_t lane_base = uint32_t(quarter * 32) << 16;
        constexpr int KLA = KL > 0 ? KL : 1;             // KL == 0: dense-output mode, no lists
        const int cstride = p.col_stride;
        uint32_t tile = 0;
        for (int u = wid; u < units; u += nworkers) {
            const UnitInfo ui = unit_info<CL>(p, u, crank, KL);
            const int rb = ui.rb;
            const int64_t row0 = ui.row0;
            const int sblk = p.sblk_base + ((CL == 2) ? rb : rb >> 1);   // 256-row super-block (symmetric tile ownership)
            const int64_t row = row0 + quarter * 32 + lane;
            const int ct0 = ui.ct0, ct1 = ui.ct1;
            float lv[KLA];
            int li[KLA];
#pragma unroll
            for (int q = 0; q < KLA; ++q) { lv[q] = INFINITY; li[q] = -1; }
            // this row's own score parameters; padded rows: (+inf, 0)
            const float2 me = (row < p.n_pad) ? __ldg(p.colp + row) : make_float2(INFINITY, 0.f);
            const float row_norm = me.x, ta = me.x, tb = me.y;
            // the accumulator holds (s_i x_i) . (s_j x_j): 1 / s_i here, 1 / s_j folded into the column multiplier
            const float rinv = (row < ui.row_end) ? op_inv_scale(__ldg(p.norms_a + row)) : 1.f;   // padding rows: dot = 0
            float4 mypos = make_float4(0.f, 0.f, 0.f, 0.f);
            if (HAS_POS && row < ui.row_end) mypos = __ldg(p.pos4 + row);
            // SYM: the pre-pass bound of this row also gates its own list (nothing above it can be in
            // the final top-k), which removes the warm-up insertions of every unit; NaN on padded rows
            const float row_bound = SYM ? __ldg(p.thr + row) : INFINITY;
            for (int ct = ct0; ct < ct1; ++ct) {
                if (!owns_tile<SYM>(sblk, ct)) continue;
                const uint32_t tbuf = DENSE ? 0u : (tile & 1u), tph = DENSE ? (tile & 1u) : ((tile >> 1) & 1u);
                const uint32_t buf = tile & 1u;                           // shared-memory column parameters alternate
                const int col0 = ct * TC_BN;
#pragma unroll
                for (int c = et; c < TC_BN; c += EPI) {
                    const int64_t src = int64_t(col0 + c) * cstride;      // operand row behind column c
                    const bool ok = src < p.n_pad;
                    float2 cpj = ok ? __ldg(p.colp + src) : make_float2(INFINITY, 0.f);
                    const float cinv = ok ? op_inv_scale(__ldg(p.norms_b + src)) : 1.f;
                    cpj.y *= cinv;                                        // exact: a power of two
                    sm->colp[buf][c] = cpj;
                    if (SYM) sm->cinv[buf][c] = cinv;
                    if (HAS_POS) sm->cpos[buf][c] = ok ? __ldg(p.pos4 + src) : make_float4(0.f, 0.f, 0.f, 0.f);
                    if (SYM) sm->thr[buf][c] = __ldg(p.thr + col0 + c);    // NaN beyond n: never admitted
                }
                ptx::named_bar_sync(1, EPI);
                ptx::mbar_wait(ptx::smem_u32(&sm->tmem_full[tbuf]), tph);
                ptx::tc_fence_after();
                const bool diag = p.exclude_self && (col0 < row0 + TC_BM) && (row0 < col0 + TC_BN) && cstride == 1;
                const int self_col = diag ? int(row - col0) : -1;     // column inside this tile, if any
                const bool transposed = SYM && ct != sblk;             // the diagonal tile serves rows only
                // ranking score of (this row, column c) and, SYM, of (column c, this row)
                auto scores = [&](int c, float acc, float& s_row, float& s_tr) {
                    const float2 cp = sm->colp[buf][c];
                    const float dot = acc * rinv;                         // exact (power of two): x_i . x_j * s_j
                    s_row = fmaf(cp.y, dot, cp.x);
                    if (SYM) s_tr = fmaf(tb * sm->cinv[buf][c], dot, ta);
                    if (HAS_POS) {
                        const float4 q = sm->cpos[buf][c];
                        const float dx = mypos.x - q.x, dy = mypos.y - q.y, dz = mypos.z - q.z;
                        const float g = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
                        s_row = fmaf(p.gscale, g, s_row);
                        if (SYM) s_tr = fmaf(p.gscale, g, s_tr);
                    }
                };
#pragma unroll 1
                for (int ch = DENSE ? chalf * 4 : 0; ch < (DENSE ? chalf * 4 + 4 : TC_BN / 32); ++ch) {
                    uint32_t r[32];
                    ptx::tmem_ld_32x32(tmem_base + lane_base + tbuf * TC_BN + ch * 32, r);
                    if constexpr (DENSE) {
                        uint32_t rx[32];                                   // the cross-term accumulator of the tile
                        ptx::tmem_ld_32x32(tmem_base + lane_base + TC_BN + ch * 32, rx);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(rx[j]));
                    } else {
                        ptx::tmem_ld_wait();
                    }
                    if constexpr (KL == 0) {
                        // dense mode: K = exp(-lambda_h (n_i + n_j - 2 dot + g d2_g)), 128 contiguous bytes per thread
                        float* dst = p.dense_out + (row - p.row_begin) * p.dense_ld + col0 + ch * 32;
                        const bool row_ok = row < p.row_end;
                        const bool full = row_ok && (col0 + ch * 32 + 32 <= p.n) && ((p.dense_ld & 3) == 0);
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            float v[4];
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                float s, unused;
                                scores(ch * 32 + j + q, __uint_as_float(r[j + q]), s, unused);
                                // a row against itself: d2 = 0 exactly, whatever the split product left of it
                                const bool self = p.dense_self && row == int64_t(col0 + ch * 32 + j + q);
                                v[q] = self ? 1.f : expf(-p.lambda_h * (row_norm + s));
                            }
                            if (full) {
                                *reinterpret_cast<float4*>(dst + j) = make_float4(v[0], v[1], v[2], v[3]);
                            } else if (row_ok) {
#pragma unroll
                                for (int q = 0; q < 4; ++q)
                                    if (col0 + ch * 32 + j + q < p.n) dst[j + q] = v[q];
                            }
                        }
                        continue;
                    }
                    // pass 1 (straight-line): bitmask of the columns that beat the row's current k-th
                    // best and, SYM, of the columns whose admission threshold this row beats
                    uint32_t hits = 0, hits_t = 0;
                    const float thr = SYM ? fminf(lv[KLA - 1], row_bound) : lv[KLA - 1];
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int c = ch * 32 + j;
                        float s, st;
                        scores(c, __uint_as_float(r[j]), s, st);
                        if (c == self_col) s = INFINITY;
                        hits |= (s < thr) ? (1u << j) : 0u;
                        if (SYM) hits_t |= (st < sm->thr[buf][c]) ? (1u << j) : 0u;
                        if (!SYM) r[j] = __float_as_uint(s);      // pass 2 picks the score itself, not the dot product
                    }
                    // pass 2 (rare): ascending column order keeps the lowest index among equal scores
                    while (hits) {
                        const int j = __ffs(hits) - 1;
                        hits &= hits - 1;
                        float s, st;
                        if (SYM) scores(ch * 32 + j, __uint_as_float(select32(r, j)), s, st);
                        else s = __uint_as_float(select32(r, j));
                        if (s < lv[KLA - 1]) list_insert<KLA>(lv, li, s, (col0 + ch * 32 + j) * cstride - ui.col_off);
                    }
                    if (SYM) {
                        // transposed survivors: (score, column, row) into this warp's queue
                        if (!transposed) hits_t = 0;
                        while (hits_t) {
                            const int j = __ffs(hits_t) - 1;
                            hits_t &= hits_t - 1;
                            float s, st;
                            scores(ch * 32 + j, __uint_as_float(select32(r, j)), s, st);
                            const int slot = atomicAdd(&sm->wq_cnt[quarter], 1);
                            if (slot < TC_WQ_CAP)
                                sm->wq[quarter][slot] = make_uint2(__float_as_uint(st),
                                                                   uint32_t((ch * 32 + j) << 8) | uint32_t(quarter * 32 + lane));
                            else
                                bucket_push(p, col0 + ch * 32 + j, int(row), __float_as_uint(st));
                        }
                    }
                }
                ptx::tc_fence_before();
                if (CG == 1) ptx::mbar_arrive(ptx::smem_u32(&sm->tmem_empty[tbuf]));
                else ptx::mbar_arrive_cluster(ptx::smem_u32(&sm->tmem_empty[tbuf]), 0);    // the leader's barrier
                if (SYM) {
                    // drain the warp's queue into the per-column buckets, one entry per lane and round
                    __syncwarp();
                    const int queued = min(sm->wq_cnt[quarter], TC_WQ_CAP);
                    for (int e = lane; e < queued; e += 32) {
                        const uint2 q = sm->wq[quarter][e];
                        bucket_push(p, col0 + int(q.y >> 8), int(row0) + int(q.y & 0xffu), q.x);
                    }
                    __syncwarp();
                    if (lane == 0) sm->wq_cnt[quarter] = 0;
                    __syncwarp();
                }
                ++tile;
            }
            // the unit's result: the threshold (pre-pass) or the candidate list of (row, split)
            if (KL > 0 && row < ui.row_end) {
                if (p.thr_out != nullptr) {
                    float t = INFINITY;                   // the list ascends: min over q >= rank is lv[rank]
#pragma unroll
                    for (int q = 0; q < KLA; ++q) t = fminf(t, (q >= p.thr_rank) ? lv[q] : INFINITY);
                    // guaranteed slack (knn_tc.cuh tc_eps): the error of this pass's score plus that of the main
                    // pass's score the bound will be compared with; Euclidean / Gaussian scores n_j - 2 dot scale
                    // with n_i + n_max, cosine scores -dot / |x_j| with |x_i|
                    const float scale = (p.thr_margin_cos != 0) ? 1.f / fmaxf(fabsf(tb), 1e-30f) : ta + __ldg(p.norm_max);
                    const float rho = p.rho_max != nullptr ? __ldg(p.rho_max) : TC_RHO_WORST;
                    p.thr_out[row] = t + tc_slack(p.d_pad, p.prods, p.pre_main_prods, rho) * scale + TC_SLACK_REL * fabsf(t);
                } else {
                    const int64_t o = ui.list_off + (row - ui.list_row0) * KL;
#pragma unroll
                    for (int q = 0; q < KLA; ++q) { p.part_val[o + q] = lv[q]; p.part_idx[o + q] = li[q]; }
                }
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (CL == 2) ptx::cluster_sync();                  // neither CTA may leave while its peer still reads / writes its smem
    if (warp == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc<CG>(tmem_base, TC_TMEM_COLS);
    }
}

// ---------------------------------------------------------------------------------------------
int tc_cta_group() {                       // 1 (default) or 2 = CTA pairs, HG_B200_TC_CG=2
    const char* e = getenv("HG_B200_TC_CG");
    return (e && atoi(e) == 2) ? 2 : 1;
}
// CTA pairs in the symmetric pass (cta_group::2: each CTA stages its own 128 A rows and half of the B tile, so a
// B tile leaves the L2 once per pair).  The pass runs against the board's power limit, and of all the variants
// measured back to back for seconds (profiles/r02_power_probe_variants.jsonl) the pairs hold the highest clock per
// unit of work from D = 1024 up: C3 30.8 ms against 36.3 (single CTAs, same tiles) and 32.5 (whole B rows, 4 stages).
// Narrow rows are bound by the epilogue, where single CTAs are 1-3 % ahead.  HG_B200_TC_SYM_CG=1/2 overrides.
static int tc_sym_cta_group(int64_t d_pad) {
    if (const char* e = getenv("HG_B200_TC_SYM_CG")) { const int v = atoi(e); if (v == 1 || v == 2) return v; }
    if (tc_cta_group() == 2) return 2;
    return d_pad >= 1024 ? 2 : 1;
}
static int tc_multicast() {                // multicast pairs (cg 1 only): HG_B200_TC_MC=0/1
    if (tc_cta_group() != 1) return 0;
    const char* e = getenv("HG_B200_TC_MC");
    return (e && atoi(e) != 0) ? 1 : 0;
}

static int list_len(int k) {            // k plus slack for the exact re-rank, in the register-array sizes compiled
    return (k + 4 <= 8) ? 8 : (k + 4 <= 16) ? 16 : (k + 8 <= 24) ? 24 : 32;
}

// Concurrency shape rc x cc = workers and the split count.  The A row blocks of the workers
// running together must stay L2-resident while B streams through (A block = rows x d_pad x
// (hi + lo)); keep them under ~40 MB of the 126 MB L2.  Then the fewest column splits (a
// multiple of cc) whose last round wastes the least.  `major` = row blocks or row-block pairs,
// `tiles` = column tiles a major unit sweeps.
static void shape_plan(TcPlan& pl, int major, int tiles, int64_t d_pad, int min_splits = 1) {
    const int workers = kNumSMs / pl.cl;               // CTAs or CTA pairs
    const int64_t a_block = int64_t(TC_BM * pl.cl) * d_pad * 4;
    int64_t a_mb = 40;
    if (const char* e = getenv("HG_B200_TC_A_MB")) { const int v = atoi(e); if (v >= 1 && v <= 512) a_mb = v; }   // tuning override
    const int64_t rc_cap = (a_mb << 20) / a_block;
    const int cc_max = 4 / pl.cl;                      // 148 = 4 * 37, 74 = 2 * 37
    int best_cc = cc_max, best_s = cc_max;
    double best_eff = -1.0;
    for (int cc = 1; cc <= cc_max; cc *= 2) {
        const int rc = workers / cc;
        if (rc > rc_cap && cc < cc_max) continue;
        for (int s = cc; s <= TC_MAX_SPLITS; s += cc) {
            if (s > tiles && s > cc) break;
            if (s < min_splits && s + cc <= TC_MAX_SPLITS && s + cc <= tiles) continue;
            const int64_t units = int64_t(major) * s;
            // a unit is ceil(tiles / s) tiles long: count the padding of the last split as waste too
            const double fill = double(tiles) / double(ceil_div(tiles, s) * s);
            const double eff = fill * double(units) / double(ceil_div(units, workers) * workers);
            if (eff > best_eff + 0.015) { best_eff = eff; best_cc = cc; best_s = s; }
        }
    }
    if (const char* ov = getenv("HG_B200_TC_PLAN")) {          // tuning override: "cc,splits"
        int cc = 0, sp = 0;
        if (sscanf(ov, "%d,%d", &cc, &sp) == 2 && (cc == 1 || cc == 2 || cc == 4) && cc <= cc_max && sp >= cc &&
            sp <= TC_MAX_SPLITS && sp % cc == 0) { best_cc = cc; best_s = sp; }
    }
    pl.cc = best_cc;
    pl.rc = workers / best_cc;
    pl.splits = best_s;
    pl.grid = int(min64(workers, int64_t(major) * pl.splits)) * pl.cl;
}

TcPlan tc_make_plan(int64_t n, int64_t n_rows, int k, int64_t d_pad, bool lists) {
    TcPlan pl{};
    pl.cg = tc_cta_group();
    // long classic launches over wide rows (the row block of a giant slide: 12.5k x 100k x 1536 runs 7.9 ms on CTA
    // pairs against 9.1) are up against the same power / L2-feed limits as the symmetric pass; short ones are not
    if (lists && !getenv("HG_B200_TC_CG") && d_pad >= 1024 && double(n) * double(n_rows) >= double(1 << 30)) pl.cg = 2;
    pl.mc = lists ? tc_multicast() : 0;
    pl.cl = (pl.cg == 2 || pl.mc) ? 2 : 1;
    pl.n_rows = n_rows;
    pl.num_rb = int(ceil_div(n_rows, TC_BM * pl.cl));
    pl.num_u = pl.num_rb;
    pl.num_ct = int(ceil_div(n, TC_BN));
    pl.kl = list_len(k);
    shape_plan(pl, pl.num_rb, pl.num_ct, d_pad);
    pl.ct_per_split = int(ceil_div(pl.num_ct, pl.splits));     // trailing splits may be empty
    return pl;
}

// Threshold pre-pass: all rows against every stride-th column.  One unit per row block when the A
// blocks of all 148 CTAs plus the column sample fit the L2 (the bound then comes straight out of the
// kernel); otherwise 2 or 4 column splits run side by side on fewer row blocks and
// thr_merge_kernel (rerank.cu) combines their lists.
TcPlan tc_make_plan_prepass(int64_t n, int k, int64_t d_pad, int stride) {
    TcPlan pl{};
    pl.cg = 1; pl.cl = 1;
    pl.n_rows = n;
    pl.num_rb = int(ceil_div(n, TC_BM));
    pl.num_u = pl.num_rb;
    pl.num_ct = int(ceil_div(ceil_div(n, stride), TC_BN));
    pl.kl = list_len(k);
    const int64_t a_block = int64_t(TC_BM) * d_pad * 4;
    const int64_t sample = ceil_div(n, stride) * d_pad * 4;
    int cc = 1;
    while (cc < 4 && (kNumSMs / cc) * a_block + sample > (int64_t(100) << 20)) cc *= 2;
    if (const char* e = getenv("HG_B200_TC_PRE_CC")) {          // tuning / test override: 1, 2 or 4
        const int v = atoi(e);
        if (v == 1 || v == 2 || v == 4) cc = v;
    }
    if (cc > pl.num_ct) cc = 1;
    pl.cc = cc; pl.rc = kNumSMs / cc; pl.splits = cc;
    pl.ct_per_split = int(ceil_div(pl.num_ct, pl.splits));
    pl.grid = int(min64(kNumSMs, int64_t(pl.num_rb) * pl.splits));
    return pl;
}

// Lists per row in a symmetric pass over a giant slide.  The wider slack of the two-product pass puts ~24 candidates
// per row inside it at 100k rows and more the denser the slide; a list that fills up inside the slack sends its row to
// the exact fallback, which at this scale costs milliseconds per thousand rows.  A row block of an 8-way partitioned
// 250k slide would get 3 lists per row from the load balance alone (C5: 20 ms of fused kernel, 37 ms of fallback);
// six spread the candidates thin enough.
static int sym_min_splits(int64_t n) { return n >= 65536 ? 6 : 1; }

// Symmetric plan over the full row range: the classic unit grid, of which each unit computes
// only the tiles it owns (about every other one).
TcPlan tc_make_plan_sym(int64_t n, int k, int64_t d_pad) {
    TcPlan pl{};
    pl.cg = tc_sym_cta_group(d_pad);
    pl.mc = pl.cg == 1 ? tc_multicast() : 0;
    pl.cl = (pl.cg == 2 || pl.mc) ? 2 : 1;
    pl.sym = 1;
    pl.n_rows = n;
    pl.num_rb = int(ceil_div(n, TC_BM * pl.cl));
    pl.num_u = pl.num_rb;
    pl.num_ct = int(ceil_div(n, TC_BN));
    pl.kl = list_len(k);
    shape_plan(pl, pl.num_rb, pl.num_ct, d_pad, sym_min_splits(n));
    pl.ct_per_split = int(ceil_div(pl.num_ct, pl.splits));
    return pl;
}

TcPlan tc_make_plan_sym_rows(int64_t n, int64_t n_rows, int k, int64_t d_pad) {
    TcPlan pl{};
    pl.cg = tc_sym_cta_group(d_pad); pl.mc = 0; pl.cl = pl.cg;
    pl.sym = 1;
    pl.n_rows = n_rows;
    pl.num_rb = int(ceil_div(n_rows, TC_BM * pl.cl));
    pl.num_u = pl.num_rb;
    pl.num_ct = int(ceil_div(n, TC_BN));
    pl.kl = list_len(k);
    shape_plan(pl, pl.num_rb, pl.num_ct, d_pad, sym_min_splits(n));
    pl.ct_per_split = int(ceil_div(pl.num_ct, pl.splits));
    return pl;
}

// one unit per row block, every stride-th column: the bound comes straight out of the kernel (thr_out)
TcPlan tc_make_plan_prepass_rows(int64_t n, int64_t n_rows, int k, int64_t d_pad, int stride) {
    TcPlan pl{};
    pl.cg = 1; pl.cl = 1;
    pl.n_rows = n_rows;
    pl.num_rb = int(ceil_div(n_rows, TC_BM));
    pl.num_u = pl.num_rb;
    pl.num_ct = int(ceil_div(ceil_div(n, stride), TC_BN));
    pl.kl = list_len(k);
    pl.cc = 1; pl.rc = kNumSMs; pl.splits = 1;
    pl.ct_per_split = pl.num_ct;
    pl.grid = int(min64(kNumSMs, pl.num_rb));
    (void)d_pad;
    return pl;
}

// Symmetric mode pays a 1/stride pre-pass and needs enough sample columns for a useful bound; it
// wins from n ~ 9k up (measured at D = 1024: 2 % at n = 8192, 21 % at n = 10240; break-even at 9k x 512).  HG_B200_TC_SYM=0 disables it,
// HG_B200_TC_SYM=<stride> forces it (any n with enough sample columns) with that pre-pass stride.
int tc_sym_stride(int64_t n, int64_t n_rows, int k) {
    if (n_rows != n) return 0;
    int stride = (n >= 32768) ? 16 : 8;
    bool forced = false;
    if (const char* e = getenv("HG_B200_TC_SYM")) {
        const int v = atoi(e);
        if (v <= 0) return 0;
        if (v >= 2 && v <= 64) { stride = v; forced = true; }
    }
    if (!forced && n < 9216) return 0;
    if (n / stride < 4 * (k + 8)) return 0;
    return stride;
}

static PFN_cuTensorMapEncodeTiled_v12000 get_encode_fn() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
    if (fn) return fn;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
        return nullptr;
    fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
    return fn;
}

// row_stride > 1: a view of every row_stride-th operand row (the pre-pass column sample)
// hi_only: the box is the first 64 bytes (32 hi values) of every 128-byte k-block, 64-byte swizzle
static int make_tmap(CUtensorMap* tm, const op16_t* ops, int64_t n_pad, int64_t d_pad, int box_rows,
                     int row_stride = 1, bool hi_only = false) {
    auto enc = get_encode_fn();
    HG_REQUIRE(enc != nullptr, HG_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
    // operand row = d_pad / 32 blocks of (32 hi | 32 lo) FP16 = 4 * d_pad bytes; one box row = one such block
    const cuuint64_t row_bytes = cuuint64_t(ops_row_elems(d_pad)) * 2;
    cuuint64_t dims[3] = {cuuint64_t(d_pad) * 2, cuuint64_t(ceil_div(n_pad, row_stride)), 1};
    cuuint64_t strides[2] = {row_bytes * cuuint64_t(row_stride), row_bytes * cuuint64_t(n_pad)};
    cuuint32_t box[3] = {cuuint32_t(hi_only ? TC_BK : 2 * TC_BK), cuuint32_t(box_rows), 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<op16_t*>(ops), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, hi_only ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    HG_REQUIRE(r == CUDA_SUCCESS, HG_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d)", int(r));
    return HG_OK;
}

template <int KL, bool HAS_POS, int CG, bool SYM, bool MC = false>
static int launch_one(const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, cudaStream_t st) {
    auto kern = knn_tc_kernel<KL, HAS_POS, CG, SYM, MC>;
    constexpr int CL = (CG == 2 || MC) ? 2 : 1;
    constexpr size_t smem = tc_smem_bytes<CG>();
    static_assert(smem <= 232448, "over the 227 KB dynamic shared memory limit");
    HG_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(unsigned(p.plan.grid));
    cfg.blockDim = dim3(KL == 0 ? TC_THREADS_DENSE : TC_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CL; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = (CL > 1) ? 1 : 0;
    HG_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, p));
    count_launches(1);
    return HG_OK;
}

template <int KL>
static int launch_kl(const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, cudaStream_t st) {
    if constexpr (KL > 0) {
        if (p.plan.sym && p.plan.mc)
            return p.pos4 ? launch_one<KL, true, 1, true, true>(ta, tb, p, st) : launch_one<KL, false, 1, true, true>(ta, tb, p, st);
        if (p.plan.sym && p.plan.cg == 2)
            return p.pos4 ? launch_one<KL, true, 2, true>(ta, tb, p, st) : launch_one<KL, false, 2, true>(ta, tb, p, st);
        if (p.plan.sym)
            return p.pos4 ? launch_one<KL, true, 1, true>(ta, tb, p, st) : launch_one<KL, false, 1, true>(ta, tb, p, st);
        if (p.plan.mc)
            return p.pos4 ? launch_one<KL, true, 1, false, true>(ta, tb, p, st) : launch_one<KL, false, 1, false, true>(ta, tb, p, st);
    }
    if (p.plan.cg == 2)
        return p.pos4 ? launch_one<KL, true, 2, false>(ta, tb, p, st) : launch_one<KL, false, 2, false>(ta, tb, p, st);
    return p.pos4 ? launch_one<KL, true, 1, false>(ta, tb, p, st) : launch_one<KL, false, 1, false>(ta, tb, p, st);
}

// ops_a / ops_b: operand buffers behind the query rows and the candidate columns -- the same buffer for a
// slide's self-kNN, two buffers for a rectangular product (rows of one matrix against another's)
static int tc_launch_impl(const op16_t* ops_a, int64_t n_pad_a, const op16_t* ops, int64_t n_pad,
                          const TcParams& p_in, cudaStream_t st) {
    TcParams p = p_in;
    p.n_pad = n_pad;
    // A row blocks are re-read for every column tile (keep them), B tiles stream through.
    // HG_B200_TC_HINTS=<a><b> overrides: n = normal, f = evict first, l = evict last
    p.hint_a = ptx::kEvictLast; p.hint_b = ptx::kEvictNormal;
    // The producer refills the four 48 KB stages two at a time: 96 KB bursts keep the CTAs that sweep the
    // same B tiles close enough for the L2 to serve them (C3: 40.8 -> 37.6 ms; unpaired, DRAM reads were
    // 85 GB per launch against 18 GB), while the MMAs still start on the first half as soon as it lands.
    // HG_B200_TC_PAIR=0 refills stage by stage.
    p.pair_stages = 1;
    if (const char* e = getenv("HG_B200_TC_PAIR")) p.pair_stages = atoi(e) != 0 ? 1 : 0;
    if (ops_a != ops) { p.hint_a = ptx::kEvictNormal; p.hint_b = ptx::kEvictLast; }   // rectangular: the (small) B side is the reused one
    if (const char* e = getenv("HG_B200_TC_HINTS")) {
        auto pick = [](char c, uint64_t dflt) {
            return c == 'n' ? ptx::kEvictNormal : c == 'f' ? ptx::kEvictFirst : c == 'l' ? ptx::kEvictLast : dflt;
        };
        if (e[0] && e[1]) { p.hint_a = pick(e[0], p.hint_a); p.hint_b = pick(e[1], p.hint_b); }
    }
    if (p.col_stride < 1) p.col_stride = 1;
    HG_REQUIRE(!p.plan.sym || (p.plan.kl > 0 && p.col_stride == 1 && p.thr &&
                               ((p.peer_chunk == 0 && p.row_begin == 0 && p.row_end == p.n && p.bkt_cnt && p.bkt && p.ovf) ||
                                (p.peer_chunk > 0 && p.ovf && p.row_begin % 256 == 0 && p.sblk_base == p.row_begin / 256))),
               HG_ERR_BAD_ARG, "tc_launch: incomplete symmetric-mode parameters");
    HG_REQUIRE(p.col_stride == 1 || (p.plan.cl == 1 && p.plan.kl > 0), HG_ERR_BAD_ARG, "tc_launch: strided columns need single CTAs");
    HG_REQUIRE(!p.plan.mc || p.plan.kl > 0, HG_ERR_BAD_ARG, "tc_launch: multicast pairs keep lists (no dense mode)");
    CUtensorMap ta, tb;
    if (p.prods == 0) p.prods = 3;
    if (p.pre_main_prods == 0) p.pre_main_prods = p.prods;
    HG_REQUIRE(p.prods >= 1 && p.prods <= 3 && (p.prods > 1 || (p.plan.cg == 1 && p.col_stride > 1)), HG_ERR_BAD_ARG,
               "tc_launch: 1 to 3 products per k-step, the single product is for the threshold pre-pass");
    HG_REQUIRE(p.prods == 3 || p.plan.kl > 0, HG_ERR_BAD_ARG, "tc_launch: dense output needs all three products");
    HG_REQUIRE(p.norms_a != nullptr, HG_ERR_BAD_ARG, "tc_launch: the row norms behind the operand scales are missing");
    if (p.norms_b == nullptr) p.norms_b = p.norms_a;
#ifdef HG_TC_FORCE_PRODS          // timing experiments only, never defined by the Makefile (results are wrong): the
                                  // one-product run behind "64 B / clock / SM" in DESIGN.md 3.0 was built with -DHG_TC_FORCE_PRODS=1
    if (p.plan.kl > 0) p.prods = HG_TC_FORCE_PRODS;
#endif
    // No product reads the lo halves of B when there are two or one: the B tiles are then 64-byte rows, a stage is
    // 32 KB (24 KB for CTA pairs) instead of 48 (32) and more of them fit.  What binds the kernel is the 64 B / clock
    // an SM takes in from the L2: with whole B rows a tile needs 1280 D bytes for 16 D cycles of MMAs (2 products).
    p.b_hi_only = p.prods <= 2 ? 1 : 0;
    if (const char* e = getenv("HG_B200_TC_B_HI")) { if (atoi(e) == 0) p.b_hi_only = 0; }    // A/B switch
    {
        const int b_rows = TC_BN / (p.plan.cg == 2 ? 2 : 1);
        const int stage_bytes = TC_BM * 128 + b_rows * (p.b_hi_only ? 64 : 128);
        const int budget = (p.plan.cg == 2 ? 6 * 32 : 4 * 48) * 1024;
        p.stages = std::min(TC_MAX_STAGES, budget / stage_bytes) & ~1;               // even: stages are refilled in pairs
    }
    int rc = make_tmap(&ta, ops_a, n_pad_a, p.d_pad, TC_BM, 1);
    if (rc != HG_OK) return rc;
    rc = make_tmap(&tb, ops, n_pad, p.d_pad, TC_BN / p.plan.cl, p.col_stride, p.b_hi_only != 0);   // a CTA of a pair stages half of the B tile
    if (rc != HG_OK) return rc;
    switch (p.plan.kl) {
        case 0:  return launch_kl<0>(ta, tb, p, st);
        case 8:  return launch_kl<8>(ta, tb, p, st);
        case 16: return launch_kl<16>(ta, tb, p, st);
        case 24: return launch_kl<24>(ta, tb, p, st);
        case 32: return launch_kl<32>(ta, tb, p, st);
        default: break;
    }
    set_error("tc_launch: unsupported list length %d", p.plan.kl);
    return HG_ERR_UNSUPPORTED;
}

int tc_launch(const op16_t* ops, int64_t n_pad, const TcParams& p, cudaStream_t st) {
    return tc_launch_impl(ops, n_pad, ops, n_pad, p, st);
}

// Grouped launch: the units of several slides that share one operand buffer (TcUnit, knn_tc.cuh).  Classic list
// mode on single CTAs; p carries colp / pos4 / gscale / exclude_self / part_val / part_idx / d_pad, and n = n_pad =
// the rows of the whole group.
int tc_launch_units(const op16_t* ops, int64_t n_pad, int k, const TcUnit* unit_tab, int num_units,
                    const TcParams& p_in, cudaStream_t st) {
    HG_REQUIRE(unit_tab != nullptr && num_units > 0 && !p_in.thr_out && p_in.col_stride <= 1, HG_ERR_BAD_ARG,
               "tc_launch_units: bad argument");
    TcParams p = p_in;
    p.plan = TcPlan{};
    p.plan.cg = 1; p.plan.cl = 1; p.plan.kl = list_len(k);
    p.plan.grid = int(min64(kNumSMs, num_units));
    p.unit_tab = unit_tab; p.num_tab_units = num_units;
    p.n = n_pad; p.row_begin = 0; p.row_end = n_pad;
    return tc_launch_impl(ops, n_pad, ops, n_pad, p, st);
}
int tc_list_len(int k) { return list_len(k); }

// Rows of matrix A (p.row_begin .. p.row_end) against the p.n rows of matrix B: classic list mode only.
// p.colp / p.n describe B; nothing is excluded by index.
int tc_launch_rect(const op16_t* ops_a, int64_t n_pad_a, const op16_t* ops_b, int64_t n_pad_b,
                   const TcParams& p, cudaStream_t st) {
    HG_REQUIRE(ops_a != ops_b && !p.plan.sym && p.plan.cl == 1 && p.plan.kl > 0 && !p.exclude_self && !p.pos4 &&
               !p.thr_out && p.col_stride <= 1 && p.row_end <= n_pad_a,
               HG_ERR_BAD_ARG, "tc_launch_rect: classic list mode without positions only");
    return tc_launch_impl(ops_a, n_pad_a, ops_b, n_pad_b, p, st);
}

// One unit per 128-row block sweeping every column tile: the B side is small (centres), every A block is
// used once per column tile, so neither column splits nor an L2 chunking of the rows buy anything.
TcPlan tc_make_plan_rect(int64_t cols, int64_t n_rows, int k, int64_t d_pad) {
    TcPlan pl{};
    pl.cg = 1; pl.mc = 0; pl.cl = 1;
    pl.n_rows = n_rows;
    pl.num_rb = int(ceil_div(n_rows, TC_BM));
    pl.num_u = pl.num_rb;
    pl.num_ct = int(ceil_div(cols, TC_BN));
    pl.kl = list_len(k);
    pl.cc = 1; pl.rc = kNumSMs; pl.splits = 1;
    pl.ct_per_split = pl.num_ct;
    pl.grid = int(min64(kNumSMs, pl.num_rb));
    (void)d_pad;
    return pl;
}

}  // namespace hg
