// Persistent warp-specialised bf16 GEMM for the T5 linears (HF:277,298-299,338 q/k/v/o,
// HF:92-103 wi/wo), hand-written for sm_100a:
//   C[M,N] = A[M,K] . B[N,K]^T      A = activations (row-major), B = nn.Linear weight [out,in]
// TMA (128-byte swizzle) -> shared-memory ring -> tcgen05.mma with fp32 accumulators in TMEM
// (double-buffered) -> tcgen05.ld epilogue straight to global.  cta_group::2 pairs two SMs on a
// 256 x BLOCK_N tile (each CTA loads its half of A and B); cta_group::1 is the single-SM variant.
// Warp roles: 0 = TMA producer, 1 = MMA issuer (leader CTA only), 2 = TMEM allocator, 4..7 = epilogue.
#include <cstdlib>
#include <type_traits>
#include "pn_common.cuh"
#include "pn_tc.cuh"
#include "pn_t5_internal.cuh"

using namespace pntc;

namespace {

constexpr int BLOCK_M = 128;  // rows per CTA
constexpr int BLOCK_K = 64;   // one 128-byte swizzle atom of bf16
constexpr int UMMA_K = 16;
constexpr int kGemmThreads = 256;
constexpr int A_STAGE_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KB
constexpr int EPI_BUF_BYTES = 32 * 128;               // one epilogue warp's 32 rows x 128 B staging tile
constexpr int EPI_BYTES = 4 * 2 * EPI_BUF_BYTES;      // 4 epilogue warps, double-buffered: 32 KB
// TMA-fed residual epilogue: per warp a ring of kResBufs fp32 tiles [32 rows x 32 columns] (residual in, updated
// residual out, in place) and two bf16 tiles [32 x 32] for the copy of the updated stream.  The path is bound by the
// latency of the residual loads: [131072 x 768] x [768 x 768] takes 205 us with 3 tiles (5 stages), 195 us with 4 or 5
// (4 stages); requesting the tiles into L2 eight chunks ahead as well makes it slower (203 -> 209 us), and so does
// requesting two column-adjacent tiles per round instead of one per chunk (194 -> 198 us).
#ifndef PN_RES_BUFS
#define PN_RES_BUFS 5
#endif
constexpr int kResBufs = PN_RES_BUFS;   // tiles are requested kResBufs - 1 chunks ahead
constexpr int RES_Y_BYTES = 32 * 64;
constexpr int RES_WARP_BYTES = kResBufs * EPI_BUF_BYTES + 2 * RES_Y_BYTES;   // 24 KB
constexpr int EPI_BYTES_TMA = 4 * RES_WARP_BYTES;                             // 96 KB: 4 pipeline stages left
static_assert(8 * (2 * 8 + 5 + 4 * kResBufs) <= 512, "barrier area");
// Masked bf16 epilogue (kTmaMask): the staging tiles + a ring of kMaskBufs mask tiles per warp, requested
// kMaskBufs - 1 chunks ahead (into the next output tile).  2 tiles (one chunk ahead) left the HBM latency of every
// mask tile exposed: [131072 x 768] x [768 x 3072] 545 us against 443 us for the same-shape forward product.
constexpr int kMaskBufs = 4;
constexpr int EPI_BYTES_MASK = EPI_BYTES + 4 * kMaskBufs * EPI_BUF_BYTES;   // 96 KB: 4 pipeline stages
static_assert(kMaskBufs <= kResBufs, "mask ring uses the residual-tile barriers");
// Norm-backward residual epilogue (kNormRes = ring depth): per warp kNormRes fp32 tiles + as many bf16 tiles of the
// normed stream + the two bf16 copy tiles.  3 buffers = 22 KB per warp, 88 KB: 4 pipeline stages, as the plain TMA
// residual path; 4 buffers = 28 KB per warp, 112 KB: 3 stages; 2 buffers = 16 KB per warp, 64 KB: 5 stages.
// PN_NORM_BUFS (environment) picks, default 2.
__host__ __device__ constexpr int norm_warp_bytes(int bufs) { return bufs * (EPI_BUF_BYTES + RES_Y_BYTES) + 2 * RES_Y_BYTES; }
static_assert(4 <= kResBufs && norm_warp_bytes(2) % 1024 == 0 && norm_warp_bytes(3) % 1024 == 0 && norm_warp_bytes(4) % 1024 == 0, "norm-backward epilogue tiles");

template <int kCtaGroup, int BLOCK_N, int kEpiBytes = EPI_BYTES>
struct GemmCfg {
  static constexpr int EPI = kEpiBytes;
  static constexpr int B_ROWS = BLOCK_N / kCtaGroup;        // B rows this CTA loads
  static constexpr int B_STAGE_BYTES = B_ROWS * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int kAvail = 227 * 1024 - kEpiBytes - 1024 /*align slack*/ - 512 /*barriers*/;
  static constexpr int kStages = kAvail / STAGE_BYTES > 8 ? 8 : kAvail / STAGE_BYTES;
  static constexpr int kAccStages = 2;
  static constexpr int TMEM_COLS = kAccStages * BLOCK_N;    // 512 or 256 (power of two)
  static constexpr int SMEM_BYTES = kStages * STAGE_BYTES + kEpiBytes + 1024 /*align slack*/ + 512 /*barriers*/;
  static constexpr uint32_t IDESC = make_idesc_bf16(BLOCK_M * kCtaGroup, BLOCK_N, 0, 0);
};

struct GemmArgs {
  int M, N, K;
  void* out;               // bf16 [M, ldc] (mode 0) or fp32 [M, ldc] (mode 1, accumulated in place)
  int ldc;
  int mode;                // PN_EPI_*
  int splits;              // k-splits per output tile (> 1 only with PN_EPI_ATOMIC_F32)
  int relu;
  const float* row_scale;  // nullable: out = acc * row_scale[row]
  // fused T5 RMSNorm of the A operand (HF:55-69): A holds the un-normalised bf16 residual stream, the norm weight is
  // folded into B, and acc is scaled by rsqrt(sum_p sumsq_in[p*M + row] / norm_dim + eps)
  const float* sumsq_in;
  int sumsq_parts;
  int norm_dim;
  float eps;
  // residual mode extras: bf16 copy of the updated residual stream and per-n-tile row sums of squares of it
  __nv_bfloat16* copy_out;  // nullable, [M, ld_copy]
  int ld_copy;
  float* sumsq_out;         // nullable, [N / BLOCK_N][ss_stride]
  long long ss_stride;      // rows of the whole sumsq_out array (= M unless the launch is one row slab of a larger call)
  // bf16 mode, nullable: ReLU backward fused into a data-gradient GEMM, out = mask[row, col] > 0 ? acc : 0
  const __nv_bfloat16* mask;
  int ld_mask;
  // block-diagonal B (kTN = false): rows [g*group_rows, (g+1)*group_rows) of A multiply the B rows starting at
  // g*group_b_rows (one small B per document); group_rows is a multiple of the M tile.  0 = one B for all rows.
  int group_rows, group_b_rows;
  // batched kTN: split s covers the K rows of batch s and writes its own output at out + s*batch_stride
  long long batch_stride;
  int batches;
  PnDrop drop;             // dropout on the produced values (thresh 0 = off), see PnGemmExtra
  float out_scale;         // scalar factor folded into the row scale
  int m_major;             // tile walk: 1 = a CTA (pair) takes whole m-blocks and walks their n-tiles back to back
  float* dot_out;          // masked bf16 epilogue, nullable: [N / BLOCK_N][ss_stride] row sums of out * mask
  const float* nb_coef;    // kNormRes: out += acc - nb_coef[row] * hb[row, col]
  const __nv_bfloat16* nb_hb;
  int ld_hb;
};

// kTN = false: C[M,N] = A[M,K] . B[N,K]^T   (both operands K-major: forward and data-gradient GEMMs)
// kTN = true : C[M,N] += A[K,M]^T . B[K,N]   (both operands MN-major: weight gradients dW = dY^T X, K = rows);
//              the K range is split over args.splits work items per tile, accumulated with fp32 atomics
// kTmaRes  : PN_EPI_RESID_F32 with the residual tiles moved by TMA (tmC = the fp32 output as 32 x 32 boxes, tmY = the
//              bf16 copy as 32 x 32 boxes); trades pipeline stages for the 64 KB of epilogue tiles
// kTmaMask : PN_EPI_BF16 with a ReLU-backward mask (tmY = the mask as 32 x 64 boxes): each epilogue warp receives the
//              mask tile of its next chunk by TMA while it works on the current one and applies it thread-per-row
//              while packing (no per-thread global loads, no second pass over the staged tile); costs one stage
// kNormRes : PN_EPI_RESID_F32 of the data gradient of a normed linear with the RMSNorm backward folded in (tmC, tmY as
//              kTmaRes, tmH = the bf16 stream the norm read as 32 x 32 boxes): out += acc - coef[row] * hb[row, col]
template <int kCtaGroup, int BLOCK_N, bool kTN, bool kTmaRes, bool kTmaMask = false, int kNormRes = 0>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmY,
                 const __grid_constant__ CUtensorMap tmH, const GemmArgs args) {
  static_assert(!(kNormRes && (kTmaRes || kTmaMask)), "one epilogue variant per instantiation");
  using Cfg = GemmCfg<kCtaGroup, BLOCK_N,
                      kNormRes ? 4 * norm_warp_bytes(kNormRes) : (kTmaRes ? EPI_BYTES_TMA : (kTmaMask ? EPI_BYTES_MASK : EPI_BYTES))>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t epi_smem = smem_base + kStages * Cfg::STAGE_BYTES;   // 1024-aligned
  const uint32_t bar_base = epi_smem + Cfg::EPI;
  // barrier layout: full[kStages] empty[kStages] tmem_full[2] tmem_empty[2] | tmem ptr | residual-tile barriers
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (kStages + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * kStages + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * kStages + 2 + s); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * kStages + 4);
  auto res_bar = [&](int w, int b) { return bar_base + 8u * (2 * kStages + 5 + w * kResBufs + b); };
  uint32_t* tmem_slot_gen =
      reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta_rank = (kCtaGroup == 2) ? cluster_ctarank() : 0;
  const bool leader = cta_rank == 0;
  const int num_groups = gridDim.x / kCtaGroup;   // CTAs (1) or CTA pairs (2)
  const int group_id = blockIdx.x / kCtaGroup;

  const int tile_m = BLOCK_M * kCtaGroup;
  const int m_tiles = (args.M + tile_m - 1) / tile_m;
  const int n_tiles = args.N / BLOCK_N;
  const int num_out_tiles = m_tiles * n_tiles;
  const int num_tiles = num_out_tiles * args.splits;            // work items: (k-split, output tile)
  const int total_kb = (args.K + BLOCK_K - 1) / BLOCK_K;
  const int kb_per_split = (total_kb + args.splits - 1) / args.splits;
  // Work item -> output tile.  Round-robin over tiles (item % num_out_tiles) hands the n-tiles of one 256 x K block
  // of A to different CTA pairs, which drift apart over the ~80 waves of an M = 512 000 launch until the block has left
  // L2 between their reads (1.38x the DRAM bytes of the K = 3072 projection, profiles/r1t_ncu_full_summary.md).
  // m_major: the it-th item of pair g is n-tile (it % n_tiles) of m-block (it / n_tiles) * groups + g, so one pair
  // reads an A block n_tiles times back to back (and all pairs sweep the B tiles in step); the m-blocks left over
  // after the full rounds are dealt out tile by tile for balance.  A bijection on [0, num_out_tiles).
  const int mm_full_rounds = m_tiles / num_groups;
  const int mm_full_items = mm_full_rounds * n_tiles;              // items per pair in the full rounds
  auto tile_of = [&](int item) -> int {
    if (!args.m_major) return item % num_out_tiles;
    const int it_ = item / num_groups, g_ = item - it_ * num_groups;
    if (it_ < mm_full_items) return ((it_ / n_tiles) * num_groups + g_) * n_tiles + it_ % n_tiles;
    return item;   // tail: item >= mm_full_items * num_groups = first tile of the left-over m-blocks
  };
  auto kb_range = [&](int item, int& kb0, int& kb1) {
    const int sp = item / num_out_tiles;
    kb0 = sp * kb_per_split;
    kb1 = kb0 + kb_per_split < total_kb ? kb0 + kb_per_split : total_kb;
  };

  asm volatile("griddepcontrol.launch_dependents;");   // the next kernel's CTAs may take SMs as this grid's CTAs exit
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
    if (kTmaRes || kTmaMask || kNormRes) tma_prefetch_desc(&tmY);
    if (kNormRes) tma_prefetch_desc(&tmH);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 4 * kCtaGroup);  // one arrive per epilogue warp per CTA
    }
    if (kTmaRes || kTmaMask || kNormRes)
      for (int w = 0; w < 4; ++w)
        for (int b = 0; b < kResBufs; ++b) mbar_init(res_bar(w, b), 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc<kCtaGroup>(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish<kCtaGroup>();
  }
  tc_fence_before();
  if (kCtaGroup == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_gen;
  // Programmatic dependent launch (launch_gemm sets the stream-serialization attribute): everything above — barrier
  // init, tensor-map prefetch, TMEM allocation — may run while the previous kernel of the stream drains its last
  // tiles; nothing below touches global memory before that kernel has completed and flushed.
  asm volatile("griddepcontrol.wait;" ::: "memory");

  if (warp == 0 && lane == 0) {
    // ===================== TMA producer =====================
    int s = 0;
    uint32_t ph = 0;
    for (int item = group_id; item < num_tiles; item += num_groups) {
      const int tile = tile_of(item);
      const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
      const int m0 = m_blk * tile_m + (int)cta_rank * BLOCK_M;
      int n0 = n_blk * BLOCK_N + (int)cta_rank * Cfg::B_ROWS;
      if (!kTN && args.group_rows > 0) n0 += (m_blk * tile_m / args.group_rows) * args.group_b_rows;
      int kb0, kb1;
      kb_range(item, kb0, kb1);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(empty_bar(s), ph ^ 1);
        const uint32_t sa = smem_base + s * Cfg::STAGE_BYTES;
        const uint32_t sb = sa + A_STAGE_BYTES;
        if (kCtaGroup == 1) mbar_arrive_expect_tx(full_bar(s), Cfg::STAGE_BYTES);
        else if (leader) mbar_arrive_expect_tx(full_bar(s), 2 * Cfg::STAGE_BYTES);  // both CTAs' halves
        if (!kTN) {
          if (kCtaGroup == 1) {
            tma_load_2d(sa, &tmA, full_bar(s), kb * BLOCK_K, m0);
            tma_load_2d(sb, &tmB, full_bar(s), kb * BLOCK_K, n0);
          } else {
            tma_load_2d_2sm(sa, &tmA, full_bar(s), kb * BLOCK_K, m0);
            tma_load_2d_2sm(sb, &tmB, full_bar(s), kb * BLOCK_K, n0);
          }
        } else {
          // MN-major tiles: boxes of 64 k-rows x 64 columns (8 KB), one per 64-wide block of M / N
#pragma unroll
          for (int j = 0; j < BLOCK_M / 64; ++j) {
            if (kCtaGroup == 1) tma_load_2d(sa + j * 8192, &tmA, full_bar(s), m0 + j * 64, kb * BLOCK_K);
            else tma_load_2d_2sm(sa + j * 8192, &tmA, full_bar(s), m0 + j * 64, kb * BLOCK_K);
          }
#pragma unroll
          for (int j = 0; j < Cfg::B_ROWS / 64; ++j) {
            if (kCtaGroup == 1) tma_load_2d(sb + j * 8192, &tmB, full_bar(s), n0 + j * 64, kb * BLOCK_K);
            else tma_load_2d_2sm(sb + j * 8192, &tmB, full_bar(s), n0 + j * 64, kb * BLOCK_K);
          }
        }
        if (++s == kStages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1 && lane == 0 && leader) {
    // ===================== MMA issuer =====================
    int s = 0;
    uint32_t ph = 0;
    int it = 0;
    constexpr uint32_t IDESC = kTN ? make_idesc_bf16(BLOCK_M * kCtaGroup, BLOCK_N, 1, 1) : Cfg::IDESC;
    for (int item = group_id; item < num_tiles; item += num_groups, ++it) {
      const int as = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      int kb0, kb1;
      kb_range(item, kb0, kb1);
      mbar_wait_cluster(tempty_bar(as), aph ^ 1);  // epilogues of both CTAs drained this stage
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BLOCK_N;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(full_bar(s), ph);
        tc_fence_after();
        const uint32_t sa = smem_base + s * Cfg::STAGE_BYTES;
        const uint32_t sb = sa + A_STAGE_BYTES;
#pragma unroll
        for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
          uint64_t adesc, bdesc;
          if (!kTN) {
            // advance 32 bytes (16 bf16) inside the swizzle atom: +2 in the 16-byte address field
            adesc = make_desc_kmajor_sw128(sa) + 2u * k;
            bdesc = make_desc_kmajor_sw128(sb) + 2u * k;
          } else {
            // 16 k-rows of 128 B further down; 64-wide MN blocks are 8 KB apart (LBO), 8-row groups 1 KB (SBO)
            adesc = make_desc_mnmajor_sw128(sa + k * 2048, 8192);
            bdesc = make_desc_mnmajor_sw128(sb + k * 2048, 8192);
          }
          umma_bf16<kCtaGroup>(d_tmem, adesc, bdesc, IDESC, ((kb - kb0) | k) ? 1u : 0u);
        }
        umma_commit<kCtaGroup>(empty_bar(s));  // frees the smem slot (in both CTAs) when MMAs retire
        if (++s == kStages) { s = 0; ph ^= 1; }
      }
      umma_commit<kCtaGroup>(tfull_bar(as));   // accumulator ready (signalled in both CTAs)
    }
  } else if (warp >= 4 && kNormRes) {
    // ===================== norm-backward residual epilogue =====================
    // dh[row, col] += acc - coef[row] * hb[row, col] and the bf16 copy of the updated dh: the RMSNorm backward of a
    // normed linear finished inside its data-gradient GEMM.  Structure of the TMA residual epilogue below; every
    // ring slot holds the fp32 dh tile and the bf16 tile of the stream the norm read, both on the slot's mbarrier.
    const int ew = warp - 4;
    constexpr int kChunks = BLOCK_N / 32;
    constexpr int RB = kNormRes ? kNormRes : 1;
    const uint32_t wbase = epi_smem + ew * norm_warp_bytes(RB);
    const uint32_t hbase = wbase + RB * EPI_BUF_BYTES;
    const uint32_t ybase = hbase + RB * RES_Y_BYTES;
    const uint32_t swz = (uint32_t)(lane & 7), yswz = (uint32_t)((lane >> 1) & 3);
    static_assert(RB - 1 <= kChunks, "prefetch distance reaches at most into the next tile");
    auto load_tile = [&](uint32_t b, int r0, int c0) {
      mbar_arrive_expect_tx(res_bar(ew, b), EPI_BUF_BYTES + RES_Y_BYTES);
      tma_load_2d(wbase + b * EPI_BUF_BYTES, &tmC, res_bar(ew, b), c0, r0);
      tma_load_2d(hbase + b * RES_Y_BYTES, &tmH, res_bar(ew, b), c0, r0);
    };
    auto tile_origin = [&](int item, int& r0, int& c0) {
      const int tile = tile_of(item);
      r0 = (tile / n_tiles) * tile_m + (int)cta_rank * BLOCK_M + ew * 32;
      c0 = (tile % n_tiles) * BLOCK_N;
    };
    if (lane == 0 && group_id < num_tiles) {
      int r0, c0;
      tile_origin(group_id, r0, c0);
      for (int i = 0; i + 1 < RB; ++i) load_tile(i, r0, c0 + i * 32);
    }
    uint32_t rb = 0, rph = 0;
    int it = 0;
    uint32_t q = 0;
    for (int item = group_id; item < num_tiles; item += num_groups, ++it) {
      const int tile = tile_of(item);
      const int as = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
      const int row0 = m_blk * tile_m + (int)cta_rank * BLOCK_M + ew * 32;
      const int row = row0 + lane;
      const int col0 = n_blk * BLOCK_N;
      const bool live = row < args.M;
      float rs = args.out_scale;
      if (live && args.row_scale) rs *= args.row_scale[row];
      const float nk = live ? -args.nb_coef[row] : 0.0f;
      const uint32_t t_row = tmem_base + ((uint32_t)(ew * 32) << 16) + as * BLOCK_N;
      int nrow0 = 0, ncol0 = 0;
      const bool nvalid = item + num_groups < num_tiles;
      if (nvalid) tile_origin(item + num_groups, nrow0, ncol0);
      mbar_wait(tfull_bar(as), aph);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < kChunks; ++c, ++q) {
        const uint32_t rbuf = wbase + rb * EPI_BUF_BYTES + lane * 128;
        const uint32_t hbuf = hbase + rb * RES_Y_BYTES + lane * 64;
        const uint32_t ybuf = ybase + (q & 1) * RES_Y_BYTES + lane * 64;
        uint32_t v[32];
        tmem_ld_32x32(t_row + c * 32, v);
        mbar_wait(res_bar(ew, rb), rph);
        tmem_ld_wait();
        __syncwarp();   // lane 0 has seen the store of chunk q - 2 drain before ybuf is rewritten
        // args.drop: dropout mask of the PREVIOUS sub-layer's output, applied to the bf16 copy only (the copy is the
        // gradient of that sub-layer's linear output; dh itself is the gradient of the stream and stays unmasked)
        PnDropRun drun = {};
        unsigned long long dbase = 0;
        if (args.drop.thresh) {
          dbase = (unsigned long long)args.drop.idx_offset + (unsigned long long)row * args.N + (col0 + c * 32);
          drun = pn_drop_run(args.drop, dbase, 16);
        }
#pragma unroll
        for (int p = 0; p < 4; ++p) {
          uint32_t hw[4];   // 8 bf16 of the normed stream: 64-byte rows, piece p of row r at p ^ ((r >> 1) & 3)
          asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(hw[0]), "=r"(hw[1]), "=r"(hw[2]), "=r"(hw[3])
                       : "r"(hbuf + ((p ^ yswz) << 4)));
#pragma unroll
          for (int jj = 0; jj < 2; ++jj) {
            const int j = 2 * p + jj;
            const uint32_t addr = rbuf + ((j ^ swz) << 4);
            float4 h;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(h.x), "=f"(h.y), "=f"(h.z), "=f"(h.w) : "r"(addr));
            const uint32_t w0 = hw[2 * jj], w1 = hw[2 * jj + 1];
            h.x = fmaf(__uint_as_float(v[j * 4 + 0]), rs, fmaf(nk, __uint_as_float(w0 << 16), h.x));
            h.y = fmaf(__uint_as_float(v[j * 4 + 1]), rs, fmaf(nk, __uint_as_float(w0 & 0xffff0000u), h.y));
            h.z = fmaf(__uint_as_float(v[j * 4 + 2]), rs, fmaf(nk, __uint_as_float(w1 << 16), h.z));
            h.w = fmaf(__uint_as_float(v[j * 4 + 3]), rs, fmaf(nk, __uint_as_float(w1 & 0xffff0000u), h.w));
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(h.x), "f"(h.y), "f"(h.z), "f"(h.w)
                         : "memory");
            if (args.drop.thresh) {
              float m0, m1, m2, m3;
              pn_drop_run_get(args.drop, drun, dbase, 2 * j, m0, m1);
              pn_drop_run_get(args.drop, drun, dbase, 2 * j + 1, m2, m3);
              h.x *= m0; h.y *= m1; h.z *= m2; h.w *= m3;
            }
            __nv_bfloat162 lo = __floats2bfloat162_rn(h.x, h.y), hi = __floats2bfloat162_rn(h.z, h.w);
            v[j * 2 + 0] = *reinterpret_cast<uint32_t*>(&lo);   // v[0 .. 15] becomes the packed bf16 row
            v[j * 2 + 1] = *reinterpret_cast<uint32_t*>(&hi);
          }
        }
#pragma unroll
        for (int p = 0; p < 4; ++p)
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(ybuf + ((p ^ yswz) << 4)), "r"(v[p * 4 + 0]),
                       "r"(v[p * 4 + 1]), "r"(v[p * 4 + 2]), "r"(v[p * 4 + 3])
                       : "memory");
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmC, wbase + rb * EPI_BUF_BYTES, col0 + c * 32, row0);
          tma_store_2d(&tmY, ybase + (q & 1) * RES_Y_BYTES, col0 + c * 32, row0);
          tma_store_commit();
          tma_store_wait_read<1>();   // chunk q - 1 has left shared memory: its slot takes chunk q + RB - 1
          const uint32_t bl = rb == 0 ? RB - 1 : rb - 1;
          const int cn = c + RB - 1;
          if (cn < kChunks) load_tile(bl, row0, col0 + cn * 32);
          else if (nvalid) load_tile(bl, nrow0, ncol0 + (cn - kChunks) * 32);
        }
        if (++rb == RB) { rb = 0; rph ^= 1; }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (kCtaGroup == 1) mbar_arrive(tempty_bar(as));
        else mbar_arrive_remote_relaxed(tempty_bar(as), 0);
      }
    }
    if (lane == 0) tma_store_wait_all<0>();
  } else if (warp >= 4 && kTmaRes) {
    // ===================== residual epilogue, tiles moved by TMA =====================
    // out[row, col] += acc * rs, plus the bf16 copy and the row sums of squares of the updated values.  Each warp
    // streams its 32 rows as 32-column chunks through a ring of fp32 tiles in shared memory: the residual tile of
    // chunk q + 2 is requested (TMA load, own mbarrier) when chunk q has been handed to its TMA store, thread r
    // updates row r of the tile in place (128-byte swizzle: conflict-free) and the same bytes go back out.  No
    // global address arithmetic, no bound checks (the tensor maps zero-fill / clip rows beyond M), and the row sum
    // of squares is a per-thread scalar.
    const int ew = warp - 4;
    constexpr int kChunks = BLOCK_N / 32;
    const uint32_t wbase = epi_smem + ew * RES_WARP_BYTES;
    const uint32_t ybase = wbase + kResBufs * EPI_BUF_BYTES;
    const uint32_t swz = (uint32_t)(lane & 7), yswz = (uint32_t)((lane >> 1) & 3);
    // lane 0 requests tiles kResBufs - 1 chunks ahead; the origin of the NEXT work item is formed once per tile, so
    // the per-chunk issue path has no divisions (it sits on the warp's critical path: the others wait at __syncwarp)
    // ring slots this instantiation cycles through (the layout always holds kResBufs of them): the prefetch distance
    // reaches at most into the next tile
    constexpr int kRing = kResBufs - 1 <= kChunks ? kResBufs : kChunks + 1;
    auto load_tile = [&](uint32_t b, int r0, int c0) {
      mbar_arrive_expect_tx(res_bar(ew, b), EPI_BUF_BYTES);
      tma_load_2d(wbase + b * EPI_BUF_BYTES, &tmC, res_bar(ew, b), c0, r0);
    };
    auto tile_origin = [&](int item, int& r0, int& c0) {
      const int tile = tile_of(item);
      r0 = (tile / n_tiles) * tile_m + (int)cta_rank * BLOCK_M + ew * 32;
      c0 = (tile % n_tiles) * BLOCK_N;
    };
    if (lane == 0 && group_id < num_tiles) {
      int r0, c0;
      tile_origin(group_id, r0, c0);
      for (int i = 0; i + 1 < kRing; ++i) load_tile(i, r0, c0 + i * 32);
    }
    uint32_t rb = 0, rph = 0;   // ring slot and mbarrier phase of the current chunk
    int it = 0;
    uint32_t q = 0;
    for (int item = group_id; item < num_tiles; item += num_groups, ++it) {
      const int tile = tile_of(item);
      const int as = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
      const int row0 = m_blk * tile_m + (int)cta_rank * BLOCK_M + ew * 32;
      const int row = row0 + lane;
      const int col0 = n_blk * BLOCK_N;
      const bool live = row < args.M;
      float rs = args.out_scale;
      if (live) {
        if (args.row_scale) rs *= args.row_scale[row];
        if (args.sumsq_in) {
          float ss = 0.0f;
          for (int p = 0; p < args.sumsq_parts; ++p) ss += args.sumsq_in[(size_t)p * args.M + row];
          rs *= rsqrtf(ss / (float)args.norm_dim + args.eps);
        }
      }
      const uint32_t t_row = tmem_base + ((uint32_t)(ew * 32) << 16) + as * BLOCK_N;
      int nrow0 = 0, ncol0 = 0;
      const bool nvalid = item + num_groups < num_tiles;
      if (nvalid) tile_origin(item + num_groups, nrow0, ncol0);
      float ssq = 0.0f;
      mbar_wait(tfull_bar(as), aph);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < kChunks; ++c, ++q) {
        const uint32_t rbuf = wbase + rb * EPI_BUF_BYTES + lane * 128;
        const uint32_t ybuf = ybase + (q & 1) * RES_Y_BYTES + lane * 64;
        uint32_t v[32];
        tmem_ld_32x32(t_row + c * 32, v);
        mbar_wait(res_bar(ew, rb), rph);
        tmem_ld_wait();
        __syncwarp();   // lane 0 has seen the store of chunk q - 2 drain (its wait below) before ybuf is rewritten
        PnDropRun drun = {};
        unsigned long long dbase = 0;
        if (args.drop.thresh) {   // the chunk's 16 index pairs share the inner hash of the mask (pn_drop_run)
          dbase = (unsigned long long)args.drop.idx_offset + (unsigned long long)row * args.N + (col0 + c * 32);
          drun = pn_drop_run(args.drop, dbase, 16);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t addr = rbuf + ((j ^ swz) << 4);
          float4 h;
          asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(h.x), "=f"(h.y), "=f"(h.z), "=f"(h.w) : "r"(addr));
          float k0 = rs, k1 = rs, k2 = rs, k3 = rs;
          if (args.drop.thresh) {   // sub-layer output dropout: the mask multiplies the product, not the residual
            float m0, m1, m2, m3;
            pn_drop_run_get(args.drop, drun, dbase, 2 * j, m0, m1);
            pn_drop_run_get(args.drop, drun, dbase, 2 * j + 1, m2, m3);
            k0 *= m0; k1 *= m1; k2 *= m2; k3 *= m3;
          }
          h.x = fmaf(__uint_as_float(v[j * 4 + 0]), k0, h.x);
          h.y = fmaf(__uint_as_float(v[j * 4 + 1]), k1, h.y);
          h.z = fmaf(__uint_as_float(v[j * 4 + 2]), k2, h.z);
          h.w = fmaf(__uint_as_float(v[j * 4 + 3]), k3, h.w);
          asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(h.x), "f"(h.y), "f"(h.z), "f"(h.w)
                       : "memory");
          ssq += h.x * h.x + h.y * h.y + h.z * h.z + h.w * h.w;
          __nv_bfloat162 lo = __floats2bfloat162_rn(h.x, h.y), hi = __floats2bfloat162_rn(h.z, h.w);
          v[j * 2 + 0] = *reinterpret_cast<uint32_t*>(&lo);   // v[0 .. 15] becomes the packed bf16 row
          v[j * 2 + 1] = *reinterpret_cast<uint32_t*>(&hi);
        }
        if (args.copy_out) {
#pragma unroll
          for (int p = 0; p < 4; ++p)   // 64-byte rows, 64-byte swizzle: piece p of row r sits at p ^ ((r >> 1) & 3)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(ybuf + ((p ^ yswz) << 4)), "r"(v[p * 4 + 0]),
                         "r"(v[p * 4 + 1]), "r"(v[p * 4 + 2]), "r"(v[p * 4 + 3])
                         : "memory");
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          tma_store_2d(&tmC, wbase + rb * EPI_BUF_BYTES, col0 + c * 32, row0);
          if (args.copy_out) tma_store_2d(&tmY, ybase + (q & 1) * RES_Y_BYTES, col0 + c * 32, row0);
          tma_store_commit();
          tma_store_wait_read<1>();   // chunk q - 1 has left shared memory: its fp32 tile takes chunk q + kRing - 1
          const uint32_t bl = rb == 0 ? kRing - 1 : rb - 1;
          const int cn = c + kRing - 1;
          if (cn < kChunks) load_tile(bl, row0, col0 + cn * 32);
          else if (nvalid) load_tile(bl, nrow0, ncol0 + (cn - kChunks) * 32);
        }
        if (++rb == kRing) { rb = 0; rph ^= 1; }
      }
      if (args.sumsq_out && live) args.sumsq_out[(size_t)n_blk * args.ss_stride + row] = ssq;
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (kCtaGroup == 1) mbar_arrive(tempty_bar(as));
        else mbar_arrive_remote_relaxed(tempty_bar(as), 0);
      }
    }
    if (lane == 0) tma_store_wait_all<0>();
  } else if (warp >= 4) {
    // ===================== epilogue: TMEM -> registers -> swizzled smem tile -> global =====================
    // Each warp owns 32 accumulator rows (TMEM lane quarter = warp % 4) and a private double-buffered 4 KB
    // staging tile [32 rows x 128 B] in the 128-byte-swizzle layout: thread r writes row r conflict-free.
    //   bf16 modes : the tile is a 32 x 64 bf16 box, written back by one TMA store (coalesced, asynchronous)
    //   fp32 modes : the tile is a 32 x 32 fp32 box; the warp re-reads it row-contiguously (8 lanes = one 128 B
    //                row) so the residual read-modify-write, the bf16 copy and the fp32 store are all coalesced;
    //                the residual rows are prefetched into registers one chunk ahead
    const int ew = warp - 4;
    const uint32_t stage0 = epi_smem + ew * (2 * EPI_BUF_BYTES);
    uint8_t* stage_gen = smem_raw + (stage0 - smem_u32(smem_raw));
    const uint32_t swz = (uint32_t)(lane & 7);
    const int rrow = lane >> 3, rj = lane & 7;        // read-back mapping: iteration i -> row 4*i + rrow, chunk rj
    int it = 0;
    uint32_t cc = 0;                                  // chunks staged so far (buffer = cc & 1)
    // kTmaMask: a ring of kMaskBufs 4 KB mask tiles per warp behind the staging tiles; the tile of chunk cc +
    // kMaskBufs - 1 is requested while chunk cc is processed (slot = chunk % kMaskBufs, own mbarrier per slot)
    const uint32_t mstage0 = epi_smem + 4 * (2 * EPI_BUF_BYTES) + ew * (kMaskBufs * EPI_BUF_BYTES);
    uint32_t mslot = 0, mph = 0;                      // ring slot and mbarrier phase of the current chunk
    auto load_mask_tile = [&](uint32_t b, int r0, int c0) {
      mbar_arrive_expect_tx(res_bar(ew, b), EPI_BUF_BYTES);
      tma_load_2d(mstage0 + b * EPI_BUF_BYTES, &tmY, res_bar(ew, b), c0, r0);
    };
    auto warp_origin = [&](int item_, int& r0, int& c0) {
      const int t_ = tile_of(item_);
      r0 = (t_ / n_tiles) * tile_m + (int)cta_rank * BLOCK_M + ew * 32;
      c0 = (t_ % n_tiles) * BLOCK_N;
    };
    static_assert(!kTmaMask || kMaskBufs - 1 <= BLOCK_N / 64, "mask prefetch reaches at most into the next tile");
    if (kTmaMask && lane == 0 && group_id < num_tiles) {
      int r0, c0;
      warp_origin(group_id, r0, c0);
      for (int i = 0; i + 1 < kMaskBufs; ++i) load_mask_tile(i, r0, c0 + i * 64);
    }
    float rs_next = 1.0f;        // row scale of the NEXT tile's row, requested a tile ahead (an L2 round trip otherwise
    bool rs_next_ok = false;     // stalls the first multiply of every tile: 4 % of the masked epilogue's samples)
    for (int item = group_id; item < num_tiles; item += num_groups, ++it) {
      const int tile = tile_of(item);
      const int as = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      const int m_blk = tile / n_tiles, n_blk = tile % n_tiles;
      const int row0 = m_blk * tile_m + (int)cta_rank * BLOCK_M + ew * 32;   // first row of this warp
      const int row = row0 + lane;
      const int col0 = n_blk * BLOCK_N;
      const bool live = row < args.M;
      float rs = args.out_scale;
      if (live && args.row_scale) rs *= rs_next_ok ? rs_next : args.row_scale[row];
      rs_next_ok = false;
      if (args.row_scale && item + num_groups < num_tiles) {
        int nr0, nc0;
        warp_origin(item + num_groups, nr0, nc0);
        if (nr0 + lane < args.M) { rs_next = args.row_scale[nr0 + lane]; rs_next_ok = true; }
      }
      if (live) {
        if (args.sumsq_in) {
          float ss = 0.0f;
          for (int p = 0; p < args.sumsq_parts; ++p) ss += args.sumsq_in[(size_t)p * args.M + row];
          rs *= rsqrtf(ss / (float)args.norm_dim + args.eps);
        }
      }
      const uint32_t t_row = tmem_base + ((uint32_t)(ew * 32) << 16) + as * BLOCK_N;
      if (kTmaMask) {
        int nrow0 = 0, ncol0 = 0;
        const bool nvalid = item + num_groups < num_tiles;
        if (nvalid) warp_origin(item + num_groups, nrow0, ncol0);
        mbar_wait(tfull_bar(as), aph);
        tc_fence_after();
        float dsum = 0.0f;   // sum over the tile's columns of out * mask (the mask holds the forward activation)
#pragma unroll 1
        for (int c = 0; c < BLOCK_N / 64; ++c, ++cc) {
          const uint32_t buf = stage0 + (cc & 1) * EPI_BUF_BYTES;
          const uint32_t mbuf = mstage0 + mslot * EPI_BUF_BYTES + lane * 128;
          uint32_t v0[32], v1[32];
          tmem_ld_32x32(t_row + c * 64, v0);
          tmem_ld_32x32(t_row + c * 64 + 32, v1);
          if (lane == 0) tma_store_wait_read<1>();   // the store that used this staging buffer two chunks ago has drained
          __syncwarp();                              // ... and every lane is done with the mask tile of chunk cc - 1
          if (lane == 0) {                           // whose slot takes chunk cc + kMaskBufs - 1
            const uint32_t bl = mslot == 0 ? kMaskBufs - 1 : mslot - 1;
            const int cn = c + kMaskBufs - 1;
            if (cn < BLOCK_N / 64) load_mask_tile(bl, row0, col0 + cn * 64);
            else if (nvalid) load_mask_tile(bl, nrow0, ncol0 + (cn - BLOCK_N / 64) * 64);
          }
          mbar_wait(res_bar(ew, mslot), mph);
          if (++mslot == kMaskBufs) { mslot = 0; mph ^= 1; }
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t(&v)[32] = j < 4 ? v0 : v1;
            uint32_t mk0, mk1, mk2, mk3;
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(mk0), "=r"(mk1), "=r"(mk2), "=r"(mk3)
                         : "r"(mbuf + ((j ^ swz) << 4)));
            const uint32_t mk[4] = {mk0, mk1, mk2, mk3};
            uint32_t pk[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float x0 = __uint_as_float(v[(j & 3) * 8 + 2 * e]) * rs;
              const float x1 = __uint_as_float(v[(j & 3) * 8 + 2 * e + 1]) * rs;
              __nv_bfloat162 h = __floats2bfloat162_rn(x0, x1);
              // the mask holds ReLU outputs (>= 0): positive <=> any magnitude bit set
              pk[e] = *reinterpret_cast<uint32_t*>(&h) &
                      (((mk[e] & 0x00007fffu) ? 0x0000ffffu : 0u) | ((mk[e] & 0x7fff0000u) ? 0xffff0000u : 0u));
              dsum = fmaf(x0, __uint_as_float(mk[e] << 16), dsum);          // a zero activation contributes nothing
              dsum = fmaf(x1, __uint_as_float(mk[e] & 0xffff0000u), dsum);
            }
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(buf + lane * 128 + ((j ^ swz) << 4)),
                         "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3])
                         : "memory");
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmC, buf, col0 + c * 64, row0);   // rows beyond M are clipped by the tensor map
            tma_store_commit();
          }
        }
        if (args.dot_out && live) args.dot_out[(size_t)n_blk * args.ss_stride + row] = dsum;
      } else if (args.mode == PN_EPI_BF16) {
        // ReLU-backward mask: read with the coalesced mapping (lanes 0..7 = the eight 16-byte pieces of one 128-byte
        // row, four rows per instruction: 4 L1 wavefronts instead of the 32 of a thread-per-row read) and applied to
        // the staged tile in shared memory just before the TMA store.  The mask of the NEXT chunk is requested a
        // chunk ahead (the first one before the accumulator is even ready).
        uint4 mk[8], mkn[8];
        auto load_mask = [&](uint4(&dst)[8], int c) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int r = row0 + 4 * i + rrow;
            dst[i] = r < args.M ? __ldg(reinterpret_cast<const uint4*>(args.mask + (size_t)r * args.ld_mask + col0 +
                                                                       c * 64 + rj * 8))
                                : make_uint4(0u, 0u, 0u, 0u);
          }
        };
        if (args.mask) load_mask(mkn, 0);
        mbar_wait(tfull_bar(as), aph);
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < BLOCK_N / 64; ++c, ++cc) {
          const uint32_t buf = stage0 + (cc & 1) * EPI_BUF_BYTES;
          uint32_t v0[32], v1[32];
          tmem_ld_32x32(t_row + c * 64, v0);
          tmem_ld_32x32(t_row + c * 64 + 32, v1);
          if (args.mask) {
#pragma unroll
            for (int j = 0; j < 8; ++j) mk[j] = mkn[j];
            if (c + 1 < BLOCK_N / 64) load_mask(mkn, c + 1);
          }
          if (lane == 0) tma_store_wait_read<1>();   // the store that used this buffer two chunks ago has drained
          __syncwarp();
          PnDropRun drun = {};
          unsigned long long dbase = 0;
          if (args.drop.thresh) {   // the chunk's 32 index pairs share the inner hash of the mask (pn_drop_run)
            dbase = (unsigned long long)args.drop.idx_offset + (unsigned long long)row * args.N + (col0 + c * 64);
            drun = pn_drop_run(args.drop, dbase, 32);
          }
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t(&v)[32] = j < 4 ? v0 : v1;
            uint32_t pk[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              float x0 = __uint_as_float(v[(j & 3) * 8 + 2 * e]) * rs;
              float x1 = __uint_as_float(v[(j & 3) * 8 + 2 * e + 1]) * rs;
              if (args.relu) { x0 = fmaxf(x0, 0.0f); x1 = fmaxf(x1, 0.0f); }
              if (args.drop.thresh) {   // FFN hidden dropout (after the activation, HF:103)
                float m0, m1;
                pn_drop_run_get(args.drop, drun, dbase, j * 4 + e, m0, m1);
                x0 *= m0; x1 *= m1;
              }
              __nv_bfloat162 h = __floats2bfloat162_rn(x0, x1);
              pk[e] = *reinterpret_cast<uint32_t*>(&h);
            }
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(buf + lane * 128 + ((j ^ swz) << 4)),
                         "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3])
                         : "memory");
          }
          if (args.mask) {
            __syncwarp();
            // the mask holds ReLU outputs (>= 0): positive <=> any magnitude bit set
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int rr = 4 * i + rrow;
              const uint32_t addr = buf + rr * 128 + ((rj ^ (rr & 7)) << 4);
              uint32_t x0, x1, x2, x3;
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(x0), "=r"(x1), "=r"(x2), "=r"(x3) : "r"(addr));
              auto keep = [](uint32_t x, uint32_t m) {
                return x & (((m & 0x00007fffu) ? 0x0000ffffu : 0u) | ((m & 0x7fff0000u) ? 0xffff0000u : 0u));
              };
              x0 = keep(x0, mk[i].x); x1 = keep(x1, mk[i].y); x2 = keep(x2, mk[i].z); x3 = keep(x3, mk[i].w);
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(x0), "r"(x1), "r"(x2), "r"(x3)
                           : "memory");
            }
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmC, buf, col0 + c * 64, row0);   // rows beyond M are clipped by the tensor map
            tma_store_commit();
          }
        }
      } else {
        const bool resid = args.mode == PN_EPI_RESID_F32;
        const bool atomic = args.mode == PN_EPI_ATOMIC_F32;
        float* outf = reinterpret_cast<float*>(args.out) + (size_t)(item / num_out_tiles) * args.batch_stride;
        float ssq[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) ssq[i] = 0.0f;
        // The four epilogue warps are one warp per scheduler, so this path is bound by its own instruction count:
        // row pointers are formed once per tile (row i of the read-back mapping = base + i * 4 rows, a 32-bit
        // multiple folded into one IMAD.WIDE per access) and a slab that lies entirely inside M takes the copy of
        // the loop without the per-row bound checks.
        float* const orow = outf + (size_t)(row0 + rrow) * args.ldc + col0 + rj * 4;
        const uint32_t ostep = 4u * (uint32_t)args.ldc;
        __nv_bfloat16* const crow =
            args.copy_out ? args.copy_out + (size_t)(row0 + rrow) * args.ld_copy + col0 + rj * 4 : nullptr;
        const uint32_t cstep = 4u * (uint32_t)args.ld_copy;
        auto run = [&](auto full_tag) {
          constexpr bool kFull = decltype(full_tag)::value;
          auto row_ok = [&](int i) { return kFull || row0 + 4 * i + rrow < args.M; };
          float4 pre[2][8];                            // residual rows of the next two chunks (register prefetch)
          auto prefetch = [&](float4(&dst)[8], int c) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              dst[i] = (resid && c < BLOCK_N / 32 && row_ok(i))
                           ? *reinterpret_cast<const float4*>(orow + (size_t)(i * ostep) + c * 32)
                           : make_float4(0.f, 0.f, 0.f, 0.f);
          };
          prefetch(pre[0], 0);                         // issued before the accumulator is even ready
          prefetch(pre[1], 1);
          mbar_wait(tfull_bar(as), aph);
          tc_fence_after();
#pragma unroll 1
          for (int c2 = 0; c2 < BLOCK_N / 32; c2 += 2) {
#pragma unroll
            for (int u = 0; u < 2; ++u, ++cc) {
              const int c = c2 + u;
              uint8_t* bufg = stage_gen + (cc & 1) * EPI_BUF_BYTES;
              uint32_t v[32];
              tmem_ld_32x32(t_row + c * 32, v);
              tmem_ld_wait();
              __syncwarp();                            // every lane finished reading this buffer two chunks ago
              PnDropRun drun = {};
              unsigned long long dbase = 0;
              if (args.drop.thresh) {   // the chunk's 16 index pairs share the inner hash of the mask (pn_drop_run)
                dbase = (unsigned long long)args.drop.idx_offset + (unsigned long long)row * args.N + (col0 + c * 32);
                drun = pn_drop_run(args.drop, dbase, 16);
              }
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                float4 f;
                f.x = __uint_as_float(v[j * 4 + 0]) * rs;
                f.y = __uint_as_float(v[j * 4 + 1]) * rs;
                f.z = __uint_as_float(v[j * 4 + 2]) * rs;
                f.w = __uint_as_float(v[j * 4 + 3]) * rs;
                if (args.drop.thresh) {   // sub-layer output dropout (staged thread-per-row: this thread owns `row`)
                  float m0, m1, m2, m3;
                  pn_drop_run_get(args.drop, drun, dbase, 2 * j, m0, m1);
                  pn_drop_run_get(args.drop, drun, dbase, 2 * j + 1, m2, m3);
                  f.x *= m0; f.y *= m1; f.z *= m2; f.w *= m3;
                }
                *reinterpret_cast<float4*>(bufg + lane * 128 + ((j ^ swz) << 4)) = f;
              }
              __syncwarp();
              // read-back: rows 4*i + rrow; (rr & 7) = ((i & 1) * 4 + rrow) & 7 takes two values per lane
              const uint32_t sw_even = (uint32_t)((rj ^ (rrow & 7)) << 4), sw_odd = (uint32_t)((rj ^ ((rrow + 4) & 7)) << 4);
              const uint8_t* rb = bufg + rrow * 128;
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const float4 acc = *reinterpret_cast<const float4*>(rb + i * 512 + ((i & 1) ? sw_odd : sw_even));
                float4 h = pre[u][i];
                h.x += acc.x; h.y += acc.y; h.z += acc.z; h.w += acc.w;
                if (row_ok(i)) {
                  float4* dst = reinterpret_cast<float4*>(orow + (size_t)(i * ostep) + c * 32);
                  if (atomic) atomicAdd(dst, h);       // red.global.add.v4.f32
                  else *dst = h;
                  if (crow) {
                    __nv_bfloat162 lo = __floats2bfloat162_rn(h.x, h.y), hi = __floats2bfloat162_rn(h.z, h.w);
                    *reinterpret_cast<uint2*>(crow + (size_t)(i * cstep) + c * 32) =
                        make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
                  }
                }
                ssq[i] += h.x * h.x + h.y * h.y + h.z * h.z + h.w * h.w;
              }
              prefetch(pre[u], c + 2);   // the registers just consumed take the rows of two chunks ahead
            }
          }
        };
        if (row0 + 32 <= args.M) run(std::true_type{});
        else run(std::false_type{});
        if (args.sumsq_out) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float t = ssq[i];
            t += __shfl_xor_sync(0xffffffffu, t, 1);
            t += __shfl_xor_sync(0xffffffffu, t, 2);
            t += __shfl_xor_sync(0xffffffffu, t, 4);
            const int r = row0 + 4 * i + rrow;
            if (rj == 0 && r < args.M) args.sumsq_out[(size_t)n_blk * args.ss_stride + r] = t;
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (kCtaGroup == 1) mbar_arrive(tempty_bar(as));
        else mbar_arrive_remote_relaxed(tempty_bar(as), 0);  // the leader's MMA warp owns the wait
      }
    }
    if (lane == 0) tma_store_wait_all<0>();            // smem must outlive the last bulk stores
  }

  // ---- teardown (single-lane role loops rejoin their warps before the aligned barrier)
  __syncwarp();
  tc_fence_before();
  if (kCtaGroup == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 2) tmem_dealloc<kCtaGroup>(tmem_base, Cfg::TMEM_COLS);
}

// ---------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

}  // namespace

// row-major [rows, cols] tensor of bf16 or fp32, box = [box_rows, box_cols]
static int make_tmap_2d(CUtensorMap* out, CUtensorMapDataType dt, int elem_bytes, CUtensorMapSwizzle swz,
                        const void* base, uint64_t rows, uint64_t cols, uint64_t row_stride_elems,
                        uint32_t box_rows, uint32_t box_cols) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return pn_fail(PN_E_LAUNCH, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_elems * (uint64_t)elem_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return pn_fail(PN_E_LAUNCH, "cuTensorMapEncodeTiled failed (%d) rows=%llu cols=%llu", (int)r,
                   (unsigned long long)rows, (unsigned long long)cols);
  return PN_OK;
}

// bf16 row-major [rows, cols] tensor, box = [box_rows, 64 cols], 128-byte swizzle
int pn_make_tmap_bf16_2d(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols,
                         uint64_t row_stride_elems, uint32_t box_rows, uint32_t box_cols) {
  return make_tmap_2d(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, CU_TENSOR_MAP_SWIZZLE_128B, base, rows, cols,
                      row_stride_elems, box_rows, box_cols);
}

template <int kCtaGroup, int BLOCK_N, bool kTN, bool kTmaRes = false, bool kTmaMask = false, int kNormRes = 0>
static int launch_gemm(const void* A, int lda, const void* B, int ldb, GemmArgs args, cudaStream_t st) {
  using Cfg = GemmCfg<kCtaGroup, BLOCK_N,
                      kNormRes ? 4 * norm_warp_bytes(kNormRes) : (kTmaRes ? EPI_BYTES_TMA : (kTmaMask ? EPI_BYTES_MASK : EPI_BYTES))>;
  CUtensorMap tmA, tmB;
  int rc;
  if (!kTN) {
    rc = pn_make_tmap_bf16_2d(&tmA, A, args.M, args.K, lda, BLOCK_M, BLOCK_K);
    if (rc) return rc;
    const uint64_t b_rows = args.group_rows > 0
                                ? (uint64_t)((args.M + args.group_rows - 1) / args.group_rows) * args.group_b_rows
                                : (uint64_t)args.N;   // block-diagonal B: reads past the last group are zero-filled
    if (args.group_rows > 0 && (args.group_rows % (BLOCK_M * kCtaGroup)))
      return pn_fail(PN_E_BADARG, "pn_gemm: group_rows=%d must be a multiple of the %d-row tile", args.group_rows,
                     BLOCK_M * kCtaGroup);
    rc = pn_make_tmap_bf16_2d(&tmB, B, b_rows, args.K, ldb, Cfg::B_ROWS, BLOCK_K);
    if (rc) return rc;
  } else {   // operands are [K rows, M | N columns] row-major: 64 x 64 boxes
    rc = pn_make_tmap_bf16_2d(&tmA, A, args.K, args.M, lda, BLOCK_K, 64);
    if (rc) return rc;
    rc = pn_make_tmap_bf16_2d(&tmB, B, args.K, args.N, ldb, BLOCK_K, 64);
    if (rc) return rc;
  }
  CUtensorMap tmC = tmA;   // only the bf16 epilogue stores through a tensor map (32-row x 64-column boxes)
  if (args.mode == PN_EPI_BF16) {
    rc = pn_make_tmap_bf16_2d(&tmC, args.out, args.M, args.N, args.ldc, 32, 64);
    if (rc) return rc;
  }
  CUtensorMap tmY = tmA, tmH = tmA;
  if (kNormRes) {
    rc = make_tmap_2d(&tmH, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, CU_TENSOR_MAP_SWIZZLE_64B, args.nb_hb, args.M, args.N,
                      args.ld_hb, 32, 32);
    if (rc) return rc;
  }
  if (kTmaRes || kNormRes) {   // fp32 residual stream and its bf16 copy as 32-row x 32-column boxes (128 / 64 bytes per box row)
    rc = make_tmap_2d(&tmC, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, CU_TENSOR_MAP_SWIZZLE_128B, args.out, args.M, args.N,
                      args.ldc, 32, 32);
    if (rc) return rc;
    if (args.copy_out) {
      rc = make_tmap_2d(&tmY, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, CU_TENSOR_MAP_SWIZZLE_64B, args.copy_out, args.M,
                        args.N, args.ld_copy, 32, 32);
      if (rc) return rc;
    }
  }
  if (kTmaMask) {   // the ReLU-backward mask as 32-row x 64-column boxes (the layout of the staged output tile)
    rc = pn_make_tmap_bf16_2d(&tmY, args.mask, args.M, args.N, args.ld_mask, 32, 64);
    if (rc) return rc;
  }
  auto kern = gemm_bf16_kernel<kCtaGroup, BLOCK_N, kTN, kTmaRes, kTmaMask, kNormRes>;
  static bool attr_done = false;
  if (!attr_done) {
    PN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES));
    attr_done = true;
  }
  const int tile_m = BLOCK_M * kCtaGroup;
  const int tiles = ((args.M + tile_m - 1) / tile_m) * (args.N / BLOCK_N);
  const int max_groups = PN_SM_COUNT / kCtaGroup;
  args.splits = 1;
  if (kTN && args.batches > 0) {
    args.splits = args.batches;   // one work item per (batch, output tile); K rows of batch s = split s
  } else if (kTN) {
    // enough k-splits to give every SM (pair) a few work items, each at least 8 k-blocks deep
    const int total_kb = (args.K + BLOCK_K - 1) / BLOCK_K;
    int want = (2 * max_groups + tiles - 1) / tiles;
    if (want > total_kb / 8) want = total_kb / 8;
    if (want < 1) want = 1;
    const int per = (total_kb + want - 1) / want;
    args.splits = (total_kb + per - 1) / per;          // no empty split
  }
  int groups = max_groups;
  if (tiles * args.splits < groups) groups = tiles * args.splits;
  // measured and left OFF (profiles/r4o_gemm_ab.txt, same box): re-rank 8 087 against 8 523 pairs/s, training 110.3
  // against 108.0 ms — with every pair sweeping the same B tile at the same moment the weight tiles become an L2 hot
  // spot, which costs more than the A re-reads it removes
  static const int want_mmajor = [] { const char* e = getenv("PN_GEMM_MMAJOR"); return e ? atoi(e) : 0; }();
  args.m_major = (!kTN && args.splits == 1 && want_mmajor) ? 1 : 0;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(groups * kCtaGroup);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCtaGroup;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  // PN_GEMM_PDL=0: plain stream order (the kernel's griddepcontrol instructions are then no-ops)
  static const int want_pdl = [] { const char* e = getenv("PN_GEMM_PDL"); return e ? atoi(e) : 1; }();
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = want_pdl ? 2 : 1;
  // launches below ~200 FLOP per byte of compulsory traffic (skinny products, the fp32 residual read-modify-write
  // of the narrow output projections) are HBM-bound: they are timed as their own class with bytes as the work
  const double flops = 2.0 * args.M * args.N * args.K;
  const double out_b = args.mode == PN_EPI_BF16 ? 2.0 : (args.mode == PN_EPI_RESID_F32 ? 8.0 : 4.0);
  const double bytes = 2.0 * ((double)args.M * args.K + (double)(args.group_rows > 0 ? args.M / args.group_rows : 1) *
                                                            args.N * args.K) +
                       (double)args.M * args.N *
                           (out_b + (args.copy_out ? 2.0 : 0.0) + (args.mask ? 2.0 : 0.0) + (kNormRes ? 2.0 : 0.0));
  const bool hbm_bound = flops < 200.0 * bytes;
  PnProfScope prof(hbm_bound ? PN_PROF_GEMM_HBM : (kTN ? PN_PROF_WGRAD : PN_PROF_GEMM), hbm_bound ? bytes : flops, st);
  PN_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmC, tmY, tmH, args));
  g_pn_launches.fetch_add(1, std::memory_order_relaxed);
  return PN_OK;
}

// Batched C_b[M,N] (fp32, += into zeroed memory) = A_b[Kper,M]^T . B_b[Kper,N]: batch b owns rows
// [b*Kper, (b+1)*Kper) of A and B and the output block out + b*M*ldc.  Kper % 64 == 0.
int pn_gemm_tn_batched(const void* A, int lda, const void* B, int ldb, float* out, int ldc, int M, int N, int Kper,
                       int batches, cudaStream_t st) {
  if (M <= 0 || N <= 0 || Kper <= 0 || batches <= 0) return pn_fail(PN_E_BADARG, "pn_gemm_tn_batched: empty problem");
  if ((N % 128) || (M % 8) || (Kper % 64))
    return pn_fail(PN_E_UNSUPPORTED, "pn_gemm_tn_batched: need N %% 128 == 0, M %% 8 == 0, Kper %% 64 == 0");
  if (M > 128) return pn_fail(PN_E_UNSUPPORTED, "pn_gemm_tn_batched: M=%d > 128", M);
  if ((lda % 8) || (ldb % 8) || (ldc % 4) || ((uintptr_t)A & 15) || ((uintptr_t)B & 15) || ((uintptr_t)out & 15))
    return pn_fail(PN_E_BADARG, "pn_gemm_tn_batched: operands must be 16-byte aligned");
  GemmArgs a = {};
  a.M = M; a.N = N; a.K = Kper * batches; a.out = out; a.ldc = ldc; a.mode = PN_EPI_ATOMIC_F32; a.norm_dim = 1;
  a.out_scale = 1.0f;
  a.batches = batches; a.batch_stride = (long long)M * ldc;
  return (N % 256) == 0 ? launch_gemm<1, 256, true>(A, lda, B, ldb, a, st)
                        : launch_gemm<1, 128, true>(A, lda, B, ldb, a, st);
}

int pn_gemm_wgrad(const void* A, int lda, const void* B, int ldb, float* out, int ldc, int M, int N, int K,
                  int cta_group, cudaStream_t st) {
  if (M <= 0 || N <= 0 || K <= 0) return pn_fail(PN_E_BADARG, "pn_gemm_wgrad: empty problem");
  if (N % 128) return pn_fail(PN_E_UNSUPPORTED, "pn_gemm_wgrad: N=%d must be a multiple of 128", N);
  if (M % 8) return pn_fail(PN_E_UNSUPPORTED, "pn_gemm_wgrad: M=%d must be a multiple of 8", M);
  if ((lda % 8) || (ldb % 8) || (ldc % 4) || ((uintptr_t)A & 15) || ((uintptr_t)B & 15) || ((uintptr_t)out & 15))
    return pn_fail(PN_E_BADARG, "pn_gemm_wgrad: operands must be 16-byte aligned");
  GemmArgs a = {};
  a.M = M; a.N = N; a.K = K; a.out = out; a.ldc = ldc; a.mode = PN_EPI_ATOMIC_F32; a.norm_dim = 1;
  a.out_scale = 1.0f;
  const bool n256 = (N % 256) == 0;
  if (cta_group == 2)
    return n256 ? launch_gemm<2, 256, true>(A, lda, B, ldb, a, st) : launch_gemm<2, 128, true>(A, lda, B, ldb, a, st);
  if (cta_group == 1)
    return n256 ? launch_gemm<1, 256, true>(A, lda, B, ldb, a, st) : launch_gemm<1, 128, true>(A, lda, B, ldb, a, st);
  return pn_fail(PN_E_BADARG, "pn_gemm_wgrad: cta_group must be 1 or 2");
}

extern "C" int pn_gemm_bf16_wgrad(const void* A, int64_t lda, const void* B, int64_t ldb, float* out, int64_t ldc,
                                  int64_t M, int64_t N, int64_t K, int cta_group, void* stream) {
  PN_CHECK_ARG(A && B && out);
  return pn_gemm_wgrad(A, (int)lda, B, (int)ldb, out, (int)ldc, (int)M, (int)N, (int)K, cta_group,
                       (cudaStream_t)stream);
}

int pn_gemm_bf16_ex(const void* A, int lda, const void* B, int ldb, void* out, int ldc, int M, int N, int K,
                    int mode, int relu, const PnGemmExtra* ex, int cta_group, cudaStream_t st) {
  if (M <= 0 || N <= 0 || K <= 0) return pn_fail(PN_E_BADARG, "pn_gemm: empty problem");
  if (K % BLOCK_K) return pn_fail(PN_E_UNSUPPORTED, "pn_gemm: K=%d must be a multiple of 64", K);
  if (N % 128) return pn_fail(PN_E_UNSUPPORTED, "pn_gemm: N=%d must be a multiple of 128", N);
  if ((lda % 8) || (ldb % 8) || (ldc % 8) || ((uintptr_t)A & 15) || ((uintptr_t)B & 15) ||
      ((uintptr_t)out & 15))
    return pn_fail(PN_E_BADARG, "pn_gemm: operands must be 16-byte aligned with ld %% 8 == 0");
  if (mode != PN_EPI_BF16 && mode != PN_EPI_RESID_F32 && mode != PN_EPI_F32)
    return pn_fail(PN_E_BADARG, "pn_gemm: mode");
  GemmArgs a = {};
  a.M = M; a.N = N; a.K = K; a.out = out; a.ldc = ldc; a.mode = mode; a.relu = relu; a.splits = 1;
  a.row_scale = ex ? ex->row_scale : nullptr;
  a.sumsq_in = ex ? ex->sumsq_in : nullptr;
  a.sumsq_parts = ex ? ex->sumsq_parts : 0;
  a.norm_dim = ex ? ex->norm_dim : 1;
  a.eps = ex ? ex->eps : 0.0f;
  a.copy_out = ex ? ex->copy_out : nullptr;
  a.ld_copy = ex ? ex->ld_copy : 0;
  a.sumsq_out = ex ? ex->sumsq_out : nullptr;
  a.group_rows = ex ? ex->group_rows : 0;
  a.group_b_rows = ex ? ex->group_b_rows : 0;
  a.mask = ex ? ex->relu_mask : nullptr;
  a.ld_mask = ex ? ex->ld_mask : 0;
  a.out_scale = ex && ex->out_scale != 0.0f ? ex->out_scale : 1.0f;
  if (ex) a.drop = ex->drop;
  a.dot_out = ex ? ex->dot_out : nullptr;
  a.nb_coef = ex ? ex->nb_coef : nullptr;
  a.nb_hb = ex ? ex->nb_hb : nullptr;
  a.ld_hb = ex ? ex->ld_hb : 0;
  if (a.drop.thresh && (mode == PN_EPI_F32 || (N & 1)))
    return pn_fail(PN_E_BADARG, "pn_gemm: dropout is built into the bf16 and the residual epilogues only");
  if (a.mask && (mode != PN_EPI_BF16 || (a.ld_mask % 8) || ((uintptr_t)a.mask & 15)))
    return pn_fail(PN_E_BADARG, "pn_gemm: relu mask needs the bf16 epilogue, 16-byte alignment and ld %% 8 == 0");
  if (a.copy_out && ((a.ld_copy % 4) || ((uintptr_t)a.copy_out & 7)))
    return pn_fail(PN_E_BADARG, "pn_gemm: bf16 copy must be 8-byte aligned with ld %% 4 == 0");
  if ((a.copy_out || a.sumsq_out) && mode == PN_EPI_BF16)
    return pn_fail(PN_E_BADARG, "pn_gemm: copy/sumsq outputs need an fp32 epilogue mode");
  a.ss_stride = M;
  const bool n256 = (N % 256) == 0;
  static const int tma_mask = [] { const char* e = getenv("PN_GEMM_TMA_MASK"); return e ? atoi(e) : 1; }();
  if (a.dot_out && !(tma_mask && a.mask && mode == PN_EPI_BF16 && cta_group == 2 && n256 && !relu && !a.drop.thresh))
    return pn_fail(PN_E_UNSUPPORTED, "pn_gemm: dot_out needs the TMA-masked bf16 epilogue (2-CTA, N %% 256 == 0)");
  if (a.nb_coef) {   // RMSNorm backward folded into the data-gradient GEMM
    if (mode != PN_EPI_RESID_F32 || cta_group != 2 || !n256 || !a.nb_hb || !a.copy_out || (N & 1) ||
        a.sumsq_in || a.sumsq_out || a.group_rows || (ldc % 4) || (a.ld_copy % 8) || ((uintptr_t)a.copy_out & 15) ||
        (a.ld_hb % 8) || ((uintptr_t)a.nb_hb & 15))
      return pn_fail(PN_E_UNSUPPORTED, "pn_gemm: the norm-backward epilogue needs the 2-CTA residual mode with "
                                       "N %% 256 == 0, a bf16 copy output and 16-byte aligned operands");
    // same-box A/B (profiles/r4x_ab.txt): 2 slots / 5 stages 100.7 and 102.0 ms per training step, 3 slots / 4 stages
    // 100.9 and 102.4 ms, 4 slots / 3 stages +1.2 ms: the kernel is main-loop-bound, stages matter more than prefetch depth
    static const int norm_bufs = [] { const char* e = getenv("PN_NORM_BUFS"); return e ? atoi(e) : 2; }();
    return norm_bufs == 4   ? launch_gemm<2, 256, false, false, false, 4>(A, lda, B, ldb, a, st)
           : norm_bufs == 2 ? launch_gemm<2, 256, false, false, false, 2>(A, lda, B, ldb, a, st)
                            : launch_gemm<2, 256, false, false, false, 3>(A, lda, B, ldb, a, st);
  }
  // Deep-K residual launches over very many rows go out as row slabs: with the static tile order the CTA pairs that
  // share an A block drift apart over the ~80 waves of an M = 512 000 launch until the block has left L2 between
  // their reads (1.45x the DRAM reads, profiles/r2b_gemm_traffic.csv); at 131 072 rows per launch they do not.
  // PN_GEMM_SLAB = rows per slab (0 = one launch).
  static const int slab_rows = [] { const char* e = getenv("PN_GEMM_SLAB"); return e ? atoi(e) : 0; }();
  if (slab_rows >= 256 && mode == PN_EPI_RESID_F32 && cta_group == 2 && K >= 2048 && M >= 2 * slab_rows &&
      !a.row_scale && !a.sumsq_in && a.group_rows == 0 && !a.drop.thresh) {
    const int step = slab_rows / 256 * 256;
    for (int r0 = 0; r0 < M; r0 += step) {
      GemmArgs b = a;
      b.M = M - r0 < step ? M - r0 : step;
      b.out = reinterpret_cast<float*>(out) + (size_t)r0 * ldc;
      if (a.copy_out) b.copy_out = a.copy_out + (size_t)r0 * a.ld_copy;
      if (a.sumsq_out) b.sumsq_out = a.sumsq_out + r0;
      const __nv_bfloat16* Ab = reinterpret_cast<const __nv_bfloat16*>(A) + (size_t)r0 * lda;
      const int rc = n256 ? launch_gemm<2, 256, false>(Ab, lda, B, ldb, b, st) : launch_gemm<2, 128, false>(Ab, lda, B, ldb, b, st);
      if (rc) return rc;
    }
    return PN_OK;
  }
  // Residual epilogue through TMA for the launches whose epilogue is the bound (the narrow output projections,
  // < 256 FLOP per byte: 4.10 -> 4.48 TB/s for that class in the re-rank step, r2y); the deep-K FFN output projection
  // hides its epilogue behind the main loop and keeps the sixth pipeline stage instead (1 237 vs 1 215 TFLOP/s).
  // PN_GEMM_TMA_RES=0 keeps the register path everywhere, =2 forces TMA for every residual launch (A/B runs).
  static const int tma_res = [] { const char* e = getenv("PN_GEMM_TMA_RES"); return e ? atoi(e) : 1; }();
  const bool epi_bound = 2.0 * N * K < 256.0 * (2.0 * K + N * (a.copy_out ? 10.0 : 8.0));
  if (tma_res && (epi_bound || tma_res == 2) && mode == PN_EPI_RESID_F32 && cta_group == 2 && n256 && (ldc % 4) == 0 &&
      (!a.copy_out || ((a.ld_copy % 8) == 0 && ((uintptr_t)a.copy_out & 15) == 0)))
    return launch_gemm<2, 256, false, true>(A, lda, B, ldb, a, st);
  // masked (ReLU-backward) epilogue with the mask tile moved by TMA: the FFN data gradient [131072 x 768] x [768 x 3072]
  // took 704 us with the register path against 443 us for the same-shape forward GEMM (profiles/r4n_train_launches*);
  // same-box A/B of the training step: 110.1 / 110.4 ms with it, 113.1 / 112.0 ms without (profiles/r4o_gemm_ab.txt).
  // PN_GEMM_TMA_MASK=0 keeps the register path.
  if (tma_mask && a.mask && mode == PN_EPI_BF16 && cta_group == 2 && n256 && !relu && !a.drop.thresh)
    return launch_gemm<2, 256, false, false, true>(A, lda, B, ldb, a, st);
  if (cta_group == 2) {
    return n256 ? launch_gemm<2, 256, false>(A, lda, B, ldb, a, st) : launch_gemm<2, 128, false>(A, lda, B, ldb, a, st);
  } else if (cta_group == 1) {
    return n256 ? launch_gemm<1, 256, false>(A, lda, B, ldb, a, st) : launch_gemm<1, 128, false>(A, lda, B, ldb, a, st);
  }
  return pn_fail(PN_E_BADARG, "pn_gemm: cta_group must be 1 or 2");
}

// number of per-n-tile partial sums the residual epilogue writes per row for an N-wide output
int pn_gemm_sumsq_parts(int N) { return (N % 256) == 0 ? N / 256 : N / 128; }

int pn_gemm_bf16_internal(const void* A, int lda, const void* B, int ldb, void* out, int ldc, int M,
                          int N, int K, int mode, int relu, const float* row_scale, int cta_group,
                          cudaStream_t st) {
  PnGemmExtra ex = {};
  ex.row_scale = row_scale;
  return pn_gemm_bf16_ex(A, lda, B, ldb, out, ldc, M, N, K, mode, relu, &ex, cta_group, st);
}

extern "C" int pn_gemm_bf16(const void* A, int64_t lda, const void* B, int64_t ldb, void* out,
                            int64_t ldc, int64_t M, int64_t N, int64_t K, int mode, int relu,
                            const float* row_scale, int cta_group, void* stream) {
  PN_CHECK_ARG(A && B && out);
  return pn_gemm_bf16_internal(A, (int)lda, B, (int)ldb, out, (int)ldc, (int)M, (int)N, (int)K, mode,
                               relu, row_scale, cta_group, (cudaStream_t)stream);
}

/* building blocks of the fused norm backward: see include/pn_b200.h */
extern "C" int pn_gemm_bf16_relu_bwd(const void* dY, int64_t ld_dy, const void* Wt, int64_t ldw, const void* act,
                                     int64_t ld_act, const float* row_scale, void* out, int64_t ldc, float* dot_out,
                                     int64_t M, int64_t N, int64_t K, void* stream) {
  PN_CHECK_ARG(dY && Wt && act && out);
  PnGemmExtra ex = {};
  ex.row_scale = row_scale; ex.relu_mask = (const __nv_bfloat16*)act; ex.ld_mask = (int)ld_act; ex.dot_out = dot_out;
  return pn_gemm_bf16_ex(dY, (int)ld_dy, Wt, (int)ldw, out, (int)ldc, (int)M, (int)N, (int)K, PN_EPI_BF16, 0, &ex, 2,
                         (cudaStream_t)stream);
}

extern "C" int pn_gemm_bf16_norm_bwd(const void* Ys, int64_t ld_ys, const void* Wt, int64_t ldw, const void* hb,
                                     int64_t ld_hb, const float* coef, float* dh, int64_t ld_dh, void* dhb,
                                     int64_t ld_dhb, int64_t M, int64_t N, int64_t K, void* stream) {
  PN_CHECK_ARG(Ys && Wt && hb && coef && dh && dhb);
  PnGemmExtra ex = {};
  ex.nb_coef = coef; ex.nb_hb = (const __nv_bfloat16*)hb; ex.ld_hb = (int)ld_hb;
  ex.copy_out = (__nv_bfloat16*)dhb; ex.ld_copy = (int)ld_dhb;
  return pn_gemm_bf16_ex(Ys, (int)ld_ys, Wt, (int)ldw, dh, (int)ld_dh, (int)M, (int)N, (int)K, PN_EPI_RESID_F32, 0, &ex,
                         2, (cudaStream_t)stream);
}
