// Backward of the T5 encoder self-attention (autograd of HF modeling_t5.py:308-336 under lceT5.py:155):
//   P = softmax(Q K^T + bias),  O = P V
//   dV = P^T dO,   dP = dO V^T,   dS = P (.) (dP - D),  D_i = sum_d dO_id O_id,
//   dQ = dS K,     dK = dS^T Q,   dbias[h, j - i] += dS_ij        (no 1/sqrt(d) scale in T5)
// Flash-style on tcgen05, P recomputed from the forward's log2-domain row logsumexp; nothing [S,S]-sized
// ever reaches HBM.  One templated kernel, two launches:
//   MODE 0 (dK, dV, dbias): a work item is (doc, 128-key block); K, V stay resident, the query blocks stream
//   MODE 1 (dQ)           : a work item is (doc, 128-query block); Q, dO stay resident, the key blocks stream
// Per inner step both modes form S = Q K^T and dP = dO V^T as 128x128 UMMAs into TMEM (rows = queries), 256
// threads (two per query row, 64 key columns each) turn them into P and dS (bf16, swizzled shared-memory
// tiles), and the accumulating UMMAs consume those tiles MN-major (P^T dO, dS^T Q) or K-major (dS K).
// A CTA serves one head for its whole life, so the per-diagonal bias gradient accumulates in shared memory
// (warp-skewed shuffle reduction: one shared atomic per diagonal per warp) and is flushed once.
// Warp roles: 0..7 compute, 8 = TMA producer, 9 = TMEM allocator + MMA issuer, 10..13 (MODE 0) = bias-gradient
// diagonal sums read back from the dS tile.
#include "pn_common.cuh"
#include <stdlib.h>
#include "pn_tc.cuh"
#include "pn_t5_internal.cuh"

using namespace pntc;

namespace {

constexpr int kThreadsKV = 512;   // MODE 0: + 4 diagonal warps (+ 2 idle: setmaxnreg works on whole warpgroups)
constexpr int kThreadsQ = 512;    // MODE 1: 6 idle warps only donate registers
constexpr int DK = 64;
constexpr int BLK = 128;                       // rows per query / key block
constexpr int TILE_BYTES = BLK * DK * 2;       // 16 KB: [128 x 64] bf16, 128-byte swizzle
constexpr int PT_BYTES = 2 * TILE_BYTES;       // [128 q x 128 k] bf16 as two 64-column sub-tiles
constexpr int MAX_S = 512;
constexpr int BIAS_W = 2 * MAX_S - 1;
constexpr int BIAS_STRIDE = 1032;
constexpr float kLog2e = 1.4426950408889634f;

// smem: resident pair x2 | streamed pair x2 | P tile | dS tile | bias copies | diag accumulators | barriers
// The resident pair is double-buffered by item: the next item's tiles arrive while this one is still being worked on
// (its first S / dP products are issued before this item's last accumulating ones, and the item epilogue may still
// read the current pair for the row dots), so no item boundary waits for a TMA load.
constexpr int SM_RES = 0;                              // 2 buffers x ([0] = Q or K, [1] = dO or V)
constexpr int SM_STR = 4 * TILE_BYTES;                 // 2 buffers x (2 tiles)
constexpr int SM_P = SM_STR + 4 * TILE_BYTES;
constexpr int SM_DS = SM_P + PT_BYTES;
constexpr int SM_BIAS = SM_DS + PT_BYTES;
constexpr int SM_DIAG = SM_BIAS + 4 * BIAS_STRIDE * 4;
constexpr int SM_BAR = SM_DIAG + 1024 * 4;
enum { B_RES_FULL0 = 0, B_RES_FULL1, B_RES_FREE0, B_RES_FREE1, B_STR_FULL0, B_STR_FULL1, B_STR_EMPTY0, B_STR_EMPTY1, B_SP_FULL, B_SP_FREE,
       B_PD_READY, B_PD_FREE, B_ACC_FULL, B_ACC_FREE, NUM_BARS };
constexpr int BWD_SMEM = SM_BAR + NUM_BARS * 8 + 16 + 1024;

constexpr uint32_t TM_S = 0, TM_DP = 128, TM_ACC0 = 256, TM_ACC1 = 320, TM_COLS = 512;
constexpr uint32_t TM_DS = 320;   // MODE 1: dS as packed bf16 (64 columns), the A operand of dQ += dS K

struct BwdArgs {
  int S, n_heads, docs, cph;
  int q_col, k_col, v_col;          // column blocks of qkv / dqkv
  const float* bias;                // [H, 1023]
  const float* lse2;                // [rows, H] log2-domain logsumexp from the forward
  const float* dvec;                // [rows, H] D_i = sum_d dO O
  const float* row_scale;           // nullable [rows]: dq/dk/dv rows are scaled (fused-norm backward)
  __nv_bfloat16* dqkv;              // [rows, ld] gradients, same layout as qkv
  int ld;
  float* dbias_diag;                // [H, 1023] += per-diagonal sums (MODE 0 only)
  PnDrop drop;                      // dropout of the probabilities in the forward (HF:331), regenerated here
  const __nv_bfloat16* qkv;         // the forward's q | k | v rows (same layout as dqkv), read only when dot_out is set
  float* dot_out;                   // nullable [4 * H][dot_stride]: partial row sums of dqkv * qkv (see internal.cuh)
  long long dot_stride;
};

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

// packed fp32 pairs (sm_100 FFMA2 / FADD2 / FMUL2)
__device__ __forceinline__ uint64_t f2pack(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void f2unpack(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}

template <int MODE>
__global__ void __maxnreg__(120)
t5_attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                   const BwdArgs a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  float* sbias = reinterpret_cast<float*>(gbase + SM_BIAS);
  float* sdiag = reinterpret_cast<float*>(gbase + SM_DIAG);
  const uint32_t bar0 = base + SM_BAR;
  auto bar = [&](int i) { return bar0 + 8u * i; };
  const uint32_t tmem_slot = bar0 + 8u * NUM_BARS;
  uint32_t* tmem_slot_gen = reinterpret_cast<uint32_t*>(gbase + SM_BAR + 8 * NUM_BARS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int head = blockIdx.x / a.cph, c_in_head = blockIdx.x % a.cph;
  const int S = a.S;
  const int nb = (S + BLK - 1) / BLK;                 // blocks per document (queries and keys alike)
  const int items_total = a.docs * nb;
  const int n_items = c_in_head < items_total ? (items_total - c_in_head + a.cph - 1) / a.cph : 0;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmDO);
    for (int b = 0; b < 2; ++b) {
      mbar_init(bar(B_RES_FULL0 + b), 1);
      mbar_init(bar(B_RES_FREE0 + b), a.dot_out ? 8 : 1);   // row dots: the compute warps read the resident tiles last
    }
    mbar_init(bar(B_STR_FULL0), 1);
    mbar_init(bar(B_STR_FULL1), 1);
    mbar_init(bar(B_STR_EMPTY0), 1);
    mbar_init(bar(B_STR_EMPTY1), 1);
    mbar_init(bar(B_SP_FULL), 1);
    mbar_init(bar(B_SP_FREE), 8);
    mbar_init(bar(B_PD_READY), 8);
    mbar_init(bar(B_PD_FREE), MODE == 0 ? 5 : 1);   // tcgen05.commit (+ the 4 diagonal warps)
    mbar_init(bar(B_ACC_FULL), 1);
    mbar_init(bar(B_ACC_FREE), 8);
    fence_mbar_init();
  }
  if (warp == 9) {
    tmem_alloc<1>(tmem_slot, TM_COLS);
    tmem_relinquish<1>();
  }
  for (int i = threadIdx.x; i < 4 * BIAS_STRIDE; i += blockDim.x) {
    const int sh = i / BIAS_STRIDE, k = i % BIAS_STRIDE;
    const int src = k + sh;
    sbias[i] = src < BIAS_W ? a.bias[(size_t)head * BIAS_W + src] * kLog2e : 0.0f;
  }
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) sdiag[i] = 0.0f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot_gen;

  // item i of this CTA -> (doc, outer block); inner step t -> inner block
  auto item_doc = [&](int i) { return (c_in_head + i * a.cph) / nb; };
  auto item_ob = [&](int i) { return (c_in_head + i * a.cph) % nb; };
  // tile sources: MODE 0 resident = (K, V) of key block ob, streamed = (Q, dO) of query block t
  //               MODE 1 resident = (Q, dO) of query block ob, streamed = (K, V) of key block t
  const int col_q = a.q_col + head * DK, col_k = a.k_col + head * DK, col_v = a.v_col + head * DK;
  const int col_do = head * DK;

  // MODE 0 registers: the 8 compute warps take what the other two warpgroups give up
  if (warp == 8) {
    reg_dealloc<40>();
    if (lane == 0) {
      // ===================== TMA producer =====================
      int g = 0;
      for (int it = 0; it < n_items; ++it) {
        const int row0 = item_doc(it) * S, ob = item_ob(it);
        const int rbuf = it & 1;
        const uint32_t rdst = base + SM_RES + rbuf * 2 * TILE_BYTES;
        mbar_wait(bar(B_RES_FREE0 + rbuf), ((it >> 1) & 1) ^ 1);
        mbar_arrive_expect_tx(bar(B_RES_FULL0 + rbuf), 2 * TILE_BYTES);
        if (MODE == 0) {
          tma_load_2d(rdst, &tmQKV, bar(B_RES_FULL0 + rbuf), col_k, row0 + ob * BLK);
          tma_load_2d(rdst + TILE_BYTES, &tmQKV, bar(B_RES_FULL0 + rbuf), col_v, row0 + ob * BLK);
        } else {
          tma_load_2d(rdst, &tmQKV, bar(B_RES_FULL0 + rbuf), col_q, row0 + ob * BLK);
          tma_load_2d(rdst + TILE_BYTES, &tmDO, bar(B_RES_FULL0 + rbuf), col_do, row0 + ob * BLK);
        }
        for (int t = 0; t < nb; ++t, ++g) {
          const int b = g & 1;
          mbar_wait(bar(B_STR_EMPTY0 + b), ((g >> 1) & 1) ^ 1);
          mbar_arrive_expect_tx(bar(B_STR_FULL0 + b), 2 * TILE_BYTES);
          const uint32_t dst = base + SM_STR + b * 2 * TILE_BYTES;
          if (MODE == 0) {
            tma_load_2d(dst, &tmQKV, bar(B_STR_FULL0 + b), col_q, row0 + t * BLK);
            tma_load_2d(dst + TILE_BYTES, &tmDO, bar(B_STR_FULL0 + b), col_do, row0 + t * BLK);
          } else {
            tma_load_2d(dst, &tmQKV, bar(B_STR_FULL0 + b), col_k, row0 + t * BLK);
            tma_load_2d(dst + TILE_BYTES, &tmQKV, bar(B_STR_FULL0 + b), col_v, row0 + t * BLK);
          }
        }
      }
    }
  } else if (warp == 9) {
    reg_dealloc<40>();
    // ===================== MMA issuer =====================
    // The whole warp runs this loop so that every shared-memory descriptor is warp-uniform (uniform registers,
    // no per-instruction vector->uniform moves: a [128 x 64 x 16] MMA is only 32 tensor-core cycles of work and
    // would otherwise be bound by its own issue); one lane issues.
    constexpr uint32_t IDESC_SP = make_idesc_bf16(128, BLK, 0, 0);
    constexpr uint32_t IDESC_TN = make_idesc_bf16(128, DK, 1, 1);    // A = P^T / dS^T (MN-major), B MN-major
    constexpr uint32_t IDESC_NN = make_idesc_bf16(128, DK, 0, 1);    // A = dS from TMEM, B = K MN-major
    const int total = n_items * nb;
    const bool leader = lane == 0;
    // K-major descriptors (S = Q K^T, dP = dO V^T) and MN-major ones (accumulating products); a 16-element K step
    // advances a K-major descriptor by 2 (32 B) and an MN-major one by 128 (16 rows x 128 B)
    const uint64_t p_mn = make_desc_mnmajor_sw128(base + SM_P, TILE_BYTES);
    const uint64_t ds_mn = make_desc_mnmajor_sw128(base + SM_DS, TILE_BYTES);
    auto issue_sp = [&](int g) {
      const int it = g / nb, t = g % nb;
      const uint32_t str = base + SM_STR + (g & 1) * 2 * TILE_BYTES;
      const uint32_t res = base + SM_RES + (it & 1) * 2 * TILE_BYTES;
      const uint64_t res0_k = make_desc_kmajor_sw128(res), res1_k = make_desc_kmajor_sw128(res + TILE_BYTES);
      const uint64_t str0_k = make_desc_kmajor_sw128(str), str1_k = make_desc_kmajor_sw128(str + TILE_BYTES);
      // MODE 0: streamed = (Q, dO), resident = (K, V);  MODE 1: resident = (Q, dO), streamed = (K, V)
      const uint64_t dq = MODE == 0 ? str0_k : res0_k, ddo = MODE == 0 ? str1_k : res1_k;
      const uint64_t dk = MODE == 0 ? res0_k : str0_k, dv = MODE == 0 ? res1_k : str1_k;
      mbar_wait(bar(B_STR_FULL0 + (g & 1)), (g >> 1) & 1);
      if (t == 0) mbar_wait(bar(B_RES_FULL0 + (it & 1)), (it >> 1) & 1);
      if (g > 0) mbar_wait(bar(B_SP_FREE), (g - 1) & 1);   // compute threads hold S, dP of the previous step
      tc_fence_after();
      if (leader) {
#pragma unroll
        for (int k = 0; k < DK / 16; ++k) umma_bf16<1>(tmem + TM_S, dq + 2u * k, dk + 2u * k, IDESC_SP, k ? 1u : 0u);
#pragma unroll
        for (int k = 0; k < DK / 16; ++k) umma_bf16<1>(tmem + TM_DP, ddo + 2u * k, dv + 2u * k, IDESC_SP, k ? 1u : 0u);
        umma_commit<1>(bar(B_SP_FULL));
      }
      __syncwarp();
    };
    auto issue_acc = [&](int g) {
      const int it = g / nb, t = g % nb;
      const uint32_t str = base + SM_STR + (g & 1) * 2 * TILE_BYTES;
      const uint64_t str0_mn = make_desc_mnmajor_sw128(str, TILE_BYTES);
      const uint64_t str1_mn = make_desc_mnmajor_sw128(str + TILE_BYTES, TILE_BYTES);
      mbar_wait(bar(B_PD_READY), g & 1);
      if (t == 0 && it > 0) mbar_wait(bar(B_ACC_FREE), (it - 1) & 1);
      tc_fence_after();
      const uint32_t first = t ? 1u : 0u;
      if (leader) {
        if (MODE == 0) {
          // dV[k, :] += sum_q P[q, k] dO[q, :]      dK[k, :] += sum_q dS[q, k] Q[q, :]
#pragma unroll
          for (int kk = 0; kk < BLK / 16; ++kk)
            umma_bf16<1>(tmem + TM_ACC0, p_mn + 128u * kk, str1_mn + 128u * kk, IDESC_TN, kk ? 1u : first);
#pragma unroll
          for (int kk = 0; kk < BLK / 16; ++kk)
            umma_bf16<1>(tmem + TM_ACC1, ds_mn + 128u * kk, str0_mn + 128u * kk, IDESC_TN, kk ? 1u : first);
        } else {
          // dQ[q, :] += sum_k dS[q, k] K[k, :]
#pragma unroll
          for (int kk = 0; kk < BLK / 16; ++kk)
            umma_bf16_ts(tmem + TM_ACC0, tmem + TM_DS + kk * 8, str0_mn + 128u * kk, IDESC_NN, kk ? 1u : first);
        }
        umma_commit<1>(bar(B_PD_FREE));
        umma_commit<1>(bar(B_STR_EMPTY0 + (g & 1)));
        if (t == nb - 1) {
          umma_commit<1>(bar(B_ACC_FULL));
          if (!a.dot_out) umma_commit<1>(bar(B_RES_FREE0 + (it & 1)));
        }
      }
      __syncwarp();
    };
    if (total > 0) issue_sp(0);
    for (int g = 0; g < total; ++g) {
      // S / dP of the next step are formed while the compute warps work on this one — across item boundaries too
      // (the next item's resident pair sits in the other buffer)
      if (g + 1 < total) issue_sp(g + 1);
      issue_acc(g);
    }
  } else if (warp < 8) {
    // ===================== compute warps: thread = (query row r, 64-key half) =====================
    if (MODE == 0) reg_alloc<184>();   // per scheduler: 2 x 184 + 72 (diagonal warp) + 40 = 4 x 120
    else reg_alloc<200>();             //                2 x 200 + 2 x 40
    const int w4 = warp & 3, half = warp >> 2;
    const int r = w4 * 32 + lane;                   // query row inside the block = TMEM lane
    const uint32_t lane_addr = (uint32_t)(w4 * 32) << 16;
    const int sh = 3 - (r & 3);                     // shifted bias copy that keeps this row 16-byte aligned
    const float* bcopy = sbias + sh * BIAS_STRIDE;
    uint8_t* prow = gbase + SM_P + half * TILE_BYTES + r * 128;
    uint8_t* dsrow = gbase + SM_DS + half * TILE_BYTES + r * 128;
    // row statistics (-lse2, D) of step g2's query row; fetched one step ahead so the loads never stall a step
    const int total = n_items * nb;
    auto row_stats = [&](int g2, float& nl, float& dd) {
      nl = 0.0f; dd = 0.0f;
      if (g2 < total) {
        const int it2 = g2 / nb;
        const int qi2 = (MODE == 0 ? g2 % nb : item_ob(it2)) * BLK + r;
        if (qi2 < S) {
          const size_t o = (size_t)(item_doc(it2) * S + qi2) * a.n_heads + head;
          nl = a.lse2[o];   // raw values: nothing may depend on these loads before the next step
          dd = a.dvec[o];
        }
      }
    };
    float lse, dv, lse_nx, dv_nx;
    row_stats(0, lse, dv);
    int g = 0;
    for (int it = 0; it < n_items; ++it) {
      const int row0 = item_doc(it) * S, ob = item_ob(it);
      for (int t = 0; t < nb; ++t, ++g) {
        const int qb = MODE == 0 ? t : ob, kb = MODE == 0 ? ob : t;
        const int qi = qb * BLK + r;
        const int k0 = kb * BLK + half * 64;
        const int kvalid = qi < S ? min(64, S - k0) : 0;   // keys of this half that exist (0 for a padding row)
        const float nlse = -lse;
        const bool full = __all_sync(0xffffffffu, kvalid == 64);   // no padding rows or keys in this warp's block
        // dropout mask of row qi: element [doc, head, qi, key] at an even row pitch.  With dropout O = (P (.) M) V:
        // dV sums P (.) M, and dS = P (.) (M (.) dP - D) with D = rowsum(dO (.) O) as before.
        const bool dropping = a.drop.thresh != 0;
        const unsigned long long dbase =
            (unsigned long long)a.drop.idx_offset +
            ((unsigned long long)(item_doc(it) * a.n_heads + head) * S + (qi < S ? qi : 0)) * (S + (S & 1)) + k0;
        PnDropRun drun = {};
        bool dfast = true;
        if (dropping) {   // the 32 index pairs of this thread's 64 keys share their inner hash (pn_drop_run)
          drun = pn_drop_run(a.drop, dbase, 32);
          dfast = __all_sync(0xffffffffu, drun.fast);
        }
        mbar_wait(bar(B_SP_FULL), g & 1);
        tc_fence_after();
        uint32_t pp[32], pd[32];                    // P and dS of this thread's 64 columns, packed bf16
        uint32_t sraw[2][32], dpraw[2][32];
        tmem_ld_32x32(tmem + lane_addr + TM_S + half * 64, sraw[0]);
        tmem_ld_32x32(tmem + lane_addr + TM_DP + half * 64, dpraw[0]);
        tmem_ld_32x32(tmem + lane_addr + TM_S + half * 64 + 32, sraw[1]);
        tmem_ld_32x32(tmem + lane_addr + TM_DP + half * 64 + 32, dpraw[1]);
        tmem_ld_wait();
        // S, dP are in registers: the tensor core forms the next step's while this one is processed
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(B_SP_FREE));
#pragma unroll
        for (int c2 = 0; c2 < 2; ++c2) {
          const uint32_t(&s)[32] = sraw[c2];
          const uint32_t(&dp)[32] = dpraw[c2];
          const float4* bp = reinterpret_cast<const float4*>(bcopy + (k0 + c2 * 32 - qi + (MAX_S - 1) - sh));
          if (full) {
            // packed fp32 pairs (FFMA2 / FADD2 / FMUL2): half the issue slots of the scalar form
            const uint64_t l2 = f2pack(kLog2e, kLog2e), nl2 = f2pack(nlse, nlse), ndv2 = f2pack(-dv, -dv);
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 b = bp[j >> 2];
#pragma unroll
              for (int h2 = 0; h2 < 2; ++h2) {
                const uint64_t sv = f2pack(__uint_as_float(s[j + 2 * h2]), __uint_as_float(s[j + 2 * h2 + 1]));
                const uint64_t bv = h2 ? f2pack(b.z, b.w) : f2pack(b.x, b.y);
                const uint64_t x = add2(fma2(sv, l2, bv), nl2);
                float x0, x1;
                f2unpack(x, x0, x1);
                const float p0 = ex2_approx(x0), p1 = ex2_approx(x1);
                uint64_t dpv = f2pack(__uint_as_float(dp[j + 2 * h2]), __uint_as_float(dp[j + 2 * h2 + 1]));
                float m0 = 1.0f, m1 = 1.0f;
                if (dropping) {
                  if (dfast) pn_drop_run_mul2(a.drop, drun, (c2 * 32 + j + 2 * h2) >> 1, m0, m1);
                  else pn_drop_mul2(a.drop, dbase + c2 * 32 + j + 2 * h2, m0, m1);
                  dpv = mul2(dpv, f2pack(m0, m1));
                }
                const uint64_t dsv = mul2(f2pack(p0, p1), add2(dpv, ndv2));
                float d0, d1;
                f2unpack(dsv, d0, d1);
                pp[c2 * 16 + (j >> 1) + h2] = pack2(p0 * m0, p1 * m1);
                pd[c2 * 16 + (j >> 1) + h2] = pack2(d0, d1);
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 b = bp[j >> 2];
              const float bb[4] = {b.x, b.y, b.z, b.w};
              float p[4], d[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float x = fmaf(__uint_as_float(s[j + e]), kLog2e, bb[e]) + nlse;
                p[e] = (c2 * 32 + j + e < kvalid) ? ex2_approx(x) : 0.0f;
                const float mk = pn_drop_mul(a.drop, dbase + c2 * 32 + j + e);
                d[e] = p[e] * (__uint_as_float(dp[j + e]) * mk - dv);
                p[e] *= mk;
              }
              pp[c2 * 16 + (j >> 1)] = pack2(p[0], p[1]);
              pp[c2 * 16 + (j >> 1) + 1] = pack2(p[2], p[3]);
              pd[c2 * 16 + (j >> 1)] = pack2(d[0], d[1]);
              pd[c2 * 16 + (j >> 1) + 1] = pack2(d[2], d[3]);
            }
          }
        }
        // next step's row statistics: requested here, where S / dP have left the registers (asked for at the top of
        // the step the two values were spilled at once, and the spill store waited for the loads: 7 % of the pass's
        // samples), and still a barrier wait, the tile writes and the next step's TMEM loads ahead of their first use
        row_stats(g + 1, lse_nx, dv_nx);
        if (g > 0) mbar_wait(bar(B_PD_FREE), (g - 1) & 1);   // previous step's accumulating MMAs read P, dS
        if (MODE == 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int ch = (j ^ (r & 7)) << 4;      // 128-byte swizzle: 16-byte chunk j of row r
            *reinterpret_cast<uint4*>(prow + ch) = make_uint4(pp[4 * j], pp[4 * j + 1], pp[4 * j + 2], pp[4 * j + 3]);
            *reinterpret_cast<uint4*>(dsrow + ch) = make_uint4(pd[4 * j], pd[4 * j + 1], pd[4 * j + 2], pd[4 * j + 3]);
          }
          fence_proxy_async();
        } else {
          // dS stays in tensor memory: rows = queries, two keys per 32-bit column (TS-mode operand of dQ += dS K)
          tc_fence_after();
          tmem_st_32x32(tmem + lane_addr + TM_DS + half * 32, pd);
          tmem_st_wait();
          tc_fence_before();
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(B_PD_READY));
        lse = lse_nx;
        dv = dv_nx;

        if (t == nb - 1) {
          // ---- item epilogue: dV, dK of key block ob (MODE 0) or dQ of query block ob (MODE 1)
          // row-dot operands: the forward's v | k (MODE 0) or q (MODE 1) values of this thread's 32 columns, read from
          // the resident tiles (row r, 16-byte chunks half*4 .. half*4+3 under the 128-byte swizzle: conflict-free,
          // where the same bytes from global memory would be 32 sectors per instruction).  With dot_out set the
          // resident pair is released by these warps instead of by the last MMA's commit.
          uint4 xq[4], yq[4];
          if (a.dot_out) {
            mbar_wait(bar(B_RES_FULL0 + (it & 1)), (it >> 1) & 1);   // completed long ago: orders these reads after the TMA writes
            const uint32_t rsrc = base + SM_RES + (it & 1) * 2 * TILE_BYTES;
            const uint32_t t0 = rsrc + (MODE == 0 ? TILE_BYTES : 0) + r * 128;
            const uint32_t t1 = rsrc + r * 128;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t ch = (uint32_t)(((half * 4 + j) ^ (r & 7)) << 4);
              asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                           : "=r"(xq[j].x), "=r"(xq[j].y), "=r"(xq[j].z), "=r"(xq[j].w) : "r"(t0 + ch));
              if (MODE == 0)
                asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                             : "=r"(yq[j].x), "=r"(yq[j].y), "=r"(yq[j].z), "=r"(yq[j].w) : "r"(t1 + ch));
            }
          }
          mbar_wait(bar(B_ACC_FULL), it & 1);
          tc_fence_after();
          uint32_t a0[32], a1[32];
          tmem_ld_32x32(tmem + lane_addr + TM_ACC0 + half * 32, a0);
          if (MODE == 0) tmem_ld_32x32(tmem + lane_addr + TM_ACC1 + half * 32, a1);
          tmem_ld_wait();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            mbar_arrive(bar(B_ACC_FREE));
            if (a.dot_out) mbar_arrive(bar(B_RES_FREE0 + (it & 1)));   // MMAs done (B_ACC_FULL) and this warp's reads too
          }
          const int oi = ob * BLK + r;
          const bool live = oi < S;
          const size_t grow = (size_t)(row0 + (live ? oi : 0));
          const float sc = (live && a.row_scale) ? a.row_scale[grow] : 1.0f;
          if (a.dot_out && live) {
            // <dq, q> (MODE 1) or <dv, v> + <dk, k> (MODE 0) over this thread's 32 columns: with the forward's
            // q | k | v = rs h W'^T these sums are the row dot <G, h> of the norm behind the projection
            float ds = 0.0f;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint4 x = xq[j];
              const uint32_t xw[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                ds = fmaf(__uint_as_float(a0[8 * j + 2 * e]), __uint_as_float(xw[e] << 16), ds);
                ds = fmaf(__uint_as_float(a0[8 * j + 2 * e + 1]), __uint_as_float(xw[e] & 0xffff0000u), ds);
              }
              if (MODE == 0) {
                const uint4 y = yq[j];
                const uint32_t yw[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  ds = fmaf(__uint_as_float(a1[8 * j + 2 * e]), __uint_as_float(yw[e] << 16), ds);
                  ds = fmaf(__uint_as_float(a1[8 * j + 2 * e + 1]), __uint_as_float(yw[e] & 0xffff0000u), ds);
                }
              }
            }
            a.dot_out[(size_t)((MODE * a.n_heads + head) * 2 + half) * a.dot_stride + grow] = ds * sc;
          }
          // The gradient rows leave through this warp's own 32 rows of the (now idle) P tile: thread-per-row writes
          // under an XOR swizzle, read back so that one store instruction covers 4 rows x 128 B (MODE 0: 64 B of dV
          // and 64 B of dK per row) or 8 rows x 64 B (MODE 1) — 8 segments per instruction instead of 32.
          uint8_t* stg = gbase + SM_P + half * TILE_BYTES + (w4 * 32) * 128;
          const int sw = lane & 7;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 u;
            u.x = pack2(__uint_as_float(a0[8 * j + 0]) * sc, __uint_as_float(a0[8 * j + 1]) * sc);
            u.y = pack2(__uint_as_float(a0[8 * j + 2]) * sc, __uint_as_float(a0[8 * j + 3]) * sc);
            u.z = pack2(__uint_as_float(a0[8 * j + 4]) * sc, __uint_as_float(a0[8 * j + 5]) * sc);
            u.w = pack2(__uint_as_float(a0[8 * j + 6]) * sc, __uint_as_float(a0[8 * j + 7]) * sc);
            *reinterpret_cast<uint4*>(stg + lane * 128 + ((j ^ sw) << 4)) = u;
            if (MODE == 0) {
              u.x = pack2(__uint_as_float(a1[8 * j + 0]) * sc, __uint_as_float(a1[8 * j + 1]) * sc);
              u.y = pack2(__uint_as_float(a1[8 * j + 2]) * sc, __uint_as_float(a1[8 * j + 3]) * sc);
              u.z = pack2(__uint_as_float(a1[8 * j + 4]) * sc, __uint_as_float(a1[8 * j + 5]) * sc);
              u.w = pack2(__uint_as_float(a1[8 * j + 6]) * sc, __uint_as_float(a1[8 * j + 7]) * sc);
              *reinterpret_cast<uint4*>(stg + lane * 128 + (((4 + j) ^ sw) << 4)) = u;
            }
          }
          __syncwarp();
          {
            const int blk_row0 = ob * BLK + w4 * 32;                     // first row of this warp inside the document
            __nv_bfloat16* gout = a.dqkv + (size_t)(row0 + blk_row0) * a.ld + head * DK + half * 32;
            if (MODE == 0) {
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                const int rl = k * 4 + (lane >> 3), c = lane & 7;
                const uint4 v = *reinterpret_cast<const uint4*>(stg + rl * 128 + ((c ^ (rl & 7)) << 4));
                if (blk_row0 + rl < S)
                  *reinterpret_cast<uint4*>(gout + (size_t)rl * a.ld + (c < 4 ? a.v_col + c * 8 : a.k_col + (c - 4) * 8)) = v;
              }
            } else {
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const int rl = k * 8 + (lane >> 2), c = lane & 3;
                const uint4 v = *reinterpret_cast<const uint4*>(stg + rl * 128 + ((c ^ (rl & 7)) << 4));
                if (blk_row0 + rl < S) *reinterpret_cast<uint4*>(gout + (size_t)rl * a.ld + a.q_col + c * 8) = v;
              }
            }
          }
          __syncwarp();   // the staging rows are this warp's P rows of the next step
        }
      }
    }
  } else {
    // ===================== diagonal warps (MODE 0): dbias[j - i] += dS_ij =====================
    // The dS tile [128 q x 128 k] is cut into 8-row x 8-diagonal parallelograms: thread (c, sub) owns the wrapped
    // diagonals (k - q) mod 128 in [8c, 8c + 8) of the row blocks a = sub and sub + 8.  Row 8a + rr reads them
    // from the two 16-byte chunks (a + c) mod 16 (elements rr..7) and (a + c + 1) mod 16 (elements 0..rr-1),
    // conflict-free, with compile-time register indices; a chunk that wrapped past column 127 belongs to the
    // diagonal 128 below.  The eight threads of a class then reduce-scatter by shuffles, so every diagonal of
    // the tile has exactly one owner: no atomics.
    if (MODE == 0 && warp < 14) reg_dealloc<72>();
    else reg_dealloc<40>();
    if (MODE == 0 && warp < 14) {
      const int u = threadIdx.x - 320;
      const int c = u >> 3, sub = u & 7;
      const uint8_t* ds = gbase + SM_DS;
      int g = 0;
      for (int it = 0; it < n_items; ++it) {
        const int ob = item_ob(it);
        for (int t = 0; t < nb; ++t, ++g) {
          mbar_wait(bar(B_PD_READY), g & 1);
          float accn[8], accw[8];        // diagonals 8c + i, and 8c + i - 128 (wrapped chunks)
#pragma unroll
          for (int i = 0; i < 8; ++i) accn[i] = accw[i] = 0.0f;
#pragma unroll
          for (int blk = 0; blk < 2; ++blk) {
            const int ab = sub + 8 * blk;
            const int b0 = (ab + c) & 15, b1 = (ab + c + 1) & 15;
            const bool wrap0 = ab + c >= 16, wrap1 = ab + c + 1 >= 16;
            float lo[8], hi[8];          // sums from the first / second chunk
#pragma unroll
            for (int i = 0; i < 8; ++i) lo[i] = hi[i] = 0.0f;
            const uint8_t* rowp = ds + ab * 8 * 128;
#pragma unroll
            for (int rr = 0; rr < 8; ++rr) {
              const uint4 v0 = *reinterpret_cast<const uint4*>(rowp + (b0 >> 3) * TILE_BYTES + rr * 128 +
                                                               (((b0 & 7) ^ rr) << 4));
              const uint32_t w0[4] = {v0.x, v0.y, v0.z, v0.w};
#pragma unroll
              for (int e = rr; e < 8; ++e)
                lo[e - rr] += __uint_as_float((e & 1) ? (w0[e >> 1] & 0xffff0000u) : (w0[e >> 1] << 16));
              if (rr > 0) {
                const uint4 v1 = *reinterpret_cast<const uint4*>(rowp + (b1 >> 3) * TILE_BYTES + rr * 128 +
                                                                 (((b1 & 7) ^ rr) << 4));
                const uint32_t w1[4] = {v1.x, v1.y, v1.z, v1.w};
#pragma unroll
                for (int e = 0; e < rr; ++e)
                  hi[e + 8 - rr] += __uint_as_float((e & 1) ? (w1[e >> 1] & 0xffff0000u) : (w1[e >> 1] << 16));
              }
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              accn[i] += (wrap0 ? 0.0f : lo[i]) + (wrap1 ? 0.0f : hi[i]);
              accw[i] += (wrap0 ? lo[i] : 0.0f) + (wrap1 ? hi[i] : 0.0f);
            }
          }
          // the tile is consumed: let the compute warps overwrite it while the partial sums are combined
          __syncwarp();
          if (lane == 0) mbar_arrive(bar(B_PD_FREE));
          // reduce-scatter over the 8 threads of the class: thread `sub` ends up with diagonal i = sub
          float n4[4], w4[4];
          {
            const bool up = sub & 4;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float sn = up ? accn[j] : accn[4 + j], sw = up ? accw[j] : accw[4 + j];
              n4[j] = (up ? accn[4 + j] : accn[j]) + __shfl_xor_sync(0xffffffffu, sn, 4);
              w4[j] = (up ? accw[4 + j] : accw[j]) + __shfl_xor_sync(0xffffffffu, sw, 4);
            }
          }
          float n2[2], w2[2];
          {
            const bool up = sub & 2;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const float sn = up ? n4[j] : n4[2 + j], sw = up ? w4[j] : w4[2 + j];
              n2[j] = (up ? n4[2 + j] : n4[j]) + __shfl_xor_sync(0xffffffffu, sn, 2);
              w2[j] = (up ? w4[2 + j] : w4[j]) + __shfl_xor_sync(0xffffffffu, sw, 2);
            }
          }
          const bool up = sub & 1;
          const float vn = (up ? n2[1] : n2[0]) + __shfl_xor_sync(0xffffffffu, up ? n2[0] : n2[1], 1);
          const float vw = (up ? w2[1] : w2[0]) + __shfl_xor_sync(0xffffffffu, up ? w2[0] : w2[1], 1);
          const int idx = (ob - t) * BLK + (MAX_S - 1) + 8 * c + sub;   // (key block - query block) * 128 + 511 + d
          if (idx >= 0 && idx < BIAS_W) sdiag[idx] += vn;
          if (idx - BLK >= 0 && idx - BLK < BIAS_W) sdiag[idx - BLK] += vw;
        }
      }
    }
  }

  // ---- teardown
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc<1>(tmem, TM_COLS);
  if (MODE == 0) {
    for (int i = threadIdx.x; i < BIAS_W; i += blockDim.x) {
      const float v = sdiag[i];
      if (v != 0.0f) atomicAdd(a.dbias_diag + (size_t)head * BIAS_W + i, v);
    }

  }
}

// D[row, h] = sum_d dO[row, h*64 + d] * O[row, h*64 + d]: eight lanes per (row, head)
__global__ void attn_dvec_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ d_o,
                                 int ld, int64_t rows, int H, float* __restrict__ dvec) {
  const int64_t gi = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t grp = gi >> 3;
  const int l8 = (int)(gi & 7);
  float acc = 0.0f;
  if (grp < rows * H) {
    const int64_t row = grp / H;
    const int h = (int)(grp % H);
    const size_t off = (size_t)row * ld + h * DK + l8 * 8;
    const uint4 x = *reinterpret_cast<const uint4*>(o + off);
    const uint4 y = *reinterpret_cast<const uint4*>(d_o + off);
    const uint32_t xs[4] = {x.x, x.y, x.z, x.w}, ys[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      acc = fmaf(__uint_as_float(xs[e] << 16), __uint_as_float(ys[e] << 16), acc);
      acc = fmaf(__uint_as_float(xs[e] & 0xffff0000u), __uint_as_float(ys[e] & 0xffff0000u), acc);
    }
  }
  acc += __shfl_xor_sync(0xffffffffu, acc, 1);
  acc += __shfl_xor_sync(0xffffffffu, acc, 2);
  acc += __shfl_xor_sync(0xffffffffu, acc, 4);
  if (l8 == 0 && grp < rows * H) dvec[grp] = acc;
}

// grad of relative_attention_bias.weight [buckets, H] += per-diagonal sums folded by bucket(j - i)
__global__ void rel_bias_grad_kernel(const float* __restrict__ diag, const int* __restrict__ bucket_of, int H,
                                     int buckets, float* __restrict__ grad) {
  __shared__ float sb[64];
  const int h = blockIdx.x;
  for (int i = threadIdx.x; i < 64; i += blockDim.x) sb[i] = 0.0f;
  __syncthreads();
  for (int i = threadIdx.x; i < BIAS_W; i += blockDim.x) {
    const float v = diag[(size_t)h * BIAS_W + i];
    if (v != 0.0f) atomicAdd(&sb[bucket_of[i]], v);
  }
  __syncthreads();
  for (int b = threadIdx.x; b < buckets; b += blockDim.x) grad[(size_t)b * H + h] += sb[b];
}

}  // namespace

int pn_t5_self_attn_bwd(const __nv_bfloat16* qkv, int ld, int q_col, int k_col, int v_col, int docs, int S,
                        int n_heads, const float* bias, const __nv_bfloat16* ctx, const __nv_bfloat16* dctx,
                        int ld_ctx, const float* lse2, float* dvec, const float* row_scale,
                        __nv_bfloat16* dqkv, float* dbias_diag, cudaStream_t st, PnDrop drop, float* dot_out) {
  if (S < 1 || S > MAX_S) return pn_fail(PN_E_UNSUPPORTED, "self-attention backward: S=%d not in [1,512]", S);
  if (docs <= 0) return PN_OK;
  const int64_t rows = (int64_t)docs * S;
  CUtensorMap tmQKV, tmDO;
  int rc = pn_make_tmap_bf16_2d(&tmQKV, qkv, (uint64_t)rows, (uint64_t)ld, ld, BLK, DK);
  if (rc) return rc;
  rc = pn_make_tmap_bf16_2d(&tmDO, dctx, (uint64_t)rows, (uint64_t)ld_ctx, ld_ctx, BLK, DK);
  if (rc) return rc;
  static bool attr_done = false;
  if (!attr_done) {
    PN_CUDA(cudaFuncSetAttribute(t5_attn_bwd_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM));
    PN_CUDA(cudaFuncSetAttribute(t5_attn_bwd_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, BWD_SMEM));
    attr_done = true;
  }
  {
    const int64_t threads = rows * n_heads * 8;
    attn_dvec_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(ctx, dctx, ld_ctx, rows, n_heads, dvec);
    PN_LAUNCHED();
  }
  const int nb = (S + BLK - 1) / BLK;
  BwdArgs a;
  a.S = S; a.n_heads = n_heads; a.docs = docs;
  a.cph = PN_SM_COUNT / n_heads;
  if (a.cph < 1) a.cph = 1;
  if (a.cph > docs * nb) a.cph = docs * nb;
  a.q_col = q_col; a.k_col = k_col; a.v_col = v_col;
  a.bias = bias; a.lse2 = lse2; a.dvec = dvec; a.row_scale = row_scale;
  a.dqkv = dqkv; a.ld = ld; a.dbias_diag = dbias_diag;
  a.drop = drop;
  a.qkv = qkv; a.dot_out = dot_out; a.dot_stride = rows;
  // five [128 x 128 x 64] products per (query block, key block) pair, S and dP formed twice
  PnProfScope prof(PN_PROF_ATTN_BWD, 14.0 * docs * n_heads * (double)S * S * DK, st);
  auto launched = [&](const void* fn, const char* which) -> int {
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) return PN_OK;
    cudaFuncAttributes fa = {};
    cudaFuncGetAttributes(&fa, fn);
    return pn_fail(PN_E_LAUNCH, "pn_t5_self_attn_bwd: %s launch failed: %s (regs %d, max threads %d, static smem %zu, "
                   "max dynamic smem %d, requested %d)", which, cudaGetErrorString(e), fa.numRegs,
                   fa.maxThreadsPerBlock, fa.sharedSizeBytes, fa.maxDynamicSharedSizeBytes, BWD_SMEM);
  };
  static int only = -2;   // PN_ATTN_BWD_ONLY=0|1 runs a single pass (timing tools only)
  if (only == -2) {
    const char* e = getenv("PN_ATTN_BWD_ONLY");
    only = e ? atoi(e) : -1;
  }
  if (only != 1) {
    t5_attn_bwd_kernel<0><<<n_heads * a.cph, kThreadsKV, BWD_SMEM, st>>>(tmQKV, tmDO, a);
    pn_count_launch();
    if (int rc2 = launched((const void*)t5_attn_bwd_kernel<0>, "dK/dV")) return rc2;
  }
  if (only != 0) {
    t5_attn_bwd_kernel<1><<<n_heads * a.cph, kThreadsQ, BWD_SMEM, st>>>(tmQKV, tmDO, a);
    pn_count_launch();
    if (int rc2 = launched((const void*)t5_attn_bwd_kernel<1>, "dQ")) return rc2;
  }
  return PN_OK;
}

int pn_t5_rel_bias_grad(const float* dbias_diag, const int* bucket_of, int n_heads, int buckets, float* grad,
                        cudaStream_t st) {
  if (buckets > 64) return pn_fail(PN_E_UNSUPPORTED, "relative bias gradient: %d buckets > 64", buckets);
  rel_bias_grad_kernel<<<n_heads, 256, 0, st>>>(dbias_diag, bucket_of, n_heads, buckets, grad);
  PN_LAUNCHED();
  return PN_OK;
}

/* test entry: see include/pn_b200.h */
extern "C" int pn_t5_self_attention_bwd(const void* qkv, int32_t docs, int32_t S, int32_t n_heads,
                                        const float* bias, const void* ctx, const void* dctx,
                                        const float* lse2, const float* row_scale, void* dqkv,
                                        float* dbias_diag, float* dvec_ws, void* stream) {
  PN_CHECK_ARG(qkv && bias && ctx && dctx && lse2 && dqkv && dbias_diag && dvec_ws && docs >= 0 && n_heads > 0);
  const int di = n_heads * DK;
  return pn_t5_self_attn_bwd((const __nv_bfloat16*)qkv, 3 * di, 0, di, 2 * di, docs, S, n_heads, bias,
                             (const __nv_bfloat16*)ctx, (const __nv_bfloat16*)dctx, di, lse2, dvec_ws,
                             row_scale, (__nv_bfloat16*)dqkv, dbias_diag, (cudaStream_t)stream);
}

extern "C" int pn_t5_self_attention_bwd_dot(const void* qkv, int32_t docs, int32_t S, int32_t n_heads,
                                            const float* bias, const void* ctx, const void* dctx,
                                            const float* lse2, const float* row_scale, void* dqkv,
                                            float* dbias_diag, float* dvec_ws, float* dot_out, void* stream) {
  PN_CHECK_ARG(qkv && bias && ctx && dctx && lse2 && dqkv && dbias_diag && dvec_ws && dot_out && docs >= 0 &&
               n_heads > 0);
  const int di = n_heads * DK;
  return pn_t5_self_attn_bwd((const __nv_bfloat16*)qkv, 3 * di, 0, di, 2 * di, docs, S, n_heads, bias,
                             (const __nv_bfloat16*)ctx, (const __nv_bfloat16*)dctx, di, lse2, dvec_ws,
                             row_scale, (__nv_bfloat16*)dqkv, dbias_diag, (cudaStream_t)stream, PnDrop(), dot_out);
}

// ---------------------------------------------------------------------------------------------------------
// Allocation-free variant of pn_t5_self_attention_bwd (no counterpart in the reference; added
// for callers that cannot allocate: CUDA graph capture, tight serving loops, harnesses that
// forbid cudaMalloc).  The D-vector workspace [docs*S, H] the two backward passes read lives in a file-scope device
// array instead of a caller buffer.  Every GPU that loads the library holds the array until exit.
constexpr long long PN_ATTN_BWD_STATIC_MAX = 1ll << 20;   // docs * S * n_heads <= 1 Mi floats (4 MiB)
static __device__ __align__(256) float g_attn_bwd_dvec_scratch[PN_ATTN_BWD_STATIC_MAX];

// In host code the array's name is a host-side stand-in, not a device address: the device address is looked up once
// per device, with that device current (cudaGetSymbolAddress allocates nothing), and cached.
static int attn_bwd_static_scratch(float** out) {
  static float* cached[64] = {nullptr};
  static std::atomic<int> ready[64];
  int dev = 0;
  PN_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return pn_fail(PN_E_UNSUPPORTED, "pn_t5_self_attention_bwd_static: device index %d >= 64", dev);
  if (!ready[dev].load(std::memory_order_acquire)) {
    float* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_attn_bwd_dvec_scratch);
    if (e != cudaSuccess || p == nullptr)
      return pn_fail(PN_E_LAUNCH, "pn_t5_self_attention_bwd_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                     cudaGetErrorString(e));
    cached[dev] = p;                       // two racing threads look up the same address
    ready[dev].store(1, std::memory_order_release);
  }
  *out = cached[dev];
  return PN_OK;
}

extern "C" int pn_t5_self_attention_bwd_static(const void* qkv, int32_t docs, int32_t S, int32_t n_heads,
                                               const float* bias, const void* ctx, const void* dctx,
                                               const float* lse2, const float* row_scale, void* dqkv,
                                               float* dbias_diag, void* stream) {
  PN_CHECK_ARG(qkv && bias && ctx && dctx && lse2 && dqkv && dbias_diag && docs >= 0 && n_heads > 0 && S >= 1);
  const long long need = (long long)docs * S * n_heads;
  if (need > PN_ATTN_BWD_STATIC_MAX)
    return pn_fail(PN_E_UNSUPPORTED,
                   "pn_t5_self_attention_bwd_static: docs*S*n_heads = %lld exceeds the built-in scratch of %lld floats; "
                   "use pn_t5_self_attention_bwd with a caller workspace",
                   need, PN_ATTN_BWD_STATIC_MAX);
  float* dvec = nullptr;
  if (int rc = attn_bwd_static_scratch(&dvec)) return rc;
  // nothing in the scratch has to start at zero: attn_dvec_kernel overwrites every entry the passes read
  const int di = n_heads * DK;
  return pn_t5_self_attn_bwd((const __nv_bfloat16*)qkv, 3 * di, 0, di, 2 * di, docs, S, n_heads, bias,
                             (const __nv_bfloat16*)ctx, (const __nv_bfloat16*)dctx, di, lse2, dvec, row_scale,
                             (__nv_bfloat16*)dqkv, dbias_diag, (cudaStream_t)stream);
}
