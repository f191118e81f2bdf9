// T5 encoder self-attention, second generation (HF modeling_t5.py:308-336): softmax(Q K^T + bias[h, j-i]) V,
// no 1/sqrt(d) scaling, additive learned relative-position bias, fp32 softmax.
//
// Persistent, warp-specialised flash attention on tcgen05 with two query blocks in flight per CTA:
//   * a CTA serves ONE head for its whole life (bias table loaded once, four 16-byte-aligned shifted
//     copies so a thread reads its Toeplitz row with LDS.128) and walks documents doc0, doc0+cph, ...
//   * K, V (S x 64 bf16 each) and all four 128-row Q blocks of a document live in shared memory; the
//     TMA producer refills each 16 KB slot for the NEXT document as soon as the MMA issuer has
//     committed its last use, so loads are hidden behind the current document's softmax
//   * two softmax warpgroups (A: query blocks 0,2  B: blocks 1,3) ping-pong on the tensor pipe:
//     S = Q K^T (UMMA 128x128x64, SS) lands in TMEM, thread r owns row r, adds bias, keeps a lazily
//     updated row max (O is rescaled in TMEM only when the max grows by > 2^8), writes P as packed bf16
//     back over S in TMEM and O += P V runs as a TS-mode UMMA (A from TMEM, V MN-major from smem)
//   * the [N,h,S,S] score tensor never exists; P never touches shared memory.
// Warp roles: 0..3 = softmax A, 4..7 = softmax B, 8 = TMA producer, 9 / 11 = MMA issuers of stream A / B,
// 10 = TMEM allocator (the utility warps carry the highest warp ids: the scheduler favours them whenever
// they wake up, so an MMA issue is never starved by the softmax warps sharing its scheduler).
#include "pn_common.cuh"
#include "pn_tc.cuh"
#include "pn_t5_internal.cuh"
#include <stdlib.h>

using namespace pntc;

namespace {

constexpr int kThreads = 384;
constexpr int DK = 64;
constexpr int QB = 128;                    // query rows per block
constexpr int KC = 128;                    // keys per chunk
constexpr int TILE_BYTES = 128 * DK * 2;   // 16 KB
constexpr int MAX_S = 512;
constexpr int BIAS_W = 2 * MAX_S - 1;      // 1023 relative positions
constexpr int BIAS_STRIDE = 1032;          // floats per shifted copy (stride/4 = 2 mod 8: LDS.128 conflict-free)
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kRescaleThreshold = 8.0f;  // log2 units: lazy O rescale
constexpr float kRedoThreshold = 60.0f;    // log2 units: stale max too far off, redo the chunk exactly

constexpr int SM_K = 0;
constexpr int SM_V = 4 * TILE_BYTES;
constexpr int SM_Q = 8 * TILE_BYTES;
constexpr int SM_BIAS = 12 * TILE_BYTES;
constexpr int SM_BAR = SM_BIAS + 4 * BIAS_STRIDE * 4;
// barriers (8 B each): kfull[4] vfull[4] kempty[4] vempty[4] qfull[4] qempty[4] sfull[2] pready[2] ofull[2]
// oempty[2] sfree[2] pfree[2]
constexpr int BAR_KFULL = 0, BAR_VFULL = 4, BAR_KEMPTY = 8, BAR_VEMPTY = 12, BAR_QFULL = 16,
              BAR_QEMPTY = 20, BAR_SFULL = 24, BAR_PREADY = 26, BAR_OFULL = 28, BAR_OEMPTY = 30,
              BAR_SFREE = 32, BAR_PFREE = 34, NUM_BARS = 36;
constexpr int SM_STAGE = SM_BAR + NUM_BARS * 8 + 16;      // 8 softmax warps x 2 KB output staging
constexpr int ATTN2_SMEM = SM_STAGE + 8 * 2048 + 1024;

// TMEM columns: S_A 0..127, S_B 128..255, O_A 256..319, O_B 320..383, P_A 384..447, P_B 448..511 (packed bf16)
constexpr uint32_t TM_S = 0, TM_O = 256, TM_P = 384, TM_COLS = 512;

struct Attn2Args {
  int S, n_heads, docs, cph;      // cph = CTAs per head
  int q_col, k_col, v_col;
  const float* bias;              // [n_heads, 1023] additive bias by (j - i + 511)
  __nv_bfloat16* out;
  int ld_out;
  float* lse;                     // optional [docs*S, n_heads]: log2-domain logsumexp of each row (training)
  const int32_t* klen;            // optional [docs]: keys >= klen[doc] are padding (HF attention_mask, right-padded)
  long long* dbg;                 // optional [2][64][8] clock64 stamps of CTA 0 (PN_ATTN_DBG)
  PnDrop drop;                    // training: dropout of the probabilities (HF:331); the row sum stays un-dropped
};

#define DBG_STAMP(who, step, ev)                                                        \
  do {                                                                                  \
    if (a.dbg && blockIdx.x == 0 && (step) < 64) a.dbg[((who)*64 + (step)) * 8 + (ev)] = clock64(); \
  } while (0)

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
// packed fp32 pairs (sm_100 FFMA2 / FADD2): one issue slot for two elements of the softmax row
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
// 2^x for a pair on the FMA pipe instead of MUFU.EX2 (the exponential unit is the busiest pipe of this kernel:
// 16 lanes / clk / SM against 128 for FFMA): Cody-Waite split x = n + f with the round-to-nearest magic constant
// (n in the low mantissa bits of t = x + 1.5 * 2^23), a cubic minimax polynomial for 2^f on [-0.5, 0.5] (relative
// error 7.5e-5, far below the bf16 rounding of P) and n added into the exponent field.  x is clamped at -125 (2^-125
// is 0 for every purpose here and keeps the exponent arithmetic in range); x <= ~+60 by the redo threshold.
#ifndef PN_ATTN_DEFER_EPI
// 1: a query block's epilogue runs one step late, after the first step of the stream's next block (see the step loop).
// Measured and left off (profiles/r5e_attn_defer_epilogue_ab.txt, same box): 508-510 TFLOP/s with the immediate epilogue,
// 469 with the deferred one (-8 %; re-rank 9 058 -> 8 915 pairs/s): the next block's first P V then waits for the O read
// (BAR_OEMPTY) and lands ~2 200 cycles after its P was ready, and the epilogue itself is no shorter.
#define PN_ATTN_DEFER_EPI 0
#endif
#ifndef PN_ATTN_POLY
// Measured (same box, profiles/r4h_attn_poly_ab.txt; 250 docs x 12 heads x 512): 0 -> 496.7 TFLOP/s, 1 -> 501.5 (+1 %),
// 2 -> 434.7 (-12 %); re-rank step 8 581 / 8 575 / 8 352 pairs/s.  The kernel is not MUFU-bound (two softmax warps per
// scheduler: issue + dependency latency), so the default stays 0.
#define PN_ATTN_POLY 0     // 0: every exponential on MUFU; 1: one pair in four on the FMA pipe; 2: one in two
#endif
__device__ __forceinline__ void ex2_poly2(float x0, float x1, float& p0, float& p1) {
  const uint64_t magic = f2pack(12582912.0f, 12582912.0f), nmagic = f2pack(-12582912.0f, -12582912.0f);
  const uint64_t x = f2pack(fmaxf(x0, -125.0f), fmaxf(x1, -125.0f));
  const uint64_t t = add2(x, magic);
  const uint64_t n = add2(t, nmagic);
  const uint64_t f = fma2(n, f2pack(-1.0f, -1.0f), x);
  uint64_t p = fma2(f, f2pack(0.055171459913253784f, 0.055171459913253784f), f2pack(0.2426108568906784f, 0.2426108568906784f));
  p = fma2(p, f, f2pack(0.6932609677314758f, 0.6932609677314758f));
  p = fma2(p, f, f2pack(0.9999281167984009f, 0.9999281167984009f));
  float t0, t1, q0, q1;
  f2unpack(t, t0, t1);
  f2unpack(p, q0, q1);
  p0 = __int_as_float(__float_as_int(q0) + (__float_as_int(t0) << 23));
  p1 = __int_as_float(__float_as_int(q1) + (__float_as_int(t1) << 23));
}

// kDrop: training forward with dropout of the probabilities; the eval instantiation carries none of that code (its
// registers are all spoken for: 168 per thread at 384 threads)
template <bool kDrop>
__global__ void __launch_bounds__(kThreads, 1)
t5_self_attn2_kernel(const __grid_constant__ CUtensorMap tmQKV, const Attn2Args a) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* gbase = smem_raw + (base - smem_u32(smem_raw));
  float* sbias = reinterpret_cast<float*>(gbase + SM_BIAS);
  const uint32_t bar0 = base + SM_BAR;
  auto bar = [&](int i) { return bar0 + 8u * i; };
  const uint32_t tmem_slot = bar0 + 8u * NUM_BARS;
  uint32_t* tmem_slot_gen = reinterpret_cast<uint32_t*>(gbase + SM_BAR + 8 * NUM_BARS);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int head = blockIdx.x / a.cph;
  const int doc0 = blockIdx.x % a.cph;
  const int S = a.S;
  const int n_qb = (S + QB - 1) / QB, n_kc = (S + KC - 1) / KC;
  const int rounds = (n_qb + 1) / 2;            // B idles in the last round when n_qb is odd
  const int n_items = doc0 < a.docs ? (a.docs - doc0 + a.cph - 1) / a.cph : 0;
  const int steps_per_item = rounds * n_kc;
  const int total_steps = n_items * steps_per_item;
  const bool b_used = n_qb > 1;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmQKV);
    for (int i = 0; i < 4; ++i) {
      mbar_init(bar(BAR_KFULL + i), 1);
      mbar_init(bar(BAR_VFULL + i), 1);
      mbar_init(bar(BAR_KEMPTY + i), b_used ? 2 : 1);   // one commit per softmax stream's last use
      mbar_init(bar(BAR_VEMPTY + i), b_used ? 2 : 1);
      mbar_init(bar(BAR_QFULL + i), 1);
      mbar_init(bar(BAR_QEMPTY + i), 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar(BAR_SFULL + i), 1);
      mbar_init(bar(BAR_PREADY + i), 4);   // one arrive per softmax warp
      mbar_init(bar(BAR_OFULL + i), 1);
      mbar_init(bar(BAR_OEMPTY + i), 4);
      mbar_init(bar(BAR_SFREE + i), 4);
      mbar_init(bar(BAR_PFREE + i), 1);
    }
    fence_mbar_init();
  }
  if (warp == 10) {
    tmem_alloc<1>(tmem_slot, TM_COLS);
    tmem_relinquish<1>();
  }
  // shifted copies of this head's bias row, pre-multiplied by log2(e): copy[sh][k] = t[k + sh]
  for (int i = threadIdx.x; i < 4 * BIAS_STRIDE; i += kThreads) {
    const int sh = i / BIAS_STRIDE, k = i % BIAS_STRIDE;
    const int src = k + sh;
    sbias[i] = src < BIAS_W ? a.bias[(size_t)head * BIAS_W + src] * kLog2e : 0.0f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot_gen;

  if (warp >= 8) {
    reg_dealloc<56>();
    if (warp == 8 && lane == 0) {
      // ===================== TMA producer: runs ahead, bounded by the slot-empty barriers =====================
      for (int it = 0; it < n_items; ++it) {
        const int row0 = (doc0 + it * a.cph) * S;
        const uint32_t ph = it & 1;
        auto load_q = [&](int qb) {
          mbar_wait(bar(BAR_QEMPTY + qb), ph ^ 1);
          mbar_arrive_expect_tx(bar(BAR_QFULL + qb), TILE_BYTES);
          tma_load_2d(base + SM_Q + qb * TILE_BYTES, &tmQKV, bar(BAR_QFULL + qb), a.q_col + head * DK,
                      row0 + qb * QB);
        };
        load_q(0);
        if (n_qb > 1) load_q(1);
        for (int c = 0; c < n_kc; ++c) {
          mbar_wait(bar(BAR_KEMPTY + c), ph ^ 1);
          mbar_arrive_expect_tx(bar(BAR_KFULL + c), TILE_BYTES);
          tma_load_2d(base + SM_K + c * TILE_BYTES, &tmQKV, bar(BAR_KFULL + c), a.k_col + head * DK,
                      row0 + c * KC);
          mbar_wait(bar(BAR_VEMPTY + c), ph ^ 1);
          mbar_arrive_expect_tx(bar(BAR_VFULL + c), TILE_BYTES);
          tma_load_2d(base + SM_V + c * TILE_BYTES, &tmQKV, bar(BAR_VFULL + c), a.v_col + head * DK,
                      row0 + c * KC);
        }
        for (int qb = 2; qb < n_qb; ++qb) load_q(qb);
      }
    } else if (warp == 9 || warp == 11) {
      // ===================== MMA issuers: warp 9 drives stream A, warp 11 stream B =====================
      // Each stream alternates QK(n+1) / PV(n); a dedicated issuing warp per stream means a slow softmax
      // warpgroup never delays the other one's MMAs (the two streams touch disjoint TMEM columns and only
      // read the shared K/V tiles; each commits its own last use of a tile to the tile's empty barrier).
      // The WHOLE warp runs the loop and one lane issues: every descriptor is then warp-uniform and sits in a uniform
      // register.  With a single divergent thread each tcgen05.mma was wrapped in its own vector->uniform waterfall
      // (4 dependent R2UR + ELECT + branch, ~80 cycles per instruction: profiles/r5b_attn_fwd_source_summary.md) and the
      // eight [128 x 64 x 16] products of P V — 32 tensor-core cycles each — took ~700 cycles to ISSUE, all of it on
      // the softmax warps' critical path (P buffer hand-back every step, O hand-over at the end of a query block).
      const bool leader = lane == 0;
      constexpr uint32_t IDESC_QK = make_idesc_bf16(128, KC, 0, 0);  // A = Q K-major, B = K K-major
      constexpr uint32_t IDESC_PV = make_idesc_bf16(128, DK, 0, 1);  // A = P (TMEM),  B = V MN-major
      // warp index and TMEM base through a lane-0 shuffle: values the compiler can prove warp-uniform
      const int wg = __shfl_sync(0xffffffffu, warp, 0) == 9 ? 0 : 1;
      const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
      const int my_rounds = wg == 0 ? rounds : n_qb / 2;             // rounds in which this stream has a block
      const int steps = n_items * my_rounds * n_kc;
      const uint32_t s_tmem = tmem_u + TM_S + wg * 128, p_tmem = tmem_u + TM_P + wg * 64,
                     o_tmem = tmem_u + TM_O + wg * DK;
      struct Cur { int it, rd, c; };
      auto advance = [&](Cur& u) {
        if (++u.c == n_kc) { u.c = 0; if (++u.rd == my_rounds) { u.rd = 0; ++u.it; } }
      };
      auto uni = [](const Cur& v) {                    // the walk's state, provably warp-uniform for the compiler
        return Cur{__shfl_sync(0xffffffffu, v.it, 0), __shfl_sync(0xffffffffu, v.rd, 0),
                   __shfl_sync(0xffffffffu, v.c, 0)};
      };
      auto issue_qk = [&](const Cur& uu, int n) {      // n = index of this step in the stream
        const Cur u = uni(uu);
        const int qb = 2 * u.rd + wg;
        const uint32_t ph = u.it & 1;
        if (n > 0) mbar_wait(bar(BAR_SFREE + wg), (n - 1) & 1);     // softmax holds S(n-1) in registers
        if (wg == 0 && leader) DBG_STAMP(1, n, 0);
        if (u.c == 0) mbar_wait(bar(BAR_QFULL + qb), ph);
        mbar_wait(bar(BAR_KFULL + u.c), ph);
        tc_fence_after();
        const uint64_t adesc = make_desc_kmajor_sw128(base + SM_Q + qb * TILE_BYTES);
        const uint64_t bdesc = make_desc_kmajor_sw128(base + SM_K + u.c * TILE_BYTES);
        if (leader) {
#pragma unroll
          for (int k = 0; k < DK / 16; ++k)
            umma_bf16<1>(s_tmem, adesc + 2u * k, bdesc + 2u * k, IDESC_QK, k ? 1u : 0u);
          umma_commit<1>(bar(BAR_SFULL + wg));
          if (wg == 0) DBG_STAMP(1, n, 1);
          if (u.c == n_kc - 1) umma_commit<1>(bar(BAR_QEMPTY + qb));
          if (u.rd == my_rounds - 1) umma_commit<1>(bar(BAR_KEMPTY + u.c));   // stream's last use of K_c
        }
        __syncwarp();
      };
      auto issue_pv = [&](const Cur& uu, int n) {
        const Cur u = uni(uu);
        mbar_wait(bar(BAR_PREADY + wg), n & 1);
        if (wg == 0 && leader) DBG_STAMP(1, n, 2);
        if (u.c == 0) mbar_wait(bar(BAR_OEMPTY + wg), ((n / n_kc) & 1) ^ 1);
        mbar_wait(bar(BAR_VFULL + u.c), u.it & 1);
        tc_fence_after();
        // a 16-key step advances the MN-major descriptor of V by 16 rows x 128 B = 128 address units of 16 B
        const uint64_t vdesc = make_desc_mnmajor_sw128(base + SM_V + u.c * TILE_BYTES, 1024);
        if (leader) {
#pragma unroll
          for (int k = 0; k < KC / 16; ++k)
            umma_bf16_ts(o_tmem, p_tmem + k * 8, vdesc + 128u * k, IDESC_PV, (u.c | k) ? 1u : 0u);
          umma_commit<1>(bar(BAR_PFREE + wg));                             // P buffer / O readable again
          if (wg == 0) DBG_STAMP(1, n, 3);
          if (u.c == n_kc - 1) umma_commit<1>(bar(BAR_OFULL + wg));
          if (u.rd == my_rounds - 1) umma_commit<1>(bar(BAR_VEMPTY + u.c));
        }
        __syncwarp();
      };
      if (steps > 0) {
        Cur pv{0, 0, 0}, qk{0, 0, 0};
        issue_qk(qk, 0);
        advance(qk);
        for (int n = 0; n < steps; ++n) {
          if (n + 1 < steps) { issue_qk(qk, n + 1); advance(qk); }   // S(n+1) is produced while softmax(n) runs
          issue_pv(pv, n);
          advance(pv);
        }
      }
    }
  } else {
    // ===================== softmax warpgroups: thread r owns query row r of its block =====================
    reg_alloc<224>();
    const int wg = warp >> 2;                       // 0 = A, 1 = B
    const int w4 = warp & 3;                        // TMEM lane quarter this warp may touch
    const int r = w4 * 32 + lane;                   // row inside the query block = TMEM lane
    const uint32_t lane_addr = (uint32_t)(w4 * 32) << 16;
    const uint32_t s_addr = tmem + lane_addr + TM_S + wg * 128;
    const uint32_t o_addr = tmem + lane_addr + TM_O + wg * DK;
    const uint32_t p_addr = tmem + lane_addr + TM_P + wg * 64;
    const int sh = 3 - (r & 3);                     // which shifted bias copy keeps this row 16-byte aligned
    const float* bcopy = sbias + sh * BIAS_STRIDE;
    float m_used = 0.0f, l = 0.0f, mx_prev = 0.0f;
    int n = 0;                                      // valid steps done by this stream
    // epilogue of query block qb_ of document doc_ (the blk_-th block this stream finishes): ctx[row, head*64..] = O / l
    auto epilogue = [&](float l_, float m_, int qb_, int doc_, int blk_) {
      mbar_wait(bar(BAR_OFULL + wg), blk_ & 1);
      tc_fence_after();
      if (warp == 0 && lane == 0) DBG_STAMP(0, (blk_ + 1) * n_kc - 1, 7);
      uint32_t o0[32], o1[32];
      tmem_ld_32x32(o_addr, o0);
      tmem_ld_32x32(o_addr + 32, o1);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(BAR_OEMPTY + wg));
      // rows -> bf16, transposed through a 2 KB per-warp staging tile so that every store instruction writes
      // 8 rows x 64 contiguous bytes instead of 32 rows x 16 bytes
      const float inv = 1.0f / l_;
      const int qi_ = qb_ * QB + r;
      if (a.lse && qi_ < S) a.lse[(size_t)(doc_ * S + qi_) * a.n_heads + head] = m_ + log2f(l_);
      uint8_t* stage = gbase + SM_STAGE + warp * 2048;
      const int doc_row0 = doc_ * S + qb_ * QB + w4 * 32;
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {
        const uint32_t(&o)[32] = h2 ? o1 : o0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          uint4 u;
          u.x = pack_bf16x2(__uint_as_float(o[j * 8 + 0]) * inv, __uint_as_float(o[j * 8 + 1]) * inv);
          u.y = pack_bf16x2(__uint_as_float(o[j * 8 + 2]) * inv, __uint_as_float(o[j * 8 + 3]) * inv);
          u.z = pack_bf16x2(__uint_as_float(o[j * 8 + 4]) * inv, __uint_as_float(o[j * 8 + 5]) * inv);
          u.w = pack_bf16x2(__uint_as_float(o[j * 8 + 6]) * inv, __uint_as_float(o[j * 8 + 7]) * inv);
          *reinterpret_cast<uint4*>(stage + ((lane * 4 + (j ^ ((lane >> 1) & 3))) << 4)) = u;
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int row = 8 * i + (lane >> 2), j = lane & 3;
          const uint4 u = *reinterpret_cast<const uint4*>(stage + ((row * 4 + (j ^ ((row >> 1) & 3))) << 4));
          if (qb_ * QB + w4 * 32 + row < S)
            *reinterpret_cast<uint4*>(a.out + (size_t)(doc_row0 + row) * a.ld_out + head * DK + h2 * 32 + j * 8) = u;
        }
        __syncwarp();
      }
    };
#if PN_ATTN_DEFER_EPI
    bool pend = false;
    float pend_l = 1.0f, pend_m = 0.0f;
    int pend_qb = 0, pend_doc = 0, pend_blk = 0;
#endif
    int it = 0, rd = 0, c = 0;
    for (int s = 0; s < total_steps; ++s, (++c == n_kc ? (c = 0, (++rd == rounds ? (rd = 0, ++it) : 0)) : 0)) {
      const int qb = 2 * rd + wg;
      if (qb >= n_qb) continue;
      const int qi = qb * QB + r;                   // query position inside the document
      // keys beyond the document's length (attention_mask == 0, or beyond S in the ragged last chunk) get a score no
      // exponential survives: -inf without a mask (bit-identical to the unmasked kernel), a finite -1e30 with one so
      // that a chunk with NO valid key leaves the running max alone (chunk 0 always holds key 0: klen >= 1)
      const int klen = a.klen ? min(S, max(1, a.klen[doc0 + it * a.cph])) : S;
      const int kvalid = min(KC, klen - c * KC);
      const float kmasked = a.klen ? -1e30f : -INFINITY;
      if (warp == 0 && lane == 0) DBG_STAMP(0, n, 0);
      mbar_wait(bar(BAR_SFULL + wg), n & 1);
      tc_fence_after();
      if (warp == 0 && lane == 0) DBG_STAMP(0, n, 1);
      float x[KC];
      {
        uint32_t(&xu)[KC] = reinterpret_cast<uint32_t(&)[KC]>(x);
        tmem_ld_32x32(s_addr + 0, reinterpret_cast<uint32_t(&)[32]>(xu[0]));
        tmem_ld_32x32(s_addr + 32, reinterpret_cast<uint32_t(&)[32]>(xu[32]));
        tmem_ld_32x32(s_addr + 64, reinterpret_cast<uint32_t(&)[32]>(xu[64]));
        tmem_ld_32x32(s_addr + 96, reinterpret_cast<uint32_t(&)[32]>(xu[96]));
        tmem_ld_wait();
      }
      // S is in registers: the tensor core may overwrite the buffer with the next chunk's scores
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(BAR_SFREE + wg));
      if (warp == 0 && lane == 0) DBG_STAMP(0, n, 2);
      const float4* bp = reinterpret_cast<const float4*>(bcopy + (c * KC - qi + (MAX_S - 1) - sh));
      if (kvalid < KC) {   // ragged last chunk: keys beyond S do not exist (-inf survives the bias FFMA)
#pragma unroll
        for (int jj = 0; jj < KC; ++jj)
          if (jj >= kvalid) x[jj] = kmasked;
      }
      // columns 0..31: x = s*log2e + bias*log2e, and their max (seeds the row max of a fresh query block)
      float m0 = -INFINITY, m1 = -INFINITY;
      const uint64_t l2e2 = f2pack(kLog2e, kLog2e);
#pragma unroll
      for (int jj = 0; jj < 32; jj += 4) {
        const float4 b = bp[jj >> 2];
        f2unpack(fma2(f2pack(x[jj + 0], x[jj + 1]), l2e2, f2pack(b.x, b.y)), x[jj + 0], x[jj + 1]);
        f2unpack(fma2(f2pack(x[jj + 2], x[jj + 3]), l2e2, f2pack(b.z, b.w)), x[jj + 2], x[jj + 3]);
        m0 = fmax3(m0, x[jj + 0], x[jj + 1]);
        m1 = fmax3(m1, x[jj + 2], x[jj + 3]);
      }
      bool pfree_seen = n == 0;
      auto rescale_o = [&](float m_new) {   // O, l: reference max m_used -> m_new (PV of the previous step retired)
        if (!pfree_seen) {
          mbar_wait(bar(BAR_PFREE + wg), (n - 1) & 1);
          tc_fence_after();
          pfree_seen = true;
        }
        const float alpha = ex2_approx(m_used - m_new);
        l *= alpha;
#pragma unroll
        for (int h2 = 0; h2 < 2; ++h2) {
          uint32_t v[32];
          tmem_ld_32x32(o_addr + h2 * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int d = 0; d < 32; ++d) v[d] = __float_as_uint(__uint_as_float(v[d]) * alpha);
          tmem_st_32x32(o_addr + h2 * 32, v);
        }
      };
      if (c == 0) {
        m_used = fmaxf(m0, m1);
        l = 0.0f;
      } else if (__any_sync(0xffffffffu, mx_prev > m_used + kRescaleThreshold)) {
        // lazy rescale: the previous chunk raised the row max by more than 2^8
        const float m_new = fmaxf(m_used, mx_prev);
        rescale_o(m_new);
        m_used = m_new;
      }
      // single pass: bias (columns 32..127), running max, p = 2^(x - m_used) against the STALE max so that the
      // exponentials do not wait for the row max; P -> its own TMEM buffer as packed bf16 (64 columns)
      // dropout mask of this row's chunk: element [doc, head, qi, c*KC + jj] at an even row pitch
      constexpr bool dropping = kDrop;
      unsigned long long dbase = 0;
      if (kDrop)
        dbase = (unsigned long long)a.drop.idx_offset +
                ((unsigned long long)((doc0 + it * a.cph) * a.n_heads + head) * S + (qi < S ? qi : 0)) * (S + (S & 1)) +
                c * KC;
      // the chunk's KC / 2 index pairs share their inner hash (pn_drop_run); one row in 2^32 pairs does not
      PnDropRun drun = {};
      bool dfast = true;
      if (kDrop) {
        drun = pn_drop_run(a.drop, dbase, KC / 2);
        dfast = __all_sync(0xffffffffu, drun.fast);
      }
      float nm = -m_used;
      uint64_t nm2 = f2pack(nm, nm);
      uint64_t lsa = f2pack(0.0f, 0.0f), lsb = lsa;
      float ls0 = 0.0f, ls1 = 0.0f;
      float m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {
        uint32_t pk[32];
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          const int jj = h2 * 64 + 2 * e;
          uint64_t xa, xb;
          if (jj >= 32) {
            const float4 b = bp[jj >> 2];
            xa = fma2(f2pack(x[jj + 0], x[jj + 1]), l2e2, f2pack(b.x, b.y));
            xb = fma2(f2pack(x[jj + 2], x[jj + 3]), l2e2, f2pack(b.z, b.w));
            f2unpack(xa, x[jj + 0], x[jj + 1]);
            f2unpack(xb, x[jj + 2], x[jj + 3]);
            m2 = fmax3(m2, x[jj + 0], x[jj + 1]);
            m3 = fmax3(m3, x[jj + 2], x[jj + 3]);
          } else {
            xa = f2pack(x[jj + 0], x[jj + 1]);
            xb = f2pack(x[jj + 2], x[jj + 3]);
          }
          float e0, e1, e2, e3;
          f2unpack(add2(xa, nm2), e0, e1);
          f2unpack(add2(xb, nm2), e2, e3);
          const float p0 = ex2_approx(e0);
          const float p1 = ex2_approx(e1);
          float p2, p3;
          if (PN_ATTN_POLY == 2 || (PN_ATTN_POLY == 1 && (e & 2))) {
            ex2_poly2(e2, e3, p2, p3);
          } else {
            p2 = ex2_approx(e2);
            p3 = ex2_approx(e3);
          }
          lsa = add2(lsa, f2pack(p0, p1));
          lsb = add2(lsb, f2pack(p2, p3));
          if (dropping) {
            float m0, m1, m2, m3;
            if (dfast) {
              pn_drop_run_mul2(a.drop, drun, jj >> 1, m0, m1);
              pn_drop_run_mul2(a.drop, drun, (jj >> 1) + 1, m2, m3);
            } else {
              pn_drop_mul2(a.drop, dbase + jj, m0, m1);
              pn_drop_mul2(a.drop, dbase + jj + 2, m2, m3);
            }
            pk[e] = pack_bf16x2(p0 * m0, p1 * m1);
            pk[e + 1] = pack_bf16x2(p2 * m2, p3 * m3);
          } else {
            pk[e] = pack_bf16x2(p0, p1);
            pk[e + 1] = pack_bf16x2(p2, p3);
          }
        }
        if (h2 == 1) {
          // the whole row is known now; if the stale max was off by more than 2^60 redo the chunk exactly
          const float mx = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
          mx_prev = mx;
          if (__any_sync(0xffffffffu, mx > m_used + kRedoThreshold)) {
            const float m_new = fmaxf(m_used, mx);
            if (c > 0) rescale_o(m_new);
            m_used = m_new;
            nm = -m_used;
            nm2 = f2pack(nm, nm);
            lsa = lsb = f2pack(0.0f, 0.0f);
            if (!pfree_seen) {
              mbar_wait(bar(BAR_PFREE + wg), (n - 1) & 1);
              tc_fence_after();
              pfree_seen = true;
            }
#pragma unroll
            for (int g = 0; g < 2; ++g) {
              uint32_t pq[32];
#pragma unroll
              for (int e = 0; e < 32; ++e) {
                const float p0 = ex2_approx(x[g * 64 + 2 * e] + nm);
                const float p1 = ex2_approx(x[g * 64 + 2 * e + 1] + nm);
                ls0 += p0;
                ls1 += p1;
                float m0 = 1.0f, m1 = 1.0f;
                if (dropping) pn_drop_mul2(a.drop, dbase + g * 64 + 2 * e, m0, m1);
                pq[e] = pack_bf16x2(p0 * m0, p1 * m1);
              }
              tmem_st_32x32(p_addr + g * 32, pq);
            }
            break;   // P fully rewritten
          }
        }
        if (!pfree_seen) {   // the previous PV still owns the P buffer until its commit lands
          if (warp == 0 && lane == 0) DBG_STAMP(0, n, 3);
          mbar_wait(bar(BAR_PFREE + wg), (n - 1) & 1);
          tc_fence_after();
          pfree_seen = true;
          if (warp == 0 && lane == 0) DBG_STAMP(0, n, 4);
        }
        tmem_st_32x32(p_addr + h2 * 32, pk);
      }
      {
        float a0, a1, b0, b1;
        f2unpack(lsa, a0, a1);
        f2unpack(lsb, b0, b1);
        l += ((a0 + a1) + (b0 + b1)) + (ls0 + ls1);
      }
      if (warp == 0 && lane == 0) DBG_STAMP(0, n, 5);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(BAR_PREADY + wg));
      if (warp == 0 && lane == 0) DBG_STAMP(0, n, 6);
      ++n;
#if PN_ATTN_DEFER_EPI
      // the epilogue of a finished query block runs one step late: after the FIRST step of the stream's next block has
      // handed its P over.  By then the block's last P V has long retired (waiting for it right away cost ~1 050
      // cycles per block, profiles/r5b_attn_source_summary.md), and a step with c == 0 never touches O itself: the
      // issuer holds that step's P V (accumulate = 0, it overwrites O) until the epilogue has read O (BAR_OEMPTY)
      if (pend) {
        epilogue(pend_l, pend_m, pend_qb, pend_doc, pend_blk);
        pend = false;
      }
      if (c == n_kc - 1) {
        pend = true;
        pend_l = l; pend_m = m_used; pend_qb = qb; pend_doc = doc0 + it * a.cph; pend_blk = n / n_kc - 1;
      }
#else
      if (c == n_kc - 1) epilogue(l, m_used, qb, doc0 + it * a.cph, n / n_kc - 1);
#endif
    }
#if PN_ATTN_DEFER_EPI
    if (pend) epilogue(pend_l, pend_m, pend_qb, pend_doc, pend_blk);
#endif
  }

  // ---- teardown
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  if (warp == 10) tmem_dealloc<1>(tmem, TM_COLS);
}

}  // namespace

int pn_t5_self_attn_v1(const __nv_bfloat16* qkv, int ld_qkv, int q_col, int k_col, int v_col, int docs,
                       int S, int n_heads, const float* bias, __nv_bfloat16* out, int ld_out,
                       cudaStream_t st);

long long* g_pn_attn_dbg = nullptr;
extern "C" int pn_attn_debug_read(long long* host_out) {
  if (!g_pn_attn_dbg) return -1;
  cudaDeviceSynchronize();
  cudaMemcpy(host_out, g_pn_attn_dbg, 2 * 64 * 8 * 8, cudaMemcpyDeviceToHost);
  return 0;
}

static int attn_impl() {
  static int impl = -1;
  if (impl < 0) {
    const char* e = getenv("PN_ATTN_IMPL");
    impl = (e && e[0] == '1') ? 1 : 2;
  }
  return impl;
}

static int self_attn_impl(const __nv_bfloat16* qkv, int ld_qkv, int q_col, int k_col, int v_col, int docs, int S,
                          int n_heads, const float* bias, const int32_t* klen, __nv_bfloat16* out, int ld_out,
                          float* lse_out, PnDrop drop, cudaStream_t st);

int pn_t5_self_attn(const __nv_bfloat16* qkv, int ld_qkv, int q_col, int k_col, int v_col, int docs,
                    int S, int n_heads, const float* bias, __nv_bfloat16* out, int ld_out, float* lse_out,
                    cudaStream_t st) {
  return self_attn_impl(qkv, ld_qkv, q_col, k_col, v_col, docs, S, n_heads, bias, nullptr, out, ld_out, lse_out,
                        PnDrop(), st);
}

int pn_t5_self_attn_drop(const __nv_bfloat16* qkv, int ld_qkv, int q_col, int k_col, int v_col, int docs,
                         int S, int n_heads, const float* bias, __nv_bfloat16* out, int ld_out, float* lse_out,
                         PnDrop drop, cudaStream_t st) {
  return self_attn_impl(qkv, ld_qkv, q_col, k_col, v_col, docs, S, n_heads, bias, nullptr, out, ld_out, lse_out, drop,
                        st);
}

int pn_t5_self_attn_masked(const __nv_bfloat16* qkv, int ld_qkv, int q_col, int k_col, int v_col, int docs,
                           int S, int n_heads, const float* bias, const int32_t* klen, __nv_bfloat16* out,
                           int ld_out, float* lse_out, cudaStream_t st) {
  return self_attn_impl(qkv, ld_qkv, q_col, k_col, v_col, docs, S, n_heads, bias, klen, out, ld_out, lse_out, PnDrop(),
                        st);
}

static int self_attn_impl(const __nv_bfloat16* qkv, int ld_qkv, int q_col, int k_col, int v_col, int docs, int S,
                          int n_heads, const float* bias, const int32_t* klen, __nv_bfloat16* out, int ld_out,
                          float* lse_out, PnDrop drop, cudaStream_t st) {
  if (attn_impl() == 1 && !lse_out && !klen && !drop.thresh)
    return pn_t5_self_attn_v1(qkv, ld_qkv, q_col, k_col, v_col, docs, S, n_heads, bias, out, ld_out, st);
  if (S < 1 || S > MAX_S) return pn_fail(PN_E_UNSUPPORTED, "self-attention: S=%d not in [1,512]", S);
  if (docs <= 0) return PN_OK;
  CUtensorMap tm;
  int rc = pn_make_tmap_bf16_2d(&tm, qkv, (uint64_t)docs * S, (uint64_t)ld_qkv, ld_qkv, 128, DK);
  if (rc) return rc;
  Attn2Args a;
  a.S = S; a.n_heads = n_heads; a.docs = docs;
  a.cph = PN_SM_COUNT / n_heads;
  if (a.cph < 1) a.cph = 1;
  if (a.cph > docs) a.cph = docs;
  a.q_col = q_col; a.k_col = k_col; a.v_col = v_col;
  a.bias = bias; a.out = out; a.ld_out = ld_out;
  a.lse = lse_out;
  a.klen = klen;
  a.drop = drop;
  a.dbg = nullptr;
  if (getenv("PN_ATTN_DBG")) {
    static long long* dbg = nullptr;
    if (!dbg) { cudaMalloc(&dbg, 2 * 64 * 8 * 8); }
    cudaMemsetAsync(dbg, 0, 2 * 64 * 8 * 8, st);
    a.dbg = dbg;
    g_pn_attn_dbg = dbg;
  }
  static bool attr_done = false;
  if (!attr_done) {
    PN_CUDA(cudaFuncSetAttribute(t5_self_attn2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 ATTN2_SMEM));
    PN_CUDA(cudaFuncSetAttribute(t5_self_attn2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 ATTN2_SMEM));
    attr_done = true;
  }
  PnProfScope prof(PN_PROF_ATTN, 4.0 * docs * n_heads * (double)S * S * DK, st);
  if (drop.thresh) t5_self_attn2_kernel<true><<<n_heads * a.cph, kThreads, ATTN2_SMEM, st>>>(tm, a);
  else t5_self_attn2_kernel<false><<<n_heads * a.cph, kThreads, ATTN2_SMEM, st>>>(tm, a);
  PN_LAUNCHED();
  return PN_OK;
}
