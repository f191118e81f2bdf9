// micro-benchmark: cycles per tcgen05.mma (cta_group::1, bf16) as a function of the tile shape, the A operand source
// (shared memory = SS, tensor memory = TS) and of how many independent accumulators a chain of products alternates
// between.  One CTA per SM; one elected lane issues CHAIN products, commits, and waits for the commit's mbarrier.
// Operands are zeros (the pipe does not care).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17
//   -I pacednegatives_b200/csrc -o tools/ubench/mma_chain tools/ubench/mma_chain.cu -lcuda
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include "pn_tc.cuh"
using namespace pntc;

constexpr int CHAIN = 64;
constexpr int SMEM = 96 * 1024 + 2048;

template <int N, bool TS, int NACC>
__global__ void __launch_bounds__(128, 1) chain_kernel(long long* out, int reps) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* g = smem_raw + (base - smem_u32(smem_raw));
  const uint32_t bar = base + 96 * 1024, slot = bar + 16;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(g)[i] = 0u;
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_mbar_init(); }
  if (warp == 0) { tmem_alloc<1>(slot, 512); tmem_relinquish<1>(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<uint32_t*>(g + 96 * 1024 + 16);
  if (warp == 0) {
    constexpr uint32_t IDESC = make_idesc_bf16(128, N, 0, TS ? 1 : 0);
    const uint64_t adesc = make_desc_kmajor_sw128(base);                   // A [128 x 64] K-major (SS)
    const uint64_t bdesc_k = make_desc_kmajor_sw128(base + 16384);         // B [N x 64] K-major (SS)
    const uint64_t bdesc_mn = make_desc_mnmajor_sw128(base + 16384, 1024); // B [16 keys x N] MN-major (TS, as P V)
    long long best = 1ll << 60;
    for (int r = 0; r < reps; ++r) {
      __syncwarp();
      const long long t0 = clock64();
      if (lane == 0) {
#pragma unroll 8
        for (int i = 0; i < CHAIN; ++i) {
          const uint32_t d = tmem + (uint32_t)((i % NACC) * N) % 512u;      // NACC accumulators of N columns
          if (TS) umma_bf16_ts(d, tmem + 448 + (i & 7) * 8, bdesc_mn + 128u * (i & 7), IDESC, i >= NACC ? 1u : 0u);
          else umma_bf16<1>(d, adesc + 2u * (i & 3), bdesc_k + 2u * (i & 3), IDESC, i >= NACC ? 1u : 0u);
        }
        umma_commit<1>(bar);
      }
      __syncwarp();
      mbar_wait(bar, r & 1);
      const long long t1 = clock64();
      if (t1 - t0 < best) best = t1 - t0;
    }
    if (lane == 0) out[blockIdx.x] = best;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<1>(tmem, 512);
}

template <int N, bool TS, int NACC>
void run(long long* d_out, const char* what) {
  auto k = chain_kernel<N, TS, NACC>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
  for (int grid : {1, 148}) {
    k<<<grid, 128, SMEM>>>(d_out, 20);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, d_out, grid * 8, cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("%-3s 128x%-3dx16  accumulators %d  grid %3d: %7.1f cycles per product (chain of %d incl. commit + wait; floor %d)  %s\n",
           TS ? "TS" : "SS", N, NACC, grid, (double)mx / CHAIN, CHAIN, 128 * N / 256, e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
}

int main() {
  long long* d_out;
  cudaMalloc(&d_out, 148 * 8);
  run<64, true, 1>(d_out, "");
  run<64, true, 2>(d_out, "");
  run<64, true, 4>(d_out, "");
  run<64, false, 1>(d_out, "");
  run<64, false, 2>(d_out, "");
  run<128, false, 1>(d_out, "");
  run<128, false, 2>(d_out, "");
  run<256, false, 1>(d_out, "");
  run<256, false, 2>(d_out, "");
  run<128, true, 1>(d_out, "");
  run<128, true, 2>(d_out, "");
  return 0;
}
