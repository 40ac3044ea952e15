// Micro-benchmark (GPU box): cycles per tcgen05.mma (M = 128, K = 16, bf16, operands in shared memory,
// K-major SWIZZLE_NONE like the convolution kernels) for N = 32..256, issued as ONE dependent chain into the
// same accumulator columns, operands in the SWIZZLE_NONE or the SWIZZLE_128B form, with one or two CTAs on the SM.
// (Alternating over 2 / 4 accumulators, or moving the A start address by a row per MMA like the taps do, changes nothing.)
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I umetrack_b200/csrc -I include \
//        tools/ubench/mma_chain.cu -o /tmp/mma_chain && /tmp/mma_chain
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "umma.cuh"

using namespace umt;

// SWIZZLE_128B K-major descriptor (rows of 128 bytes = 64 bf16 of K, 8-row groups 1024 bytes apart); only the
// timing matters here, the operand values are zeros
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)1 << 16) | ((uint64_t)((1024u >> 4) & 0x3FFF) << 32) |
         (1ull << 46) | (2ull << 61);
}

__global__ void __launch_bounds__(128) chain_kernel(int n_cols, int n_acc, int n_mma, int a_step_rows, int layout, long long* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint32_t tmem_slot;
  __shared__ uint64_t bar;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(256u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int i = threadIdx.x; i < 99 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp == 0) {
    const bool leader = elect_one();
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n_cols >> 3) << 17) | ((128u >> 4) << 24);
    // A: 160 rows x 2 planes (plane distance 160 * 16 B); B: n_cols rows x 2 planes
    const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem + 8192);
    const uint32_t a1 = (a0 + 1023u) & ~1023u, b1 = a1 + 32768u;   // 1024-byte aligned tiles for the swizzled forms
    const uint64_t adesc = (layout & 1) ? umma_desc_sw128(a1) : umma_desc(a0, 160 * 16, 128);
    const uint64_t bdesc = (layout & 2) ? umma_desc_sw128(b1) : umma_desc(b0, (uint32_t)n_cols * 16, 128);
    long long t0 = 0, t1 = 0;
    if (leader) {
      t0 = clock64();
      for (int i = 0; i < n_mma; ++i) {
        const uint32_t d = tmem + (uint32_t)((i % n_acc) * n_cols);
        // a_step_rows: consecutive MMAs read A from a start address moved by that many rows (like the taps)
        tc_mma_bf16(d, adesc + (uint64_t)((i % 9) * a_step_rows), bdesc, idesc, i >= n_acc ? 1u : 0u);
      }
      tc_commit(smem_u32(&bar));
    }
    __syncwarp();
    mbar_wait(smem_u32(&bar), 0);
    t1 = clock64();
    if (leader && blockIdx.x == 0) out[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
}

// The same for CTA pairs: tcgen05.mma.cta_group::2, M = 256 over two SMs, each CTA holding its 128 rows of A and
// half of the N rows of B; issued by the leader CTA, completion multicast to both.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128) pair_chain_kernel(int n_cols, int n_mma, long long* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint32_t tmem_slot;
  __shared__ uint64_t bar;
  const int warp = threadIdx.x >> 5;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(256u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  for (int i = threadIdx.x; i < 48 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp == 0) {
    long long t0 = 0, t1 = 0;
    if (rank == 0) {
      const bool leader = elect_one();
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n_cols >> 3) << 17) | ((256u >> 4) << 24);
      const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem + 8192);
      const uint64_t adesc = umma_desc(a0, 160 * 16, 128), bdesc = umma_desc(b0, (uint32_t)(n_cols / 2) * 16, 128);
      if (leader) {
        t0 = clock64();
        for (int i = 0; i < n_mma; ++i) {
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                       "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                       ::"r"(tmem), "l"(adesc + (uint64_t)(i % 9)), "l"(bdesc), "r"(idesc), "r"(i ? 1u : 0u) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                     ::"r"(smem_u32(&bar)), "h"((uint16_t)3) : "memory");
      }
      __syncwarp();
    }
    mbar_wait(smem_u32(&bar), 0);
    t1 = clock64();
    if (rank == 0 && blockIdx.x == 0 && (threadIdx.x & 31) == 0) out[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
}

int main() {
  long long* d_out;
  cudaMalloc(&d_out, 8);
  cudaFuncSetAttribute(chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  const int n_mma = 2048;
  printf("cycles per MMA (M=128, K=16, bf16, SWIZZLE_NONE K-major), %d MMAs per CTA\n", n_mma);
  printf("%6s %8s %8s %12s %12s %14s\n", "N", "A", "B", "1 CTA / SM", "2 CTAs / SM", "math (cycles)");
  for (int n_cols : {32, 64, 128, 256})
    for (int layout = 0; layout < 4; ++layout) {
      long long c[2];
      for (int two = 0; two < 2; ++two) {
        const int grid = two ? 2 * 148 : 148;
        const size_t smem = 100 * 1024;   // two CTAs fit an SM; 148 CTAs spread one per SM
        for (int rep = 0; rep < 2; ++rep) {
          chain_kernel<<<grid, 128, smem>>>(n_cols, 1, n_mma, 1, layout, d_out);
          if (cudaDeviceSynchronize() != cudaSuccess) { printf("error %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
        }
        cudaMemcpy(&c[two], d_out, 8, cudaMemcpyDeviceToHost);
      }
      printf("%6d %8s %8s %12.1f %12.1f %14.1f\n", n_cols, (layout & 1) ? "SW128" : "none", (layout & 2) ? "SW128" : "none",
             (double)c[0] / n_mma, (double)c[1] / n_mma, 128.0 * n_cols * 16 / 4096);
    }
  cudaFuncSetAttribute(pair_chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  printf("\nCTA pairs (cta_group::2, M = 256): cycles per MMA\n%6s %14s %14s %16s\n", "N", "1 CTA / SM", "2 CTAs / SM", "math per SM");
  for (int n_cols : {64, 128, 256}) {
    long long c[2];
    for (int two = 0; two < 2; ++two) {
      const int grid = two ? 2 * 148 : 148;
      for (int rep = 0; rep < 2; ++rep) {
        pair_chain_kernel<<<grid, 128, 100 * 1024>>>(n_cols, n_mma, d_out);
        if (cudaDeviceSynchronize() != cudaSuccess) { printf("error %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
      }
      cudaMemcpy(&c[two], d_out, 8, cudaMemcpyDeviceToHost);
    }
    printf("%6d %14.1f %14.1f %16.1f\n", n_cols, (double)c[0] / n_mma, (double)c[1] / n_mma, 256.0 * n_cols * 16 / 8192);
  }
  return 0;
}
