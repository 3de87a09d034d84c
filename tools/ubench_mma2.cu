// ubench_mma2.cu -- cycles per tcgen05.mma.cta_group::2 (kind::f16, M = 256 over a CTA pair, K = 16, operands in
// shared memory, no-swizzle K-major descriptors) as a function of N, beside the cta_group::1 figure of ubench_mma.cu.
// Question it answers: is the ~60-clk floor of a short (N <= 128) MMA per INSTRUCTION (then M = 256 halves the cost per
// FLOP of the N' = 96 / N = 128 conv layers) or per operand byte read by each SM (then it does not).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_mma2 tools/ubench_mma2.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t make_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((addr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.b32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ uint32_t cluster_ctarank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync() { asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory"); }

// N = accumulator width of the pair MMA: D is 256 x N; each CTA supplies 128 rows of A and N/2 rows of B.
template <int N>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) k_mma2(int reps, int nacc, int spread, long long* out, int a_off) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  const uint32_t rank = cluster_ctarank();
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = slot;
  constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
  if (warp == 1 && rank == 0) {
    const uint32_t a0 = smem_u32(smem), b0 = a0 + 64 * 1024;
    long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      if (elect_one()) {
#pragma unroll
        for (int j = 0; j < 9; ++j) {
          const uint64_t ad = make_desc(a0 + ((r * 9 + j) % 8) * spread + (j % 3) * a_off, 4096, 128);
          const uint64_t bd = make_desc(b0 + ((r * 9 + j) % 8) * spread, (N / 2) * 16, 128);
          const uint32_t d = tmem + ((r % nacc) * N);
          const uint32_t acc = j != 0;
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(ad), "l"(bd), "r"(IDESC), "r"(acc) : "memory");
        }
      }
      __syncwarp();
    }
    if (elect_one()) asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "h"((uint16_t)1) : "memory");
    __syncwarp();
    asm volatile("{\n\t.reg .pred p;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}" ::"r"(smem_u32(&bar)) : "memory");
    long long t1 = clock64();
    if ((threadIdx.x & 31) == 0) out[blockIdx.x >> 1] = t1 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync();
  if (warp == 0) { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512)); }
}
template <int N>
int run(int grid, int nacc, int spread, int a_off = 0) {
  long long* d; CK(cudaMalloc(&d, sizeof(long long) * grid));
  CK(cudaFuncSetAttribute(k_mma2<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  const int reps = 2000;
  k_mma2<N><<<grid, 128, 200 * 1024>>>(reps, nacc, spread, d, a_off);
  CK(cudaDeviceSynchronize());
  k_mma2<N><<<grid, 128, 200 * 1024>>>(reps, nacc, spread, d, a_off);
  CK(cudaDeviceSynchronize());
  const int np = grid / 2;
  long long h[74]; CK(cudaMemcpy(h, d, sizeof(long long) * np, cudaMemcpyDeviceToHost));
  double avg = 0; for (int i = 0; i < np; ++i) avg += h[i]; avg /= np;
  printf("cta_group::2 M=256 N=%3d grid=%3d nacc=%d spread=%5d a_off=%2d: %.1f clk per MMA (floor %d), %.0f FLOP/clk/SM\n", N, grid, nacc, spread, a_off,
         avg / (reps * 9.0), N / 2, 2.0 * 128 * N * 16 * reps * 9.0 / avg);
  cudaFree(d);
  return 0;
}
int main() {
  const int grid = 148;
  run<256>(grid, 2, 8192); run<192>(grid, 2, 8192); run<128>(grid, 2, 8192); run<96>(grid, 4, 8192); run<64>(grid, 4, 8192); run<32>(grid, 4, 8192);
  printf("-- misaligned A (+16 B per x-tap)\n");
  run<256>(grid, 2, 8192, 16); run<192>(grid, 2, 8192, 16); run<128>(grid, 2, 8192, 16); run<96>(grid, 4, 8192, 16);
  return 0;
}
