// ubench_mma.cu -- cycles per tcgen05.mma (cta_group::1, kind::f16, M=128, K=16, operands in shared memory,
// no-swizzle K-major descriptors as the conv kernels build them) as a function of N, with nothing else running.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_mma tools/ubench_mma.cu
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
template <int N>
__global__ void __launch_bounds__(128, 1) k_mma(int reps, int nacc, int spread, long long* out, int a_off, int fill_kb, const uint8_t* gsrc, int rnd) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = rnd ? (0x3c003c00u ^ ((i * 2654435761u) & 0x03ff03ffu)) : 0;   // fp16 values in [1,2)
  __shared__ uint64_t fbar; __shared__ volatile int stop;
  if (threadIdx.x == 0) { stop = 0; asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&fbar))); }
  if (threadIdx.x == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = slot;
  constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  if (warp == 1) {
    const uint32_t a0 = smem_u32(smem), b0 = a0 + 64 * 1024;
    long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      if (elect_one()) {
#pragma unroll
        for (int j = 0; j < 9; ++j) {
          // A: 128 rows x 16 k: rows 16 B apart, chunks 4 KB apart; B: N rows x 16 k, chunks N*16 apart; operands move by `spread` bytes per MMA
          const uint64_t ad = make_desc(a0 + ((r * 9 + j) % 8) * spread + (j % 3) * a_off, 4096, 128);
          const uint64_t bd = make_desc(b0 + ((r * 9 + j) % 8) * spread, N * 16, 128);
          const uint32_t d = tmem + ((r % nacc) * N);
          const uint32_t acc = j != 0;
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(ad), "l"(bd), "r"(IDESC), "r"(acc) : "memory");
        }
      }
      __syncwarp();
    }
    if (elect_one()) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    __syncwarp();
    asm volatile("{\n\t.reg .pred p;\n\tW:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}" ::"r"(smem_u32(&bar)) : "memory");
    long long t1 = clock64();
    if ((threadIdx.x & 31) == 0) { out[blockIdx.x] = t1 - t0; stop = 1; }
  } else if (warp == 2 && fill_kb > 0) {
    // concurrent shared-memory fill traffic: bulk copies of fill_kb KB from global (L2-resident) into a scratch area, back to back
    if ((threadIdx.x & 31) == 0) {
      uint32_t ph = 0;
      const uint32_t dst = smem_u32(smem) + 128 * 1024, fb = smem_u32(&fbar);
      const uint8_t* src = gsrc + (size_t)blockIdx.x * 64 * 1024;
      while (!stop) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fb), "r"(fill_kb * 1024) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(fill_kb * 1024), "r"(fb) : "memory");
        asm volatile("{\n\t.reg .pred p;\n\tW2:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D2;\n\tbra W2;\n\tD2:\n\t}" ::"r"(fb), "r"(ph) : "memory");
        ph ^= 1;
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512)); }
}
template <int N>
int run(int grid, int nacc, int spread, int a_off = 0, int fill_kb = 0, int rnd = 0) {
  static uint8_t* gsrc = nullptr; if (!gsrc) { cudaMalloc(&gsrc, 148 * 64 * 1024); cudaMemset(gsrc, 0x3c, 148 * 64 * 1024); }
  long long* d; CK(cudaMalloc(&d, sizeof(long long) * grid));
  CK(cudaFuncSetAttribute(k_mma<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  const int reps = 2000;
  k_mma<N><<<grid, 128, 200 * 1024>>>(reps, nacc, spread, d, a_off, fill_kb, gsrc, rnd);
  CK(cudaDeviceSynchronize());
  k_mma<N><<<grid, 128, 200 * 1024>>>(reps, nacc, spread, d, a_off, fill_kb, gsrc, rnd);
  CK(cudaDeviceSynchronize());
  long long h[148]; CK(cudaMemcpy(h, d, sizeof(long long) * (grid < 148 ? grid : 148), cudaMemcpyDeviceToHost));
  double avg = 0; for (int i = 0; i < (grid < 148 ? grid : 148); ++i) avg += h[i]; avg /= (grid < 148 ? grid : 148);
  printf("N=%3d grid=%3d nacc=%d spread=%5d a_off=%2d fill=%2dKB rnd=%d: %.1f clk per MMA (floor %d), %.0f FLOP/clk/SM\n", N, grid, nacc, spread, a_off, fill_kb, rnd, avg / (reps * 9.0), N / 2,
         2.0 * 128 * N * 16 * reps * 9.0 / avg);
  cudaFree(d);
  return 0;
}
int main() {
  const int grid = 148;
  run<256>(grid, 2, 8192); run<192>(grid, 2, 8192); run<128>(grid, 2, 8192); run<96>(grid, 4, 8192); run<32>(grid, 4, 8192);
  printf("-- misaligned A (+16 B per x-tap)\n");
  run<256>(grid, 2, 8192, 16); run<128>(grid, 2, 8192, 16); run<96>(grid, 4, 8192, 16);
  printf("-- random fp16 data\n");
  run<256>(grid, 2, 8192, 0, 0, 1); run<96>(grid, 4, 8192, 0, 0, 1);
  printf("-- concurrent bulk-copy fill of shared memory\n");
  run<256>(grid, 2, 8192, 0, 16); run<256>(grid, 2, 8192, 0, 48); run<256>(grid, 2, 8192, 16, 48, 1); run<96>(grid, 4, 8192, 0, 32); run<96>(grid, 4, 8192, 16, 32, 1);
  return 0;
}
