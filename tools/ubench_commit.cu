// ubench_commit.cu -- what do tcgen05.commit and the accumulator hand-shake cost the MMA-issuing thread?
// Batches of MB tcgen05.mma (M=128, K=16, N template) into a ring of 4 accumulators, 1 CTA per SM, all SMs.
//   mode 0: MMAs only                         mode 1: + one commit per batch (nobody waits on it)
//   mode 2: + two commits per batch           mode 3: one commit per batch, a consumer warp per accumulator waits for it and
//   hands the accumulator back (the issuer waits for that before reusing it) -- the conv_rows protocol
//   mode 4: mode 3 + a second commit per batch (the conv_rows "box empty" commit)
// issuers = 1 or 2 (alternate batches).  Prints clk per batch next to the pipe time MB x clk-per-MMA.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_commit tools/ubench_commit.cu
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
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile("{\n\t.reg .pred p;\n\tWL:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra WD;\n\tbra WL;\n\tWD:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void commit(uint32_t bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory"); }

template <int N, int MB>
__global__ void __launch_bounds__(320, 1) k_commit(int reps, int mode, int issuers, long long* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bars[16];          // 0-3 accumulator full, 4-7 accumulator empty, 8-15 second commit target
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 16; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bars[i])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = slot;
  constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
  const uint32_t bfull = smem_u32(&bars[0]), bempty = smem_u32(&bars[4]), bsec = smem_u32(&bars[8]);
  const bool handshake = mode >= 3;
  if (warp >= 1 && warp <= issuers) {
    const int wi = warp - 1;
    const uint32_t a0 = smem_u32(smem), b0 = a0 + 32 * 1024;
    long long t_issue = 0, t_c1 = 0, t_c2 = 0, t_wait = 0;
    const long long t0 = clock64();
    for (int r = wi; r < reps; r += issuers) {
      const int buf = r & 3, use = r >> 2;
      long long ta = clock64();
      if (handshake) {
        mbar_wait(bempty + 8 * buf, (use & 1) ^ 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      }
      long long tb = clock64();
      if (elect_one()) {
#pragma unroll
        for (int j = 0; j < MB; ++j) {
          const uint64_t ad = make_desc(a0 + (j % 3) * 512 + (j / 3) * 8192, 4096, 128);
          const uint64_t bd = make_desc(b0 + j * 4096, N * 16, 128);
          const uint32_t acc = j != 0;
          asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem + buf * 128), "l"(ad), "l"(bd), "r"(IDESC), "r"(acc) : "memory");
        }
      }
      __syncwarp();
      long long tc = clock64();
      if (mode >= 1) { if (elect_one()) commit(bfull + 8 * buf); __syncwarp(); }
      long long td = clock64();
      if (mode == 2 || mode == 4) { if (elect_one()) commit(bsec + 8 * (r & 7)); __syncwarp(); }
      long long te = clock64();
      t_wait += tb - ta; t_issue += tc - tb; t_c1 += td - tc; t_c2 += te - td;
    }
    // drain
    if (elect_one()) commit(bsec + 8 * wi);
    __syncwarp();
    const long long t1 = clock64();
    if (lane == 0) { long long* o = out + (blockIdx.x * 2 + wi) * 5; o[0] = t1 - t0; o[1] = t_wait; o[2] = t_issue; o[3] = t_c1; o[4] = t_c2; }
  } else if (warp >= 4 && warp < 8 && handshake) {
    // consumer of accumulator (warp-4): wait for its commit, hand it back
    const int buf = warp - 4;
    for (int r = buf, use = 0; r < reps; r += 4, ++use) {
      mbar_wait(bfull + 8 * buf, use & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bempty + 8 * buf) : "memory");
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512)); }
}

template <int N, int MB>
int run(int mode, int issuers, double clk_per_mma) {
  const int grid = 148, reps = 4000;
  long long* d; CK(cudaMalloc(&d, sizeof(long long) * grid * 10));
  CK(cudaMemset(d, 0, sizeof(long long) * grid * 10));
  CK(cudaFuncSetAttribute(k_commit<N, MB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  for (int i = 0; i < 2; ++i) { k_commit<N, MB><<<grid, 320, 100 * 1024>>>(reps, mode, issuers, d); CK(cudaDeviceSynchronize()); }
  long long h[148 * 10]; CK(cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost));
  double a[5] = {0, 0, 0, 0, 0};
  for (int i = 0; i < grid; ++i) for (int k = 0; k < 5; ++k) a[k] += (double)h[i * 10 + k] / grid;
  const double per_issuer = (double)reps / issuers;
  printf("N=%3d MB=%d mode %d issuers %d: %7.1f clk per batch (pipe %4.0f) | issuer 0 per own batch: wait %5.1f issue %6.1f commit1 %6.1f commit2 %6.1f\n",
         N, MB, mode, issuers, a[0] / reps, MB * clk_per_mma, a[1] / per_issuer, a[2] / per_issuer, a[3] / per_issuer, a[4] / per_issuer);
  cudaFree(d);
  return 0;
}
int main() {
  for (int issuers = 1; issuers <= 2; ++issuers)
    for (int mode = 0; mode <= 4; ++mode) run<96, 6>(mode, issuers, 72.1);
  for (int issuers = 1; issuers <= 2; ++issuers)
    for (int mode = 0; mode <= 4; ++mode) run<96, 3>(mode, issuers, 72.1);
  for (int issuers = 1; issuers <= 2; ++issuers)
    for (int mode = 0; mode <= 4; ++mode) run<128, 9>(mode, issuers, 76.8);
  return 0;
}
