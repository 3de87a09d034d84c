// ubench_shfl.cu -- issue rate of the instructions the conv_rows / conv_block epilogues are made of, per SM, with the
// kernels' own occupancy (16 warps of 1 CTA per SM): SHFL (x-tap combine: 64 per warp and M-tile), packed f32x2 adds,
// F2FP pack, and the mix of one epilogue iteration.  Question: are the level-0/1 layers (~600 clk per M-tile whatever
// their HBM traffic, profiles/r02_conv.md) bound by the shuffle unit?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_shfl tools/ubench_shfl.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) { uint64_t r; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

template <int MODE>
__global__ void __launch_bounds__(512, 1) k(int reps, long long* out, uint32_t* sink) {
  uint32_t v[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = threadIdx.x * 16 + i;
  uint64_t a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = ((uint64_t)__float_as_uint(1.0f + i) << 32) | __float_as_uint(0.5f * threadIdx.x);
  __syncthreads();
  const long long t0 = clock64();
  for (int r = 0; r < reps; ++r) {
    if (MODE == 0 || MODE == 2) {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = __shfl_up_sync(0xffffffffu, v[i], 1);          // 16 independent shuffles
    }
    if (MODE == 1 || MODE == 2) {
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = add2(a[i], a[(i + 1) & 7]);                        // 8 packed adds
    }
    if (MODE == 3) {                                                                       // F2FP packs
#pragma unroll
      for (int i = 0; i < 16; i += 2) {
        uint32_t p;
        asm volatile("cvt.rn.relu.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(p) : "f"(__uint_as_float(v[i])), "f"(__uint_as_float(v[i + 1])));
        v[i] ^= p;
      }
    }
  }
  const long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s ^= v[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) s ^= (uint32_t)a[i] ^ (uint32_t)(a[i] >> 32);
  if (s == 0x12345678u) sink[0] = s;
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
}

template <int MODE>
int run(const char* what, double per_rep_warp_instr) {
  long long* d; uint32_t* sink;
  CK(cudaMalloc(&d, sizeof(long long) * 148)); CK(cudaMalloc(&sink, 4));
  const int reps = 4000;
  k<MODE><<<148, 512>>>(reps, d, sink); CK(cudaDeviceSynchronize());
  k<MODE><<<148, 512>>>(reps, d, sink); CK(cudaDeviceSynchronize());
  long long h[148]; CK(cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost));
  double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
  const double instr = per_rep_warp_instr * reps * 16;     // warp-instructions per SM
  printf("%-44s 16 warps/SM: %.2f warp-instr per clk per SM (%.1f clk per warp-instr)\n", what, instr / avg, avg / instr);
  cudaFree(d); cudaFree(sink);
  return 0;
}
int main() {
  run<0>("SHFL.UP (32-bit)", 16);
  run<1>("FADD2 (add.rn.f32x2)", 8);
  run<2>("16 SHFL + 8 FADD2 interleaved", 24);
  run<3>("F2FP.SATFINITE.RELU.F16.F32.PACK + LOP", 16);
  return 0;
}
