// ubench_tma.cu -- how fast can one SM pull DRAM-resident activation tiles into shared memory?
// Persistent kernel, 1 CTA/SM, S-stage ring; a consumer thread only waits and releases.
//   mode 0: TMA 4-D box {32 pos x 16 B, 6 rows, CPK chunks} over the PP layout (conv_rows pattern)
//   mode 1: 1-D bulk copies, one contiguous run of the same byte count per stage
//   mode 2: mode 0 + LSU prefetch.global.L2 of the box PFD items ahead, issued by idle warps
//   mode 3: mode 0 but tiles ordered so that one CTA walks ALONG a row group (x fastest per CTA)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_tma tools/ubench_tma.cu -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count)); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile("{\n\t.reg .pred p;\n\tWL:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra WD;\n\tbra WL;\n\tWD:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const void* tmap, int c0, int c1, int c2, int c3, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

struct Params {
  alignas(64) CUtensorMap tm;
  const uint8_t* base;
  int bc, br, nlw, mode, S, cpk, tiles_x, tiles_y, B, nchunks, pitch, Ppl, G, pfd, kbn;
  long long total; size_t buf_bytes;
  uint8_t* out; int write_bytes;     // optional concurrent write stream per item
};

__global__ void __launch_bounds__(448, 1) k_pull(const __grid_constant__ Params P) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t bar_full = smem_u32(smem), bar_empty = bar_full + 8 * P.S;
  const uint32_t a_base = smem_u32(smem) + 1024;
  const int stage_bytes = P.cpk * P.br * P.bc * 16;
  if (threadIdx.x == 0) {
    *reinterpret_cast<volatile int*>(smem + 512) = 0;
    for (int i = 0; i < P.S; ++i) { mbar_init(bar_full + 8 * i, (P.mode == 5 || P.mode == 6) ? P.nlw * 32 : 1); mbar_init(bar_empty + 8 * i, 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int tpi = P.tiles_x * P.tiles_y;
  auto coords = [&](long long item, int& b, int& ty, int& tx) {
    if (P.mode == 3) {   // each CTA owns contiguous runs of tiles_x tiles
      long long rg = item / P.tiles_x; tx = (int)(item - rg * P.tiles_x);
      b = (int)(rg / P.tiles_y); ty = (int)(rg - (long long)b * P.tiles_y);
    } else {
      b = (int)(item / tpi); int t = (int)(item - (long long)b * tpi); ty = t / P.tiles_x; tx = t - ty * P.tiles_x;
    }
  };
  auto next_item = [&](long long it) -> long long {   // it = CTA-local iteration index
    if (P.mode == 3) { long long run = it / P.tiles_x, k = it - run * P.tiles_x; return (run * gridDim.x + blockIdx.x) * P.tiles_x + k; }
    return blockIdx.x + it * gridDim.x;
  };
  if (warp >= 8 && warp < 8 + P.nlw && P.mode < 5) {
    if (lane == 0) {
      int s = 0; uint32_t ph = 0; long long gstage = 0;
      for (long long it = 0;; ++it) {
        long long item = next_item(it);
        if (item >= P.total) break;
        int b, ty, tx; coords(item, b, ty, tx);
        for (int kb = 0; kb < P.kbn; ++kb, ++gstage) {
          if (gstage % P.nlw != warp - 8) { if (++s == P.S) { s = 0; ph ^= 1; } continue; }
          if (P.pfd != 1) mbar_wait(bar_empty + 8 * s, ph ^ 1);
          mbar_expect_tx(bar_full + 8 * s, stage_bytes);
          if (P.mode == 1) {          // one contiguous run (wraps inside the buffer)
            const size_t off = (((size_t)item * P.kbn + kb) * stage_bytes) % (P.buf_bytes - stage_bytes);
            bulk_g2s(a_base + s * stage_bytes, P.base + (off & ~(size_t)15), stage_bytes, bar_full + 8 * s);
          } else if (P.mode == 4) {   // the box as 6*cpk bulk copies of 512 B, one thread
            for (int cc = 0; cc < P.cpk; ++cc)
              for (int row = 0; row < 6; ++row) {
                const size_t pos = (size_t)P.G + (size_t)(ty * 4 + row) * P.pitch + tx * 30;
                bulk_g2s(a_base + s * stage_bytes + (cc * 6 + row) * 512,
                         P.base + (((size_t)b * P.nchunks + kb * P.cpk + cc) * P.Ppl + pos) * 16, 512, bar_full + 8 * s);
              }
          } else {
            tma_load_4d(a_base + s * stage_bytes, &P.tm, tx * (P.pfd == 2 ? P.bc : P.bc - 2) * 2, ty * (P.br - 2), kb * P.cpk, b, bar_full + 8 * s);
          }
          if (++s == P.S) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int s = 0; uint32_t ph = 0;
      for (long long it = 0;; ++it) {
        long long item = next_item(it);
        if (item >= P.total) break;
        for (int kb = 0; kb < P.kbn; ++kb) {
          mbar_wait(bar_full + 8 * s, ph);
          mbar_arrive(bar_empty + 8 * s);
          if (++s == P.S) { s = 0; ph ^= 1; }
        }
        *reinterpret_cast<volatile int*>(smem + 512) = (int)it + 1;
      }
    }
  } else if (warp == 2 && P.mode == 2) {
    // LSU L2 prefetch of the boxes PFD iterations ahead of the consumer (paced by its progress counter)
    volatile int* prog = reinterpret_cast<volatile int*>(smem + 512);
    const int nlines = 6 * P.cpk * P.kbn * 5;
    for (long long it = 0;; ++it) {
      long long item = next_item(it);
      if (item >= P.total) break;
      while (it > (long long)*prog + P.pfd) { }
      int b, ty, tx; coords(item, b, ty, tx);
      for (int r = lane; r < nlines; r += 32) {
        const int run = r / 5, line = r - run * 5;
        const int cc = run / 6, row = run - cc * 6;
        const size_t pos = (size_t)P.G + (size_t)(ty * 4 + row) * P.pitch + tx * 30;
        const uint8_t* a = P.base + (((size_t)b * P.nchunks + cc) * P.Ppl + pos) * 16;
        const uint8_t* al = (const uint8_t*)(((uintptr_t)a & ~(uintptr_t)127) + line * 128);
        if (al < a + 512) asm volatile("prefetch.global.L2 [%0];" ::"l"(al));
      }
    }
  } else if (warp >= 3 && warp < 3 + P.nlw && (P.mode == 5 || P.mode == 6)) {
    const int t = threadIdx.x - 96, nt = P.nlw * 32;
    int s = 0; uint32_t ph = 0;
    for (long long it = 0;; ++it) {
      long long item = next_item(it);
      if (item >= P.total) break;
      int b, ty, tx; coords(item, b, ty, tx);
      for (int kb = 0; kb < P.kbn; ++kb) {
        mbar_wait(bar_empty + 8 * s, ph ^ 1);
        const int nitems = P.cpk * 6 * 32;
        if (P.mode == 5) {
          for (int i = t; i < nitems; i += nt) {
            const int l = i & 31, j = i >> 5, cc = j / 6, row = j - cc * 6;
            const size_t pos = (size_t)P.G + (size_t)(ty * 4 + row) * P.pitch + tx * 30 + l;
            const uint8_t* src = P.base + (((size_t)b * P.nchunks + kb * P.cpk + cc) * P.Ppl + pos) * 16;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(a_base + s * stage_bytes + i * 16), "l"(src) : "memory");
          }
          asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar_full + 8 * s) : "memory");
        } else {
          uint4 v[6];
          int k = 0;
          for (int i = t; i < nitems && k < 6; i += nt, ++k) {
            const int l = i & 31, j = i >> 5, cc = j / 6, row = j - cc * 6;
            const size_t pos = (size_t)P.G + (size_t)(ty * 4 + row) * P.pitch + tx * 30 + l;
            v[k] = __ldg(reinterpret_cast<const uint4*>(P.base + (((size_t)b * P.nchunks + kb * P.cpk + cc) * P.Ppl + pos) * 16));
          }
          k = 0;
          for (int i = t; i < nitems && k < 6; i += nt, ++k)
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(a_base + s * stage_bytes + i * 16), "r"(v[k].x), "r"(v[k].y), "r"(v[k].z), "r"(v[k].w) : "memory");
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          mbar_arrive(bar_full + 8 * s);
        }
        if (++s == P.S) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp >= 3 + 8 && P.out) {
    // concurrent write stream (the epilogue's stores): write_bytes per item, 16 B per thread
    const int t = (warp - 11) * 32 + lane;
    for (long long it = 0;; ++it) {
      long long item = next_item(it);
      if (item >= P.total) break;
      uint4 v = make_uint4((uint32_t)item, 1, 2, 3);
      for (int o = t * 16; o < P.write_bytes; o += 96 * 16)
        *reinterpret_cast<uint4*>(P.out + (size_t)item * P.write_bytes + o) = v;
    }
  }
}

int main(int argc, char** argv) {
  const int H = 512, W = 512; const int B = argc > 3 ? atoi(argv[3]) : 64;
  const int pitch = W + 2, G = ((W + 3) + 7) & ~7, npos = (H + 2) * pitch;
  const int Ppl = ((G + npos + 400 + (W + 3)) + 7) & ~7;
  int nchunks = argc > 1 ? atoi(argv[1]) : 4;     // channels/8 of the tensor
  int cpk = argc > 2 ? atoi(argv[2]) : 4;         // chunks per box
  const size_t bytes = (size_t)B * nchunks * Ppl * 16;
  uint8_t *d, *dout;
  CK(cudaMalloc(&d, bytes)); CK(cudaMemset(d, 1, bytes));
  const int write_bytes = 30 * 4 * 4 * 16;        // 120 positions x 4 chunks x 16 B (what the epilogue stores)
  const int tiles_x = (W + 29) / 30, tiles_y = H / 4;
  const long long total = (long long)B * tiles_x * tiles_y;
  CK(cudaMalloc(&dout, (size_t)total * write_bytes));
  typedef CUresult (*Fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* fn; cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  CK(cudaFuncSetAttribute(k_pull, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int geo[3][2] = {{32, 6}, {32, 10}, {32, 18}};
  for (auto& g : geo) {
    Params P{};
    const int bc = g[0], br = g[1];
    // 64-bit elements: the inner box dimension (<= 256 elements) can then span up to 2 KB = 128 positions
    cuuint64_t gd[4] = {(cuuint64_t)pitch * 2, (cuuint64_t)H + 2, (cuuint64_t)nchunks, (cuuint64_t)B};
    cuuint64_t gs[3] = {(cuuint64_t)pitch * 16, (cuuint64_t)Ppl * 16, (cuuint64_t)nchunks * Ppl * 16};
    cuuint32_t bx[4] = {(cuuint32_t)bc * 2, (cuuint32_t)br, (cuuint32_t)cpk, 1}, es[4] = {1, 1, 1, 1};
    CUresult r = ((Fn)fn)(&P.tm, CU_TENSOR_MAP_DATA_TYPE_UINT64, 4, d + (size_t)G * 16, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
    P.base = d; P.cpk = cpk; P.bc = bc; P.br = br; P.B = B; P.nchunks = nchunks; P.pitch = pitch; P.Ppl = Ppl; P.G = G;
    P.tiles_x = (W + bc - 3) / (bc - 2); P.tiles_y = H / (br - 2);
    P.total = (long long)B * P.tiles_x * P.tiles_y; P.kbn = nchunks / cpk; P.buf_bytes = bytes; P.mode = 0; P.nlw = 4; P.pfd = 4;
    const int stage = cpk * br * bc * 16;
    for (int var : {0, 1})                       // 1: with the epilogue's store stream beside the loads
    for (int nprod : {1})
    for (int S : {3, 5}) {
      if (1024 + S * stage > 227 * 1024) continue;
      P.S = S; P.out = var ? dout : nullptr; P.write_bytes = write_bytes * ((br - 2) / 4); P.pfd = 0; P.nlw = nprod;
      float best = 1e9;
      for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0);
        k_pull<<<148, 448, 1024 + S * stage>>>(P);
        cudaEventRecord(e1);
        CK(cudaDeviceSynchronize());
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
      }
      const double rd = (double)P.total * P.kbn * stage;
      const double outpix = (double)B * H * W;
      printf("writes %d nprod %d box %3d pos x %2d rows x %d chunks (%5d B) S %d: %.3f ms  smem-fill %.2f TB/s (%.1f B/clk/SM @1.9GHz)  %.0f clk per 128 output pixels per SM\n",
             var, nprod, bc, br, cpk, stage, S, best, rd / best / 1e9, rd / (best * 1e-3) / 148 / 1.9e9, best * 1e-3 * 1.9e9 * 148 / (outpix / 128));
    }
  }
  return 0;
}
