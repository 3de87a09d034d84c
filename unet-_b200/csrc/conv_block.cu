// conv_block.cu -- the first encoder block as ONE kernel: ConvBlock.forward of conv0_0 (src/models/unetpp.py:23-26:
// conv1+bn1+relu -> conv2+bn2+relu) followed by MaxPool2d(2,2) (unetpp.py:75,105).
//
// Why: at 512x512 the two layers are HBM passes, not tensor work (profiles/r02_conv.md: conv0_0.conv1 moves 1.56 GB,
// conv0_0.conv2 2.38 GB per 64 frames, 14 / 26 % tensor-pipe use): the 32-channel intermediate (1.07 GB) is written and
// read straight back.  Here it never leaves the SM: the first conv's accumulators go TMEM -> registers (x-tap combine,
// bias, ReLU, 16-bit pack) -> shared memory IN THE SECOND CONV'S A-OPERAND LAYOUT, and the second conv's MMAs read them
// there.  Same arithmetic in the same order as the two k_conv_rows launches it replaces (the output is bit-identical,
// tests/test_gpu_net.py::test_fused_block_equals_two_layers).
//
// Geometry (both convs use the k_conv_rows scheme: M-tile = 4 image rows x 32 columns, x-taps folded into N' = 96, an
// output column needs the partial sums of its two lane neighbours).  A work item is 16 output rows x 28 output columns:
//   second conv: 4 M-tiles; lane l <-> image column x = 28*tx - 2 + l, lanes 2..29 valid, needs the intermediate at
//                lanes 1..30 and rows y0-1 .. y0+16 (18 rows)
//   first conv:  5 M-tiles over rows y0-1 .. y0+18 (the last two unused), same lane <-> column map, lanes 1..30 valid,
//                fed by ONE TMA box of the network input: 22 rows x 32 columns x 2 chunks (zero fill outside the plane)
// Recompute: 20/16 rows of the (cheap, K = 16) first conv, 32/28 columns of both.
//
// Warp roles (608 threads): warps 0-7 first-conv epilogue (two warpgroups on alternate M-tiles -> shared memory),
// warps 8-15 second-conv epilogue (two warpgroups on alternate M-tiles -> global, + fused 2x2 max-pool), warp 16 TMA
// producer (+ TMEM alloc, resident weights), warp 17 first-conv MMA issuer, warp 18 second-conv MMA issuer.
// TMEM: four accumulators of 96 columns at stride 128 (0,1 first conv; 2,3 second conv).  The intermediate tile is double
// buffered in shared memory, so the first conv of item i+1 overlaps the second conv of item i.
#include <stdlib.h>
#include <string.h>
#include "tc_ptx.cuh"
#include "tmap.cuh"

namespace {
using namespace tcp;

constexpr int NTHREADS = 19 * 32;
constexpr int G_ROWS = 16, G_COLS = 28;            // output rows / valid columns per work item
constexpr int X_ROWS = 22, T_ROWS = 20;            // staged rows of the input box / of the intermediate tile
constexpr int SEG = 32 * 16;                       // one row of one chunk: 32 positions x 16 B
constexpr int NP = 96;                             // folded N of both convs (3 x 32)
constexpr int X_CHUNK = X_ROWS * SEG, X_STAGE = 2 * X_CHUNK;         // 16 input channels = 2 chunks
constexpr int T_CHUNK = T_ROWS * SEG, T_BUF = 4 * T_CHUNK;           // 32 intermediate channels = 4 chunks
constexpr int W1_KY = 16 * NP * 2, W1_BYTES = 3 * W1_KY;             // [ky][2 chunks][96][8]
constexpr int W2_KY = 32 * NP * 2, W2_BYTES = 3 * W2_KY;             // [ky][4 chunks][96][8]
constexpr int XS = 3;                                                // input stages
constexpr int HDR = 1024, POOL_SMEM = 16 * 1024;
constexpr int OFF_W1 = HDR, OFF_W2 = OFF_W1 + W1_BYTES, OFF_X = OFF_W2 + W2_BYTES, OFF_T = OFF_X + XS * X_STAGE,
              OFF_POOL = (OFF_T + 2 * T_BUF + 1023) & ~1023, SMEM_BYTES = OFF_POOL + POOL_SMEM;
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");

struct BlockParams {
  alignas(64) CUtensorMap tmx;                 // network input (PP, 16 channels): box = 22 rows x 32 columns x 2 chunks
  const act_t* w1; const act_t* w2;            // folded weights of the two convs (ConvLayer::w of layers 0 and 1)
  act_t* out; act_t* pool_out;                 // block output (PP, 32 channels) and its 2x2 max-pool
  int B, tiles_x, tiles_y;
  long long total_items;
  PPLayout L, Lp;
  float bias1[32], bias2[32];                  // folded BN biases, read through the constant bank
};

__global__ void __launch_bounds__(NTHREADS, 1) k_conv_block(const __grid_constant__ BlockParams P) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sb = smem_u32(smem);
  // barriers (8 B each): xfull[3] xempty[3] | t1full[2] t1empty[2] | t0full[2] t0empty[2] | t2full[2] t2empty[2] | wbar
  const uint32_t bar_xfull = sb, bar_xempty = sb + 24, bar_t1full = sb + 48, bar_t1empty = sb + 64, bar_t0full = sb + 80,
                 bar_t0empty = sb + 96, bar_t2full = sb + 112, bar_t2empty = sb + 128, bar_w = sb + 144;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 160);
  const PPLayout& L = P.L;
  const int tiles_per_img = P.tiles_x * P.tiles_y;

  if (threadIdx.x == 0) {
    for (int i = 0; i < XS; ++i) { mbar_init(bar_xfull + 8 * i, 1); mbar_init(bar_xempty + 8 * i, 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_t1full + 8 * i, 1); mbar_init(bar_t1empty + 8 * i, 4);
      mbar_init(bar_t0full + 8 * i, 20); mbar_init(bar_t0empty + 8 * i, 1);      // 5 M-tiles x 4 warps fill one intermediate tile
      mbar_init(bar_t2full + 8 * i, 1); mbar_init(bar_t2empty + 8 * i, 4);
    }
    mbar_init(bar_w, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 16) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 16) {
    // ===================== producer: resident weights once, then one input box per work item =====================
    if (lane == 0) {
      tma_prefetch_desc(&P.tmx);
      mbar_expect_tx(bar_w, W1_BYTES + W2_BYTES);
      bulk_g2s(sb + OFF_W1, P.w1, W1_BYTES, bar_w);
      bulk_g2s(sb + OFF_W2, P.w2, W2_BYTES, bar_w);
      int s = 0; uint32_t ph = 0;
      for (long long grp = blockIdx.x; grp < P.total_items; grp += gridDim.x) {
        const int b = (int)(grp / tiles_per_img), t = (int)(grp - (long long)b * tiles_per_img);
        const int ty = t / P.tiles_x, tx = t - ty * P.tiles_x;
        mbar_wait(bar_xempty + 8 * s, ph ^ 1);
        mbar_expect_tx(bar_xfull + 8 * s, X_STAGE);
        // box origin: image row 16*ty - 2, image column 28*tx - 2 (padded coordinates + 1); outside the plane -> zeros
        tma_load_4d(sb + OFF_X + s * X_STAGE, &P.tmx, (tx * G_COLS - 1) * 8, ty * G_ROWS - 1, 0, b, bar_xfull + 8 * s);
        if (++s == XS) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 17) {
    // ===================== first-conv MMA issuer: 5 M-tiles x 3 tap rows, K = 16 =====================
    constexpr uint32_t IDESC = (1u << 4) | (UB_MMA_FMT << 7) | (UB_MMA_FMT << 10) | ((uint32_t)(NP >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    mbar_wait(bar_w, 0);
    int s = 0; uint32_t ph = 0; int n1 = 0;
    for (long long grp = blockIdx.x; grp < P.total_items; grp += gridDim.x) {
      mbar_wait(bar_xfull + 8 * s, ph);
      tc_fence_after();
      const uint32_t x_st = sb + OFF_X + s * X_STAGE;
      for (int t = 0; t < 5; ++t, ++n1) {
        const int buf = n1 & 1;
        mbar_wait(bar_t1empty + 8 * buf, ((n1 >> 1) & 1) ^ 1);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int ky = 0; ky < 3; ++ky) {
            const uint64_t ad = make_desc(x_st + (4 * t + ky) * SEG, X_CHUNK, 128);
            const uint64_t bd = make_desc(sb + OFF_W1 + ky * W1_KY, NP * 16, 128);
            tc_mma(tmem_base + buf * 128, ad, bd, IDESC, ky != 0);
          }
          tc_commit(bar_t1full + 8 * buf);
          if (t == 4) tc_commit(bar_xempty + 8 * s);           // the box is free once these MMAs have read it
        }
        __syncwarp();
      }
      if (++s == XS) { s = 0; ph ^= 1; }
    }
  } else if (warp == 18) {
    // ===================== second-conv MMA issuer: 4 M-tiles x 3 tap rows x 2 K steps =====================
    constexpr uint32_t IDESC = (1u << 4) | (UB_MMA_FMT << 7) | (UB_MMA_FMT << 10) | ((uint32_t)(NP >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    mbar_wait(bar_w, 0);
    int g = 0, n2 = 0;
    for (long long grp = blockIdx.x; grp < P.total_items; grp += gridDim.x, ++g) {
      const int tb = g & 1;
      mbar_wait(bar_t0full + 8 * tb, (g >> 1) & 1);            // the intermediate tile is complete in shared memory
      tc_fence_after();
      const uint32_t t_st = sb + OFF_T + tb * T_BUF;
      for (int m = 0; m < 4; ++m, ++n2) {
        const int buf = n2 & 1;
        mbar_wait(bar_t2empty + 8 * buf, ((n2 >> 1) & 1) ^ 1);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int k16 = 0; k16 < 2; ++k16) {
              const uint64_t ad = make_desc(t_st + 2 * k16 * T_CHUNK + (4 * m + ky) * SEG, T_CHUNK, 128);
              const uint64_t bd = make_desc(sb + OFF_W2 + ky * W2_KY + 2 * k16 * NP * 16, NP * 16, 128);
              tc_mma(tmem_base + (2 + buf) * 128, ad, bd, IDESC, (ky | k16) != 0);
            }
          tc_commit(bar_t2full + 8 * buf);
          if (m == 3) tc_commit(bar_t0empty + 8 * tb);
        }
        __syncwarp();
      }
    }
  } else {
    const int wq = warp & 3;                                   // TMEM lane quarter = row inside the M-tile
    const int wg = (warp >> 2) & 1;                            // which of the two warpgroups of this role
    const bool second = warp >= 8;
    // (b, ty, tx) of this CTA's work items, advanced by gridDim.x per item without divisions in the loop
    const int step_b = (int)gridDim.x / tiles_per_img, step_r = (int)gridDim.x - step_b * tiles_per_img;
    const int step_y = step_r / P.tiles_x, step_x = step_r - step_y * P.tiles_x;
    int b = (int)blockIdx.x / tiles_per_img, ty, tx;
    {
      const int t = (int)blockIdx.x - b * tiles_per_img;
      ty = t / P.tiles_x; tx = t - ty * P.tiles_x;
    }
    auto advance = [&]() {
      tx += step_x;
      if (tx >= P.tiles_x) { tx -= P.tiles_x; ++ty; }
      ty += step_y;
      if (ty >= P.tiles_y) { ty -= P.tiles_y; ++b; }
      b += step_b;
    };
    if (!second) {
      // ===================== first-conv epilogue: TMEM -> bias, ReLU, pack -> the second conv's A operand =====================
      int g = 0, n1 = 0;
      for (long long grp = blockIdx.x; grp < P.total_items; grp += gridDim.x, ++g) {
        const int tb = g & 1;
        const int x = tx * G_COLS - 2 + lane;
        const bool col_ok = lane >= 1 && lane <= 30 && x >= 0 && x < L.W;
        const uint32_t dst0 = sb + OFF_T + tb * T_BUF + lane * 16;
        for (int t = 0; t < 5; ++t, ++n1) {
          if ((n1 & 1) != wg) continue;
          const int buf = wg, use = n1 >> 1;
          const int i = 4 * t + wq;                            // row of the intermediate tile; image row y0 - 1 + i
          const int y = ty * G_ROWS - 1 + i;
          mbar_wait(bar_t1full + 8 * buf, use & 1);
          tc_fence_after();
          // the tile buffer was last read by the second conv of item g-2
          mbar_wait(bar_t0empty + 8 * tb, ((g >> 1) & 1) ^ 1);
          const bool keep = col_ok && y >= 0 && y < L.H;       // outside the image the intermediate is the conv's zero padding
          const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + buf * 128;
          if (i < 18) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
              const int c0 = half * 16;
              uint32_t k0[16], k1[16], k2[16];
              TMEM_LD16(taddr + c0, k0);
              TMEM_LD16(taddr + 32 + c0, k1);
              TMEM_LD16(taddr + 64 + c0, k2);
              tmem_ld_wait();
              if (half == 1) {                                 // accumulator fully read: hand it back to the issuer
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_t1empty + 8 * buf);
              }
              uint32_t hv[8];
#pragma unroll
              for (int j = 0; j < 16; j += 2) {
                const uint32_t l0 = __shfl_up_sync(0xffffffffu, k0[j], 1), l1 = __shfl_up_sync(0xffffffffu, k0[j + 1], 1);
                const uint32_t r0 = __shfl_down_sync(0xffffffffu, k2[j], 1), r1 = __shfl_down_sync(0xffffffffu, k2[j + 1], 1);
                uint64_t v = add2(pack2(l0, l1), pack2(k1[j], k1[j + 1]));
                v = add2(v, pack2(r0, r1));
                v = add2(v, pack2(__float_as_uint(P.bias1[c0 + j]), __float_as_uint(P.bias1[c0 + j + 1])));
                hv[j >> 1] = keep ? relu_pack(v) : 0u;
              }
              sts_v4(dst0 + (2 * half) * T_CHUNK + i * SEG, hv[0], hv[1], hv[2], hv[3]);
              sts_v4(dst0 + (2 * half + 1) * T_CHUNK + i * SEG, hv[4], hv[5], hv[6], hv[7]);
            }
          } else {                                             // rows 18, 19 of the last M-tile are never read
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_t1empty + 8 * buf);
          }
          fence_proxy_async_smem();                            // generic-proxy writes -> visible to the tensor core's reads
          __syncwarp();
          if (lane == 0) mbar_arrive(bar_t0full + 8 * tb);
        }
        advance();
      }
    } else {
      // ===================== second-conv epilogue: TMEM -> bias, ReLU, pack -> global, + 2x2 max-pool =====================
      const size_t plane = (size_t)L.Ppl * 8;
      uint32_t pbuf = 0;
      const int pw = warp - 8;                                 // 0..7
      const uint32_t stage_mine = sb + OFF_POOL + pw * 1024 + (lane >> 1) * 16;
      const uint32_t stage_peer = sb + OFF_POOL + (pw ^ 1) * 1024 + (lane >> 1) * 16;
      int n2 = 0;
      for (long long grp = blockIdx.x; grp < P.total_items; grp += gridDim.x) {
        const int x = tx * G_COLS - 2 + lane;
        const bool valid = lane >= 2 && lane <= 29 && x < L.W;
        const bool edge = (x == -1 && lane == 1) || (x == L.W && lane <= 30);     // left / right zero border of the PP plane
        act_t* const o_grp = P.out + ((size_t)b * 4) * plane + (size_t)(L.G + x + 1) * 8;
        for (int m = 0; m < 4; ++m, ++n2) {
          if ((n2 & 1) != wg) continue;
          const int buf = wg, use = n2 >> 1;
          const int y = ty * G_ROWS + 4 * m + wq;
          const int q = (y + 1) * L.pitch;
          mbar_wait(bar_t2full + 8 * buf, use & 1);
          tc_fence_after();
          const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + (2 + buf) * 128;
          uint32_t mine[8];
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            const int c0 = half * 16;
            uint32_t k0[16], k1[16], k2[16];
            TMEM_LD16(taddr + c0, k0);
            TMEM_LD16(taddr + 32 + c0, k1);
            TMEM_LD16(taddr + 64 + c0, k2);
            tmem_ld_wait();
            if (half == 1) {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(bar_t2empty + 8 * buf);
            }
            uint32_t hv[8];
#pragma unroll
            for (int j = 0; j < 16; j += 2) {
              const uint32_t l0 = __shfl_up_sync(0xffffffffu, k0[j], 1), l1 = __shfl_up_sync(0xffffffffu, k0[j + 1], 1);
              const uint32_t r0 = __shfl_down_sync(0xffffffffu, k2[j], 1), r1 = __shfl_down_sync(0xffffffffu, k2[j + 1], 1);
              uint64_t v = add2(pack2(l0, l1), pack2(k1[j], k1[j + 1]));
              v = add2(v, pack2(r0, r1));
              v = add2(v, pack2(__float_as_uint(P.bias2[c0 + j]), __float_as_uint(P.bias2[c0 + j + 1])));
              hv[j >> 1] = valid ? relu_pack(v) : 0u;
            }
            if (valid || edge) {
              act_t* o = o_grp + (size_t)(half * 2) * plane + (size_t)q * 8;
              *reinterpret_cast<uint4*>(o) = make_uint4(hv[0], hv[1], hv[2], hv[3]);
              *reinterpret_cast<uint4*>(o + plane) = make_uint4(hv[4], hv[5], hv[6], hv[7]);
              if (y == 0 || y == L.H - 1) {                    // top / bottom zero border rows of the plane
                act_t* z = y == 0 ? o - (size_t)L.pitch * 8 : o + (size_t)L.pitch * 8;
                *reinterpret_cast<uint4*>(z) = make_uint4(0, 0, 0, 0);
                *reinterpret_cast<uint4*>(z + plane) = make_uint4(0, 0, 0, 0);
              }
            }
            // column pairs (x, x+1), x even = even lanes; activations are >= 0, so the 16-bit integer maximum is the maximum
            uint32_t cp[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) cp[j] = __vmaxu2(hv[j], __shfl_down_sync(0xffffffffu, hv[j], 1));
            if (half == (wq & 1)) {                            // even tile rows keep channels 0-15, odd rows 16-31
#pragma unroll
              for (int j = 0; j < 8; ++j) mine[j] = cp[j];
            } else if (!(lane & 1)) {
              st_shared_v4(stage_mine + pbuf * 512, cp[0], cp[1], cp[2], cp[3]);
              st_shared_v4(stage_mine + pbuf * 512 + 256, cp[4], cp[5], cp[6], cp[7]);
            }
          }
          asm volatile("bar.sync %0, 64;" ::"r"(1 + (pw >> 1)) : "memory");        // the two warps of rows 2r / 2r+1
          if (!(lane & 1) && valid) {
            uint32_t o0, o1, o2, o3, o4, o5, o6, o7;
            ld_shared_v4(stage_peer + pbuf * 512, o0, o1, o2, o3);
            ld_shared_v4(stage_peer + pbuf * 512 + 256, o4, o5, o6, o7);
            const int py = y >> 1, px = x >> 1;
            act_t* po = P.pool_out + (((size_t)b * 4 + (wq & 1) * 2) * P.Lp.Ppl + P.Lp.G + (size_t)(py + 1) * P.Lp.pitch + px + 1) * 8;
            *reinterpret_cast<uint4*>(po) = make_uint4(__vmaxu2(mine[0], o0), __vmaxu2(mine[1], o1), __vmaxu2(mine[2], o2), __vmaxu2(mine[3], o3));
            *reinterpret_cast<uint4*>(po + (size_t)P.Lp.Ppl * 8) =
                make_uint4(__vmaxu2(mine[4], o4), __vmaxu2(mine[5], o5), __vmaxu2(mine[6], o6), __vmaxu2(mine[7], o7));
          }
          pbuf ^= 1;
        }
        advance();
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 16) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512));
  }
}

}  // namespace

int conv_block_init(ub_ctx* ctx) {
  UB_CUDA(ctx, cudaFuncSetAttribute(k_conv_block, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
  return UB_OK;
}

// The shapes the kernel is written for: conv0_0 = 16(3) -> 32 -> 32 folded layers, rows in groups of 16, even width.
bool conv_block_can_fuse(const ub_ctx* ctx, const ConvLayer& l1, const ConvLayer& l2, int H, int W) {
  return ctx->fuse_block0 && ctx->conv_impl == 0 && l1.fold && l2.fold && l1.cin0 == 16 && l1.cin1 == 0 && l1.kc == 16 &&
         l1.cout == 32 && l2.cin0 == 32 && l2.cin1 == 0 && l2.kc == 32 && l2.cout == 32 && H % G_ROWS == 0 && !(W & 1);
}

int conv_block0(ub_ctx* ctx, const ConvLayer& l1, const ConvLayer& l2, const act_t* src, act_t* out, act_t* pool_out,
                int B, int H, int W, cudaStream_t st) {
  BlockParams P{};
  P.w1 = l1.w; P.w2 = l2.w; P.out = out; P.pool_out = pool_out;
  memcpy(P.bias1, l1.bias_host, sizeof P.bias1);
  memcpy(P.bias2, l2.bias_host, sizeof P.bias2);
  P.B = B; P.L = pp_layout(H, W); P.Lp = pp_layout(H / 2, W / 2);
  P.tiles_x = cdiv(W, G_COLS); P.tiles_y = H / G_ROWS;
  P.total_items = (long long)B * P.tiles_x * P.tiles_y;
  int r = tmap_pp_rows(ctx, &P.tmx, src, B, 2, P.L, 32, X_ROWS, 2);
  if (r != UB_OK) return r;
  TimeScope ts(ctx, UB_T_CONV, st);
  const long long grid = P.total_items < ctx->sm_count ? P.total_items : ctx->sm_count;
  k_conv_block<<<(unsigned)grid, NTHREADS, SMEM_BYTES, st>>>(P);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}
