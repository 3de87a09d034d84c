// conv_block.cu -- the first encoder block as ONE kernel: ConvBlock.forward of conv0_0 (src/models/unetpp.py:23-26:
// conv1+bn1+relu -> conv2+bn2+relu) followed by MaxPool2d(2,2) (unetpp.py:75,105).
//
// Why: at 512x512 the two layers write and read back a 32-channel intermediate (1.07 GB per 64 frames) and -- what
// actually paces them -- spend their time in the epilogue's x-tap combine: the folded-N scheme of k_conv_rows costs 64
// SHFL per warp and M-tile and the SM retires one SHFL per clock (tools/ubench_shfl.cu, profiles/r02_conv.md).  Here
//  * the intermediate never leaves the SM: the first conv's accumulators go TMEM -> registers (bias, ReLU, 16-bit pack)
//    -> shared memory IN THE SECOND CONV'S A-OPERAND LAYOUT, and the second conv's MMAs read them there;
//  * the first conv has no x-tap combine at all: its 3 input channels leave 13 of the 16 K slots of an MMA empty, so the
//    network input is written in "x-im2col" form (channel k = kx*3 + c holds colour c of the pixel at x + kx - 1;
//    k_preprocess<..., XI> / k_nchw_to_xi) and the three x-taps ride in K.  One N = 32 MMA per tap row, every lane valid.
//
// Geometry.  A work item is 16 output rows x 30 output columns; lane l <-> image column x = 30*tx - 1 + l.
//   first conv:  5 M-tiles (4 rows x 32 columns) over rows y0-1 .. y0+18 (the last two unused), fed by ONE TMA box of the
//                input: 22 rows x 32 columns x 2 chunks (zero fill outside the plane)
//   second conv: 4 M-tiles, k_conv_rows scheme (x-taps folded into N' = 96, an output needs the partial sums of its lane
//                neighbours: lanes 1..30 valid), A operand = the intermediate tile in shared memory (20 rows x 32 x 32 ch)
// Recompute: 20/16 rows of the (cheap, K = 16) first conv, 32/30 columns.
//
// Warp roles (608 threads): warps 0-3 first-conv epilogue (-> shared memory), warps 4-15 second-conv epilogue (three
// warpgroups on M-tiles round robin -> global, + fused 2x2 max-pool), warp 16 TMA producer (+ TMEM alloc, resident
// weights), warp 17 first-conv MMA issuer, warp 18 second-conv MMA issuer.  TMEM: three second-conv accumulators of 96
// columns, two first-conv accumulators of 32.  The intermediate tile is double buffered, so the first conv of item i+1
// overlaps the second conv of item i.
#include <stdlib.h>
#include <string.h>
#include "tc_ptx.cuh"
#include "tmap.cuh"

namespace {
using namespace tcp;

constexpr int NTHREADS = 19 * 32;
constexpr int G_ROWS = 16, G_COLS = 30;            // output rows / valid columns per work item
constexpr int X_ROWS = 22, T_ROWS = 20;            // staged rows of the input box / of the intermediate tile
constexpr int SEG = 32 * 16;                       // one row of one chunk: 32 positions x 16 B
constexpr int NP = 96;                             // folded N of the second conv (3 x 32)
constexpr int X_CHUNK = X_ROWS * SEG, X_STAGE = 2 * X_CHUNK;         // 16 input channels = 2 chunks
constexpr int T_CHUNK = T_ROWS * SEG, T_BUF = 4 * T_CHUNK;           // 32 intermediate channels = 4 chunks
constexpr int W1_KY = 16 * 32 * 2, W1_BYTES = 3 * W1_KY;             // [ky][2 chunks][32][8], k = kx*3 + c
constexpr int W2_KY = 32 * NP * 2, W2_BYTES = 3 * W2_KY;             // [ky][4 chunks][96][8]
constexpr int XS = 3;                                                // input stages
constexpr int HDR = 1024, POOL_SMEM = 12 * 1024;
constexpr int TM_C2 = 0, TM_C1 = 384;              // TMEM columns: second-conv accumulators at 0/128/256, first-conv at 384/416
constexpr int OFF_W1 = HDR, OFF_W2 = OFF_W1 + W1_BYTES, OFF_X = OFF_W2 + W2_BYTES, OFF_T = OFF_X + XS * X_STAGE,
              OFF_POOL = (OFF_T + 2 * T_BUF + 1023) & ~1023, SMEM_BYTES = OFF_POOL + POOL_SMEM;
static_assert(SMEM_BYTES <= 227 * 1024, "shared memory budget");

struct BlockParams {
  alignas(64) CUtensorMap tmx;                 // network input (PP, x-im2col, 16 channels): box = 22 rows x 32 columns x 2 chunks
  const act_t* w1; const act_t* w2;            // ConvLayer::wx of layer 0 (x-taps in K), ConvLayer::w of layer 1 (x-taps in N)
  act_t* out; act_t* pool_out;                 // block output (PP, 32 channels) and its 2x2 max-pool
  int B, tiles_x, tiles_y;
  long long total_items;
  PPLayout L, Lp;
  unsigned long long* sat;                     // non-null: count outputs (of both convs) that saturate the 16-bit activation type
  float bias1[32], bias2[32];                  // folded BN biases, read through the constant bank
};

// outputs of a packed row that hit the largest finite 16-bit value (diagnostic of ub_set_option "check_saturation")
__device__ __forceinline__ void count_saturated(unsigned long long* sat, const uint32_t* hv, int n, int lane) {
#ifdef UB_ACT_BF16
  constexpr uint32_t MAXW = 0x7F7F7F7Fu;
#else
  constexpr uint32_t MAXW = 0x7BFF7BFFu;
#endif
  int cnt = 0;
  for (int j = 0; j < n; ++j) cnt += __popc(__vcmpeq2(hv[j], MAXW) & 0x00010001u);
  cnt = __reduce_add_sync(0xffffffffu, cnt);
  if (lane == 0 && cnt) atomicAdd(sat, (unsigned long long)cnt);
}

__global__ void __launch_bounds__(NTHREADS, 1) k_conv_block(const __grid_constant__ BlockParams P) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t sb = smem_u32(smem);
  // barriers (8 B each): xfull[3] xempty[3] | t1full[2] t1empty[2] | t0full[2] t0empty[2] | t2full[3] t2empty[3] | wbar
  const uint32_t bar_xfull = sb, bar_xempty = sb + 24, bar_t1full = sb + 48, bar_t1empty = sb + 64, bar_t0full = sb + 80,
                 bar_t0empty = sb + 96, bar_t2full = sb + 112, bar_t2empty = sb + 136, bar_w = sb + 160;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 176);
  const PPLayout& L = P.L;
  const int tiles_per_img = P.tiles_x * P.tiles_y;

  if (threadIdx.x == 0) {
    for (int i = 0; i < XS; ++i) { mbar_init(bar_xfull + 8 * i, 1); mbar_init(bar_xempty + 8 * i, 1); }
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_t1full + 8 * i, 1); mbar_init(bar_t1empty + 8 * i, 4);
      mbar_init(bar_t0full + 8 * i, 20); mbar_init(bar_t0empty + 8 * i, 1);      // 5 M-tiles x 4 warps fill one intermediate tile
    }
    for (int i = 0; i < 3; ++i) { mbar_init(bar_t2full + 8 * i, 1); mbar_init(bar_t2empty + 8 * i, 4); }
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
        // box origin: image row 16*ty - 2, image column 30*tx - 1 (padded coordinates + 1); outside the plane -> zeros
        tma_load_4d(sb + OFF_X + s * X_STAGE, &P.tmx, tx * G_COLS * 8, ty * G_ROWS - 1, 0, b, bar_xfull + 8 * s);
        if (++s == XS) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 17) {
    // ===================== first-conv MMA issuer: 5 M-tiles x 3 tap rows, K = 16 (9 used: 3 x-taps x 3 colours), N = 32 =====================
    constexpr uint32_t IDESC = (1u << 4) | (UB_MMA_FMT << 7) | (UB_MMA_FMT << 10) | ((uint32_t)(32 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
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
            const uint64_t bd = make_desc(sb + OFF_W1 + ky * W1_KY, 32 * 16, 128);
            tc_mma(tmem_base + TM_C1 + buf * 32, ad, bd, IDESC, ky != 0);
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
    int g = 0, buf = 0; uint32_t use = 0;                      // accumulator of M-tile n2: buffer n2 % 3, use n2 / 3
    for (long long grp = blockIdx.x; grp < P.total_items; grp += gridDim.x, ++g) {
      const int tb = g & 1;
      mbar_wait(bar_t0full + 8 * tb, (g >> 1) & 1);            // the intermediate tile is complete in shared memory
      tc_fence_after();
      const uint32_t t_st = sb + OFF_T + tb * T_BUF;
      for (int m = 0; m < 4; ++m) {
        mbar_wait(bar_t2empty + 8 * buf, (use & 1) ^ 1);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int k16 = 0; k16 < 2; ++k16) {
              const uint64_t ad = make_desc(t_st + 2 * k16 * T_CHUNK + (4 * m + ky) * SEG, T_CHUNK, 128);
              const uint64_t bd = make_desc(sb + OFF_W2 + ky * W2_KY + 2 * k16 * NP * 16, NP * 16, 128);
              tc_mma(tmem_base + TM_C2 + buf * 128, ad, bd, IDESC, (ky | k16) != 0);
            }
          tc_commit(bar_t2full + 8 * buf);
          if (m == 3) tc_commit(bar_t0empty + 8 * tb);
        }
        __syncwarp();
        if (++buf == 3) { buf = 0; ++use; }
      }
    }
  } else {
    const int wq = warp & 3;                                   // TMEM lane quarter = row inside the M-tile
    const bool second = warp >= 4;
    const int wg = (warp >> 2) - 1;                            // second-conv epilogue: which of its three warpgroups
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
        const int x = tx * G_COLS - 1 + lane;
        const bool col_ok = x >= 0 && x < L.W;
        const uint32_t dst0 = sb + OFF_T + tb * T_BUF + lane * 16;
        // the tile buffer was last read by the second conv of item g-2
        mbar_wait(bar_t0empty + 8 * tb, ((g >> 1) & 1) ^ 1);
        for (int t = 0; t < 5; ++t, ++n1) {
          const int buf = n1 & 1, use = n1 >> 1;
          const int i = 4 * t + wq;                            // row of the intermediate tile; image row y0 - 1 + i
          const int y = ty * G_ROWS - 1 + i;
          mbar_wait(bar_t1full + 8 * buf, use & 1);
          tc_fence_after();
          const bool keep = col_ok && y >= 0 && y < L.H;       // outside the image the intermediate is the conv's zero padding
          const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + TM_C1 + buf * 32;
          if (i < 18) {
            uint32_t k[32];
            TMEM_LD32(taddr, k);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_t1empty + 8 * buf); // accumulator read: hand it back to the issuer
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              uint32_t hv[4];
#pragma unroll
              for (int j = 0; j < 8; j += 2) {
                const uint64_t v = add2(pack2(k[8 * c + j], k[8 * c + j + 1]),
                                        pack2(__float_as_uint(P.bias1[8 * c + j]), __float_as_uint(P.bias1[8 * c + j + 1])));
                hv[j >> 1] = keep ? relu_pack(v) : 0u;
              }
              sts_v4(dst0 + c * T_CHUNK + i * SEG, hv[0], hv[1], hv[2], hv[3]);
              if (P.sat != nullptr) count_saturated(P.sat, hv, 4, lane);
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
      const int pw = warp - 4;                                 // 0..11
      const uint32_t stage_mine = sb + OFF_POOL + pw * 1024 + (lane >> 1) * 16;
      const uint32_t stage_peer = sb + OFF_POOL + (pw ^ 1) * 1024 + (lane >> 1) * 16;
      int turn = 0; uint32_t use = 0;                          // M-tile n2 belongs to warpgroup n2 % 3 (= its accumulator), use n2 / 3
      for (long long grp = blockIdx.x; grp < P.total_items; grp += gridDim.x) {
        const int x = tx * G_COLS - 1 + lane;
        const bool valid = lane >= 1 && lane <= G_COLS && x < L.W;
        const bool edge = x == -1 || x == L.W;                 // left / right zero border of the PP plane
        act_t* const o_grp = P.out + ((size_t)b * 4) * plane + (size_t)(L.G + x + 1) * 8;
        for (int m = 0; m < 4; ++m) {
          const bool mine_turn = turn == wg;
          const uint32_t use_now = use;
          if (++turn == 3) { turn = 0; ++use; }
          if (!mine_turn) continue;
          const int buf = wg;
          const int y = ty * G_ROWS + 4 * m + wq;
          const int q = (y + 1) * L.pitch;
          mbar_wait(bar_t2full + 8 * buf, use_now & 1);
          tc_fence_after();
          const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + TM_C2 + buf * 128;
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
            if (P.sat != nullptr) count_saturated(P.sat, hv, 8, lane);
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
            // column pairs (x, x+1), x even = odd lanes; activations are >= 0, so the 16-bit integer maximum is the maximum
            uint32_t cp[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) cp[j] = __vmaxu2(hv[j], __shfl_down_sync(0xffffffffu, hv[j], 1));
            if (half == (wq & 1)) {                            // even tile rows keep channels 0-15, odd rows 16-31
#pragma unroll
              for (int j = 0; j < 8; ++j) mine[j] = cp[j];
            } else if (lane & 1) {
              st_shared_v4(stage_mine + pbuf * 512, cp[0], cp[1], cp[2], cp[3]);
              st_shared_v4(stage_mine + pbuf * 512 + 256, cp[4], cp[5], cp[6], cp[7]);
            }
          }
          asm volatile("bar.sync %0, 64;" ::"r"(1 + (pw >> 1)) : "memory");        // the two warps of rows 2r / 2r+1
          if ((lane & 1) && valid) {
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

// The shapes the kernel is written for: conv0_0 = 3 -> 32 -> 32, rows in groups of 16, even width.
bool conv_block_can_fuse(const ub_ctx* ctx, const ConvLayer& l1, const ConvLayer& l2, int H, int W) {
  return ctx->fuse_block0 && ctx->conv_impl == 0 && l1.wx && l2.fold && l1.cin0 == 16 && l1.cin1 == 0 && l1.kc == 16 &&
         l1.cout == 32 && l2.cin0 == 32 && l2.cin1 == 0 && l2.kc == 32 && l2.cout == 32 && H % G_ROWS == 0 && !(W & 1);
}

int conv_block0(ub_ctx* ctx, const ConvLayer& l1, const ConvLayer& l2, const act_t* src, act_t* out, act_t* pool_out,
                int B, int H, int W, unsigned long long* sat, cudaStream_t st) {
  BlockParams P{};
  P.sat = sat;
  P.w1 = l1.wx; P.w2 = l2.w; P.out = out; P.pool_out = pool_out;
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
