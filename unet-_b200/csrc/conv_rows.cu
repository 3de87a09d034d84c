// conv_rows.cu -- 3x3 conv (+folded BN bias, ReLU, optional fused 1x1 head + argmax) for the
// two high-resolution pyramid levels (Cout = 32 / 64), the layers that hold the pixels.
//
// Replaces ConvBlock.forward's conv+bn+relu (src/models/unetpp.py:23-26) and, for the last
// layer, `final` (unetpp.py:119) + softmax/argmax (infer_two_stage_burr.py:299-300).
//
// GEMM view.  The kernel's x-taps are folded into the GEMM N dimension (N' = 3*Cout, so one
// activation tile feeds all three x-taps and the shared-memory operand path is not the limit:
// 128 B/clk feeds N' = 96 at 86 % and N' = 192 at 100 % of the tensor pipe, N = 32 only 40 %):
//   D[(r,l), kx*Cout+co] = sum_{ky,ci} X[y0+r+ky-1, x0-1+l, ci] * W[co,ky,kx,ci]
//   out[y0+r, x0-1+l, co] = D[(r,l-1), 0*Cout+co] + D[(r,l), 1*Cout+co] + D[(r,l+1), 2*Cout+co]
// One work item is a 2-D tile of 4 image rows x 32 columns (M = 128: TMEM lane = r*32 + l, so
// warp = image row, lane = column).  The column neighbours l-1 / l+1 are lanes of the same
// warp -> two shuffles, no cross-warp exchange; columns 1..30 of a tile are valid and tiles
// advance by 30 columns.  The six input rows y0-1..y0+4 of all KC/8 chunks are staged ONCE per
// k-block by a single TMA box (cp.async.bulk.tensor -> UTMALDG; per-lane bulk copies would be
// serialised at ~100 clk each, profiles/r01_conv.md); the A operand of tap row ky is segments
// ky..ky+3, i.e. a +512 B descriptor offset -- half the L2->shared traffic of staging three
// shifted copies.  Weights stay resident in shared memory when they fit.
//
// Warp roles (608 threads): warps 0-15 epilogue, warp 16 bulk-copy producer (+TMEM alloc),
// warps 17-18 MMA issuers (alternate M-tiles, one elected lane each).  TMEM holds four accumulators of 96 columns (Cout 32,
// one per epilogue warpgroup) or two of 192 columns (Cout 64, two warpgroups per accumulator,
// 32 channels each), so the MMAs of tiles i+1.. overlap the epilogue of tile i.
// Epilogue arithmetic is packed (FADD2, F2FP.RELU.SATFINITE.PACK) -- it is the instruction-
// bound part of these layers (profiles/r01_conv.md).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "tc_ptx.cuh"
#include "tmap.cuh"

namespace {
using namespace tcp;

constexpr int T_ROWS = 4, T_COLS = 32, T_VALID = 30, N_SEG = T_ROWS + 2;
constexpr int SEG_BYTES = T_COLS * 16;            // one staged row segment of one 8-channel chunk

struct RowsParams {
  alignas(64) CUtensorMap tm0;                 // source 0 (skip) and source 1 (upsampled): tmap_pp_rows views,
  alignas(64) CUtensorMap tm1;                 // box = 6 rows x 32 columns x KC/8 chunks
  act_t* out;
  const act_t* w; const float* bias;
  int B, kb0, kbn, num_stages, b_resident;
  int tiles_x, tiles_y;                        // tiles_y counts tile GROUPS: mt vertically adjacent 4-row M-tiles share one box
  int mt, a_chunk, a_stage;                    // M-tiles per group; bytes of one chunk / one stage of the box ((4*mt+2) rows)
  long long total_items;
  PPLayout L;
  int ncls; float* logits; uint8_t* pred;      // fused 1x1 head (ncls > 0), COUT == 32
  act_t* pool_out; PPLayout Lp; int pool_smem;  // fused MaxPool2d(2,2) (POOL kernels): pooled PP plane, its layout, staging offset
  // fused Upsample(x2, bilinear, align_corners) of source 1 (UPS kernels): tm1 views the HALF-resolution tensor, box =
  // low_rows x UP_LC columns x KC/8 chunks, staged in a ring of low_slots boxes at shared-memory offset low_off
  float up_sy, up_sx; int low_rows, low_chunk, low_stage, low_slots, low_off;
  unsigned long long* sat;                     // non-null: count outputs that saturate the 16-bit activation type
  int wbytes;                                  // CG2 kernels: bytes of this CTA's half of the weights (all k-blocks, resident)
  float head[8 * 33];                          // [ncls][32] weights then [ncls] biases
  float biasv[64];                             // folded BN bias per output channel: read through the constant bank (the
                                               // shared-memory pipe is what bounds these kernels, profiles/r02_conv.md)
#ifdef UB_ROWS_TRACE
  long long* trace;                            // [8 roles][TR_N][8] clock64 stamps of block 0 (tools/rows_trace.py)
#endif
};

// Pipeline trace (debug builds only, tools/build_trace_lib.sh): block 0 stamps clock64 at the hand-over points of every
// role.  Roles: 0 producer (index = box), 1 MMA issuer (index = M-tile), 2-5 epilogue warpgroups (index = use),
// 6 interpolation warp 0 (index = source-1 stage).
#ifdef UB_ROWS_TRACE
constexpr int TR_N = 1024;
int g_trace_dump = 0;
#define UB_TR(role, idx, ev)                                                                          \
  do {                                                                                                \
    if (blockIdx.x == 0 && (idx) < TR_N) P.trace[(((role) * TR_N) + (idx)) * 8 + (ev)] = clock64();   \
  } while (0)
#else
#define UB_TR(role, idx, ev) do { } while (0)
#endif

constexpr int NTHREADS = 19 * 32;        // 16 epilogue warps, producer, two MMA issuers
constexpr int UP_WARPS = 8;               // UPS kernels: interpolation warps 19..26 (two per 8-channel chunk of a k-block)
constexpr int UP_LC = 18;                 // staged low-resolution columns: 32 output columns touch at most 17 + 1
constexpr int UP_BAR = 9;                 // named barrier of the interpolation warps (1..8 belong to the fused pool)

template <int COUT, int KC, bool CG2 = false>
struct RCfg {
  static constexpr int NP = 3 * COUT;                          // folded N
  static constexpr int NPH = CG2 ? NP / 2 : NP;                // rows of B this CTA holds (a CTA pair splits N between its two CTAs)
  static constexpr int CPK = KC / 8;
  static constexpr int B_KY = KC * NPH * 2;                    // one ky of one k-block
  static constexpr int B_STAGE = 3 * B_KY;
  static constexpr bool SPLIT = COUT == 64;                    // two warpgroups share one accumulator
  static constexpr int NBUF = SPLIT ? 2 : 4;
  static constexpr int T_STRIDE = SPLIT ? 256 : 128;           // TMEM columns per accumulator slot
  static constexpr int WARPS_PER_BUF = SPLIT ? 8 : 4;
  static constexpr int HDR = 1024;                             // barriers, TMEM slot, bias
};

// Interpolation warp of a UPS kernel.  The eight warps form two teams that take alternate source-1 stages (so two stages
// are interpolated concurrently); inside a team a warp = (pair of chunks of the k-block, half of the box rows), lane = box
// column, NR = rows per warp = 2*mt+1.  A thread runs the two chunks as two independent dependency chains (the walk is
// latency-bound: load -> x-blend -> y-blend -> store).  The arithmetic is k_upsample2's: source index =
// dst * (in-1)/(out-1) in fp32, x-blend of every low-resolution row once, one y-blend per output, all as packed 16-bit
// a + l*(c-a).  Per group a thread sets up its column terms and a row table (low-resolution row offset, packed y weight)
// in registers.  Each warp arrives on the stage's barriers by itself (full / lempty count UP_TEAM arrivals per CTA).
constexpr int UP_TEAM = UP_WARPS / 2;     // warps per team = arrivals per stage and CTA
template <int NR, bool CG2>
__device__ __forceinline__ void ups_warp(const RowsParams& P, const int iw, const int lane, const uint32_t bar_full,
                                         const uint32_t bar_empty, const uint32_t bar_lfull, const uint32_t bar_lempty,
                                         const uint32_t a_base, const uint32_t low_base) {
  const int team = iw >> 2, cp = iw & 1, r_begin = ((iw >> 1) & 1) * NR;   // chunks 2*cp, 2*cp+1
  const int S = P.num_stages, LS = P.low_slots, kb0 = P.kb0, kbn = P.kbn, W = P.L.W, H = P.L.H, mtr = T_ROWS * P.mt;
  const int a_stage = P.a_stage, low_stage = P.low_stage, tiles_x = P.tiles_x, tiles_y = P.tiles_y;
  const uint32_t a_chunk = P.a_chunk, low_chunk = P.low_chunk;
  const float sx = P.up_sx, sy = P.up_sy;
  const long long total = P.total_items;
  const int tiles_per_img = tiles_x * tiles_y;
  const int step_b = (int)gridDim.x / tiles_per_img, step_r = (int)gridDim.x - step_b * tiles_per_img;
  const int step_y = step_r / tiles_x, step_x = step_r - step_y * tiles_x;
  int tgy, tx;
  {
    const int t = (int)blockIdx.x % tiles_per_img;
    tgy = t / tiles_x; tx = t - tgy * tiles_x;
  }
  const uint32_t dst0 = a_base + 2 * cp * a_chunk + lane * 16 + r_begin * SEG_BYTES;
  const uint32_t lo0 = low_base + 2 * cp * low_chunk;
  int s = 0; uint32_t ph = 0; int ls = 0; uint32_t lph = 0;
  int nst = 0;                                                  // source-1 stages seen so far (team = parity)
  const long long rank = CG2 ? (long long)cluster_ctarank() : 0;
  for (long long grp = blockIdx.x; grp - rank < total; grp += gridDim.x) {
    if (CG2 && grp >= total) { tx = tiles_x - 1; tgy = tiles_y - 1; }       // odd tail: the peer repeats the last group
    const int xb = tx * T_VALID - 1, yb = tgy * mtr - 1;
    const int xl0 = (int)(sx * (float)(xb < 0 ? 0 : xb)), yl0 = (int)(sy * (float)(yb < 0 ? 0 : yb));
    const int x = xb + lane;
    const float fx = sx * (float)(x < 0 ? 0 : (x >= W ? W - 1 : x));
    const int x0 = (int)fx;
    const float lxf = fx - (float)x0;
    const act2_t lx2 = f2a2(lxf, lxf);
    const uint32_t lo_col = lo0 + (uint32_t)(x0 - xl0) * 16;
    int roff[NR]; uint32_t lyp[NR]; uint32_t ok = 0;            // row table of this group
#pragma unroll
    for (int i = 0; i < NR; ++i) {
      const int y = yb + r_begin + i;
      const float fy = sy * (float)(y < 0 ? 0 : (y >= H ? H - 1 : y));
      const int y0 = (int)fy;
      const float lyf = fy - (float)y0;
      const act2_t l2 = f2a2(lyf, lyf);
      roff[i] = (y0 - yl0) * (UP_LC * 16);
      lyp[i] = *reinterpret_cast<const uint32_t*>(&l2);
      if (y >= 0 && y < H && x >= 0 && x < W) ok |= 1u << i;      // outside: zero border (or never-consumed columns)
    }
    for (int kb = 0; kb < kbn; ++kb) {
      if (kb >= kb0) {
        if ((nst & 1) == team) {
          if (iw == 0 && lane == 0) UB_TR(6, nst >> 1, 0);
          mbar_wait_sleep(bar_lfull + 8 * ls, lph);              // low-resolution patch landed
          if (iw == 0 && lane == 0) UB_TR(6, nst >> 1, 1);
          mbar_wait_sleep(bar_empty + 8 * s, ph ^ 1);            // the stage's previous MMAs have retired
          if (iw == 0 && lane == 0) UB_TR(6, nst >> 1, 2);
          const uint32_t lo = lo_col + ls * low_stage, dst = dst0 + s * a_stage;
          act2_t top[2][4], bot[2][4];
          auto blend_row = [&](uint32_t addr, act2_t (*r)[4]) {  // both chunks of this thread: r[0], r[1]
            uint32_t av[2][4], cv[2][4];
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              lds_v4(addr + c * low_chunk, av[c][0], av[c][1], av[c][2], av[c][3]);
              lds_v4(addr + c * low_chunk + 16, cv[c][0], cv[c][1], cv[c][2], cv[c][3]);
            }
#pragma unroll
            for (int c = 0; c < 2; ++c)
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const act2_t ha = *reinterpret_cast<const act2_t*>(&av[c][j]), hc = *reinterpret_cast<const act2_t*>(&cv[c][j]);
                r[c][j] = __hfma2(lx2, __hsub2(hc, ha), ha);
              }
          };
          int cur = -(1 << 20);
#pragma unroll
          for (int i = 0; i < NR; ++i) {
            if (roff[i] != cur) {                                // warp-uniform
              if (roff[i] == cur + UP_LC * 16) {
#pragma unroll
                for (int c = 0; c < 2; ++c)
#pragma unroll
                  for (int j = 0; j < 4; ++j) top[c][j] = bot[c][j];
              } else {
                blend_row(lo + roff[i], top);
              }
              blend_row(lo + roff[i] + UP_LC * 16, bot);
              cur = roff[i];
            }
            const act2_t ly2 = *reinterpret_cast<const act2_t*>(&lyp[i]);
            const bool keep = (ok >> i) & 1u;
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              uint32_t o[4];
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const act2_t v = __hfma2(ly2, __hsub2(bot[c][j], top[c][j]), top[c][j]);
                o[j] = keep ? *reinterpret_cast<const uint32_t*>(&v) : 0u;
              }
              sts_v4(dst + c * a_chunk + i * SEG_BYTES, o[0], o[1], o[2], o[3]);
            }
          }
          fence_proxy_async_smem();                              // generic-proxy writes -> visible to the tensor core's reads
          __syncwarp();
          if (lane == 0) {                                       // CG2: bar_full is the leader CTA's barrier (shared::cluster address)
            if (CG2) mbar_arrive_cluster(bar_full + 8 * s); else mbar_arrive(bar_full + 8 * s);
            mbar_arrive(bar_lempty + 8 * ls);
          }
          if (iw == 0 && lane == 0) UB_TR(6, nst >> 1, 3);
        }
        ++nst;
        if (++ls == LS) { ls = 0; lph ^= 1; }
      }
      if (++s == S) { s = 0; ph ^= 1; }
    }
    tx += step_x;
    if (tx >= tiles_x) { tx -= tiles_x; ++tgy; }
    tgy += step_y;
    if (tgy >= tiles_y) tgy -= tiles_y;
  }
}

// NCLS: 0 = plain layer; 3 = fused head with exactly 3 classes (the shipped model); 8 = fused head with a run-time
// class count <= 8.  The class count must be a compile-time constant on the hot configuration: with the run-time
// bound the head cost 8x32 predicated FFMAs plus local-memory traffic per pixel and made the last layer issue-bound
// (845 instructions per warp and tile instead of ~300, profiles/r01_conv.md).
//
// POOL: the epilogue also writes MaxPool2d(2,2) of its output (src/models/unetpp.py:75,105-108) into a second PP plane,
// so the next level's input never costs a re-read of this layer's output.  Columns pair up inside a warp (one shuffle),
// rows pair up across the two warps that hold tile rows 2r / 2r+1: they swap column-pooled halves through a 16 KB
// double-buffered staging area and a 64-thread named barrier, then each stores 16 of the 32 channels.  Only the
// interior of the pooled plane is written; its zero border is set once by the host (net_glue.cu).
constexpr int POOL_SMEM = 16 * 1024;      // 16 warps x 2 buffers x [2 x 16 slots] x 16 B

//
// UPS: source 1 is given at HALF resolution and `Upsample(scale_factor=2, mode='bilinear', align_corners=True)`
// (src/models/unetpp.py:76,111-116) happens on the way into shared memory: for a k-block of source 1 the producer
// stages the low-resolution patch under the box (TMA), and eight interpolation warps write the K-major A operand of
// the stage -- the same bytes the TMA box of the materialised tensor would have delivered (same packed 16-bit
// a + l*(c-a) blends as k_upsample2, so the layer's output is bit-identical).  The TMA / MMA / epilogue warps do not
// change: the stage's `full` barrier simply gets its arrival from the interpolation warps.  Removes the upsampled
// tensor (2 x 16-bit x Cin1 per pixel written and read back) from HBM.
//
// CG2: the CTAs run as pairs (clusters of two on one TPC) and the leader issues tcgen05.mma.cta_group::2 with M = 256:
// rows 0-127 are the leader's M-tile, rows 128-255 the peer's (each from its own shared memory, into its own TMEM), and
// each CTA holds half of the N' weight rows.  Measured (tools/ubench_mma2.cu): an N' = 96 MMA costs 72 clk per 128 rows
// with cta_group::1 (the ~60-clk floor is per instruction) but 49.4 clk per 256 rows as a pair -- 5450 -> 7970
// FLOP/clk/SM -- and the B operand is read from shared memory once per 256 rows.  Both CTAs walk the same sequence of
// stages; everything that tells the issuing thread "ready" (TMA complete_tx, interpolation warps, epilogue warps
// handing accumulators back) arrives on the LEADER's barriers, and the leader's commits are multicast to both CTAs.
// Weights are resident: every CTA loads its half once before the loop.
template <int COUT, int KC, int NCLS, bool POOL, bool UPS, bool CG2>
__global__ void __launch_bounds__(UPS ? NTHREADS + UP_WARPS * 32 : NTHREADS, 1) k_conv_rows(const __grid_constant__ RowsParams P) {
  using C = RCfg<COUT, KC, CG2>;
  const uint32_t rank = CG2 ? cluster_ctarank() : 0;
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = P.num_stages;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  const uint32_t bar_full = smem_u32(bars), bar_empty = bar_full + 8 * S;
  const uint32_t bar_tfull = bar_empty + 8 * S, bar_tempty = bar_tfull + 8 * 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 8 * (2 * S + 8));
  [[maybe_unused]] const uint32_t bar_lfull = smem_u32(smem) + 256, bar_lempty = bar_lfull + 64;   // UPS: low-resolution ring
  [[maybe_unused]] const uint32_t low_base = smem_u32(smem) + P.low_off;
  const uint32_t a_base = smem_u32(smem) + C::HDR;
  const uint32_t b_base = a_base + S * P.a_stage;
  const PPLayout& L = P.L;
  const int tiles_per_img = P.tiles_x * P.tiles_y;

  constexpr int NCTA = CG2 ? 2 : 1;
  const uint32_t bar_w = smem_u32(smem) + 384;                 // CG2: this CTA's resident weights have landed
  if (threadIdx.x == 0) {
    // CG2: the leader's full / tempty barriers also count the peer's interpolation / epilogue warps
    for (int i = 0; i < S; ++i) { mbar_init(bar_full + 8 * i, UPS ? NCTA * UP_TEAM : 1); mbar_init(bar_empty + 8 * i, 2); }
    for (int i = 0; i < C::NBUF; ++i) { mbar_init(bar_tfull + 8 * i, 1); mbar_init(bar_tempty + 8 * i, NCTA * C::WARPS_PER_BUF); }
    if (UPS) for (int i = 0; i < P.low_slots; ++i) { mbar_init(bar_lfull + 8 * i, 1); mbar_init(bar_lempty + 8 * i, UP_TEAM); }
    if (CG2) mbar_init(bar_w, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 16) {
    if (CG2) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(512));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(512));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // where "ready" signals go: the leader CTA's barriers (shared::cluster addresses), or this CTA's own without pairs
  [[maybe_unused]] const uint32_t lead_full = CG2 ? mapa_u32(bar_full, 0) : bar_full;
  [[maybe_unused]] const uint32_t lead_tempty = CG2 ? mapa_u32(bar_tempty, 0) : bar_tempty;
  if constexpr (CG2) {
    // resident weights: this CTA's half of every k-block, loaded once; the pair starts when both halves are in
    if (threadIdx.x == 0) {
      mbar_expect_tx(bar_w, P.wbytes);
      const uint8_t* src = reinterpret_cast<const uint8_t*>(P.w) + (size_t)rank * P.wbytes;
      for (int o = 0; o < P.wbytes; o += 32768) bulk_g2s(b_base + o, src + o, P.wbytes - o < 32768 ? P.wbytes - o : 32768, bar_w);
    }
    mbar_wait(bar_w, 0);
    cluster_sync();                                            // also: the peer's barriers exist before anything arrives on them
  }

  if (warp == 16) {
    // ===================== producer: one TMA box (+ one weight copy) per k-block =====================
    if (lane == 0) {
      tma_prefetch_desc(&P.tm0);
      tma_prefetch_desc(&P.tm1);
      int s = 0; uint32_t ph = 0; bool first = true;
      [[maybe_unused]] int ls = 0; [[maybe_unused]] uint32_t lph = 0;
      [[maybe_unused]] int nbox = 0, nlow = 0;
      // UPS: the ring of low-resolution patches has its own cursor and runs ahead of the stage ring (a box needs ~2000 clk
      // from issue to landing, profiles/r02_conv.md): a patch is requested as soon as a ring slot is free, not when the
      // in-order walk over the stages reaches its k-block.
      [[maybe_unused]] long long lgrp = blockIdx.x, low_issued = 0, low_needed = 0;
      [[maybe_unused]] int lkb = P.kb0;
      [[maybe_unused]] auto issue_low = [&](bool blocking) -> bool {
        if (lgrp - rank >= P.total_items) return false;
        UB_TR(7, nlow, 0);
        if (blocking) mbar_wait(bar_lempty + 8 * ls, lph ^ 1);
        else if (!mbar_test(bar_lempty + 8 * ls, lph ^ 1)) return false;
        UB_TR(7, nlow, 1);
        ++nlow;
        const long long g = lgrp < P.total_items ? lgrp : P.total_items - 1;
        const int b = (int)(g / tiles_per_img), t = (int)(g - (long long)b * tiles_per_img);
        const int ty = t / P.tiles_x, tx = t - ty * P.tiles_x;
        // rows yl0.., columns xl0.. of the half-size image (padded coordinates +1) lie under the box
        const int xb = tx * T_VALID - 1, yb = ty * T_ROWS * P.mt - 1;
        const int xl0 = (int)(P.up_sx * (float)(xb < 0 ? 0 : xb)), yl0 = (int)(P.up_sy * (float)(yb < 0 ? 0 : yb));
        mbar_expect_tx(bar_lfull + 8 * ls, P.low_stage);
        tma_load_4d(low_base + ls * P.low_stage, &P.tm1, (xl0 + 1) * 8, yl0 + 1, (lkb - P.kb0) * C::CPK, b, bar_lfull + 8 * ls);
        if (++ls == P.low_slots) { ls = 0; lph ^= 1; }
        if (++lkb == P.kbn) { lkb = P.kb0; lgrp += gridDim.x; }
        ++low_issued;
        return true;
      };
      for (long long grp0 = blockIdx.x; grp0 - rank < P.total_items; grp0 += gridDim.x) {
        const long long grp = grp0 < P.total_items ? grp0 : P.total_items - 1;    // CG2 odd tail: the peer repeats the last group
        const int b = (int)(grp / tiles_per_img), t = (int)(grp - (long long)b * tiles_per_img);
        const int ty = t / P.tiles_x, tx = t - ty * P.tiles_x;
        const bool load_b = !CG2 && (!P.b_resident || first);
        for (int kb = 0; kb < P.kbn; ++kb) {
          const uint32_t full = bar_full + 8 * s;
          if (UPS && kb >= P.kb0) {
            while (issue_low(false)) { }
            if (low_issued <= low_needed) issue_low(true);       // normally the ring is already ahead of the stages
            ++low_needed;
            if (load_b) {                                      // weights of this k-block ride on the stage's `full` barrier
              mbar_wait(bar_empty + 8 * s, ph ^ 1);            // (its previous phase is over once the slot has been released)
              mbar_expect_tx_only(full, C::B_STAGE);
              bulk_g2s(b_base + (P.b_resident ? kb : s) * C::B_STAGE, P.w + (size_t)kb * 3 * (size_t)(KC * C::NP), C::B_STAGE, full);
            }
            if (++s == S) { s = 0; ph ^= 1; }
            continue;
          }
          UB_TR(0, nbox, 0);
          if (UPS) { while (!mbar_test(bar_empty + 8 * s, ph ^ 1)) issue_low(false); }
          else mbar_wait(bar_empty + 8 * s, ph ^ 1);
          UB_TR(0, nbox, 1);
          // box origin: padded row 4*mt*ty (= image row y0-1), padded column 30*tx (= image column x0-1)
          if constexpr (CG2) {
            // both CTAs' boxes complete on the leader's barrier; the leader expects the bytes of both
            if (rank == 0) {
              mbar_expect_tx(full, 2 * P.a_stage);
              if (UPS) mbar_arrive_n(full, 2 * UP_TEAM - 1);
            }
            const CUtensorMap* tm = kb < P.kb0 ? &P.tm0 : &P.tm1;
            tma_load_4d_pair(a_base + s * P.a_stage, tm, tx * T_VALID * 8, ty * T_ROWS * P.mt, (kb < P.kb0 ? kb : kb - P.kb0) * C::CPK, b,
                             lead_full + 8 * s);
          } else {
          mbar_expect_tx(full, P.a_stage + (load_b ? C::B_STAGE : 0));
          if (UPS) mbar_arrive_n(full, UP_TEAM - 1);             // UPS kernels: `full` counts the four warps of an interpolation team
          if (kb < P.kb0) tma_load_4d(a_base + s * P.a_stage, &P.tm0, tx * T_VALID * 8, ty * T_ROWS * P.mt, kb * C::CPK, b, full);
          else tma_load_4d(a_base + s * P.a_stage, &P.tm1, tx * T_VALID * 8, ty * T_ROWS * P.mt, (kb - P.kb0) * C::CPK, b, full);
          }
          if (load_b)
            bulk_g2s(b_base + (P.b_resident ? kb : s) * C::B_STAGE, P.w + (size_t)kb * 3 * (size_t)(KC * C::NP), C::B_STAGE, full);
          UB_TR(0, nbox, 2);
          ++nbox;
          if (++s == S) { s = 0; ph ^= 1; }
        }
        first = false;
      }
    }
  } else if ((warp == 17 || warp == 18) && rank != 0) {
    // CG2 peer: its tensor core is driven by the leader's instructions
  } else if (warp == 17 || warp == 18) {
    // ===================== MMA issuers (warps 17, 18) =====================
    // Two warps issue alternate M-tiles (tile parity = accumulator parity).  For a lone issuer the barrier waits and the
    // two tcgen05.commit per tile (~90 clk each) do not overlap its own MMAs: tools/ubench_commit.cu measures 647 clk per
    // 6-MMA batch with one issuer against 336 with two (profiles/r01_ubench_commit.log).  In this kernel the change is
    // neutral on the kbn = 1 layers, which are bound by their load + store streams (profiles/r01_conv.md, findings 10-11);
    // it keeps the tensor pipe off the critical path.  Each whole warp walks the loop, one elected lane issues; no
    // divisions in here.
    {
      constexpr uint32_t IDESC = (1u << 4) | (UB_MMA_FMT << 7) | (UB_MMA_FMT << 10) |
                                 ((uint32_t)(C::NP >> 3) << 17) | ((uint32_t)((CG2 ? 256 : 128) >> 4) << 24);
      const int wi = warp - 17;
      [[maybe_unused]] const uint32_t peer_empty = CG2 ? mapa_u32(bar_empty, 1) : 0;
      int it = 0;                                                 // counts M-tiles (accumulator uses)
      int s = 0; uint32_t ph = 0;
      for (long long grp = blockIdx.x; grp - rank < P.total_items; grp += gridDim.x, it += P.mt) {
        for (int kb = 0; kb < P.kbn; ++kb) {
          if (lane == 0 && kb == 0) UB_TR(1, it + wi, 4);
          mbar_wait(bar_full + 8 * s, ph);
          tc_fence_after();
          if (lane == 0 && kb == 0) UB_TR(1, it + wi, 3);          // first box of the group landed
          const uint32_t a_st = a_base + s * P.a_stage;
          const uint32_t b_st = b_base + ((CG2 || P.b_resident) ? kb : s) * C::B_STAGE;
          bool mine = false;
          for (int m = 0; m < P.mt; ++m) {
            if (((it + m) & 1) != wi) continue;
            mine = true;
            const int buf = (it + m) % C::NBUF, use = (it + m) / C::NBUF;
            if (kb == 0) {                                     // first touch of this accumulator: the epilogue must have drained it
              if (lane == 0) UB_TR(1, it + m, 0);
              mbar_wait(bar_tempty + 8 * buf, (use & 1) ^ 1);
              tc_fence_after();
              if (lane == 0) UB_TR(1, it + m, 1);
            }
            const uint32_t d_tmem = tmem_base + buf * C::T_STRIDE;
            if (elect_one()) {
#pragma unroll
              for (int ky = 0; ky < 3; ++ky) {
#pragma unroll
                for (int k16 = 0; k16 < KC / 16; ++k16) {
                  // A: M-tile m, tap row ky = box rows 4m+ky .. 4m+ky+3; rows 16 B apart, 8-row groups 128 B apart,
                  // the two K chunks a_chunk apart
                  const uint64_t ad = make_desc(a_st + 2 * k16 * P.a_chunk + (m * T_ROWS + ky) * SEG_BYTES, P.a_chunk, 128);
                  const uint64_t bd = make_desc(b_st + ky * C::B_KY + 2 * k16 * C::NPH * 16, C::NPH * 16, 128);
                  if (CG2) tc_mma_pair(d_tmem, ad, bd, IDESC, (kb | ky | k16) != 0);
                  else tc_mma(d_tmem, ad, bd, IDESC, (kb | ky | k16) != 0);
                }
              }
              if (kb == P.kbn - 1) {                                                         // accumulator of M-tile m complete (in both CTAs)
                if (CG2) tc_commit_pair(bar_tfull + 8 * buf, 3);
                else tc_commit(bar_tfull + 8 * buf);
              }
            }
            __syncwarp();
            if (lane == 0 && kb == P.kbn - 1) UB_TR(1, it + m, 2);
          }
          // hand the box back once this warp's MMAs have read it (plain arrive if it had no M-tile in this group)
          if (elect_one()) {
            if (CG2) {
              if (mine) tc_commit_pair(bar_empty + 8 * s, 3);
              else { mbar_arrive(bar_empty + 8 * s); mbar_arrive_cluster(peer_empty + 8 * s); }
            } else {
              if (mine) tc_commit(bar_empty + 8 * s);
              else mbar_arrive(bar_empty + 8 * s);
            }
          }
          __syncwarp();
          if (++s == S) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (UPS && warp >= 19) {
    // ===================== interpolation warps (UPS): A operand of the source-1 k-blocks =====================
    if constexpr (UPS) {
      if (P.mt == 4) ups_warp<9, CG2>(P, warp - 19, lane, lead_full, bar_empty, bar_lfull, bar_lempty, a_base, low_base);
      else if (P.mt == 2) ups_warp<5, CG2>(P, warp - 19, lane, lead_full, bar_empty, bar_lfull, bar_lempty, a_base, low_base);
      else ups_warp<3, CG2>(P, warp - 19, lane, lead_full, bar_empty, bar_lfull, bar_lempty, a_base, low_base);
    }
  } else {
    // ===================== epilogue (warps 0-15) =====================
    const int wq = warp & 3;                                   // TMEM lane quarter = tile row
    const int grp_w = warp >> 2;                               // epilogue warpgroup
    const int my_buf = C::SPLIT ? (grp_w >> 1) : grp_w;
    const int ch0 = C::SPLIT ? (grp_w & 1) * 32 : 0;             // first of this warp's 32 channels
    constexpr int NCHUNK = COUT / 8;
    // The 16 epilogue warps share four issue ports with ~360 instructions per warp and tile: per-tile index
    // arithmetic is kept out of the loop.  (b, tgy, tx) of the CTA's next group = current + gridDim.x items, carried
    // digit by digit; the divisions happen once here.
    const int step_b = (int)gridDim.x / tiles_per_img, step_r = (int)gridDim.x - step_b * tiles_per_img;
    const int step_y = step_r / P.tiles_x, step_x = step_r - step_y * P.tiles_x;
    int b = (int)blockIdx.x / tiles_per_img, tgy, tx;
    {
      const int t = (int)blockIdx.x - b * tiles_per_img;
      tgy = t / P.tiles_x; tx = t - tgy * P.tiles_x;
    }
    const size_t plane = (size_t)L.Ppl * 8;
    [[maybe_unused]] uint32_t pbuf = 0;                        // POOL: staging buffer of this tile
    [[maybe_unused]] const uint32_t stage_mine = smem_u32(smem) + P.pool_smem + warp * 1024 + (lane >> 1) * 16;
    [[maybe_unused]] const uint32_t stage_peer = smem_u32(smem) + P.pool_smem + (warp ^ 1) * 1024 + (lane >> 1) * 16;
    int it = 0;
    for (long long grp = blockIdx.x; grp - rank < P.total_items; grp += gridDim.x) {
    if (CG2 && grp >= P.total_items) { b = P.B - 1; tgy = P.tiles_y - 1; tx = P.tiles_x - 1; }   // odd tail: the peer repeats the last group
    if (lane == 0 && wq == 0) UB_TR(2 + grp_w, it / C::NBUF, 4);
    const int x = tx * T_VALID - 1 + lane;
    const bool valid = lane >= 1 && lane <= T_VALID && x < L.W;
    const bool edge = x == -1 || x == L.W;                     // left / right zero border of the PP plane
    act_t* const o_grp = P.out + ((size_t)b * NCHUNK + (ch0 >> 3)) * plane + (size_t)(L.G + x + 1) * 8;
    for (int m = 0; m < P.mt; ++m, ++it) {
      if (it % C::NBUF != my_buf) continue;
      const int use = it / C::NBUF;
      const int ty = tgy * P.mt + m;                           // 4-row M-tile index inside the image
      const int y = ty * T_ROWS + wq;
      const int q = (y + 1) * L.pitch;                         // row offset relative to G
      if (lane == 0 && wq == 0) UB_TR(2 + grp_w, use, 0);
      mbar_wait(bar_tfull + 8 * my_buf, use & 1);
      tc_fence_after();
      if (lane == 0 && wq == 0) UB_TR(2 + grp_w, use, 1);
      const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + my_buf * C::T_STRIDE;
      constexpr bool head = NCLS > 0;
      constexpr int NL = head ? NCLS : 1;
      const int ncls = NCLS == 8 ? P.ncls : NCLS;
      float lg[NL];
      [[maybe_unused]] uint32_t mine[8];                       // POOL: column-pooled channels this warp will store
      [[maybe_unused]] uint64_t lg2[3];                        // NCLS == 3: (bias, 0) + packed partial sums of even / odd channels
      if (head) {
#pragma unroll
        for (int c = 0; c < NL; ++c) lg[c] = c < ncls ? P.head[ncls * 32 + c] : 0.f;
        if constexpr (NCLS == 3) {
#pragma unroll
          for (int c = 0; c < 3; ++c) lg2[c] = pack2(__float_as_uint(lg[c]), 0u);
        }
      }
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int c0 = ch0 + half * 16;
        uint32_t k0[16], k1[16], k2[16];
        TMEM_LD16(taddr + c0, k0);
        TMEM_LD16(taddr + COUT + c0, k1);
        TMEM_LD16(taddr + 2 * COUT + c0, k2);
        tmem_ld_wait();
        if (half == 0 && lane == 0 && wq == 0) UB_TR(2 + grp_w, use, 5);
        if (half == 1) {                                       // accumulator fully read: hand it back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if (CG2) mbar_arrive_cluster(lead_tempty + 8 * my_buf); else mbar_arrive(bar_tempty + 8 * my_buf);
          }
          if (lane == 0 && wq == 0) UB_TR(2 + grp_w, use, 2);
        }
        uint32_t hv[8];
#pragma unroll
        for (int j = 0; j < 16; j += 2) {
          const uint32_t l0 = __shfl_up_sync(0xffffffffu, k0[j], 1), l1 = __shfl_up_sync(0xffffffffu, k0[j + 1], 1);
          const uint32_t r0 = __shfl_down_sync(0xffffffffu, k2[j], 1), r1 = __shfl_down_sync(0xffffffffu, k2[j + 1], 1);
          const float2 bb = make_float2(P.biasv[c0 + j], P.biasv[c0 + j + 1]);
          uint64_t v = add2(pack2(l0, l1), pack2(k1[j], k1[j + 1]));
          v = add2(v, pack2(r0, r1));
          v = add2(v, pack2(__float_as_uint(bb.x), __float_as_uint(bb.y)));
          if (head) {
            uint32_t lo, hi;
            asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v));
            const float f0 = fmaxf(__uint_as_float(lo), 0.f), f1 = fmaxf(__uint_as_float(hi), 0.f);   // ReLU; fp32 feeds the head directly
            if constexpr (NCLS == 3) {
              // packed f32x2 FMAs: the (even, odd) channel pair accumulates into the two halves of lg2[c], summed at the end
              const uint64_t f01 = pack2(__float_as_uint(f0), __float_as_uint(f1));
#pragma unroll
              for (int c = 0; c < 3; ++c)
                lg2[c] = fma2(pack2(__float_as_uint(P.head[c * 32 + ((c0 + j) & 31)]), __float_as_uint(P.head[c * 32 + ((c0 + j + 1) & 31)])), f01, lg2[c]);
            } else {
#pragma unroll
              for (int c = 0; c < NL; ++c)
                if (NCLS != 8 || c < ncls) lg[c] = fmaf(P.head[c * 32 + ((c0 + j + 1) & 31)], f1, fmaf(P.head[c * 32 + ((c0 + j) & 31)], f0, lg[c]));
            }
          } else {
            hv[j >> 1] = valid ? relu_pack(v) : 0u;
          }
        }
        if (!head && P.sat != nullptr) {                       // diagnostic: outputs that hit the largest finite value
#ifdef UB_ACT_BF16
          constexpr uint32_t MAXW = 0x7F7F7F7Fu;
#else
          constexpr uint32_t MAXW = 0x7BFF7BFFu;
#endif
          int cnt = 0;
#pragma unroll
          for (int j = 0; j < 8; ++j) cnt += __popc(__vcmpeq2(hv[j], MAXW) & 0x00010001u);
          cnt = __reduce_add_sync(0xffffffffu, cnt);
          if (lane == 0 && cnt) atomicAdd(P.sat, (unsigned long long)cnt);
        }
        if (!head && (valid || edge)) {
          act_t* o = o_grp + (size_t)(half * 2) * plane + (size_t)q * 8;
          *reinterpret_cast<uint4*>(o) = make_uint4(hv[0], hv[1], hv[2], hv[3]);
          *reinterpret_cast<uint4*>(o + plane) = make_uint4(hv[4], hv[5], hv[6], hv[7]);
          // top / bottom zero border rows of the plane
          if ((ty == 0 && wq == 0) || (ty == P.tiles_y * P.mt - 1 && wq == T_ROWS - 1)) {
            act_t* z = wq == 0 ? o - (size_t)L.pitch * 8 : o + (size_t)L.pitch * 8;
            *reinterpret_cast<uint4*>(z) = make_uint4(0, 0, 0, 0);
            *reinterpret_cast<uint4*>(z + plane) = make_uint4(0, 0, 0, 0);
          }
        }
        if constexpr (POOL) {
          // column pairs (x, x+1), x even = odd lanes; activations are >= 0, so the 16-bit integer maximum is the maximum
          uint32_t cp[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) cp[j] = __vmaxu2(hv[j], __shfl_down_sync(0xffffffffu, hv[j], 1));
          if (half == (wq & 1)) {                                // even tile rows keep channels 0-15, odd rows 16-31
#pragma unroll
            for (int j = 0; j < 8; ++j) mine[j] = cp[j];
          } else if (lane & 1) {
            st_shared_v4(stage_mine + pbuf * 512, cp[0], cp[1], cp[2], cp[3]);
            st_shared_v4(stage_mine + pbuf * 512 + 256, cp[4], cp[5], cp[6], cp[7]);
          }
        }
      }
      if constexpr (POOL) {
        asm volatile("bar.sync %0, 64;" ::"r"(1 + (warp >> 1)) : "memory");       // the two warps of rows 2r / 2r+1
        if ((lane & 1) && valid) {
          uint32_t o0, o1, o2, o3, o4, o5, o6, o7;
          ld_shared_v4(stage_peer + pbuf * 512, o0, o1, o2, o3);
          ld_shared_v4(stage_peer + pbuf * 512 + 256, o4, o5, o6, o7);
          const int py = y >> 1, px = tx * (T_VALID / 2) + (lane >> 1);
          act_t* po = P.pool_out + (((size_t)b * NCHUNK + (ch0 >> 3) + (wq & 1) * 2) * P.Lp.Ppl + P.Lp.G +
                                    (size_t)(py + 1) * P.Lp.pitch + px + 1) * 8;
          *reinterpret_cast<uint4*>(po) = make_uint4(__vmaxu2(mine[0], o0), __vmaxu2(mine[1], o1), __vmaxu2(mine[2], o2), __vmaxu2(mine[3], o3));
          *reinterpret_cast<uint4*>(po + (size_t)P.Lp.Ppl * 8) =
              make_uint4(__vmaxu2(mine[4], o4), __vmaxu2(mine[5], o5), __vmaxu2(mine[6], o6), __vmaxu2(mine[7], o7));
        }
        pbuf ^= 1;
      }
      if constexpr (NCLS == 3) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          uint32_t lo, hi;
          asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(lg2[c]));
          lg[c] = __uint_as_float(lo) + __uint_as_float(hi);
        }
      }
      if (head && valid) {
        const size_t hw = (size_t)L.H * L.W, pix = (size_t)y * L.W + x;
        int best = 0;
        float bestv = lg[0];
#pragma unroll
        for (int c = 0; c < NL; ++c)
          if (NCLS != 8 || c < ncls) {
            if (P.logits) P.logits[((size_t)b * ncls + c) * hw + pix] = lg[c];
            if (c > 0 && lg[c] > bestv) { best = c; bestv = lg[c]; }      // first maximum wins, as np.argmax
          }
        if (P.pred) P.pred[(size_t)b * hw + pix] = (uint8_t)best;
      }
      if (lane == 0 && wq == 0) UB_TR(2 + grp_w, use, 3);
    }
    tx += step_x;
    if (tx >= P.tiles_x) { tx -= P.tiles_x; ++tgy; }
    tgy += step_y;
    if (tgy >= P.tiles_y) { tgy -= P.tiles_y; ++b; }
    b += step_b;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (CG2) cluster_sync();                   // no CTA leaves (or frees its TMEM) while the pair's MMAs / remote arrivals may still touch it
  if (warp == 16) {
    tc_fence_after();
    if (CG2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512));
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512));
  }
}

constexpr int SMEM_BUDGET = 227 * 1024;

template <int COUT, int KC, int NCLS, bool POOL, bool UPS, bool CG2>
int launch_rows_cg(ub_ctx* ctx, RowsParams& P, const ConvArgs& a, const ConvLayer& Lr, cudaStream_t st, bool init_only) {
  using C = RCfg<COUT, KC, CG2>;
  static_assert(!UPS || KC == 32, "fused upsample: the interpolation warps are laid out for four chunks per k-block");
  auto kern = k_conv_rows<COUT, KC, NCLS, POOL, UPS, CG2>;
  constexpr int LOW_SLOTS = CG2 ? 4 : 2;      // pair kernels hold half the weights: room for a deeper ring
  // UPS: room for the ring of low-resolution patches; a patch under a (4*mt+2)-row box has at most 2*mt+3 rows
  // (the source index advances by < 0.5 per output row), sized here for the largest mt
  constexpr int BUDGET = SMEM_BUDGET - (POOL ? POOL_SMEM : 0) - (UPS ? LOW_SLOTS * C::CPK * (2 * C::NBUF + 3) * UP_LC * 16 : 0);
  if (init_only) {
    UB_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET));
    return UB_OK;
  }
  // One TMA box costs ~0.42 us of the SM's TMA unit whatever its size (tools/ubench_tma.cu): make it big.
  // mt vertically adjacent M-tiles share one box of 4*mt+2 rows.  With several k-blocks all mt accumulators are live
  // at once, so mt = NBUF gives up the TMEM double buffering across groups -- measured faster anyway (UB_ROWS_MT sweep,
  // profiles/r01_conv.md).
  const int resident_bytes = P.kbn * C::B_STAGE;
  static const int mt_env = getenv("UB_ROWS_MT") ? atoi(getenv("UB_ROWS_MT")) : 4;
  const int mt_max = mt_env < 1 ? 1 : (mt_env < C::NBUF ? mt_env : C::NBUF);
  int mt = 1, stages = 0;
  for (int cand = mt_max; cand >= 1; cand >>= 1) {
    if (a.H % (T_ROWS * cand)) continue;
    const int a_stage = C::CPK * (T_ROWS * cand + 2) * SEG_BYTES;
    int st_n;
    if (C::HDR + resident_bytes + 2 * a_stage <= BUDGET) { P.b_resident = 1; st_n = (BUDGET - C::HDR - resident_bytes) / a_stage; }
    else if (CG2) continue;                                    // pair kernels keep their (half-size) weights resident
    else { P.b_resident = 0; st_n = (BUDGET - C::HDR) / (a_stage + C::B_STAGE); }
    if (st_n >= 2 || cand == 1) { mt = cand; stages = st_n; break; }
  }
  if (stages > 8) stages = 8;
  if (stages < 2) return ub_fail(ctx, UB_ERR_ARG, "conv_rows: COUT=%d KC=%d does not fit 2 stages", COUT, KC);
  P.num_stages = stages;
  P.mt = mt;
  P.a_chunk = (T_ROWS * mt + 2) * SEG_BYTES;
  P.a_stage = C::CPK * P.a_chunk;
  P.tiles_y = a.H / (T_ROWS * mt);
  P.total_items = (long long)a.B * P.tiles_x * P.tiles_y;
  int r = tmap_pp_rows(ctx, &P.tm0, a.src0, a.B, Lr.cin0 / 8, P.L, T_COLS, T_ROWS * mt + 2, C::CPK);
  int smem = C::HDR + stages * P.a_stage + (P.b_resident ? resident_bytes : stages * C::B_STAGE);
  if (UPS) {
    const int h = a.H / 2, w = a.W / 2;
    P.up_sy = (float)(h - 1) / (float)(a.H - 1); P.up_sx = (float)(w - 1) / (float)(a.W - 1);   // as upsample2_pp
    P.low_rows = 2 * mt + 3; P.low_chunk = P.low_rows * UP_LC * 16; P.low_stage = C::CPK * P.low_chunk;
    P.low_slots = LOW_SLOTS; P.low_off = smem; smem += LOW_SLOTS * P.low_stage;
    if (r == UB_OK) r = tmap_pp_rows(ctx, &P.tm1, a.src1, a.B, Lr.cin1 / 8, pp_layout(h, w), UP_LC, P.low_rows, C::CPK);
  } else if (r == UB_OK) {
    r = Lr.cin1 ? tmap_pp_rows(ctx, &P.tm1, a.src1, a.B, Lr.cin1 / 8, P.L, T_COLS, T_ROWS * mt + 2, C::CPK)
                : tmap_pp_rows(ctx, &P.tm1, a.src0, a.B, Lr.cin0 / 8, P.L, T_COLS, T_ROWS * mt + 2, C::CPK);
  }
  if (r != UB_OK) return r;
  if (POOL) { P.pool_smem = (smem + 1023) & ~1023; smem = P.pool_smem + POOL_SMEM; }
  long long grid = P.total_items < ctx->sm_count ? P.total_items : ctx->sm_count;
  if (CG2) {
    // clusters of two; an odd last group is computed by both CTAs of its pair (identical stores)
    const long long pairs = (P.total_items + 1) / 2, max_pairs = ctx->sm_count / 2;
    grid = 2 * (pairs < max_pairs ? pairs : max_pairs);
    P.w = Lr.w2; P.wbytes = resident_bytes;
  }
#ifdef UB_ROWS_TRACE
  static long long* d_trace = nullptr;
  const size_t tr_bytes = sizeof(long long) * 8 * TR_N * 8;
  if (!d_trace) UB_CUDA(ctx, cudaMalloc(&d_trace, tr_bytes));
  UB_CUDA(ctx, cudaMemsetAsync(d_trace, 0, tr_bytes, st));
  P.trace = d_trace;
#endif
  if (CG2) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(UPS ? NTHREADS + UP_WARPS * 32 : NTHREADS);
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    UB_CUDA(ctx, cudaLaunchKernelEx(&cfg, kern, P));
  } else {
    kern<<<(unsigned)grid, UPS ? NTHREADS + UP_WARPS * 32 : NTHREADS, smem, st>>>(P);
  }
  UB_CHECK_LAUNCH(ctx);
#ifdef UB_ROWS_TRACE
  if (const char* dir = getenv("UB_ROWS_TRACE_DIR")) {
    std::vector<long long> h(8 * TR_N * 8);
    UB_CUDA(ctx, cudaStreamSynchronize(st));
    UB_CUDA(ctx, cudaMemcpy(h.data(), d_trace, tr_bytes, cudaMemcpyDeviceToHost));
    char path[512];
    snprintf(path, sizeof path, "%s/rows_trace_%02d.bin", dir, g_trace_dump++);
    if (FILE* f = fopen(path, "wb")) {
      const long long hdr[8] = {COUT, KC, NCLS, P.kbn, P.mt, P.num_stages, P.total_items, grid};
      fwrite(hdr, sizeof hdr, 1, f);
      fwrite(h.data(), tr_bytes, 1, f);
      fclose(f);
    }
  }
#endif
  return UB_OK;
}

// CTA pairs (cta_group::2 MMAs) unless UB_ROWS_PAIR=0; init registers both instantiations
bool rows_pair_enabled() {
  static const bool on = !(getenv("UB_ROWS_PAIR") && atoi(getenv("UB_ROWS_PAIR")) == 0);
  return on;
}
template <int COUT, int KC, int NCLS, bool POOL = false, bool UPS = false>
int launch_rows(ub_ctx* ctx, RowsParams& P, const ConvArgs& a, const ConvLayer& Lr, cudaStream_t st, bool init_only) {
  if (init_only) {
    int r = launch_rows_cg<COUT, KC, NCLS, POOL, UPS, false>(ctx, P, a, Lr, st, true);
    return r != UB_OK ? r : launch_rows_cg<COUT, KC, NCLS, POOL, UPS, true>(ctx, P, a, Lr, st, true);
  }
  if (rows_pair_enabled() && Lr.w2 && ctx->sm_count >= 2) return launch_rows_cg<COUT, KC, NCLS, POOL, UPS, true>(ctx, P, a, Lr, st, false);
  return launch_rows_cg<COUT, KC, NCLS, POOL, UPS, false>(ctx, P, a, Lr, st, false);
}

int dispatch_rows(ub_ctx* ctx, int cout, int kc, RowsParams& P, const ConvArgs& a, const ConvLayer& Lr, cudaStream_t st, bool init_only) {
  if (P.ncls > 0) {
    if (cout != 32 || kc != 32) return ub_fail(ctx, UB_ERR_ARG, "fused head needs COUT=32 KC=32 (got %d, %d)", cout, kc);
    return P.ncls == 3 ? launch_rows<32, 32, 3>(ctx, P, a, Lr, st, init_only) : launch_rows<32, 32, 8>(ctx, P, a, Lr, st, init_only);
  }
  if (a.up1 || (init_only && P.low_slots)) {
    if (cout == 32 && kc == 32) return launch_rows<32, 32, 0, false, true>(ctx, P, a, Lr, st, init_only);
    if (cout == 64 && kc == 32) return launch_rows<64, 32, 0, false, true>(ctx, P, a, Lr, st, init_only);
    return ub_fail(ctx, UB_ERR_ARG, "conv_rows: no fused upsample for COUT=%d KC=%d", cout, kc);
  }
  if (P.pool_out) {
    if (cout == 32 && kc == 32) return launch_rows<32, 32, 0, true>(ctx, P, a, Lr, st, init_only);
    if (cout == 64 && kc == 64) return launch_rows<64, 64, 0, true>(ctx, P, a, Lr, st, init_only);
    return ub_fail(ctx, UB_ERR_ARG, "conv_rows: no fused pool for COUT=%d KC=%d", cout, kc);
  }
  if (cout == 32 && kc == 16) return launch_rows<32, 16, 0>(ctx, P, a, Lr, st, init_only);
  if (cout == 32 && kc == 32) return launch_rows<32, 32, 0>(ctx, P, a, Lr, st, init_only);
  if (cout == 64 && kc == 32) return launch_rows<64, 32, 0>(ctx, P, a, Lr, st, init_only);
  if (cout == 64 && kc == 64) return launch_rows<64, 64, 0>(ctx, P, a, Lr, st, init_only);
  return ub_fail(ctx, UB_ERR_ARG, "conv_rows: unsupported COUT=%d KC=%d", cout, kc);
}

}  // namespace

int conv_rows_init(ub_ctx* ctx) {
  static RowsParams P{};
  const int cfgs[10][4] = {{32, 16, 0, 0}, {32, 32, 0, 0}, {64, 32, 0, 0}, {64, 64, 0, 0}, {32, 32, 3, 0}, {32, 32, 8, 0},
                           {32, 32, 0, 1}, {64, 64, 0, 1}, {32, 32, 0, 2}, {64, 32, 0, 2}};
  ConvArgs a{}; ConvLayer Lr{};
  for (auto& c : cfgs) {
    P.ncls = c[2];
    P.pool_out = c[3] == 1 ? reinterpret_cast<act_t*>(&P) : nullptr;     // any non-null value: selects the POOL instantiation
    P.low_slots = c[3] == 2;                                             // ... and the UPS instantiation
    int r = dispatch_rows(ctx, c[0], c[1], P, a, Lr, nullptr, true);
    if (r != UB_OK) return r;
  }
  return UB_OK;
}

bool conv_rows_can_upsample(const ub_ctx* ctx, const ConvLayer& L) {
  static const bool off = getenv("UB_FUSE_UP") && atoi(getenv("UB_FUSE_UP")) == 0;
  return !off && ctx->conv_impl != 1 && L.fold && L.cin1 > 0 && L.kc == 32 && (L.cout == 32 || L.cout == 64);
}

bool conv_rows_can_pool(const ub_ctx* ctx, const ConvLayer& L) {
  static const bool off = getenv("UB_FUSE_POOL") && atoi(getenv("UB_FUSE_POOL")) == 0;
  return !off && ctx->conv_impl != 1 && L.fold && L.cin1 == 0 && ((L.cout == 32 && L.kc == 32) || (L.cout == 64 && L.kc == 64));
}

int conv3x3_rows(ub_ctx* ctx, const ConvLayer& Lr, const ConvArgs& a, cudaStream_t st) {
  if (a.H % T_ROWS) return ub_fail(ctx, UB_ERR_ARG, "conv_rows: H=%d must be a multiple of %d", a.H, T_ROWS);
  RowsParams P{};
  P.out = a.out; P.w = Lr.w; P.bias = Lr.bias;
  memcpy(P.biasv, Lr.bias_host, sizeof(float) * (size_t)(Lr.cout < 64 ? Lr.cout : 64));
  P.B = a.B; P.kb0 = Lr.cin0 / Lr.kc; P.kbn = (Lr.cin0 + Lr.cin1) / Lr.kc;
  P.L = pp_layout(a.H, a.W);
  P.tiles_x = cdiv(a.W, T_VALID);
  P.ncls = 0; P.logits = a.logits; P.pred = a.pred; P.sat = a.sat;
  if (a.pool_out) {
    if (a.head_w || (a.H & 1) || (a.W & 1)) return ub_fail(ctx, UB_ERR_ARG, "conv_rows: fused pool needs even H, W and no head");
    P.pool_out = a.pool_out; P.Lp = pp_layout(a.H / 2, a.W / 2);
  }
  if (a.up1 && (!Lr.cin1 || a.pool_out || a.head_w || (a.H & 1) || (a.W & 1)))
    return ub_fail(ctx, UB_ERR_ARG, "conv_rows: fused upsample needs a second source, even H, W, no pool and no head");
  if (a.head_w) {
    if (Lr.cout != 32) return ub_fail(ctx, UB_ERR_ARG, "fused head needs cout == 32");
    P.ncls = a.num_classes;
    // head weights travel in the kernel parameters (constant bank): copy from the host mirror
    memcpy(P.head, ctx->head_host, sizeof(float) * (size_t)a.num_classes * 33);
  }
  TimeScope ts(ctx, UB_T_CONV, st);
  return dispatch_rows(ctx, Lr.cout, Lr.kc, P, a, Lr, st, false);
}
