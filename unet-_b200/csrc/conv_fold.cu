// conv_fold.cu -- 3x3 conv for narrow layers (Cout = 32 / 64) with the kernel's x-taps folded
// into the GEMM N dimension.
//
// Why: with N = Cout = 32 the nine-shifted-GEMM form (conv_tc.cu) re-reads the 4 KB activation
// tile from shared memory for every tap; ncu showed the UMMA operand path ("tc" pipe) 73 %
// busy while the tensor pipe was only 16-19 % busy (profiles/r01_conv_baseline.md).  Here
//   D[p, kx*Cout+co] = sum_{ky,ci} X[p + (ky-1)*pitch, ci] * W[co,ky,kx,ci]        (N' = 3*Cout)
// so one activation tile feeds all three x-taps (3x less A traffic, every MAC useful), and
//   out[p, co] = D[p-1, 0*Cout+co] + D[p, 1*Cout+co] + D[p+1, 2*Cout+co]
// is resolved in the epilogue: rows are TMEM lanes = threads, so row +-1 is a warp shuffle;
// the 2x3 warp-edge rows go through a 1-2 KB shared-memory exchange, and tiles overlap by two
// rows (126 valid rows per 128-row tile).
//
// Warp roles (320 threads): warps 0-3 epilogue group 0 (TMEM buffer 0, even tiles),
// warps 4-7 epilogue group 1 (buffer 1, odd tiles), warp 8 bulk-copy producer (+TMEM alloc),
// warp 9 MMA issuer.  Weights stay resident in shared memory when they fit.
#include <string.h>
#include "tc_ptx.cuh"

namespace {
using namespace tcp;

constexpr int TILE_M = 128, TILE_VALID = 126;

struct FoldParams {
  const act_t* src0; const act_t* src1; act_t* out;
  const act_t* w; const float* bias;
  int B, kb0, kbn, nch0, nch1, num_stages, b_resident;
  long long total_items;
  PPLayout L;
  int ncls; float* logits; uint8_t* pred;      // fused 1x1 head (ncls > 0), COUT == 32
  float head[8 * 33];                          // [ncls][32] weights then [ncls] biases
};

template <int COUT, int KC>
struct FCfg {
  static constexpr int NP = 3 * COUT;                         // folded N
  static constexpr int CPK = KC / 8;
  static constexpr int A_SEG = TILE_M * 16;                   // one chunk, one ky: 2048 B
  static constexpr int A_STAGE = 3 * CPK * A_SEG;
  static constexpr int B_KY = KC * NP * 2;                    // one ky of one k-block
  static constexpr int B_STAGE = 3 * B_KY;
  static constexpr int TMEM_COLS = NP <= 128 ? 256 : 512;     // two accumulators of NP columns
  static constexpr int TMEM_STRIDE = NP <= 128 ? 128 : 256;
  static constexpr int XCH = 2 * 2 * 4 * 2 * COUT * 4;        // [group][set][warp][side][COUT] fp32
  static constexpr int HDR = 1024 + XCH;                      // barriers, bias, exchange
};

template <int COUT, int KC>
__global__ void __launch_bounds__(320, 1) k_conv_fold(const __grid_constant__ FoldParams P) {
  using C = FCfg<COUT, KC>;
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = P.num_stages;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  const uint32_t bar_full = smem_u32(bars), bar_empty = bar_full + 8 * S;
  const uint32_t bar_tfull = bar_empty + 8 * S, bar_tempty = bar_tfull + 16;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 8 * (2 * S + 4));
  float* s_bias = reinterpret_cast<float*>(smem + 512);        // [COUT] (<= 64 floats)
  float* s_xch = reinterpret_cast<float*>(smem + 1024);
  const uint32_t a_base = smem_u32(smem) + C::HDR;
  const uint32_t b_base = a_base + S * C::A_STAGE;
  const PPLayout& L = P.L;

  if (threadIdx.x == 0) {
    for (int i = 0; i < S; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_tfull + 8 * i, 1); mbar_init(bar_tempty + 8 * i, 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < COUT) s_bias[threadIdx.x] = P.bias[threadIdx.x];
  if (warp == 8) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "n"(C::TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 8) {
    // ===================== producer =====================
    int s = 0; uint32_t ph = 0; bool first = true;
    for (long long item = blockIdx.x; item < P.total_items; item += gridDim.x) {
      const int mt = (int)(item % L.ntiles_fold), b = (int)(item / L.ntiles_fold);
      const int p0 = L.G + mt * TILE_VALID - 1;                  // position of tile row 0
      const bool load_b = !P.b_resident || first;
      for (int kb = 0; kb < P.kbn; ++kb) {
        const uint32_t full = bar_full + 8 * s;
        if (lane == 0) {
          mbar_wait(bar_empty + 8 * s, ph ^ 1);
          mbar_expect_tx(full, C::A_STAGE + (load_b ? C::B_STAGE : 0));
        }
        __syncwarp();
        const act_t* plane0 =
            kb < P.kb0 ? P.src0 + ((size_t)b * P.nch0 + (size_t)kb * C::CPK) * L.Ppl * 8
                       : P.src1 + ((size_t)b * P.nch1 + (size_t)(kb - P.kb0) * C::CPK) * L.Ppl * 8;
        const uint32_t a_st = a_base + s * C::A_STAGE;
        const uint32_t b_st = b_base + (P.b_resident ? kb : s) * C::B_STAGE;
        constexpr int NA = 3 * C::CPK;
        const int ncopy = NA + (load_b ? 3 : 0);
        for (int c = lane; c < ncopy; c += 32) {
          if (c < NA) {
            const int ky = c / C::CPK, cc = c - ky * C::CPK;
            bulk_g2s(a_st + c * C::A_SEG, plane0 + ((size_t)cc * L.Ppl + p0 + (ky - 1) * L.pitch) * 8, C::A_SEG, full);
          } else {
            const int ky = c - NA;
            bulk_g2s(b_st + ky * C::B_KY, P.w + ((size_t)kb * 3 + ky) * (size_t)(KC * C::NP), C::B_KY, full);
          }
        }
        if (++s == S) { s = 0; ph ^= 1; }
      }
      first = false;
    }
  } else if (warp == 9) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t IDESC = (1u << 4) | (UB_MMA_FMT << 7) | (UB_MMA_FMT << 10) |
                                 ((uint32_t)(C::NP >> 3) << 17) | ((uint32_t)(TILE_M >> 4) << 24);
      int s = 0; uint32_t ph = 0; int it = 0;
      for (long long item = blockIdx.x; item < P.total_items; item += gridDim.x, ++it) {
        const int buf = it & 1;
        mbar_wait(bar_tempty + 8 * buf, ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * C::TMEM_STRIDE;
        for (int kb = 0; kb < P.kbn; ++kb) {
          mbar_wait(bar_full + 8 * s, ph);
          tc_fence_after();
          const uint32_t a_st = a_base + s * C::A_STAGE;
          const uint32_t b_st = b_base + (P.b_resident ? kb : s) * C::B_STAGE;
#pragma unroll
          for (int ky = 0; ky < 3; ++ky) {
#pragma unroll
            for (int k16 = 0; k16 < KC / 16; ++k16) {
              const uint64_t ad = make_desc(a_st + (ky * C::CPK + 2 * k16) * C::A_SEG, C::A_SEG, 128);
              const uint64_t bd = make_desc(b_st + ky * C::B_KY + 2 * k16 * C::NP * 16, C::NP * 16, 128);
              tc_mma(d_tmem, ad, bd, IDESC, (kb | ky | k16) != 0);
            }
          }
          tc_commit(bar_empty + 8 * s);
          if (++s == S) { s = 0; ph ^= 1; }
        }
        tc_commit(bar_tfull + 8 * buf);
      }
    }
  } else {
    // ===================== epilogue (two groups of four warps) =====================
    const int grp = warp >> 2, wq = warp & 3;
    const int row = wq * 32 + lane;
    constexpr int NCHUNK = COUT / 8;
    int it = 0;
    for (long long item = blockIdx.x; item < P.total_items; item += gridDim.x, ++it) {
      if ((it & 1) != grp) continue;
      const int mt = (int)(item % L.ntiles_fold), b = (int)(item / L.ntiles_fold);
      const int q = mt * TILE_VALID - 1 + row;                   // position index relative to G
      const int yy = q / L.pitch, xx = q - yy * L.pitch;
      const bool in_img = q >= 0 && q < L.npos;
      const bool interior = in_img && yy >= 1 && yy <= L.H && xx >= 1 && xx <= L.W;
      const bool valid_row = row >= 1 && row <= TILE_VALID && in_img;
      const int use = it >> 1;                                   // per-group iteration
      float* xs = s_xch + (size_t)((grp * 2 + (use & 1)) * 4) * 2 * COUT;
      mbar_wait(bar_tfull + 8 * grp, use & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + grp * C::TMEM_STRIDE;
      // pass 1: publish the warp-edge rows (lane 31's kx=0 block, lane 0's kx=2 block)
#pragma unroll 1
      for (int c0 = 0; c0 < COUT; c0 += 16) {
        uint32_t e0[16], e2[16];
        TMEM_LD16(taddr + c0, e0);
        TMEM_LD16(taddr + 2 * COUT + c0, e2);
        tmem_ld_wait();
        if (lane == 31) {
#pragma unroll
          for (int j = 0; j < 16; j += 4)
            *reinterpret_cast<uint4*>(xs + (wq * 2 + 0) * COUT + c0 + j) = make_uint4(e0[j], e0[j + 1], e0[j + 2], e0[j + 3]);
        }
        if (lane == 0) {
#pragma unroll
          for (int j = 0; j < 16; j += 4)
            *reinterpret_cast<uint4*>(xs + (wq * 2 + 1) * COUT + c0 + j) = make_uint4(e2[j], e2[j + 1], e2[j + 2], e2[j + 3]);
        }
      }
      named_bar_sync(1 + grp, 128);
      // pass 2: combine the three x-taps, bias, ReLU, store (or fused head)
      float lg[8];
#pragma unroll
      for (int c = 0; c < 8; ++c) lg[c] = c < P.ncls ? P.head[P.ncls * 32 + c] : 0.f;
#pragma unroll
      for (int c0 = 0; c0 < COUT; c0 += 16) {
        uint32_t k0[16], k1[16], k2[16];
        TMEM_LD16(taddr + c0, k0);
        TMEM_LD16(taddr + COUT + c0, k1);
        TMEM_LD16(taddr + 2 * COUT + c0, k2);
        tmem_ld_wait();
        float f[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          float l = __shfl_up_sync(0xffffffffu, __uint_as_float(k0[j]), 1);
          float r = __shfl_down_sync(0xffffffffu, __uint_as_float(k2[j]), 1);
          if (lane == 0 && wq > 0) l = xs[((wq - 1) * 2 + 0) * COUT + c0 + j];
          if (lane == 31 && wq < 3) r = xs[((wq + 1) * 2 + 1) * COUT + c0 + j];
          const float v = l + __uint_as_float(k1[j]) + r + s_bias[c0 + j];
          f[j] = interior ? fminf(fmaxf(v, 0.f), UB_ACT_MAX) : 0.f;
        }
        if (P.ncls > 0) {
#pragma unroll
          for (int c = 0; c < 8; ++c)
            if (c < P.ncls)
#pragma unroll
              for (int j = 0; j < 16; ++j) lg[c] = fmaf(P.head[c * 32 + c0 + j], f[j], lg[c]);
        } else if (valid_row) {
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            act2_t h[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) h[j] = f2a2(f[g * 8 + 2 * j], f[g * 8 + 2 * j + 1]);
            *reinterpret_cast<uint4*>(P.out + (((size_t)b * NCHUNK + (c0 >> 3) + g) * L.Ppl + L.G + q) * 8) =
                *reinterpret_cast<uint4*>(h);
          }
        }
      }
      if (P.ncls > 0 && valid_row && interior) {
        const size_t hw = (size_t)L.H * L.W, pix = (size_t)(yy - 1) * L.W + (xx - 1);
        int best = 0;
#pragma unroll
        for (int c = 0; c < 8; ++c)
          if (c < P.ncls) {
            if (P.logits) P.logits[((size_t)b * P.ncls + c) * hw + pix] = lg[c];
            if (c > 0 && lg[c] > lg[best]) best = c;
          }
        if (P.pred) P.pred[(size_t)b * hw + pix] = (uint8_t)best;
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tempty + 8 * grp);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(C::TMEM_COLS));
  }
}

constexpr int SMEM_BUDGET = 227 * 1024;

template <int COUT, int KC>
int launch_fold(ub_ctx* ctx, FoldParams& P, cudaStream_t st, bool init_only) {
  using C = FCfg<COUT, KC>;
  auto kern = k_conv_fold<COUT, KC>;
  if (init_only) {
    UB_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET));
    return UB_OK;
  }
  const int resident_bytes = P.kbn * C::B_STAGE;
  int stages;
  if (C::HDR + resident_bytes + 3 * C::A_STAGE <= SMEM_BUDGET) {
    P.b_resident = 1;
    stages = (SMEM_BUDGET - C::HDR - resident_bytes) / C::A_STAGE;
  } else {
    P.b_resident = 0;
    stages = (SMEM_BUDGET - C::HDR) / (C::A_STAGE + C::B_STAGE);
  }
  if (stages > 8) stages = 8;
  if (stages < 2) return ub_fail(ctx, UB_ERR_ARG, "conv_fold: COUT=%d KC=%d does not fit 2 stages", COUT, KC);
  P.num_stages = stages;
  const int smem = C::HDR + stages * C::A_STAGE + (P.b_resident ? resident_bytes : stages * C::B_STAGE);
  long long grid = P.total_items < ctx->sm_count ? P.total_items : ctx->sm_count;
  kern<<<(unsigned)grid, 320, smem, st>>>(P);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

int dispatch_fold(ub_ctx* ctx, int cout, int kc, FoldParams& P, cudaStream_t st, bool init_only) {
  if (cout == 32 && kc == 16) return launch_fold<32, 16>(ctx, P, st, init_only);
  if (cout == 32 && kc == 32) return launch_fold<32, 32>(ctx, P, st, init_only);
  if (cout == 64 && kc == 32) return launch_fold<64, 32>(ctx, P, st, init_only);
  return ub_fail(ctx, UB_ERR_ARG, "conv_fold: unsupported COUT=%d KC=%d", cout, kc);
}

}  // namespace

int conv_fold_init(ub_ctx* ctx) {
  static FoldParams P{};
  const int cfgs[3][2] = {{32, 16}, {32, 32}, {64, 32}};
  for (auto& c : cfgs) {
    int r = dispatch_fold(ctx, c[0], c[1], P, nullptr, true);
    if (r != UB_OK) return r;
  }
  return UB_OK;
}

int conv3x3_fold(ub_ctx* ctx, const ConvLayer& Lr, const ConvArgs& a, cudaStream_t st) {
  FoldParams P{};
  P.src0 = a.src0; P.src1 = a.src1; P.out = a.out; P.w = Lr.w; P.bias = Lr.bias;
  P.B = a.B; P.kb0 = Lr.cin0 / Lr.kc; P.kbn = (Lr.cin0 + Lr.cin1) / Lr.kc;
  P.nch0 = Lr.cin0 / 8; P.nch1 = Lr.cin1 / 8;
  P.L = pp_layout(a.H, a.W);
  P.total_items = (long long)a.B * P.L.ntiles_fold;
  P.ncls = 0; P.logits = a.logits; P.pred = a.pred;
  if (a.head_w) {
    if (Lr.cout != 32) return ub_fail(ctx, UB_ERR_ARG, "fused head needs cout == 32");
    P.ncls = a.num_classes;
    // head weights travel in the kernel parameters (constant bank): copy from the host mirror
    memcpy(P.head, ctx->head_host, sizeof(float) * (size_t)a.num_classes * 33);
  }
  TimeScope ts(ctx, UB_T_CONV, st);
  return dispatch_fold(ctx, Lr.cout, Lr.kc, P, st, false);
}
