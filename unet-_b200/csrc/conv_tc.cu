// conv_tc.cu -- 3x3 conv (+folded BN bias, ReLU, optional fused 1x1 head + argmax) as an
// implicit GEMM on the 5th-gen tensor cores (tcgen05.mma, fp32 accumulators in TMEM),
// fed by 1-D bulk async copies (cp.async.bulk -> UBLKCP) with an mbarrier pipeline.
//
// Replaces ConvBlock.forward's conv+bn+relu (src/models/unetpp.py:23-26) and, for the
// last layer, `final` (unetpp.py:119) + softmax/argmax (infer_two_stage_burr.py:299-300).
//
// GEMM view (see common.cuh for the PP layout):
//   D[p, co] = sum_{tap, ci} X[p + off(tap), ci] * Wt[tap, ci, co]
//   M = 128 consecutive flat positions of one image, N = BN output channels,
//   K = 9 taps x Cin, walked in k-blocks of KC channels x 9 taps.
//   A operand = activations, K-major, no-swizzle core matrices (8 positions x 16 B):
//               rows 16 B apart inside a chunk plane, K chunks LBO apart.  With SHIFT_A the
//               three taps of a kernel row share ONE staged segment of 130(+6) positions
//               and differ only by +16 B in the descriptor start address.
//   B operand = weights, K-major, [tap][chunk][BN][8] in shared memory.
// Warp roles (192 threads): warps 0-3 epilogue (TMEM lane quarter = warp id),
// warp 4 bulk-copy producer (+TMEM alloc), warp 5 MMA issuer (one elected lane).
// Persistent CTAs, one per SM; TMEM accumulators double-buffered so the epilogue of item i
// overlaps the MMAs of item i+1.
#include <stdlib.h>
#include "tc_ptx.cuh"

namespace {
using namespace tcp;

constexpr int SEG_POS = 136;                 // staged positions per shifted segment (130 used)
constexpr int SEG_BYTES = SEG_POS * 16;      // 2176
constexpr int TILE_M = 128;

struct TcParams {
  const act_t* src0; const act_t* src1; act_t* out;
  const act_t* w; const float* bias;
  int B, kb0, kbn, nch0, nch1, cout, n_ntiles, num_stages, b_resident;
  long long total_items;
  PPLayout L;
  const float* head_w; const float* head_b; int ncls; float* logits; uint8_t* pred;
  unsigned long long* sat;     // non-null: count outputs clamped to the largest finite 16-bit value
};

template <int BN, int KC, bool SHIFT_A>
struct Cfg {
  static constexpr int KYS = BN >= 256 ? 1 : 3;                          // kernel rows per pipeline stage
  static constexpr int CPK = KC / 8;                                   // chunks per k-block
  static constexpr int A_SEGS = SHIFT_A ? 3 : 9;
  static constexpr int A_SEG_BYTES = SHIFT_A ? SEG_BYTES : TILE_M * 16;  // per chunk
  static constexpr int B_TAP = KC * BN * 2;
  static constexpr int B_KB = 9 * B_TAP;                                 // weights of one k-block (all taps)
  // One pipeline stage = KYS kernel rows of one k-block: their activation segment(s) and taps of weights.  N = 256:
  // one row per stage -- with whole k-blocks only two 87 KB stages fit and the tensor pipe waited for the refill of
  // the other one; 7 stages of 29 KB hide it better (250 -> 238 us per 309-GFLOP layer).  N <= 128: whole k-blocks --
  // the three MMAs of one row (~230 clk) are shorter than the producer lane's three bulk-copy issues (~100 clk each),
  // so row-sized stages made those layers producer-bound (253 -> 422 us).
  static constexpr int A_KY = (A_SEGS / 3) * CPK * A_SEG_BYTES;          // one kernel row
  static constexpr int B_KY = 3 * B_TAP;
  static constexpr int A_STAGE = KYS * A_KY;
  static constexpr int B_STAGE = KYS * B_KY;
  static constexpr int TMEM_COLS = (2 * BN < 32) ? 32 : 2 * BN;        // power of two for BN in {32..256}
  static constexpr int HDR = 2048;                                     // barriers + head weights
};

// Work item i of this CTA.  Unpaired: items (image, m-tile, n-tile) round-robin over the grid.  PAIR: the two CTAs of
// a cluster walk the SAME (n-tile, k-block) sequence on two neighbouring m-tiles, so each loads half of every weight
// stage and multicasts it to both (the deep layers stream 74 KB of weights per 13 KB of activations and were bound
// by L2->SM bandwidth, profiles/r01_conv.md).  An odd tail item is computed by both CTAs (identical stores).
template <int PAIR>
__device__ __forceinline__ bool tc_item(const TcParams& P, int i, uint32_t rank, int& nt, int& mt, int& b) {
  if (!PAIR) {
    const long long item = blockIdx.x + (long long)i * gridDim.x;
    if (item >= P.total_items) return false;
    nt = (int)(item % P.n_ntiles);
    const long long r = item / P.n_ntiles;
    mt = (int)(r % P.L.ntiles); b = (int)(r / P.L.ntiles);
    return true;
  }
  const long long M = (long long)P.B * P.L.ntiles, ppn = (M + 1) / 2;
  const long long j = (blockIdx.x >> 1) + (long long)i * (gridDim.x >> 1);
  if (j >= ppn * P.n_ntiles) return false;
  nt = (int)(j / ppn);
  long long m = 2 * (j - (long long)nt * ppn) + rank;
  if (m > M - 1) m = M - 1;
  mt = (int)(m % P.L.ntiles); b = (int)(m / P.L.ntiles);
  return true;
}

// PAIR: 0 = independent CTAs; 1 = clusters of two sharing every weight stage by multicast (M = 128 MMAs in each CTA);
// 2 = clusters of two driven by the leader's tcgen05.mma.cta_group::2: M = 256 over both CTAs' m-tiles, each CTA stages its
// own activation segments and HALF of the N weight rows (ConvLayer::w2 layout), so a weight stage costs each SM half the
// shared-memory fill and half the operand reads per FLOP.  1-D bulk copies cannot signal a barrier in the peer CTA, so the
// peer's (otherwise idle) MMA warp relays: it waits for the local `full` barrier and arrives on the leader's `pfull`.
template <int BN, int KC, bool SHIFT_A, int PAIR>
__global__ void __launch_bounds__(192, 1) k_conv_tc(const __grid_constant__ TcParams P) {
  using C = Cfg<BN, KC, SHIFT_A>;
  constexpr bool CG2 = PAIR == 2;
  constexpr int B_ST = CG2 ? C::B_STAGE / 2 : C::B_STAGE;              // bytes of weights per stage in THIS CTA
  const uint32_t cta_rank = PAIR ? cluster_ctarank() : 0;
  extern __shared__ __align__(128) uint8_t smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int S = P.num_stages;
  // header: full[S] | empty[S] | tfull[2] | tempty[2] | tmem ptr | head weights (fp32)
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  const uint32_t bar_full = smem_u32(bars), bar_empty = bar_full + 8 * S;
  const uint32_t bar_tfull = bar_empty + 8 * S, bar_tempty = bar_tfull + 16;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 8 * (2 * S + 4));
  float* s_head = reinterpret_cast<float*>(smem + 512);               // [ncls*32 + ncls] <= 8*33
  [[maybe_unused]] const uint32_t bar_pfull = smem_u32(smem) + 1600;   // CG2 leader: the peer's stage has landed (S <= 24)
  const uint32_t a_base = smem_u32(smem) + C::HDR;
  const uint32_t b_base = a_base + S * C::A_STAGE;
  const PPLayout& L = P.L;

  if (threadIdx.x == 0) {
    // PAIR: a stage is free when the MMAs of BOTH CTAs have retired (the peer writes half of it)
    for (int i = 0; i < S; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, PAIR == 1 ? 2 : 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_tfull + 8 * i, 1); mbar_init(bar_tempty + 8 * i, CG2 ? 8 : 4); }
    if (CG2) for (int i = 0; i < S; ++i) mbar_init(bar_pfull + 8 * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (P.head_w != nullptr) {
    const int n = P.ncls * 32;
    for (int i = threadIdx.x; i < P.ncls * 33; i += blockDim.x)
      s_head[i] = i < n ? P.head_w[i] : P.head_b[i - n];
  }
  if (warp == 4) {
    if (CG2) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(C::TMEM_COLS));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(C::TMEM_COLS));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (PAIR) cluster_sync();                  // the peer's barriers exist before anything is multicast to them
  const uint32_t tmem_base = *tmem_slot;
  [[maybe_unused]] const uint32_t lead_tempty = CG2 ? mapa_u32(bar_tempty, 0) : bar_tempty;

  if (warp == 4) {
    // ===================== producer: contiguous bulk copies, issued by ONE lane =====================
    // (UBLKCP takes uniform operands: per-lane addresses would be serialised by a divergent loop that
    //  costs ~100 cycles per copy -- profiles/r01_conv.md)
    if (lane == 0) {
      int s = 0; uint32_t ph = 0; bool first = true;
      int nt, mt, b;
      for (int i = 0; tc_item<PAIR>(P, i, cta_rank, nt, mt, b); ++i) {
        const int p0 = L.G + mt * TILE_M;
        const bool load_b = !P.b_resident || first;
        for (int kb = 0; kb < P.kbn; ++kb) {
          const act_t* plane0 =
              kb < P.kb0 ? P.src0 + ((size_t)b * P.nch0 + (size_t)kb * C::CPK) * L.Ppl * 8
                         : P.src1 + ((size_t)b * P.nch1 + (size_t)(kb - P.kb0) * C::CPK) * L.Ppl * 8;
          const act_t* wsrc_kb = P.w + ((size_t)nt * P.kbn + kb) * 9 * (size_t)(KC * BN);
#pragma unroll 1
          for (int ky0 = 0; ky0 < 3; ky0 += C::KYS) {
            const uint32_t full = bar_full + 8 * s;
            mbar_wait(bar_empty + 8 * s, ph ^ 1);
            mbar_expect_tx(full, C::A_STAGE + (load_b ? B_ST : 0));
            const uint32_t a_st = a_base + s * C::A_STAGE;
#pragma unroll
            for (int kyi = 0; kyi < C::KYS; ++kyi)
#pragma unroll
              for (int sg = 0; sg < C::A_SEGS / 3; ++sg) {
                const int ky = ky0 + kyi;
                const int off = SHIFT_A ? (ky - 1) * L.pitch - 1 : (ky - 1) * L.pitch + (sg - 1);
#pragma unroll
                for (int cc = 0; cc < C::CPK; ++cc)
                  bulk_g2s(a_st + kyi * C::A_KY + (sg * C::CPK + cc) * C::A_SEG_BYTES, plane0 + ((size_t)cc * L.Ppl + p0 + off) * 8,
                           C::A_SEG_BYTES, full);
              }
            if (load_b) {
              if constexpr (CG2) {
                // w2: [n-tile][k-block][stage][rank][...]: this CTA's half of the stage is one contiguous piece
                const size_t st_idx = ((size_t)nt * P.kbn + kb) * (3 / C::KYS) + ky0 / C::KYS;
                bulk_g2s(b_base + (P.b_resident ? (kb * (3 / C::KYS) + ky0 / C::KYS) : s) * B_ST,
                         P.w + (st_idx * 2 + cta_rank) * (size_t)(B_ST / 2), B_ST, full);
              } else {
                const act_t* wsrc = wsrc_kb + (size_t)ky0 * 3 * (KC * BN);
                const uint32_t wdst = b_base + (P.b_resident ? kb * C::B_KB + ky0 * C::B_KY : s * C::B_STAGE);
                if (PAIR) bulk_g2s_mcast(wdst + cta_rank * (C::B_STAGE / 2), wsrc + cta_rank * (C::B_STAGE / 4), C::B_STAGE / 2, full, 3);
                else bulk_g2s(wdst, wsrc, C::B_STAGE, full);
              }
            }
            if (++s == S) { s = 0; ph ^= 1; }
          }
        }
        first = false;
      }
    }
  } else if (CG2 && warp == 5 && cta_rank != 0) {
    // ===================== CG2 peer: relay "my stage has landed" to the leader, whose MMAs read it =====================
    const uint32_t lead_pfull = mapa_u32(bar_pfull, 0);
    int s = 0; uint32_t ph = 0;
    int nt, mt, b;
    for (int it = 0; tc_item<PAIR>(P, it, cta_rank, nt, mt, b); ++it)
      for (int i = 0; i < P.kbn * (3 / C::KYS); ++i) {
        mbar_wait(bar_full + 8 * s, ph);
        if (lane == 0) mbar_arrive_cluster(lead_pfull + 8 * s);
        __syncwarp();
        if (++s == S) { s = 0; ph ^= 1; }
      }
  } else if (warp == 5) {
    // ===================== MMA issuer: the whole warp walks the loop, one elected lane issues =====================
    // (under a plain `lane == 0` branch every UTCHMMA is wrapped in an elect-and-branch loop of ~16 dependent
    //  uniform-datapath instructions, and this warp paces the tensor pipe)
    {
      constexpr uint32_t IDESC = (1u << 4) | (UB_MMA_FMT << 7) | (UB_MMA_FMT << 10) | ((uint32_t)(BN >> 3) << 17) |
                                 ((uint32_t)((CG2 ? 2 * TILE_M : TILE_M) >> 4) << 24);
      constexpr int BNH = CG2 ? BN / 2 : BN;                      // rows of B in this CTA's shared memory
      int s = 0; uint32_t ph = 0;
      int nt, mt, b;
      for (int it = 0; tc_item<PAIR>(P, it, cta_rank, nt, mt, b); ++it) {
        const int buf = it & 1;
        mbar_wait(bar_tempty + 8 * buf, ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        for (int kb = 0; kb < P.kbn; ++kb) {
#pragma unroll 1
          for (int ky0 = 0; ky0 < 3; ky0 += C::KYS) {
            mbar_wait(bar_full + 8 * s, ph);
            if (CG2) mbar_wait(bar_pfull + 8 * s, ph);     // (a cluster-scope acquire here costs ~500 clk per stage in the issuing thread: 2.4x slower layers)
            tc_fence_after();
            const uint32_t a_st = a_base + s * C::A_STAGE;
            const uint32_t b_st = CG2 ? b_base + (P.b_resident ? (kb * (3 / C::KYS) + ky0 / C::KYS) : s) * B_ST
                                      : b_base + (P.b_resident ? kb * C::B_KB + ky0 * C::B_KY : s * C::B_STAGE);
            if (elect_one()) {
#pragma unroll
              for (int kyi = 0; kyi < C::KYS; ++kyi)
#pragma unroll
                for (int kx = 0; kx < 3; ++kx) {
                  const uint32_t a_tap = a_st + kyi * C::A_KY + (SHIFT_A ? kx * 16 : kx * C::CPK * C::A_SEG_BYTES);
#pragma unroll
                  for (int k16 = 0; k16 < KC / 16; ++k16) {
                    const uint64_t ad = make_desc(a_tap + 2 * k16 * C::A_SEG_BYTES, C::A_SEG_BYTES, 128);
                    const uint64_t bd = make_desc(b_st + (kyi * 3 + kx) * (KC * BNH * 2) + 2 * k16 * BNH * 16, BNH * 16, 128);
                    if (CG2) tc_mma_pair(d_tmem, ad, bd, IDESC, (kb | ky0 | kyi | kx | k16) != 0);
                    else tc_mma(d_tmem, ad, bd, IDESC, (kb | ky0 | kyi | kx | k16) != 0);
                  }
                }
              if (CG2) tc_commit_pair(bar_empty + 8 * s, 3);     // frees the stage in both CTAs when these MMAs retire
              else if (PAIR) tc_commit_mcast(bar_empty + 8 * s, 3);
              else tc_commit(bar_empty + 8 * s);
              if (kb == P.kbn - 1 && ky0 + C::KYS == 3) {        // accumulator ready for the epilogue (CG2: of both CTAs)
                if (CG2) tc_commit_pair(bar_tfull + 8 * buf, 3);
                else tc_commit(bar_tfull + 8 * buf);
              }
            }
            __syncwarp();
            if (++s == S) { s = 0; ph ^= 1; }
          }
        }
      }
    }
  } else {
    // ===================== epilogue: TMEM -> registers -> global =====================
    const int row = warp * 32 + lane;
    const int nco_chunks = P.cout / 8;
    int nt, mt, b;
    for (int it = 0; tc_item<PAIR>(P, it, cta_rank, nt, mt, b); ++it) {
      const int buf = it & 1;
      const int q = mt * TILE_M + row;
      const int yy = q / L.pitch, xx = q - yy * L.pitch;
      const bool interior = (q < L.npos) && yy >= 1 && yy <= L.H && xx >= 1 && xx <= L.W;
      mbar_wait(bar_tfull + 8 * buf, (it >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + buf * BN;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t v[32];
        TMEM_LD32(taddr + c0, v);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        const int co0 = nt * BN + c0;
        float f[32];
#pragma unroll
        for (int j = 0; j < 32; ++j)
          f[j] = interior ? fminf(fmaxf(__uint_as_float(v[j]) + __ldg(P.bias + co0 + j), 0.f), UB_ACT_MAX) : 0.f;
        if (P.sat != nullptr && P.head_w == nullptr) {         // diagnostic: outputs that hit the largest finite value
          int cnt = 0;
#pragma unroll
          for (int j = 0; j < 32; ++j) cnt += f[j] >= UB_ACT_MAX;
          cnt = __reduce_add_sync(0xffffffffu, cnt);
          if (lane == 0 && cnt) atomicAdd(P.sat, (unsigned long long)cnt);
        }
        if (P.head_w != nullptr) {
          if (interior) {
            float lg[8];
            int best = 0;
            const size_t hw = (size_t)L.H * L.W, pix = (size_t)(yy - 1) * L.W + (xx - 1);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              if (c < P.ncls) {
                float acc = s_head[P.ncls * 32 + c];
#pragma unroll
                for (int j = 0; j < 32; ++j) acc = fmaf(s_head[c * 32 + j], f[j], acc);
                lg[c] = acc;
                if (P.logits) P.logits[((size_t)b * P.ncls + c) * hw + pix] = acc;
                if (c > 0 && acc > lg[best]) best = c;
              }
            }
            if (P.pred) P.pred[(size_t)b * hw + pix] = (uint8_t)best;
          }
        } else {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            act2_t h[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) h[j] = f2a2(f[g * 8 + 2 * j], f[g * 8 + 2 * j + 1]);
            *reinterpret_cast<uint4*>(P.out + (((size_t)b * nco_chunks + (co0 >> 3) + g) * L.Ppl + L.G + q) * 8) =
                *reinterpret_cast<uint4*>(h);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CG2) mbar_arrive_cluster(lead_tempty + 8 * buf); else mbar_arrive(bar_tempty + 8 * buf);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync();                  // no CTA leaves while its peer may still multicast to it
  if (warp == 4) {
    tc_fence_after();
    if (CG2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(C::TMEM_COLS));
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(C::TMEM_COLS));
  }
}

constexpr int SMEM_BUDGET = 227 * 1024;

// UB_TC_PAIR: 0 = no clusters, 1 = multicast weight stages, 2 (default) = cta_group::2 MMAs over CTA pairs
int tc_pair_mode() {
  static const int mode = getenv("UB_TC_PAIR") ? atoi(getenv("UB_TC_PAIR")) : 2;
  return mode;
}

template <int BN, int KC, bool SHIFT_A>
int launch_cfg(ub_ctx* ctx, TcParams& P, const act_t* w2, cudaStream_t st, bool init_only) {
  using C = Cfg<BN, KC, SHIFT_A>;
  auto kern = k_conv_tc<BN, KC, SHIFT_A, 0>;
  auto kern_pair = k_conv_tc<BN, KC, SHIFT_A, 1>;
  auto kern_cg2 = k_conv_tc<BN, KC, SHIFT_A, 2>;
  if (init_only) {
    UB_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET));
    UB_CUDA(ctx, cudaFuncSetAttribute(kern_pair, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET));
    UB_CUDA(ctx, cudaFuncSetAttribute(kern_cg2, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET));
    return UB_OK;
  }
  const long long M = (long long)P.B * P.L.ntiles;
  const bool cg2 = tc_pair_mode() == 2 && w2 != nullptr && BN >= 64 && M >= 2 && ctx->sm_count >= 2;
  const int b_stage = cg2 ? C::B_STAGE / 2 : C::B_STAGE;            // per CTA
  // weights resident in shared memory when they fit beside >= 3 activation stages
  const int resident_bytes = P.kbn * (cg2 ? C::B_KB / 2 : C::B_KB);
  int stages;
  if (P.n_ntiles == 1 && C::HDR + resident_bytes + 3 * (3 / C::KYS) * C::A_STAGE <= SMEM_BUDGET) {
    P.b_resident = 1;
    stages = (SMEM_BUDGET - C::HDR - resident_bytes) / C::A_STAGE;
  } else {
    P.b_resident = 0;
    stages = (SMEM_BUDGET - C::HDR) / (C::A_STAGE + b_stage);
  }
  if (stages > 24) stages = 24;                // barrier header: 8 * (2 * S + 4) bytes below the head weights at +512
  if (stages < 2) return ub_fail(ctx, UB_ERR_ARG, "conv_tc: BN=%d KC=%d does not fit 2 stages", BN, KC);
  P.num_stages = stages;
  const int smem = C::HDR + stages * C::A_STAGE + (P.b_resident ? resident_bytes : stages * b_stage);
  // CTA pairs (clusters of 2): cta_group::2 MMAs, or (streamed weights only) every weight stage shared by multicast
  if (cg2 || (!P.b_resident && tc_pair_mode() == 1 && M >= 2 && ctx->sm_count >= 2)) {
    const long long pairs = (M + 1) / 2 * P.n_ntiles, max_pairs = ctx->sm_count / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * (pairs < max_pairs ? pairs : max_pairs)));
    cfg.blockDim = dim3(192);
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (cg2) { P.w = w2; UB_CUDA(ctx, cudaLaunchKernelEx(&cfg, kern_cg2, P)); }
    else UB_CUDA(ctx, cudaLaunchKernelEx(&cfg, kern_pair, P));
    UB_CHECK_LAUNCH(ctx);
    return UB_OK;
  }
  long long grid = P.total_items < ctx->sm_count ? P.total_items : ctx->sm_count;
  kern<<<(unsigned)grid, 192, smem, st>>>(P);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

template <bool SHIFT_A>
int dispatch(ub_ctx* ctx, int bn, int kc, TcParams& P, const act_t* w2, cudaStream_t st, bool init_only) {
#define UB_CASE(BN_, KC_) \
  if (bn == BN_ && kc == KC_) return launch_cfg<BN_, KC_, SHIFT_A>(ctx, P, w2, st, init_only);
  UB_CASE(32, 16) UB_CASE(32, 32) UB_CASE(64, 16) UB_CASE(64, 32) UB_CASE(128, 16) UB_CASE(128, 32)
  UB_CASE(256, 16)
#undef UB_CASE
  return ub_fail(ctx, UB_ERR_ARG, "conv_tc: unsupported tile BN=%d KC=%d", bn, kc);
}

}  // namespace

static const int kTileTable[][2] = {{32, 16}, {32, 32}, {64, 16}, {64, 32}, {128, 16}, {128, 32}, {256, 16}};

int conv_tc_init(ub_ctx* ctx) {
  TcParams P{};
  for (auto& t : kTileTable) {
    int r = dispatch<true>(ctx, t[0], t[1], P, nullptr, nullptr, true);
    if (r != UB_OK) return r;
    r = dispatch<false>(ctx, t[0], t[1], P, nullptr, nullptr, true);
    if (r != UB_OK) return r;
  }
  return UB_OK;
}

int conv3x3_tc(ub_ctx* ctx, const ConvLayer& Lr, const ConvArgs& a, cudaStream_t st) {
  TcParams P{};
  P.src0 = a.src0; P.src1 = a.src1; P.out = a.out; P.w = Lr.w; P.bias = Lr.bias;
  P.B = a.B; P.kb0 = Lr.cin0 / Lr.kc; P.kbn = (Lr.cin0 + Lr.cin1) / Lr.kc;
  P.nch0 = Lr.cin0 / 8; P.nch1 = Lr.cin1 / 8; P.cout = Lr.cout; P.n_ntiles = Lr.cout / Lr.bn;
  P.L = pp_layout(a.H, a.W);
  P.total_items = (long long)a.B * P.L.ntiles * P.n_ntiles;
  P.head_w = a.head_w; P.head_b = a.head_b; P.ncls = a.num_classes; P.logits = a.logits; P.pred = a.pred; P.sat = a.sat;
  if (a.head_w && (Lr.cout != 32 || Lr.bn != 32)) return ub_fail(ctx, UB_ERR_ARG, "fused head needs cout == bn == 32");
  TimeScope ts(ctx, UB_T_CONV, st);
  // conv_impl 0: shifted-descriptor A operand (3 staged segments); 2: nine aligned segments
  // (the cta_group::2 layout of the weights exists for the shifted-descriptor kernel only)
  return ctx->conv_impl == 2 ? dispatch<false>(ctx, Lr.bn, Lr.kc, P, nullptr, st, false)
                             : dispatch<true>(ctx, Lr.bn, Lr.kc, P, a.head_w ? nullptr : Lr.w2, st, false);
}
