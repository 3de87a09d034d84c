// cols_kernels.cuh -- "column-walking" forms of the HBM-side stencils: a warp owns a strip of the image (one aligned
// 32-bit word of a row per lane) and walks DOWN it, keeping the vertical window in registers (blur) or in a lane-private
// shared-memory ring (morphology); horizontal neighbours are the neighbouring lanes' words (SHFL).  No block-level
// synchronisation, no staging of the tile in shared memory, every global access a full coalesced row segment, each
// input row loaded once per strip (the tiled kernels re-read a 4..24-row halo per 32-row tile and spend most of their
// issue slots moving the tile through shared memory three times).
//
//   k_blur5_cols   cv2.GaussianBlur(gray,(5,5),1.0)  (infer_two_stage_burr.py:85; oracle/cvnp.py gaussian_blur5)
//   k_morph_cols   cv2.dilate / erode / morphologyEx(CLOSE | OPEN) with a MORPH_ELLIPSE footprint on a 0 / non-zero
//                  u8 plane (infer_two_stage_burr.py:78-97; oracle/cvnp.py dilate / erode)
// Included by stage2.cu after pack32 / inside32 / store_bits_as_bytes.
#pragma once
#include "flood.cuh"

namespace s2 {

// ------------------------------------------------------------------------------------------------------------------
// 5x5 Gaussian, four pixels per lane.  Lane l of strip s owns column word cw = 30 s - 1 + l (pixels 4cw .. 4cw+3);
// lanes 1..30 write, lanes 0 / 31 only supply the two neighbouring pixels.  Vertical pass on (c0,c1),(c2,c3) u16x2
// lanes from a five-row register window (the loop is unrolled by five so the window never moves), horizontal pass as
// twelve 16x8-bit dot products with the rounding constant as the seed -- the arithmetic of k_edge_fused<2>.
// REFLECT_101: rows by index; columns in the V domain (the vertical pass is column-wise, so V of a reflected column is
// V of its source column): the word left of x = 0 is (V2, V1) of the first word, the word right of x = W-1 likewise
// of the last one.  Needs W % 4 == 0 and 4-byte aligned planes (edge_fused_ok).
// ------------------------------------------------------------------------------------------------------------------
constexpr int BC_VALID = 30;

__device__ __forceinline__ uint32_t bc_hpass(uint32_t a, uint32_t mx, uint32_t my, uint32_t n) {
  // a = (v[-2], v[-1]), mx = (v0, v1), my = (v2, v3), n = (v4, v5): pixel j needs v[j-2 .. j+2]
  const uint32_t qa = __funnelshift_r(a, mx, 16), qb = __funnelshift_r(mx, my, 16), qc = __funnelshift_r(my, n, 16);
  constexpr uint32_t K1 = 14u | (62u << 8), K2 = 104u | (62u << 8), K3 = 14u, K4 = 14u << 8;
  const uint32_t p0 = __dp2a_lo(my, K3, __dp2a_lo(mx, K2, __dp2a_lo(a, K1, 32768u)));
  const uint32_t p1 = __dp2a_lo(qc, K3, __dp2a_lo(qb, K2, __dp2a_lo(qa, K1, 32768u)));
  const uint32_t p2 = __dp2a_lo(n, K3, __dp2a_lo(my, K2, __dp2a_lo(mx, K1, 32768u)));
  const uint32_t p3 = __dp2a_lo(n, K4, __dp2a_lo(qc, K2, __dp2a_lo(qb, K1, 32768u)));
  return __byte_perm(__byte_perm(p0, p1, 0x0062), __byte_perm(p2, p3, 0x0062), 0x5410);     // bits 16..23 of each
}

__global__ void __launch_bounds__(256) k_blur5_cols(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                    int H, int W, int nstrips, int R) {
  const int lane = threadIdx.x, s = blockIdx.x * blockDim.y + threadIdx.y;
  if (s >= nstrips) return;                                   // whole warps leave; no block-level barrier below
  const int Ww = W >> 2;
  const int cw = BC_VALID * s - 1 + lane;
  const bool lb = cw == 0, rb = cw == Ww - 1;
  const bool wr = cw >= 0 && cw < Ww && lane >= 1 && lane <= BC_VALID;
  const int cwc = min(max(cw, 0), Ww - 1);                    // lanes outside the image load a valid word nobody uses
  const int y0 = blockIdx.y * R, y1 = min(y0 + R, H);
  const size_t fbase = (size_t)blockIdx.z * H * W;
  const uint32_t* img = reinterpret_cast<const uint32_t*>(src + fbase) + cwc;
  uint32_t* out = reinterpret_cast<uint32_t*>(dst + fbase) + cwc;
  const int refl = 2 * (H - 1);
  uint32_t lo[5], hi[5], cur[5];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t w = __ldg(img + (uint32_t)(reflect101(y0 - 2 + k, H) * Ww));
    lo[k] = __byte_perm(w, 0u, 0x4140);                       // (c0, c1) as u16x2
    hi[k] = __byte_perm(w, 0u, 0x4342);                       // (c2, c3)
  }
  // rows from y0 + 2 on are never above the image: REFLECT_101 is min(y, 2(H-1) - y) there (max 0: rows nobody uses)
#define BC_LD(y) __ldg(img + (uint32_t)(max(min((y), refl - (y)), 0) * Ww))
#pragma unroll
  for (int k = 0; k < 5; ++k) cur[k] = BC_LD(y0 + 2 + k);
  for (int yb = y0; yb < y1; yb += 5) {
    uint32_t nxt[5];
#pragma unroll
    for (int k = 0; k < 5; ++k) nxt[k] = BC_LD(yb + 7 + k);   // the next group's rows are in flight while this one computes
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      lo[(k + 4) % 5] = __byte_perm(cur[k], 0u, 0x4140);
      hi[(k + 4) % 5] = __byte_perm(cur[k], 0u, 0x4342);
      // rows y-2 .. y+2 sit in slots k, k+1, .. (mod 5); every lane <= 255 * 256: no carry between the u16 lanes
      const uint32_t ox = 14u * (lo[k % 5] + lo[(k + 4) % 5]) + 62u * (lo[(k + 1) % 5] + lo[(k + 3) % 5]) + 104u * lo[(k + 2) % 5];
      const uint32_t oy = 14u * (hi[k % 5] + hi[(k + 4) % 5]) + 62u * (hi[(k + 1) % 5] + hi[(k + 3) % 5]) + 104u * hi[(k + 2) % 5];
      uint32_t a = __shfl_up_sync(0xffffffffu, oy, 1), n = __shfl_down_sync(0xffffffffu, ox, 1);
      const uint32_t rfl = __byte_perm(ox, oy, 0x3254);       // (V2, V1)
      if (lb) a = rfl;
      if (rb) n = rfl;
      const uint32_t word = bc_hpass(a, ox, oy, n);
      const int y = yb + k;
      if (wr && y < y1) out[(uint32_t)(y * Ww)] = word;
    }
#pragma unroll
    for (int k = 0; k < 5; ++k) cur[k] = nxt[k];
  }
#undef BC_LD
}

// ------------------------------------------------------------------------------------------------------------------
// Binary morphology, 32 pixels (one bit-packed word) per lane.  Lane l of strip s owns row word wq = 30 s - 1 + l;
// lanes 1..30 write (the horizontal reach of one or two passes of a footprint up to 25 wide is <= 24 < 32 pixels, so
// what lane 0 / 31 get wrong for want of a neighbour never reaches lane 1 / 30).
// An operator is evaluated in SCATTER form: when input row r arrives, the word dilated / eroded horizontally by the run
// of footprint row j is OR-ed / AND-ed into the accumulator of output row r - dy_j; output row t is complete once input
// row t + bottom has been seen.  Accumulators live in a ring of RS rows x 32 lanes per warp in shared memory, each
// lane touching only its own column (lane-private: no synchronisation).  The footprint rows arrive sorted by run
// inclusion (ellipse runs are nested), so the horizontal pass extends one accumulated word run by run: one funnel
// shift + one logic op per footprint COLUMN, not per cell.  CLOSE / OPEN chain two operators through a second ring.
// Out-of-image samples are ignored as cv2 does: rows outside the image are skipped, columns outside are forced to the
// operator's identity (0 before a dilation, 1 before an erosion).
// ------------------------------------------------------------------------------------------------------------------
constexpr int MC_VALID = 30;

// 32 pixels (bytes) -> one word, bit i = pixel i non-zero: per four pixels three ops for "byte != 0" -> bit 7 of the byte,
// one multiply that gathers bits 7/15/23/31 into the top nibble, one shift, one multiply-add that places the nibble.
// The 32 bytes are loaded one row AHEAD of their use (McRaw travels through the loop), so the walk never waits on memory.
struct McRaw { uint4 a, b; };
__device__ __forceinline__ McRaw mc_load(const uint8_t* __restrict__ rowx) {       // rowx = &img[y][x], 16-byte aligned
  McRaw r;
  r.a = __ldg(reinterpret_cast<const uint4*>(rowx));
  r.b = __ldg(reinterpret_cast<const uint4*>(rowx + 16));
  return r;
}
__device__ __forceinline__ uint32_t mc_pack_raw(const McRaw& r) {
  const uint32_t v[8] = {r.a.x, r.a.y, r.a.z, r.a.w, r.b.x, r.b.y, r.b.z, r.b.w};
  uint32_t word = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const uint32_t t = (((v[k] & 0x7f7f7f7fu) + 0x7f7f7f7fu) | v[k]) & 0x80808080u;
    word = ((t * 0x00204081u) >> 28) * (1u << (4 * k)) + word;
  }
  return word;
}
// may this lane use mc_load on every row ?  (its word lies inside the row, rows keep the 16-byte alignment)
__device__ __forceinline__ bool mc_fast(const uint8_t* img, int W, int x, bool inw) {
  return inw && x + 32 <= W && (W & 15) == 0 && ((reinterpret_cast<uintptr_t>(img + x) & 15) == 0);
}

// cv2.getStructuringElement(MORPH_ELLIPSE, (K, K)) for the sizes the reference uses, anchor K/2, as
// rows (dy, xa, xb) ordered so that every run contains the one before it.  The host checks each table against
// make_footprint() + sort_footprint() before it picks a specialised kernel (fp_table_matches).
template <int K> struct FpT;
#define UB_FPT(K_, N_, TOP_, BOT_, DY, XA, XB)                                                              \
  template <> struct FpT<K_> {                                                                             \
    static constexpr int N = N_, TOP = TOP_, BOT = BOT_;                                                   \
    __host__ __device__ static constexpr int dy(int j) { constexpr int t[N_] = DY; return t[j]; }          \
    __host__ __device__ static constexpr int xa(int j) { constexpr int t[N_] = XA; return t[j]; }          \
    __host__ __device__ static constexpr int xb(int j) { constexpr int t[N_] = XB; return t[j]; }          \
  };
#define UB_L(...) {__VA_ARGS__}
UB_FPT(2, 2, 1, 0, UB_L(-1, 0), UB_L(0, -1), UB_L(0, 0))
UB_FPT(3, 3, 1, 1, UB_L(-1, 1, 0), UB_L(0, 0, -1), UB_L(0, 0, 1))
UB_FPT(5, 5, 2, 2, UB_L(-2, 2, -1, 0, 1), UB_L(0, 0, -2, -2, -2), UB_L(0, 0, 2, 2, 2))
UB_FPT(8, 8, 4, 3, UB_L(-4, -3, -2, 2, 3, -1, 0, 1), UB_L(0, -3, -3, -3, -3, -4, -4, -4), UB_L(0, 3, 3, 3, 3, 3, 3, 3))
UB_FPT(25, 25, 12, 12,
       UB_L(-12, 12, -11, 11, -10, 10, -9, 9, -8, 8, -7, 7, -6, 6, -5, 5, -4, 4, -3, 3, -2, 2, -1, 1, 0),
       UB_L(0, 0, -5, -5, -7, -7, -8, -8, -9, -9, -10, -10, -10, -10, -11, -11, -11, -11, -12, -12, -12, -12, -12, -12, -12),
       UB_L(0, 0, 5, 5, 7, 7, 8, 8, 9, 9, 10, 10, 10, 10, 11, 11, 11, 11, 12, 12, 12, 12, 12, 12, 12))
#undef UB_L
#undef UB_FPT
__host__ __device__ constexpr int mc_ring_rows(int span) { return span <= 4 ? 4 : (span <= 8 ? 8 : (span <= 16 ? 16 : 32)); }

// mc_scatter with the footprint known at compile time: every shift an immediate, no loop, no table look-up
template <bool DILATE, int K>
__device__ __forceinline__ void mc_scatter_t(uint32_t w, int lane, uint32_t* __restrict__ ring, int rmask, int yi, int tmin, int tmax) {
  using F = FpT<K>;
  const uint32_t ident = DILATE ? 0u : 0xffffffffu;
  uint32_t lo = __shfl_up_sync(0xffffffffu, w, 1), hi = __shfl_down_sync(0xffffffffu, w, 1);
  if (lane == 0) lo = ident;
  if (lane == 31) hi = ident;
  uint32_t acc = ident;
#pragma unroll
  for (int j = 0; j < F::N; ++j) {
#pragma unroll
    for (int s = F::xa(j); s <= F::xb(j); ++s)
      if (j == 0 || s < F::xa(j ? j - 1 : 0) || s > F::xb(j ? j - 1 : 0)) {
        const uint32_t v = s < 0 ? __funnelshift_r(lo, w, 32 + s) : (s > 0 ? __funnelshift_r(w, hi, s) : w);
        acc = DILATE ? (acc | v) : (acc & v);
      }
    const int t = yi - F::dy(j);
    if (t >= tmin && t <= tmax) {
      uint32_t* p = ring + ((t & rmask) << 5) + lane;
      *p = DILATE ? (*p | acc) : (*p & acc);
    }
  }
}

template <bool DILATE>
__device__ __forceinline__ void mc_scatter(uint32_t w, int lane, const Footprint& fp, uint32_t* __restrict__ ring, int rmask,
                                           int yi, int tmin, int tmax) {
  const uint32_t ident = DILATE ? 0u : 0xffffffffu;
  uint32_t lo = __shfl_up_sync(0xffffffffu, w, 1), hi = __shfl_down_sync(0xffffffffu, w, 1);
  if (lane == 0) lo = ident;
  if (lane == 31) hi = ident;
  uint32_t acc = ident;
  int ca = 1, cb = 0;                                          // columns accumulated so far: [ca, cb] (empty)
  for (int j = 0; j < fp.nrows; ++j) {
    const int xa = fp.xa[j], xb = fp.xb[j];
    int s = xa, e = cb < ca ? xb : ca - 1;                      // first the columns left of the accumulated run (or all of them)
    for (int pass = 0; pass < 2; ++pass) {
      for (; s <= e; ++s) {
        const uint32_t v = s < 0 ? __funnelshift_r(lo, w, 32 + s) : __funnelshift_r(w, hi, s);
        acc = DILATE ? (acc | v) : (acc & v);
      }
      if (cb < ca) break;
      s = cb + 1; e = xb;                                       // then the columns right of it
    }
    ca = xa; cb = xb;
    const int t = yi - fp.dy[j];
    if (t >= tmin && t <= tmax) {
      uint32_t* p = ring + ((t & rmask) << 5) + lane;
      *p = DILATE ? (*p | acc) : (*p & acc);
    }
  }
}

// OP: 0 dilate, 1 erode, 2 close (dilate -> erode), 3 open (erode -> dilate).  Dynamic shared memory:
// warps x (OP >= 2 ? 2 : 1) rings x RS x 32 words, RS a power of two >= top + bottom + 1.
// K: 0 = footprint given at run time, else one of the FpT sizes (then `fp` and RS are ignored)
template <int OP, int K>
__global__ void __launch_bounds__(256, 5) k_morph_cols(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int H, int W,
                                                    const __grid_constant__ Footprint fp, uint32_t value, int nstrips, int nblk, int R, int RSr) {
  constexpr bool TWO = OP >= 2;
  constexpr bool DIL1 = OP == 0 || OP == 2, DIL2 = OP == 3;
  extern __shared__ uint32_t mc_rings[];
  const int lane = threadIdx.x, wid = threadIdx.y;
  const int task = blockIdx.x * blockDim.y + wid;
  if (task >= nstrips * nblk) return;                           // whole warps leave; no block-level barrier below
  const int rb = task / nstrips, strip = task - rb * nstrips;
  const int wq = MC_VALID * strip - 1 + lane, x = 32 * wq;
  const int Wd = (W + 31) >> 5;
  const bool inw = wq >= 0 && wq < Wd;
  const uint32_t inx = inw ? (x + 32 <= W ? 0xffffffffu : (0xffffffffu >> (x + 32 - W))) : 0u;     // columns inside the image
  const bool wr = inw && lane >= 1 && lane <= MC_VALID;
  const uint32_t id1 = DIL1 ? 0u : 0xffffffffu, id2 = DIL2 ? 0u : 0xffffffffu;
  using F = FpT<K ? K : 3>;
  const int top = K ? F::TOP : fp.top, bot = K ? F::BOT : fp.bottom;
  const int RS = K ? mc_ring_rows(F::TOP + F::BOT + 1) : RSr, rmask = RS - 1;
  const int y0 = rb * R, y1 = min(y0 + R, H);
  uint32_t* A = mc_rings + (size_t)wid * (TWO ? 2 : 1) * RS * 32;
  uint32_t* Bq = A + RS * 32;
  for (int i = 0; i < RS; ++i) {
    A[(i << 5) + lane] = id1;
    if (TWO) Bq[(i << 5) + lane] = id2;
  }
  const size_t fbase = (size_t)blockIdx.z * H * W;
  const uint8_t* img = src + fbase;
  uint8_t* out = dst + fbase;
  // rows of the first operator's output that are needed, and the input rows those need
  const int tmin = TWO ? y0 - top : y0, tmax = TWO ? y1 - 1 + bot : y1 - 1;
  const bool fast = mc_fast(img, W, x, inw);
  const int ylast = min(tmax + bot, H - 1);
  McRaw nx = {};
  if (fast && max(tmin - top, 0) <= ylast) nx = mc_load(img + (size_t)max(tmin - top, 0) * W + x);
  for (int yi = tmin - top; yi <= tmax + bot; ++yi) {
    if (yi >= 0 && yi < H) {
      const McRaw cur = nx;
      if (fast && yi + 1 <= ylast) nx = mc_load(img + (size_t)(yi + 1) * W + x);
      uint32_t w = fast ? mc_pack_raw(cur) : (inw ? pack32(img, H, W, yi, x) : 0u);
      w = DIL1 ? w : (w | ~inx);
      if (K) mc_scatter_t<DIL1, K ? K : 3>(w, lane, A, rmask, yi, tmin, tmax);
      else mc_scatter<DIL1>(w, lane, fp, A, rmask, yi, tmin, tmax);
    }
    const int tm = yi - bot;                                    // this row of the first operator is complete now
    if (tm < tmin) continue;
    uint32_t* pa = A + ((tm & rmask) << 5) + lane;
    uint32_t m = *pa;
    *pa = id1;
    if (!TWO) {
      if (wr) store_bits_as_bytes(out + (size_t)tm * W + x, m & inx, W - x, value);
      continue;
    }
    if (tm >= 0 && tm < H) {
      m = DIL2 ? (m & inx) : (m | ~inx);
      if (K) mc_scatter_t<DIL2, K ? K : 3>(m, lane, Bq, rmask, tm, y0, y1 - 1);
      else mc_scatter<DIL2>(m, lane, fp, Bq, rmask, tm, y0, y1 - 1);
    }
    const int to = tm - bot;
    if (to < y0) continue;
    uint32_t* pb = Bq + ((to & rmask) << 5) + lane;
    const uint32_t o = *pb;
    *pb = id2;
    if (wr) store_bits_as_bytes(out + (size_t)to * W + x, o & inx, W - x, value);
  }
}

// ------------------------------------------------------------------------------------------------------------------
// The morphology chain of the burr stage (infer_two_stage_burr.py:78-97 / infer_enhanced_burr.py:82-116) as five chained
// scatter operators on one walk down the strip -- the column-walking form of k_burr_morph:
//   0  band = dilate(mask, E_band);              row t complete: c = band & ~mask[t] & edges[t]
//   1  dilate(c, E_close)      2  erode(., E_close)      3  erode(., E_open)      4  dilate(., E_open)
// Between operators the columns outside the image are forced to the next operator's identity and rows outside the image
// are skipped (cv2 ignores out-of-image samples).  Six rings per warp: the five accumulators and the packed mask rows
// (row t of the mask is needed again when row t of the band completes).  Total horizontal reach <= 18 pixels < one word,
// so lanes 1..30 are exact.  Output: the packed candidate rows `bits` and the union-find entries of their run starts
// (what k_cclb_init would produce from a u8 plane -- the plane itself is never written), or bytes {0,1} when out != null.
// ------------------------------------------------------------------------------------------------------------------
struct ChainRows { int lo[5], hi[5]; };      // rows of operator s's output that are needed

// KB / KC / KO: the three ellipse sizes at compile time (FpT), or all 0 = footprints given at run time in `mc`
template <int KB, int KC, int KO>
__global__ void __launch_bounds__(256) k_burr_chain_cols(const uint8_t* __restrict__ mask, const uint8_t* __restrict__ edges,
                                                         uint8_t* __restrict__ out, uint32_t* __restrict__ bits, int* __restrict__ lab_all,
                                                         int H, int W, const __grid_constant__ MorphChain mc, int nstrips, int nblk, int R, int RSr) {
  constexpr bool T = KB != 0;
  using FB = FpT<T ? KB : 3>; using FC = FpT<T ? KC : 3>; using FO = FpT<T ? KO : 3>;
  const int RS = T ? mc_ring_rows(FB::TOP + FB::BOT + 1) : RSr;
  const int top_s[5] = {T ? FB::TOP : mc.band.top, T ? FC::TOP : mc.close.top, T ? FC::TOP : mc.close.top, T ? FO::TOP : mc.open.top, T ? FO::TOP : mc.open.top};
  const int bot_s[5] = {T ? FB::BOT : mc.band.bottom, T ? FC::BOT : mc.close.bottom, T ? FC::BOT : mc.close.bottom, T ? FO::BOT : mc.open.bottom, T ? FO::BOT : mc.open.bottom};
  extern __shared__ uint32_t mc_rings[];
  const int lane = threadIdx.x, wid = threadIdx.y;
  const int task = blockIdx.x * blockDim.y + wid;
  if (task >= nstrips * nblk) return;                           // whole warps leave; no block-level barrier below
  const int rb = task / nstrips, strip = task - rb * nstrips;
  const int wq = MC_VALID * strip - 1 + lane, x = 32 * wq;
  const int Wd = (W + 31) >> 5;
  const bool inw = wq >= 0 && wq < Wd;
  const uint32_t inx = inw ? (x + 32 <= W ? 0xffffffffu : (0xffffffffu >> (x + 32 - W))) : 0u;
  const bool wr = inw && lane >= 1 && lane <= MC_VALID;
  const int rmask = RS - 1, y0 = rb * R, y1 = min(y0 + R, H);
  uint32_t* rings = mc_rings + (size_t)wid * 6 * RS * 32;      // ring r, row t, this lane: rings[((r * RS + (t & rmask)) << 5) + lane]
#define RG(r, t) rings[(((r) * RS + ((t) & rmask)) << 5) + lane]
  for (int i = 0; i < RS; ++i) {
    RG(0, i) = 0u; RG(1, i) = 0u; RG(2, i) = 0xffffffffu; RG(3, i) = 0xffffffffu; RG(4, i) = 0u; RG(5, i) = 0u;
  }
  int lo[5], hi[5];
  lo[4] = y0; hi[4] = y1 - 1;
#pragma unroll
  for (int s = 4; s > 0; --s) { lo[s - 1] = lo[s] - top_s[s]; hi[s - 1] = hi[s] + bot_s[s]; }
#define SC(DIL, KK, FPR, W_, R_, Y_, S_)                                                           \
  do {                                                                                              \
    if (T) mc_scatter_t<DIL, T ? KK : 3>(W_, lane, rings + (R_) * RS * 32, rmask, Y_, lo[S_], hi[S_]); \
    else mc_scatter<DIL>(W_, lane, FPR, rings + (R_) * RS * 32, rmask, Y_, lo[S_], hi[S_]);         \
  } while (0)
  const size_t fbase = (size_t)blockIdx.z * H * W;
  const uint8_t* mimg = mask + fbase;
  const uint8_t* eimg = edges + fbase;
  const bool fast = mc_fast(mimg, W, x, inw) && mc_fast(eimg, W, x, inw);
  const int ylast = min(hi[0] + bot_s[0], H - 1), tlast = min(hi[0], H - 1);
  McRaw nm = {}, ne = {};
  if (fast && max(lo[0] - top_s[0], 0) <= ylast) nm = mc_load(mimg + (size_t)max(lo[0] - top_s[0], 0) * W + x);
  if (fast && max(lo[0], 0) <= tlast) ne = mc_load(eimg + (size_t)max(lo[0], 0) * W + x);
  for (int yi = lo[0] - top_s[0]; yi <= hi[0] + bot_s[0]; ++yi) {
    if (yi >= 0 && yi < H) {
      const McRaw cur = nm;
      if (fast && yi + 1 <= ylast) nm = mc_load(mimg + (size_t)(yi + 1) * W + x);
      const uint32_t wm = fast ? mc_pack_raw(cur) : (inw ? pack32(mimg, H, W, yi, x) : 0u);
      RG(5, yi) = wm;
      SC(true, KB, mc.band, wm, 0, yi, 0);
    }
    int t = yi - bot_s[0];                                      // band row t complete
    if (t < lo[0]) continue;
    uint32_t v = RG(0, t);
    RG(0, t) = 0u;
    if (t >= 0 && t < H) {
      const McRaw cur = ne;
      if (fast && t + 1 <= tlast) ne = mc_load(eimg + (size_t)(t + 1) * W + x);
      const uint32_t we = fast ? mc_pack_raw(cur) : (inw ? pack32(eimg, H, W, t, x) : 0u);
      SC(true, KC, mc.close, v & ~RG(5, t) & we, 1, t, 1);
    }
    t -= bot_s[1];                                              // dilation of CLOSE, row t complete
    if (t < lo[1]) continue;
    v = RG(1, t);
    RG(1, t) = 0u;
    if (t >= 0 && t < H) SC(false, KC, mc.close, v | ~inx, 2, t, 2);
    t -= bot_s[2];                                              // erosion of CLOSE
    if (t < lo[2]) continue;
    v = RG(2, t);
    RG(2, t) = 0xffffffffu;
    if (t >= 0 && t < H) SC(false, KO, mc.open, v | ~inx, 3, t, 3);
    t -= bot_s[3];                                              // erosion of OPEN
    if (t < lo[3]) continue;
    v = RG(3, t);
    RG(3, t) = 0xffffffffu;
    if (t >= 0 && t < H) SC(true, KO, mc.open, v & inx, 4, t, 4);
    t -= bot_s[4];                                              // dilation of OPEN = the candidate row
    if (t < y0) continue;
    const uint32_t o = RG(4, t) & inx;
    RG(4, t) = 0u;
    if (!wr) continue;
    if (out) store_bits_as_bytes(out + fbase + (size_t)t * W + x, o, W - x, 1u);
    if (bits) {
      bits[((size_t)blockIdx.z * H + t) * Wd + wq] = o;
      if (lab_all) {                                            // union-find entries of the run starts (k_cclb_init's job)
        int* l = lab_all + fbase + (size_t)t * W + x;
        uint32_t c = o;
        while (c) {
          int a, len;
          c &= ~first_run(c, a, len);
          l[a] = t * W + x + a;
        }
      }
    }
  }
#undef RG
#undef SC
}

// ------------------------------------------------------------------------------------------------------------------
// Canny hysteresis (second half of cv2.Canny; oracle/cvnp.py canny: "every candidate whose 8-connected component of
// candidates holds a strong pixel") as a flood of the strong bits through the candidate bits, one CTA per frame with
// both bit planes of the frame in shared memory (448 x 800: 2 x 48.6 KB, zero-padded by one word / one row all round).
// A visit of a candidate word seeds it with its own kept bits and the 3x3 dilation of the kept bits around it and fills the
// whole runs of candidates that hold a seed (Kogge-Stone); a thread visits a segment of consecutive rows of one word
// column in order, down then up, in place (flood.cuh).  Sweeps repeat until none changes a word.  No labels, no atomics,
// no per-pixel scratch plane: 2 x P/8 bytes in, P bytes out.  Frames whose planes do not fit take the union-find path.
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) k_hyst_frame(const uint32_t* __restrict__ bits, const uint32_t* __restrict__ strong,
                                                     uint8_t* __restrict__ dst, int H, int W, int Wd) {
  extern __shared__ uint32_t hf_smem[];
  const int Wp = Wd + 2, NP = (H + 2) * Wp;
  uint32_t* C = hf_smem;
  volatile uint32_t* S = hf_smem + NP;
  const int tid = threadIdx.x, nthr = blockDim.x;
  for (int i = tid; i < 2 * NP; i += nthr) hf_smem[i] = 0u;
  __syncthreads();
  const size_t wb = (size_t)blockIdx.x * H * Wd;
  for (int i = tid; i < H * Wd; i += nthr) {
    const int r = i / Wd, w = i - r * Wd, idx = (r + 1) * Wp + w + 1;
    C[idx] = bits[wb + i];
    S[idx] = strong[wb + i];
  }
  __syncthreads();
  fl::flood<true>(C, S, H, Wd);
  uint8_t* out = dst + (size_t)blockIdx.x * H * W;
  for (int i = tid; i < H * Wd; i += nthr) {
    const int r = i / Wd, w = i - r * Wd;
    store_bits_as_bytes(out + (size_t)r * W + 32 * w, S[(r + 1) * Wp + w + 1], W - 32 * w, 255u);
  }
}

}  // namespace s2
