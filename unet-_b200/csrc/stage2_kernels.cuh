// stage2_kernels.cuh -- integer stencil / morphology / labelling kernels of the burr stage.
// Every kernel is the device form of one function in oracle/cvnp.py (the bit-exact
// restatement of the OpenCV call the reference makes); batched over frames (blockIdx.z).
#pragma once
#include "common.cuh"

namespace s2 {

__device__ __forceinline__ int reflect101(int i, int n) {
  if (i < 0) i = -i;
  if (i >= n) i = 2 * (n - 1) - i;
  return i < 0 ? 0 : (i >= n ? n - 1 : i);     // clamp only matters for n < 3
}
__device__ __forceinline__ int clampi(int i, int n) { return i < 0 ? 0 : (i >= n ? n - 1 : i); }

// ------------------------------------------------------------------------------
// cv2.GaussianBlur(gray,(5,5),1.0) (infer_two_stage_burr.py:85): separable fixed-point
// [14,62,104,62,14], REFLECT_101, one rounding (acc + 2^15) >> 16.
// Tile 64x16 outputs per block of 256 threads; shared-memory staged, halo 2.
// ------------------------------------------------------------------------------
constexpr int GT_W = 64, GT_H = 16;           // blocks are (64, 4) threads: column = threadIdx.x, rows strided by 4 (no div/mod)
__global__ void __launch_bounds__(256) k_gauss5(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int H, int W) {
  __shared__ uint8_t s_in[GT_H + 4][GT_W + 4];
  __shared__ int s_h[GT_H + 4][GT_W];
  const int x0 = blockIdx.x * GT_W, y0 = blockIdx.y * GT_H;
  const int tx = threadIdx.x, ty = threadIdx.y;
  const uint8_t* img = src + (size_t)blockIdx.z * H * W;
  const int gx0 = reflect101(x0 + tx - 2, W), gx1 = reflect101(x0 + tx + GT_W - 2, W);
  for (int r = ty; r < GT_H + 4; r += 4) {
    const uint8_t* row = img + (size_t)reflect101(y0 + r - 2, H) * W;
    s_in[r][tx] = row[gx0];
    if (tx < 4) s_in[r][tx + GT_W] = row[gx1];
  }
  __syncthreads();
  for (int r = ty; r < GT_H + 4; r += 4)
    s_h[r][tx] = 14 * (s_in[r][tx] + s_in[r][tx + 4]) + 62 * (s_in[r][tx + 1] + s_in[r][tx + 3]) + 104 * s_in[r][tx + 2];
  __syncthreads();
  const int x = x0 + tx;
  if (x < W)
    for (int r = ty; r < GT_H; r += 4) {
      const int y = y0 + r;
      if (y < H) {
        const int acc = 14 * (s_h[r][tx] + s_h[r + 4][tx]) + 62 * (s_h[r + 1][tx] + s_h[r + 3][tx]) + 104 * s_h[r + 2][tx];
        dst[(size_t)blockIdx.z * H * W + (size_t)y * W + x] = (uint8_t)((acc + (1 << 15)) >> 16);
      }
    }
}

// ------------------------------------------------------------------------------
// First half of cv2.Canny(img, low, high) (aperture 3, L1 norm): Sobel with
// BORDER_REPLICATE, |gx|+|gy| (0 outside the image), fixed-point NMS (TG22 = 13573 <<15).
// Output: 0 none, 1 weak candidate, 2 strong.  (oracle/cvnp.py: canny_candidates)
// ------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_canny_nms(const uint8_t* __restrict__ src, uint8_t* __restrict__ cand,
                                                   int H, int W, int low, int high) {
  __shared__ uint8_t s_in[GT_H + 4][GT_W + 4];
  __shared__ int s_mag[GT_H + 2][GT_W + 2];
  __shared__ short s_gx[GT_H + 2][GT_W + 2], s_gy[GT_H + 2][GT_W + 2];
  const int x0 = blockIdx.x * GT_W, y0 = blockIdx.y * GT_H;
  const int tx = threadIdx.x, ty = threadIdx.y;
  const uint8_t* img = src + (size_t)blockIdx.z * H * W;
  const int gx0 = clampi(x0 + tx - 2, W), gx1 = clampi(x0 + tx + GT_W - 2, W);
  for (int r = ty; r < GT_H + 4; r += 4) {
    const uint8_t* row = img + (size_t)clampi(y0 + r - 2, H) * W;
    s_in[r][tx] = row[gx0];
    if (tx < 4) s_in[r][tx + GT_W] = row[gx1];
  }
  __syncthreads();
  // magnitude / gradients at the (GT_H+2) x (GT_W+2) samples around the tile; sample (r, c) = image (y0+r-1, x0+c-1)
  for (int r = ty; r < GT_H + 2; r += 4)
    for (int c = tx; c < GT_W + 2; c += GT_W) {
      const int x = x0 + c - 1, y = y0 + r - 1;
      int m = 0, gx = 0, gy = 0;
      if (x >= 0 && x < W && y >= 0 && y < H) {
        const int a = s_in[r][c], b = s_in[r][c + 1], cc = s_in[r][c + 2];
        const int d = s_in[r + 1][c], f = s_in[r + 1][c + 2];
        const int g = s_in[r + 2][c], h = s_in[r + 2][c + 1], k = s_in[r + 2][c + 2];
        gx = (cc + 2 * f + k) - (a + 2 * d + g);
        gy = (g + 2 * h + k) - (a + 2 * b + cc);
        m = abs(gx) + abs(gy);
      }
      s_mag[r][c] = m; s_gx[r][c] = (short)gx; s_gy[r][c] = (short)gy;
    }
  __syncthreads();
  const int x = x0 + tx;
  if (x >= W) return;
  for (int r = ty; r < GT_H; r += 4) {
    const int y = y0 + r;
    if (y >= H) break;
    const int m = s_mag[r + 1][tx + 1];
    uint8_t out = 0;
    if (m > low) {
      const int gx = s_gx[r + 1][tx + 1], gy = s_gy[r + 1][tx + 1];
      // |gx|, |gy| <= 1020: y << 15 and x * 13573 + (x << 16) fit 32 bits
      const int ax = abs(gx), ay = abs(gy) << 15;
      const int tg22 = ax * 13573, tg67 = tg22 + (ax << 16);
      bool keep;
      if (ay < tg22) keep = m > s_mag[r + 1][tx] && m >= s_mag[r + 1][tx + 2];
      else if (ay > tg67) keep = m > s_mag[r][tx + 1] && m >= s_mag[r + 2][tx + 1];
      else {
        const int sg = ((gx ^ gy) < 0) ? -1 : 1;
        keep = m > s_mag[r][tx + 1 - sg] && m > s_mag[r + 2][tx + 1 + sg];
      }
      if (keep) out = m > high ? 2 : 1;
    }
    cand[(size_t)blockIdx.z * H * W + (size_t)y * W + x] = out;
  }
}

// ------------------------------------------------------------------------------
// Binary morphology with a cv2 MORPH_ELLIPSE footprint given as row runs.
// The tile is bit-packed in shared memory (one ballot per 32 pixels); a footprint row is
// then a <=32-bit window test on a 64-bit funnel of two packed words.
// Out-of-image samples are ignored (identity: 0 for dilate, 1 for erode).
// ------------------------------------------------------------------------------
struct Footprint {
  int nrows;
  int8_t dy[25], xa[25], xb[25];      // row r covers offsets x+xa[r] .. x+xb[r] at y+dy[r]
  int top, bottom;                    // max upward / downward reach (>= 0)
};

constexpr int MT_W = 128, MT_H = 16, M_HALO = 12;     // halo covers footprints up to 25x25
constexpr int M_WORDS = (MT_W + 64) / 32;             // packed words per tile row: [x0-32, x0+MT_W+32)
constexpr int M_ROWS = MT_H + 2 * M_HALO;

// EXTRA: 0 none; 1: dst = (result && !mask) (detection band, infer_two_stage_burr.py:81)
template <bool DILATE>
__global__ void __launch_bounds__(256) k_morph(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int H, int W,
                                               Footprint fp, int value, const uint8_t* __restrict__ notmask) {
  __shared__ uint32_t s_bits[M_ROWS][M_WORDS + 1];   // +1: funnel reads word wd+1
  const int x0 = blockIdx.x * MT_W, y0 = blockIdx.y * MT_H;
  const size_t base = (size_t)blockIdx.z * H * W;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r_lo = M_HALO - fp.top, r_hi = M_HALO + MT_H + fp.bottom;     // rows actually needed
  for (int t = warp; t < (r_hi - r_lo) * M_WORDS; t += 8) {
    int row = r_lo + t / M_WORDS, wd = t % M_WORDS;
    int y = y0 + row - M_HALO, x = x0 - 32 + wd * 32 + lane;
    bool in = y >= 0 && y < H && x >= 0 && x < W;
    bool bit = in ? (src[base + (size_t)y * W + x] != 0) : !DILATE;
    uint32_t m = __ballot_sync(0xffffffffu, bit);
    if (lane == 0) s_bits[row][wd] = m;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < MT_W * MT_H; i += 256) {
    int ty = i / MT_W, tx = i - ty * MT_W;
    int x = x0 + tx, y = y0 + ty;
    if (x >= W || y >= H) continue;
    bool res = !DILATE;
    for (int r = 0; r < fp.nrows; ++r) {
      int row = ty + M_HALO + fp.dy[r];
      int start = tx + 32 + fp.xa[r];                    // bit index in the packed row
      int len = fp.xb[r] - fp.xa[r] + 1;
      int wd = start >> 5, sh = start & 31;
      uint64_t w64 = ((uint64_t)s_bits[row][wd + 1] << 32) | s_bits[row][wd];
      uint32_t win = (uint32_t)(w64 >> sh);
      uint32_t msk = len >= 32 ? 0xffffffffu : ((1u << len) - 1u);
      if (DILATE) { if (win & msk) { res = true; break; } }
      else        { if ((win & msk) != msk) { res = false; break; } }
    }
    size_t o = base + (size_t)y * W + x;
    if (notmask) res = res && (notmask[o] == 0);
    dst[o] = res ? (uint8_t)value : 0;
  }
}

// ------------------------------------------------------------------------------
// The whole morphology chain of the burr stage in ONE kernel on bit-packed rows
// (infer_two_stage_burr.py:78-97 / infer_enhanced_burr.py:82-116):
//   band = dilate(mask, E_band) & ~mask;  cand = edges & band;
//   cand = erode(dilate(cand, E_close), E_close);  cand = dilate(erode(cand, E_open), E_open)
// One thread owns a 32-pixel word: a footprint row [xa, xb] is an OR / AND of funnel shifts of
// the 96-bit window (w-1, w, w+1), so an operator costs ~2 instructions per footprint cell per
// 32 pixels instead of a per-pixel loop, and the five intermediate images never leave shared
// memory (six launches with 46 MB of traffic each before).  Out-of-image samples are ignored
// as cv2 does: forced to 0 before a dilation and to 1 before an erosion.
// Tile = 128 x 32 pixels, halo FM_HALO rows / 32 columns: enough for 2*(open/2) + 2*(close/2)
// + band/2 <= FM_HALO (standard 8, enhanced 18); other parameter sets take the unfused path.
// ------------------------------------------------------------------------------
constexpr int FM_W = 128, FM_TH = 32, FM_HALO = 20;
constexpr int FM_WORDS = FM_W / 32 + 2, FM_ROWS = FM_TH + 2 * FM_HALO;
struct MorphChain { Footprint band, close, open; };

// 32 pixels starting at x (multiple of 32... relative to the tile) of row y -> bit i = pixel x+i non-zero; 0 outside
// (xorv != 0: bit i = pixel != xorv's byte, e.g. 0x02020202 -> "pixel != 2")
__device__ __forceinline__ uint32_t pack32(const uint8_t* __restrict__ img, int H, int W, int y, int x, uint32_t xorv = 0u) {
  if (y < 0 || y >= H || x >= W || x + 32 <= 0) return 0u;
  const uint8_t* p = img + (size_t)y * W + x;
  uint32_t word = 0;
  if (x >= 0 && x + 32 <= W && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
    const uint4 a = *reinterpret_cast<const uint4*>(p), b = *reinterpret_cast<const uint4*>(p + 16);
    const uint32_t v[8] = {a.x ^ xorv, a.y ^ xorv, a.z ^ xorv, a.w ^ xorv, b.x ^ xorv, b.y ^ xorv, b.z ^ xorv, b.w ^ xorv};
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t t = ((((v[k] & 0x7f7f7f7fu) + 0x7f7f7f7fu) | v[k]) & 0x80808080u) >> 7;   // bit 8j = byte j non-zero
      word |= (((t * 0x00204081u) >> 21) & 15u) << (4 * k);
    }
  } else {
    for (int i = 0; i < 32; ++i)
      if (x + i >= 0 && x + i < W && (p[i] ^ (uint8_t)xorv)) word |= 1u << i;
  }
  return word;
}
__device__ __forceinline__ uint32_t inside32(int H, int W, int y, int x) {
  if (y < 0 || y >= H) return 0u;
  uint32_t m = 0xffffffffu;
  if (x < 0) m = x <= -32 ? 0u : m << (-x);
  if (x + 32 > W) m = x >= W ? 0u : m & (0xffffffffu >> (x + 32 - W));
  return m;
}

// out[r][w] = dilate / erode of plane `in` with footprint fp at word (r, w)
template <bool DILATE, int ROWS>
__device__ __forceinline__ uint32_t fm_apply(const uint32_t (*in)[FM_WORDS], int r, int w, const Footprint& fp) {
  const uint32_t ident = DILATE ? 0u : 0xffffffffu;
  uint32_t acc = ident;
  for (int j = 0; j < fp.nrows; ++j) {
    const int rr = r + fp.dy[j];
    if (rr < 0 || rr >= ROWS) continue;
    const uint32_t lo = w > 0 ? in[rr][w - 1] : ident, mid = in[rr][w], hi = w + 1 < FM_WORDS ? in[rr][w + 1] : ident;
    // the run [xa, xb] in two loops (left of the word, then from the word on): one funnel shift + one logic op per cell
    const int xa = fp.xa[j], xb = fp.xb[j];
    int sft = xa;
#pragma unroll 4
    for (; sft < 0 && sft <= xb; ++sft) { const uint32_t v = __funnelshift_r(lo, mid, 32 + sft); acc = DILATE ? (acc | v) : (acc & v); }
#pragma unroll 4
    for (; sft <= xb; ++sft) { const uint32_t v = __funnelshift_r(mid, hi, sft); acc = DILATE ? (acc | v) : (acc & v); }
  }
  return acc;
}

template <int HALO>     // rows of halo above / below the 32-row tile: >= 2*(open/2) + 2*(close/2) + band/2 (8 standard, 20 enhanced)
__global__ void __launch_bounds__(256) k_burr_morph(const uint8_t* __restrict__ mask, const uint8_t* __restrict__ edges,
                                                    uint8_t* __restrict__ out, int H, int W, MorphChain mc) {
  constexpr int ROWS = FM_TH + 2 * HALO;
  __shared__ uint32_t s_a[ROWS][FM_WORDS], s_b[ROWS][FM_WORDS], s_in[ROWS][FM_WORDS];
  const int x0 = blockIdx.x * FM_W, y0 = blockIdx.y * FM_TH;
  const size_t base = (size_t)blockIdx.z * H * W;
  constexpr int NW = ROWS * FM_WORDS;
  uint32_t e_bits[(NW + 255) / 256], m_bits[(NW + 255) / 256];
  // load: s_a = mask bits, registers keep the edge / mask words of this thread's positions
#pragma unroll
  for (int k = 0; k < (NW + 255) / 256; ++k) {
    const int i = threadIdx.x + k * 256;
    if (i < NW) {
      const int r = i / FM_WORDS, w = i - r * FM_WORDS;
      const int y = y0 - HALO + r, x = x0 - 32 + 32 * w;
      m_bits[k] = pack32(mask + base, H, W, y, x);
      e_bits[k] = pack32(edges + base, H, W, y, x);
      s_a[r][w] = m_bits[k];
      s_in[r][w] = inside32(H, W, y, x);
    }
  }
  __syncthreads();
  // cand = edges & dilate(mask, band) & ~mask                          -> s_b
#pragma unroll
  for (int k = 0; k < (NW + 255) / 256; ++k) {
    const int i = threadIdx.x + k * 256;
    if (i < NW) { const int r = i / FM_WORDS, w = i - r * FM_WORDS; s_b[r][w] = fm_apply<true, ROWS>(s_a, r, w, mc.band) & ~m_bits[k] & e_bits[k]; }
  }
  __syncthreads();
  // CLOSE: dilate -> s_a (forced to 1 outside the image for the erosion that follows), erode -> s_b
  for (int i = threadIdx.x; i < NW; i += 256) { const int r = i / FM_WORDS, w = i - r * FM_WORDS; s_a[r][w] = fm_apply<true, ROWS>(s_b, r, w, mc.close) | ~s_in[r][w]; }
  __syncthreads();
  for (int i = threadIdx.x; i < NW; i += 256) { const int r = i / FM_WORDS, w = i - r * FM_WORDS; s_b[r][w] = fm_apply<false, ROWS>(s_a, r, w, mc.close) | ~s_in[r][w]; }
  __syncthreads();
  // OPEN: erode -> s_a (forced to 0 outside for the dilation), dilate -> output
  for (int i = threadIdx.x; i < NW; i += 256) { const int r = i / FM_WORDS, w = i - r * FM_WORDS; s_a[r][w] = fm_apply<false, ROWS>(s_b, r, w, mc.open) & s_in[r][w]; }
  __syncthreads();
  for (int i = threadIdx.x; i < FM_TH * (FM_W / 32); i += 256) {
    const int ty = i / (FM_W / 32), tw = i - ty * (FM_W / 32);
    const int y = y0 + ty, x = x0 + 32 * tw;
    if (y >= H || x >= W) continue;
    const uint32_t word = fm_apply<true, ROWS>(s_a, ty + HALO, tw + 1, mc.open);
    uint8_t* o = out + base + (size_t)y * W + x;
    if (x + 32 <= W && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
      uint32_t v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = (((word >> (4 * k)) & 15u) * 0x00204081u) & 0x01010101u;
      *reinterpret_cast<uint4*>(o) = make_uint4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<uint4*>(o + 16) = make_uint4(v[4], v[5], v[6], v[7]);
    } else {
      for (int k = 0; k < 32 && x + k < W; ++k) o[k] = (word >> k) & 1u;
    }
  }
}

// ------------------------------------------------------------------------------
// One morphology operator (ub_morph: dilate / erode / close / open with a cv2 ellipse) on bit-packed shared-memory
// rows, one 32-pixel word per thread and stage -- the per-operator form of k_burr_morph.  HALO rows above and below
// the 32-row tile (template: 4 / 12 / 24 covers one or two passes of footprints up to 25), one word left and right.
// Out-of-image samples are ignored as cv2 does (0 before a dilation, 1 before an erosion).
// ------------------------------------------------------------------------------
template <bool DILATE, int ROWS>
__device__ __forceinline__ uint32_t fmw_apply(const uint32_t (*in)[FM_WORDS], int r, int w, const Footprint& fp) {
  const uint32_t ident = DILATE ? 0u : 0xffffffffu;
  uint32_t acc = ident;
  for (int j = 0; j < fp.nrows; ++j) {
    const int rr = r + fp.dy[j];
    if (rr < 0 || rr >= ROWS) continue;
    const uint32_t lo = w > 0 ? in[rr][w - 1] : ident, mid = in[rr][w], hi = w + 1 < FM_WORDS ? in[rr][w + 1] : ident;
    // the run [xa, xb] in two loops (left of the word, then from the word on): one funnel shift + one logic op per cell
    const int xa = fp.xa[j], xb = fp.xb[j];
    int sft = xa;
#pragma unroll 4
    for (; sft < 0 && sft <= xb; ++sft) { const uint32_t v = __funnelshift_r(lo, mid, 32 + sft); acc = DILATE ? (acc | v) : (acc & v); }
#pragma unroll 4
    for (; sft <= xb; ++sft) { const uint32_t v = __funnelshift_r(mid, hi, sft); acc = DILATE ? (acc | v) : (acc & v); }
  }
  return acc;
}

// OP: 0 dilate, 1 erode, 2 close (dilate -> erode), 3 open (erode -> dilate)
template <int HALO, int OP>
__global__ void __launch_bounds__(256) k_morph_words(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int H, int W,
                                                     Footprint fp, uint32_t value) {
  constexpr int ROWS = FM_TH + 2 * HALO, NW = ROWS * FM_WORDS;
  constexpr bool FIRST_DIL = OP == 0 || OP == 2;
  __shared__ uint32_t s_a[ROWS][FM_WORDS], s_b[OP >= 2 ? ROWS : 1][FM_WORDS], s_in[OP >= 1 ? ROWS : 1][FM_WORDS];
  const int x0 = blockIdx.x * FM_W, y0 = blockIdx.y * FM_TH;
  const size_t base = (size_t)blockIdx.z * H * W;
  for (int i = threadIdx.x; i < NW; i += 256) {
    const int r = i / FM_WORDS, w = i - r * FM_WORDS;
    const int y = y0 - HALO + r, x = x0 - 32 + 32 * w;
    const uint32_t b = pack32(src + base, H, W, y, x);
    if (OP >= 1) {
      const uint32_t in = inside32(H, W, y, x);
      s_in[r][w] = in;
      s_a[r][w] = FIRST_DIL ? b : (b | ~in);
    } else {
      s_a[r][w] = b;
    }
  }
  __syncthreads();
  if (OP >= 2) {
    for (int i = threadIdx.x; i < NW; i += 256) {
      const int r = i / FM_WORDS, w = i - r * FM_WORDS;
      s_b[r][w] = FIRST_DIL ? (fmw_apply<true, ROWS>(s_a, r, w, fp) | ~s_in[r][w]) : (fmw_apply<false, ROWS>(s_a, r, w, fp) & s_in[r][w]);
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < FM_TH * (FM_W / 32); i += 256) {
    const int ty = i / (FM_W / 32), tw = i - ty * (FM_W / 32);
    const int y = y0 + ty, x = x0 + 32 * tw;
    if (y >= H || x >= W) continue;
    uint32_t word;
    if (OP == 0) word = fmw_apply<true, ROWS>(s_a, ty + HALO, tw + 1, fp);
    else if (OP == 1) word = fmw_apply<false, ROWS>(s_a, ty + HALO, tw + 1, fp);
    else if (OP == 2) word = fmw_apply<false, ROWS>(s_b, ty + HALO, tw + 1, fp);
    else word = fmw_apply<true, ROWS>(s_b, ty + HALO, tw + 1, fp);
    uint8_t* o = dst + base + (size_t)y * W + x;
    if (x + 32 <= W && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
      uint32_t v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = ((((word >> (4 * k)) & 15u) * 0x00204081u) & 0x01010101u) * value;
      *reinterpret_cast<uint4*>(o) = make_uint4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<uint4*>(o + 16) = make_uint4(v[4], v[5], v[6], v[7]);
    } else {
      for (int k = 0; k < 32 && x + k < W; ++k) o[k] = ((word >> k) & 1u) ? (uint8_t)value : 0;
    }
  }
}

// out = (a != 0 && b != 0) ? 1 : 0     (edges & detection_band, infer_two_stage_burr.py:89)
__global__ void k_and01(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, uint8_t* __restrict__ out, size_t n) {
  size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i + 4 <= n) {
    uint32_t va = *reinterpret_cast<const uint32_t*>(a + i), vb = *reinterpret_cast<const uint32_t*>(b + i), o = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) o |= (uint32_t)((((va >> (8 * k)) & 255) && ((vb >> (8 * k)) & 255)) ? 1 : 0) << (8 * k);
    *reinterpret_cast<uint32_t*>(out + i) = o;
  } else {
    for (; i < n; ++i) out[i] = (a[i] && b[i]) ? 1 : 0;
  }
}

// ------------------------------------------------------------------------------
// Enhanced-variant edge maps (infer_enhanced_burr.py:93-102), REFLECT_101 borders.
// ------------------------------------------------------------------------------
__device__ __forceinline__ void sobel_lap_at(const uint8_t* img, int H, int W, int y, int x, int& s2, int& lap) {
  int ym = reflect101(y - 1, H), yp = reflect101(y + 1, H), xm = reflect101(x - 1, W), xp = reflect101(x + 1, W);
  int a = img[(size_t)ym * W + xm], b = img[(size_t)ym * W + x], c = img[(size_t)ym * W + xp];
  int d = img[(size_t)y * W + xm], e = img[(size_t)y * W + x], f = img[(size_t)y * W + xp];
  int g = img[(size_t)yp * W + xm], h = img[(size_t)yp * W + x], k = img[(size_t)yp * W + xp];
  int gx = (c + 2 * f + k) - (a + 2 * d + g);
  int gy = (g + 2 * h + k) - (a + 2 * b + c);
  s2 = gx * gx + gy * gy;
  lap = b + h + d + f - 4 * e;
}

// per-frame max of gx^2+gy^2 (the `sobel_mag.max()` of :96)
__global__ void __launch_bounds__(256) k_sobel_max(const uint8_t* __restrict__ gray, int H, int W, int* __restrict__ fmax) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  const uint8_t* img = gray + (size_t)blockIdx.z * H * W;
  int s2 = 0, lap;
  if (x < W) sobel_lap_at(img, H, W, y, x, s2, lap);
  for (int o = 16; o; o >>= 1) s2 = max(s2, __shfl_xor_sync(0xffffffffu, s2, o));
  if ((threadIdx.x & 31) == 0 && s2 > 0) atomicMax(fmax + blockIdx.z, s2);
}

// mode bit 0: Sobel edges, bit 1: Laplacian edges; result OR-ed with `base` (nullable), values {0,255}
__global__ void __launch_bounds__(256) k_enh_edges(const uint8_t* __restrict__ gray, int H, int W, const int* __restrict__ fmax,
                                                   int mode, int sobel_t, int lap_t, const uint8_t* __restrict__ base_edges,
                                                   uint8_t* __restrict__ out) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  size_t o = (size_t)blockIdx.z * H * W + (size_t)y * W + x;
  int s2, lap;
  sobel_lap_at(gray + (size_t)blockIdx.z * H * W, H, W, y, x, s2, lap);
  bool e = base_edges ? base_edges[o] != 0 : false;
  if (mode & 1) {
    int mx = fmax[blockIdx.z];
    if (mx > 0) {
      // np.uint8(mag / mag.max() * 255): two correctly rounded f64 ops, then truncation
      double m = __dmul_rn(__ddiv_rn(sqrt((double)s2), sqrt((double)mx)), 255.0);
      e = e || ((int)m > sobel_t);
    }
  }
  if (mode & 2) e = e || ((abs(lap) & 255) > lap_t);     // np.abs(lap).astype(np.uint8) wraps mod 256
  out[o] = e ? 255 : 0;
}

}  // namespace s2
