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
constexpr int GT_W = 64, GT_H = 16;
__global__ void __launch_bounds__(256) k_gauss5(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int H, int W) {
  __shared__ uint8_t s_in[GT_H + 4][GT_W + 4];
  __shared__ int s_h[GT_H + 4][GT_W];
  const int x0 = blockIdx.x * GT_W, y0 = blockIdx.y * GT_H;
  const uint8_t* img = src + (size_t)blockIdx.z * H * W;
  for (int i = threadIdx.x; i < (GT_H + 4) * (GT_W + 4); i += 256) {
    int ty = i / (GT_W + 4), tx = i - ty * (GT_W + 4);
    s_in[ty][tx] = img[(size_t)reflect101(y0 + ty - 2, H) * W + reflect101(x0 + tx - 2, W)];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < (GT_H + 4) * GT_W; i += 256) {
    int ty = i / GT_W, tx = i - ty * GT_W;
    s_h[ty][tx] = 14 * (s_in[ty][tx] + s_in[ty][tx + 4]) + 62 * (s_in[ty][tx + 1] + s_in[ty][tx + 3]) + 104 * s_in[ty][tx + 2];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < GT_H * GT_W; i += 256) {
    int ty = i / GT_W, tx = i - ty * GT_W;
    int x = x0 + tx, y = y0 + ty;
    if (x < W && y < H) {
      int acc = 14 * (s_h[ty][tx] + s_h[ty + 4][tx]) + 62 * (s_h[ty + 1][tx] + s_h[ty + 3][tx]) + 104 * s_h[ty + 2][tx];
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
  __shared__ short s_gx[GT_H][GT_W], s_gy[GT_H][GT_W];
  const int x0 = blockIdx.x * GT_W, y0 = blockIdx.y * GT_H;
  const uint8_t* img = src + (size_t)blockIdx.z * H * W;
  for (int i = threadIdx.x; i < (GT_H + 4) * (GT_W + 4); i += 256) {
    int ty = i / (GT_W + 4), tx = i - ty * (GT_W + 4);
    s_in[ty][tx] = img[(size_t)clampi(y0 + ty - 2, H) * W + clampi(x0 + tx - 2, W)];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < (GT_H + 2) * (GT_W + 2); i += 256) {
    int ty = i / (GT_W + 2), tx = i - ty * (GT_W + 2);
    int x = x0 + tx - 1, y = y0 + ty - 1;                  // image coords of this mag sample
    int m = 0;
    if (x >= 0 && x < W && y >= 0 && y < H) {
      // s_in index of (y,x) is [ty+1][tx+1]
      int a = s_in[ty][tx], b = s_in[ty][tx + 1], c = s_in[ty][tx + 2];
      int d = s_in[ty + 1][tx], f = s_in[ty + 1][tx + 2];
      int g = s_in[ty + 2][tx], h = s_in[ty + 2][tx + 1], k = s_in[ty + 2][tx + 2];
      int gx = (c + 2 * f + k) - (a + 2 * d + g);
      int gy = (g + 2 * h + k) - (a + 2 * b + c);
      m = abs(gx) + abs(gy);
      if (ty >= 1 && ty <= GT_H && tx >= 1 && tx <= GT_W) { s_gx[ty - 1][tx - 1] = (short)gx; s_gy[ty - 1][tx - 1] = (short)gy; }
    }
    s_mag[ty][tx] = m;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < GT_H * GT_W; i += 256) {
    int ty = i / GT_W, tx = i - ty * GT_W;
    int x = x0 + tx, y = y0 + ty;
    if (x >= W || y >= H) continue;
    int m = s_mag[ty + 1][tx + 1];
    uint8_t out = 0;
    if (m > low) {
      int gx = s_gx[ty][tx], gy = s_gy[ty][tx];
      long long ax = abs(gx), ay = (long long)abs(gy) << 15;
      long long tg22 = ax * 13573, tg67 = tg22 + (ax << 16);
      bool keep;
      if (ay < tg22) keep = m > s_mag[ty + 1][tx] && m >= s_mag[ty + 1][tx + 2];
      else if (ay > tg67) keep = m > s_mag[ty][tx + 1] && m >= s_mag[ty + 2][tx + 1];
      else {
        int s = ((gx ^ gy) < 0) ? -1 : 1;
        keep = m > s_mag[ty][tx + 1 - s] && m > s_mag[ty + 2][tx + 1 + s];
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
