// edge_fused.cuh -- GaussianBlur(5x5, 1.0) + the candidate half of cv2.Canny in ONE pass over the gray
// plane, four pixels per thread with packed 16-bit arithmetic, emitting what the run-based
// hysteresis consumes directly (bit-packed candidate / strong rows and the union-find entries of
// the run starts).  Replaces k_gauss5 -> k_canny_nms -> k_cclb_init (three passes, two u8 planes
// through memory, ~300 byte-granular instructions per pixel).
//
// Reference: infer_two_stage_burr.py:85-86 (GaussianBlur + Canny), infer_enhanced_burr.py:89-90.
// Arithmetic as in oracle/cvnp.py (gaussian_blur5, canny_candidates): separable [14,62,104,62,14],
// REFLECT_101, one rounding (acc + 2^15) >> 16; Sobel 3x3 on the blurred plane with
// BORDER_REPLICATE, |gx|+|gy| (0 outside the image), fixed-point NMS with TG22 = 13573 (<<15).
//
// Tile: EF_TW x EF_TH (160 x 24) outputs per CTA of 256 threads.  A "column word" cw covers the four pixels
// x0 - 4 + 4cw .. +3, so every stage works on aligned 32-bit words:
//   L  gray words with a 4-pixel halo                         g  [TH+8][42]   u8x4
//   V  vertical 5-tap on (c0,c1),(c2,c3) u16x2 lanes           vv [TH+4][44]   2 x u16x2
//   H  horizontal 5-tap as IDP.2A dot products, round, pack    bb [TH+4][44]   u8x4
//   F  (edge tiles) replicate the blurred plane one pixel outwards
//   S  Sobel on even/odd u16x2 lanes (p0,p2),(p1,p3): biased gx, gy and |gx|+|gy|
//   N  non-maximum suppression + thresholds (scalar, only where a word exceeds `low`)
//   P  32 pixels -> one candidate word + one strong word, run starts -> union-find entries
#pragma once
#include "stage2_kernels.cuh"

namespace s2 {

#ifndef UB_EF_TH
#define UB_EF_TH 24
#endif
// tile height: (TH + 4) % 4 == 0 and TH even (the V / S stages walk rows in groups of 4 / 2).  24 rows = 37.6 KB of shared memory, six
// CTAs per SM; 32 rows (48.5 KB, four CTAs) were 10 % slower, 28 / 20 / 16 rows in between (ub_blur_canny per 512 frames: 0.907 ms at
// 32, 0.840 at 28, 0.810 at 24, 0.893 at 20, 0.898 at 16; profiles/r03n_edge_tile_height.log)
constexpr int EF_TW = 160, EF_TH = UB_EF_TH;
static_assert((EF_TH + 4) % 4 == 0 && EF_TH % 2 == 0, "tile height");
constexpr int EF_CW = EF_TW / 4 + 2;            // 42 column words (one word of halo each side)
constexpr int EF_TR = 256 / EF_CW;               // 6 row lanes
constexpr int EF_CWP = EF_CW + 2;               // + one pad word each side (read by the neighbours' taps, never used)
constexpr int EF_SMEM_A = (EF_TH + 4) * EF_CWP * 8;   // vv; later gx
constexpr int EF_SMEM_G = (EF_TH + 8) * EF_CW * 4;    // g (6720 B, a multiple of 16)
constexpr int EF_SMEM_B = EF_SMEM_G + (EF_TH + 4) * EF_CWP * 4;   // g | bb; cc reuses the front after S
constexpr int EF_SMEM_M = (EF_TH + 2) * EF_CW * 8;    // mag, gy

__device__ __forceinline__ uint32_t ef_bias_sub(uint32_t a, uint32_t b) { return a - b + 0x04000400u; }   // lanes a-b+1024 (> 0: no borrow)
__device__ __forceinline__ uint32_t ef_abs_biased(uint32_t g) { return __vmaxu2(g, 0x08000800u - g); }   // lanes |g-1024|+1024

// MODE 0: Canny candidates of `src` as it is (ub_canny).  MODE 1: blur, then candidates.  MODE 2: blur only -> blur_out.
template <int MODE>
__global__ void __launch_bounds__(256) k_edge_fused(const uint8_t* __restrict__ src, uint8_t* __restrict__ blur_out,
                                                    uint32_t* __restrict__ bits, uint32_t* __restrict__ strong, int* __restrict__ lab_all,
                                                    int H, int W, int Wd, int low, int high) {
  __shared__ __align__(16) uint8_t s_a[EF_SMEM_A];       // vv -> gx
  __shared__ __align__(16) uint8_t s_b[EF_SMEM_B];       // g | bb -> cc
  __shared__ __align__(16) uint8_t s_m[EF_SMEM_M];       // mag
  __shared__ __align__(16) uint8_t s_y[EF_SMEM_M];       // gy
  uint32_t (*g)[EF_CW] = reinterpret_cast<uint32_t (*)[EF_CW]>(s_b);                 // [TH+8][42]  (6720 B)
  uint2 (*vv)[EF_CWP] = reinterpret_cast<uint2 (*)[EF_CWP]>(s_a);                    // [TH+4][44]
  uint32_t (*bb)[EF_CWP] = reinterpret_cast<uint32_t (*)[EF_CWP]>(s_b + EF_SMEM_G);  // [TH+4][44], behind g
  uint2 (*gxs)[EF_CW] = reinterpret_cast<uint2 (*)[EF_CW]>(s_a);                     // [TH+2][42]
  uint2 (*gys)[EF_CW] = reinterpret_cast<uint2 (*)[EF_CW]>(s_y);
  uint2 (*mm)[EF_CW] = reinterpret_cast<uint2 (*)[EF_CW]>(s_m);
  uint32_t (*cc)[EF_TW / 4] = reinterpret_cast<uint32_t (*)[EF_TW / 4]>(s_b);        // [TH][40]  (5120 B, after S)
  static_assert(EF_SMEM_G % 16 == 0 && EF_TH * EF_TW <= EF_SMEM_G && (EF_TH + 2) * EF_CW * 8 <= EF_SMEM_A &&
                EF_SMEM_A + EF_SMEM_B + 2 * EF_SMEM_M <= 48 * 1024, "smem carve-up");

  const int tid = threadIdx.x;
  const int x0 = blockIdx.x * EF_TW, y0 = blockIdx.y * EF_TH;
  const size_t fbase = (size_t)blockIdx.z * H * W;
  const uint8_t* img = src + fbase;
  // (tr, cw): a thread keeps its column word and walks rows with stride EF_TR -- no divisions, column terms once
  const int tr = tid / EF_CW, cw = tid - tr * EF_CW;         // tr < EF_TR for tid < 252
  const bool act = tr < EF_TR;
  const int xw = x0 - 4 + 4 * cw;                            // image x of the word's first pixel
  const bool xin = xw >= 0 && xw + 3 < W;

  if (MODE != 0) {
    // ---- L: gray words, REFLECT_101 ----
    if (act && xin) {                                // interior words: one aligned 32-bit load per row
      const uint8_t* p = img + xw;
      if (y0 >= 4 && y0 + EF_TH + 3 < H) {
        p += (y0 - 4) * W;
        for (int r = tr; r < EF_TH + 8; r += EF_TR) g[r][cw] = *reinterpret_cast<const uint32_t*>(p + r * W);
      } else {
        for (int r = tr; r < EF_TH + 8; r += EF_TR) g[r][cw] = *reinterpret_cast<const uint32_t*>(p + reflect101(y0 - 4 + r, H) * W);
      }
    } else if (act) {                                // words that straddle the left / right image border
      int xs[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) xs[k] = reflect101(xw + k, W);
      for (int r = tr; r < EF_TH + 8; r += EF_TR) {
        const uint8_t* row = img + reflect101(y0 - 4 + r, H) * W;
        g[r][cw] = (uint32_t)row[xs[0]] | ((uint32_t)row[xs[1]] << 8) | ((uint32_t)row[xs[2]] << 16) | ((uint32_t)row[xs[3]] << 24);
      }
    }
    __syncthreads();
    // ---- V: vertical taps, four output rows per thread (eight input rows in registers) ----
    for (int grp = tr; grp < (EF_TH + 4) / 4 && act; grp += EF_TR) {
      uint32_t lo[8], hi[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t w = g[4 * grp + k][cw];
        lo[k] = __byte_perm(w, 0u, 0x4140);          // (c0, c1) as u16x2
        hi[k] = __byte_perm(w, 0u, 0x4342);          // (c2, c3)
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        uint2 o;                                     // every lane <= 255*256: no carry between lanes
        o.x = 14u * (lo[k] + lo[k + 4]) + 62u * (lo[k + 1] + lo[k + 3]) + 104u * lo[k + 2];
        o.y = 14u * (hi[k] + hi[k + 4]) + 62u * (hi[k + 1] + hi[k + 3]) + 104u * hi[k + 2];
        vv[4 * grp + k][cw + 1] = o;
      }
    }
    __syncthreads();
    // ---- H: horizontal taps as 16-bit x 8-bit dot products; pixel j of the word needs v[j-2..j+2] ----
    for (int r = tr; r < EF_TH + 4 && act; r += EF_TR) {
      const uint32_t a = vv[r][cw].y;                // v[-2], v[-1]   (pad words hold garbage: only unused outputs see it)
      const uint2 m = vv[r][cw + 1];                 // v[0..1], v[2..3]
      const uint32_t n = vv[r][cw + 2].x;            // v[4], v[5]
      const uint32_t qa = __funnelshift_r(a, m.x, 16), qb = __funnelshift_r(m.x, m.y, 16), qc = __funnelshift_r(m.y, n, 16);
      constexpr uint32_t K1 = 14u | (62u << 8), K2 = 104u | (62u << 8), K3 = 14u, K4 = 14u << 8;
      const uint32_t p0 = __dp2a_lo(m.y, K3, __dp2a_lo(m.x, K2, __dp2a_lo(a, K1, 32768u)));
      const uint32_t p1 = __dp2a_lo(qc, K3, __dp2a_lo(qb, K2, __dp2a_lo(qa, K1, 32768u)));
      const uint32_t p2 = __dp2a_lo(n, K3, __dp2a_lo(m.y, K2, __dp2a_lo(m.x, K1, 32768u)));
      const uint32_t p3 = __dp2a_lo(n, K4, __dp2a_lo(qc, K2, __dp2a_lo(qb, K1, 32768u)));
      bb[r][cw + 1] = __byte_perm(__byte_perm(p0, p1, 0x0062), __byte_perm(p2, p3, 0x0062), 0x5410);   // bits 16..23 of each
    }
    __syncthreads();
    if (MODE == 2) {
      if (act && cw >= 1 && cw <= EF_TW / 4 && xw < W)
        for (int r = tr; r < EF_TH && y0 + r < H; r += EF_TR)
          *reinterpret_cast<uint32_t*>(blur_out + fbase + (uint32_t)((y0 + r) * W + xw)) = bb[r + 2][cw + 1];
      return;
    }
    // ---- F: Sobel wants BORDER_REPLICATE of the BLURRED plane (block-uniform conditions) ----
    const bool top = y0 == 0, bot = y0 + EF_TH + 1 >= H, lef = x0 == 0, rig = x0 + EF_TW + 1 >= W;
    if (top || bot) {
      const int rl = H - 1 - (y0 - 2);               // tile row of image row H-1
      for (int i = tid; i < 4 * EF_CW; i += 256) {
        const int k = i / EF_CW, cw = i - k * EF_CW;
        if (top && k == 0) bb[1][cw + 1] = bb[2][cw + 1];
        if (bot && k >= 1 && rl + k < EF_TH + 4 && rl >= 0) bb[rl + k][cw + 1] = bb[rl][cw + 1];
      }
      __syncthreads();
    }
    if (lef || rig) {
      const int cwr = (W - (x0 - 4)) / 4;            // column word that starts at x = W
      for (int r = tid; r < EF_TH + 4; r += 256) {
        if (lef) bb[r][1] = (bb[r][1] & 0x00ffffffu) | (bb[r][2] << 24);                 // x = -1 <- x = 0
        if (rig && cwr < EF_CW) bb[r][cwr + 1] = (bb[r][cwr + 1] & 0xffffff00u) | (bb[r][cwr] >> 24);   // x = W <- x = W-1
      }
      __syncthreads();
    }
  } else {
    // ---- MODE 0: the plane as given, BORDER_REPLICATE ----
    if (act && xin) {
      const uint8_t* p = img + xw;
      for (int r = tr; r < EF_TH + 4; r += EF_TR) bb[r][cw + 1] = *reinterpret_cast<const uint32_t*>(p + clampi(y0 - 2 + r, H) * W);
    } else if (act) {
      int xs[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) xs[k] = clampi(xw + k, W);
      for (int r = tr; r < EF_TH + 4; r += EF_TR) {
        const uint8_t* row = img + clampi(y0 - 2 + r, H) * W;
        bb[r][cw + 1] = (uint32_t)row[xs[0]] | ((uint32_t)row[xs[1]] << 8) | ((uint32_t)row[xs[2]] << 16) | ((uint32_t)row[xs[3]] << 24);
      }
    }
    __syncthreads();
  }

  // ---- S: Sobel, two magnitude rows per thread.  Lanes: E = (p0, p2), O = (p1, p3). ----
  for (int grp = tr; grp < (EF_TH + 2) / 2 && act; grp += EF_TR) {
    uint32_t w0[4], w1[4], w2[4], w3[4], rse[4], rso[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t L = bb[2 * grp + k][cw], C = bb[2 * grp + k][cw + 1], R = bb[2 * grp + k][cw + 2];
      w0[k] = __funnelshift_r(L, C, 24) & 0x00ff00ffu;        // (L3, C1): left neighbours of E
      w1[k] = C & 0x00ff00ffu;                                 // (C0, C2)
      w2[k] = __byte_perm(C, 0u, 0x4341);                      // (C1, C3)
      w3[k] = __funnelshift_r(C, R, 16) & 0x00ff00ffu;         // (C2, R0): right neighbours of O
      rse[k] = w0[k] + 2u * w1[k] + w2[k];                     // horizontal smooth, lanes <= 1020
      rso[k] = w1[k] + 2u * w2[k] + w3[k];
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int mr = 2 * grp + k, y = y0 - 1 + mr;
      const uint32_t c0 = w0[k] + 2u * w0[k + 1] + w0[k + 2], c1 = w1[k] + 2u * w1[k + 1] + w1[k + 2];
      const uint32_t c2 = w2[k] + 2u * w2[k + 1] + w2[k + 2], c3 = w3[k] + 2u * w3[k + 1] + w3[k + 2];
      const uint32_t gxe = ef_bias_sub(c2, c0), gxo = ef_bias_sub(c3, c1);             // gx + 1024
      const uint32_t gye = ef_bias_sub(rse[k + 2], rse[k]), gyo = ef_bias_sub(rso[k + 2], rso[k]);
      uint32_t me = ef_abs_biased(gxe) + ef_abs_biased(gye) - 0x08000800u;             // |gx| + |gy| <= 2040
      uint32_t mo = ef_abs_biased(gxo) + ef_abs_biased(gyo) - 0x08000800u;
      if (!xin || y < 0 || y >= H) me = mo = 0u;
      mm[mr][cw] = make_uint2(me, mo);
      gxs[mr][cw] = make_uint2(gxe, gxo);
      gys[mr][cw] = make_uint2(gye, gyo);
    }
  }
  __syncthreads();

  // ---- N: non-maximum suppression -> 0 none / 1 weak / 2 strong, one byte per pixel in cc ----
  for (int r = tr; r < EF_TH && act && cw >= 1 && cw <= EF_TW / 4; r += EF_TR) {
    const int mr = r + 1, c = cw - 1;
    const uint2 own = mm[mr][cw];
    uint32_t out = 0u;
    const uint32_t mx = __vmaxu2(own.x, own.y);
    if ((int)max(mx & 0xffffu, mx >> 16) > low) {
      int M[3][6];                                   // magnitudes at x-1 .. x+4 of rows y-1, y, y+1
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const uint2 o = mm[r + k][cw];
        M[k][0] = (int)(mm[r + k][cw - 1].y >> 16);
        M[k][1] = (int)(o.x & 0xffffu); M[k][2] = (int)(o.y & 0xffffu); M[k][3] = (int)(o.x >> 16); M[k][4] = (int)(o.y >> 16);
        M[k][5] = (int)(mm[r + k][cw + 1].x & 0xffffu);
      }
      const uint2 gxw = gxs[mr][cw], gyw = gys[mr][cw];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int m = M[1][j + 1];
        if (m > low) {
          const uint32_t gxl = (j & 1) ? gxw.y : gxw.x, gyl = (j & 1) ? gyw.y : gyw.x;
          const int gx = (int)((j & 2) ? gxl >> 16 : gxl & 0xffffu) - 1024;
          const int gy = (int)((j & 2) ? gyl >> 16 : gyl & 0xffffu) - 1024;
          const int ax = abs(gx), ay = abs(gy) << 15;             // <= 1020: everything fits 32 bits
          const int tg22 = ax * 13573, tg67 = tg22 + (ax << 16);
          bool keep;
          if (ay < tg22) keep = m > M[1][j] && m >= M[1][j + 2];
          else if (ay > tg67) keep = m > M[0][j + 1] && m >= M[2][j + 1];
          else if ((gx ^ gy) < 0) keep = m > M[0][j + 2] && m > M[2][j];
          else keep = m > M[0][j] && m > M[2][j + 2];
          if (keep) out |= (m > high ? 2u : 1u) << (8 * j);
        }
      }
    }
    cc[r][c] = out;
  }
  __syncthreads();

  // ---- P: candidate / strong words and the union-find entries of the run starts (k_cclb_init's job) ----
  for (int i = tid; i < EF_TH * (EF_TW / 32); i += 256) {
    const int r = i / (EF_TW / 32), w = i - r * (EF_TW / 32);
    const int y = y0 + r, wg = x0 / 32 + w;
    if (y >= H || wg >= Wd) continue;
    const uint4 a = *reinterpret_cast<const uint4*>(&cc[r][8 * w]), b = *reinterpret_cast<const uint4*>(&cc[r][8 * w + 4]);
    const uint32_t v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    uint32_t cand = 0u, str = 0u;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t s = (v[k] >> 1) & 0x01010101u, nz = (v[k] & 0x01010101u) | s;      // byte j: == 2, != 0
      cand |= ((nz * 0x01020408u) >> 24) << (4 * k);
      str |= ((s * 0x01020408u) >> 24) << (4 * k);
    }
    const int rem = W - 32 * wg;
    if (rem < 32) { cand &= (1u << rem) - 1u; str &= cand; }
    const size_t wb = ((size_t)blockIdx.z * H + y) * Wd + wg;
    bits[wb] = cand;
    strong[wb] = str;
    if (!lab_all) continue;                       // hysteresis by flood (k_hyst_frame): no union-find entries
    int* l = lab_all + fbase + (size_t)y * W + 32 * wg;
    uint32_t cbits = cand;
    while (cbits) {
      int s0, len;
      cbits &= ~first_run(cbits, s0, len);
      l[s0] = y * W + 32 * wg + s0;
    }
  }
}

}  // namespace s2
