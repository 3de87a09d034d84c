// stage2.cu -- connected-component labelling (warp-level union-find), Canny hysteresis,
// component statistics / filter, and the burr-stage schedule
// (detect_burrs_on_cable, infer_two_stage_burr.py:50-119; detect_burrs_enhanced,
// infer_enhanced_burr.py:69-138).
#include <limits.h>
#include <stdlib.h>
#include <math.h>
#include <utility>
#include "stage2_kernels.cuh"

namespace s2 {

// ------------------------------------------------------------------------------
// Union-find over pixel indices (per frame, p = y*W + x).  Links always point from the
// larger index to the smaller, so a component's root is its first pixel in raster order.
// ------------------------------------------------------------------------------
__device__ __forceinline__ int uf_find(const int* lab, int x) {
  while (true) {
    int p = __ldcg(lab + x);
    if (p == x) return x;
    x = p;
  }
}
// find with path splitting: every node on the way is re-pointed to its grandparent.  Parents only ever decrease (atomicMin)
// and a grandparent is an ancestor, so the forest stays valid under concurrent unions; on solid masks (every word of a row
// chained to the word before it, every row to the one above) this keeps the chains short instead of hundreds of hops.
__device__ __forceinline__ int uf_find_split(int* lab, int x) {
  int p = __ldcg(lab + x);
  while (p != x) {
    const int g = __ldcg(lab + p);
    if (g != p) atomicMin(lab + x, g);
    x = p;
    p = g;
  }
  return x;
}
__device__ __forceinline__ void uf_union(int* lab, int a, int b) {
  while (true) {
    a = uf_find_split(lab, a);
    b = uf_find_split(lab, b);
    if (a == b) return;
    if (a < b) { int t = a; a = b; b = t; }
    int old = atomicMin(lab + a, b);
    if (old == a) return;
    a = old;
  }
}

// run start / length of the lane's foreground run inside its 32-pixel warp segment
__device__ __forceinline__ int run_start(uint32_t mask, int lane) {
  uint32_t inv = ~mask & ((1u << lane) - 1u);
  return inv ? 32 - __clz(inv) : 0;
}
__device__ __forceinline__ int run_end(uint32_t mask, int lane) {       // inclusive
  uint32_t inv = ~mask & ~((2u << lane) - 1u);                          // zeros above lane
  return inv ? __ffs(inv) - 2 : 31;
}

// Pass 1: label = index of the start of the pixel's run within its warp segment (ballot).
__global__ void __launch_bounds__(256) k_ccl_init(const uint8_t* __restrict__ src, int* __restrict__ lab, int H, int W) {
  const int segs = (W + 31) >> 5;
  const int warp = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= H * segs) return;
  const int y = warp / segs, x = (warp - y * segs) * 32 + lane;
  const size_t base = (size_t)blockIdx.z * H * W;
  const bool fg = x < W && src[base + (size_t)y * W + x] != 0;
  const uint32_t mask = __ballot_sync(0xffffffffu, fg);
  if (x < W) lab[base + (size_t)y * W + x] = fg ? y * W + (x - lane) + run_start(mask, lane) : -1;
}

// Pass 2: merge with the row above (8-connectivity) and across warp-segment boundaries.
__global__ void __launch_bounds__(256) k_ccl_merge(int* __restrict__ lab_all, int H, int W) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  int* lab = lab_all + (size_t)blockIdx.z * H * W;
  const int p = y * W + x;
  if (lab[p] < 0) return;
  const bool w = x > 0 && lab[p - 1] >= 0;
  if (w && (x & 31) == 0) uf_union(lab, p, p - 1);
  if (y == 0) return;
  const int up = p - W;
  const bool n = lab[up] >= 0;
  const bool nw = x > 0 && lab[up - 1] >= 0;
  const bool ne = x + 1 < W && lab[up + 1] >= 0;
  if (n) {
    if (!(w && nw)) uf_union(lab, p, up);
  } else {
    if (nw && !w) uf_union(lab, p, up - 1);
    if (ne) uf_union(lab, p, up + 1);
  }
}

// Pass 3: full path compression; roots initialise their accumulators when `comp` is given.
// comp planes (each B*H*W): 0 area, 1 min x, 2 max x, 3 min y, 4 max y, 5 first-appearance key
__global__ void __launch_bounds__(256) k_ccl_flatten(int* __restrict__ lab_all, int H, int W, int* __restrict__ comp, size_t plane) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  const size_t base = (size_t)blockIdx.z * H * W;
  int* lab = lab_all + base;
  const int p = y * W + x;
  if (lab[p] < 0) return;
  const int r = uf_find(lab, p);
  lab[p] = r;
  if (comp && r == p) {
    int* c = comp + base + p;
    c[0] = 0; c[plane] = INT_MAX; c[2 * plane] = -1; c[3 * plane] = INT_MAX; c[4 * plane] = -1; c[5 * plane] = INT_MAX;
  }
}

// Canny hysteresis (second half of cv2.Canny): a candidate survives iff its 8-connected
// component of candidates holds a strong pixel.  `out` doubles as the per-root flag.
__global__ void k_hyst_mark(const uint8_t* __restrict__ cand, const int* __restrict__ lab, uint8_t* __restrict__ out, int H, int W) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  const size_t base = (size_t)blockIdx.z * H * W;
  const int p = y * W + x;
  if (cand[base + p] == 2) out[base + lab[base + p]] = 255;
}
__global__ void k_hyst_final(const int* __restrict__ lab, uint8_t* __restrict__ out, int H, int W) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  const size_t base = (size_t)blockIdx.z * H * W;
  const int p = y * W + x;
  const int r = lab[base + p];
  if (r >= 0 && r != p) out[base + p] = out[base + r];
}

// Component statistics: one atomic set per run (warp-aggregated), not per pixel.
// bgacc (nullable) [B][8]: background area / bbox for stats row 0.
__global__ void __launch_bounds__(256) k_ccl_stats(const int* __restrict__ lab_all, int H, int W, int* __restrict__ comp, size_t plane,
                                                   int* __restrict__ bgacc) {
  const int segs = (W + 31) >> 5;
  const int warp = (blockIdx.x * 256 + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= H * segs) return;
  const int y = warp / segs, x = (warp - y * segs) * 32 + lane;
  const size_t base = (size_t)blockIdx.z * H * W;
  const int r = x < W ? lab_all[base + (size_t)y * W + x] : -2;
  const bool fg = r >= 0;
  const uint32_t mask = __ballot_sync(0xffffffffu, fg);
  if (fg && run_start(mask, lane) == lane) {
    const int xe = x + (run_end(mask, lane) - lane);
    int* c = comp + base + r;
    atomicAdd(c, xe - x + 1);
    atomicMin(c + plane, x); atomicMax(c + 2 * plane, xe);
    atomicMin(c + 3 * plane, y); atomicMax(c + 4 * plane, y);
    // first appearance in cv2's 2x2-block raster scan; the key grows with x inside a row,
    // so the minimum over a run is at its first pixel.
    const int key = ((((y >> 1) * ((W + 1) >> 1)) + (x >> 1)) << 2) | ((y & 1) << 1) | (x & 1);
    atomicMin(c + 5 * plane, key);
  }
  if (bgacc) {
    const bool bg = r == -1;
    const uint32_t bm = __ballot_sync(0xffffffffu, bg);
    if (bm && lane == 0) {
      int* a = bgacc + blockIdx.z * 8;
      const int xs = (warp - y * segs) * 32;
      atomicAdd(a, __popc(bm));
      atomicMin(a + 1, xs + __ffs(bm) - 1); atomicMax(a + 2, xs + 31 - __clz(bm));
      atomicMin(a + 3, y); atomicMax(a + 4, y);
    }
  }
}

// Component filter (infer_two_stage_burr.py:103-117): burr[p] = 1 iff the component passes.
__global__ void __launch_bounds__(256) k_ccl_filter(const int* __restrict__ lab_all, int H, int W, const int* __restrict__ comp, size_t plane,
                                                    ub_burr_params P, uint8_t* __restrict__ burr, int* __restrict__ px, int px_stride) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  const size_t base = (size_t)blockIdx.z * H * W;
  bool keep = false;
  if (x < W) {
    const int r = lab_all[base + (size_t)y * W + x];
    if (r >= 0) {
      const int* c = comp + base + r;
      const int area = c[0], w = c[2 * plane] - c[plane] + 1, h = c[4 * plane] - c[3 * plane] + 1;
      const int mx = max(w, h), mn = min(w, h);
      keep = area >= P.min_area && area <= P.max_area && mx <= P.max_aspect * mn && w >= P.min_side && h >= P.min_side;
    }
    burr[base + (size_t)y * W + x] = keep ? 1 : 0;
  }
  const uint32_t m = __ballot_sync(0xffffffffu, keep);
  if (px && m && (threadIdx.x & 31) == 0) atomicAdd(px + blockIdx.z * px_stride, __popc(m));
}

// ==============================================================================
// Run-based labelling on bit-packed rows (the path of canny() and burr_stage()).
// A word holds 32 pixels of a row; a RUN is a maximal group of set bits inside one word and is
// represented by its first pixel.  Only run starts carry union-find entries, unions are issued
// per pair of touching runs, and every pass is one thread per word: ~32x fewer threads and
// memory operations than the per-pixel kernels above (which ub_ccl's OpenCV-numbered export
// still uses).  Roots are the smallest pixel index of a component, as before.
// ==============================================================================
__device__ __forceinline__ uint32_t bits_range(int a, int b) {            // bits a..b inclusive, 0 <= a <= b <= 31
  return (0xffffffffu >> (31 - b)) & (0xffffffffu << a);
}
// first run of m: returns its mask; a = first bit, len = length
__device__ __forceinline__ uint32_t first_run(uint32_t m, int& a, int& len) {
  a = __ffs(m) - 1;
  const uint32_t nt = ~(m >> a);
  len = nt ? __ffs(nt) - 1 : 32 - a;
  return bits_range(a, a + len - 1);
}
// start bit of the run of word u that contains set bit i
__device__ __forceinline__ int run_start_of(uint32_t u, int i) {
  return i == 0 ? 0 : i - __clz(~(u << (32 - i)));
}

}  // namespace s2
#include "edge_fused.cuh"
namespace s2 {

// pack + one union-find entry per run.  strong (nullable): bits of pixels == 2 (Canny strong candidates)
__global__ void __launch_bounds__(256) k_cclb_init(const uint8_t* __restrict__ src, uint32_t* __restrict__ bits, uint32_t* __restrict__ strong,
                                                   int* __restrict__ lab, int H, int W, int Wd) {
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (w >= Wd || y >= H) return;
  const size_t fb = (size_t)blockIdx.z * H * W, wb = ((size_t)blockIdx.z * H + y) * Wd + w;
  uint32_t c = pack32(src + fb, H, W, y, 32 * w);
  bits[wb] = c;
  if (strong) strong[wb] = c & ~pack32(src + fb, H, W, y, 32 * w, 0x02020202u);
  int* l = lab + fb + (size_t)y * W + 32 * w;
  while (c) {
    int a, len;
    c &= ~first_run(c, a, len);
    l[a] = y * W + 32 * w + a;
  }
}

__global__ void __launch_bounds__(256) k_cclb_merge(const uint32_t* __restrict__ bits, int* __restrict__ lab_all, int H, int W, int Wd) {
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (w >= Wd || y >= H) return;
  const uint32_t* row = bits + ((size_t)blockIdx.z * H + y) * Wd;
  uint32_t c = row[w];
  if (!c) return;
  int* lab = lab_all + (size_t)blockIdx.z * H * W;
  const int p0 = y * W + 32 * w;
  if ((c & 1u) && w > 0) {                                   // the run continues from the previous word
    const uint32_t L = row[w - 1];
    if (L >> 31) uf_union(lab, p0, p0 - 32 + run_start_of(L, 31));
  }
  if (y == 0) return;
  const uint32_t* up = row - Wd;
  const uint32_t u = up[w], ul = w > 0 ? up[w - 1] : 0u, ur = w + 1 < Wd ? up[w + 1] : 0u;
  if (!(u | (ul >> 31) | (ur & 1u))) return;
  const int q0 = p0 - W;                                     // pixel index of bit 0 of the word above
  while (c) {
    int a, len;
    c &= ~first_run(c, a, len);
    const int b = a + len - 1, me = p0 + a;
    uint32_t um = u & bits_range(a > 0 ? a - 1 : 0, b < 31 ? b + 1 : 31);    // 8-connected neighbours in the word above
    while (um) {
      int ua, ulen;
      um &= ~first_run(um, ua, ulen);
      uf_union(lab, me, q0 + run_start_of(u, ua));
    }
    if (a == 0 && (ul >> 31)) uf_union(lab, me, q0 - 32 + run_start_of(ul, 31));
    if (b == 31 && (ur & 1u)) uf_union(lab, me, q0 + 32);
  }
}

// path compression of the run starts; roots initialise their accumulators when `comp` is given
__global__ void __launch_bounds__(256) k_cclb_flatten(const uint32_t* __restrict__ bits, int* __restrict__ lab_all, int H, int W, int Wd,
                                                      int* __restrict__ comp, size_t plane) {
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (w >= Wd || y >= H) return;
  uint32_t c = bits[((size_t)blockIdx.z * H + y) * Wd + w];
  const size_t base = (size_t)blockIdx.z * H * W;
  int* lab = lab_all + base;
  while (c) {
    int a, len;
    c &= ~first_run(c, a, len);
    const int p = y * W + 32 * w + a;
    const int r = uf_find(lab, p);
    lab[p] = r;
    if (comp && r == p) {
      int* k = comp + base + p;
      k[0] = 0; k[plane] = INT_MAX; k[2 * plane] = -1; k[3 * plane] = INT_MAX; k[4 * plane] = -1; k[5 * plane] = INT_MAX;
    }
  }
}

// expand a word to 32 bytes of {0, value} and store (vector store when aligned)
__device__ __forceinline__ void store_bits_as_bytes(uint8_t* __restrict__ o, uint32_t word, int nvalid, uint32_t value) {
  if (nvalid >= 32 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
    uint32_t v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = ((((word >> (4 * k)) & 15u) * 0x00204081u) & 0x01010101u) * value;
    *reinterpret_cast<uint4*>(o) = make_uint4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<uint4*>(o + 16) = make_uint4(v[4], v[5], v[6], v[7]);
  } else {
    for (int k = 0; k < nvalid && k < 32; ++k) o[k] = ((word >> k) & 1u) ? (uint8_t)value : 0;
  }
}

}  // namespace s2
#include "cols_kernels.cuh"
namespace s2 {

// Canny hysteresis: flag[root] = 1 for components holding a strong pixel, then keep flagged runs
__global__ void __launch_bounds__(256) k_cclb_hyst_mark(const uint32_t* __restrict__ bits, const uint32_t* __restrict__ strong,
                                                        const int* __restrict__ lab_all, uint8_t* __restrict__ flag_all, int H, int W, int Wd) {
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (w >= Wd || y >= H) return;
  const size_t wb = ((size_t)blockIdx.z * H + y) * Wd + w;
  const uint32_t sg = strong[wb];
  if (!sg) return;
  uint32_t c = bits[wb];
  const size_t base = (size_t)blockIdx.z * H * W;
  while (c) {
    int a, len;
    const uint32_t rm = first_run(c, a, len);
    c &= ~rm;
    if (sg & rm) flag_all[base + lab_all[base + y * W + 32 * w + a]] = 1;
  }
}
__global__ void __launch_bounds__(256) k_cclb_hyst_final(const uint32_t* __restrict__ bits, const int* __restrict__ lab_all,
                                                         const uint8_t* __restrict__ flag_all, uint8_t* __restrict__ out, int H, int W, int Wd) {
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (w >= Wd || y >= H) return;
  uint32_t c = bits[((size_t)blockIdx.z * H + y) * Wd + w], keep = 0;
  const size_t base = (size_t)blockIdx.z * H * W;
  while (c) {
    int a, len;
    const uint32_t rm = first_run(c, a, len);
    c &= ~rm;
    if (flag_all[base + lab_all[base + y * W + 32 * w + a]]) keep |= rm;
  }
  store_bits_as_bytes(out + base + (size_t)y * W + 32 * w, keep, W - 32 * w, 255u);
}

// component statistics, one atomic set per run (flattened labels)
template <bool KEY>
__global__ void __launch_bounds__(256) k_cclb_stats(const uint32_t* __restrict__ bits, const int* __restrict__ lab_all, int H, int W, int Wd,
                                                    int* __restrict__ comp, size_t plane) {
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (w >= Wd || y >= H) return;
  uint32_t c = bits[((size_t)blockIdx.z * H + y) * Wd + w];
  const size_t base = (size_t)blockIdx.z * H * W;
  while (c) {
    int a, len;
    c &= ~first_run(c, a, len);
    const int x = 32 * w + a, xe = x + len - 1;
    int* k = comp + base + lab_all[base + y * W + x];
    atomicAdd(k, len);
    atomicMin(k + plane, x); atomicMax(k + 2 * plane, xe);
    atomicMin(k + 3 * plane, y); atomicMax(k + 4 * plane, y);
    // first appearance in cv2's 2x2-block raster scan (the component's OpenCV label order), as in k_ccl_stats
    if (KEY) atomicMin(k + 5 * plane, ((((y >> 1) * ((W + 1) >> 1)) + (x >> 1)) << 2) | ((y & 1) << 1) | (x & 1));
  }
}

// component filter (infer_two_stage_burr.py:103-117) -> burr bytes {0,1} + per-frame pixel count
__global__ void __launch_bounds__(256) k_cclb_filter(const uint32_t* __restrict__ bits, const int* __restrict__ lab_all, int H, int W, int Wd,
                                                     const int* __restrict__ comp, size_t plane, ub_burr_params P,
                                                     uint8_t* __restrict__ burr, int* __restrict__ px, int px_stride) {
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  int n = 0;
  if (w < Wd && y < H) {
    uint32_t c = bits[((size_t)blockIdx.z * H + y) * Wd + w], keep = 0;
    const size_t base = (size_t)blockIdx.z * H * W;
    while (c) {
      int a, len;
      const uint32_t rm = first_run(c, a, len);
      c &= ~rm;
      const int* k = comp + base + lab_all[base + y * W + 32 * w + a];
      const int area = k[0], cw = k[2 * plane] - k[plane] + 1, ch = k[4 * plane] - k[3 * plane] + 1;
      const int mx = max(cw, ch), mn = min(cw, ch);
      if (area >= P.min_area && area <= P.max_area && mx <= P.max_aspect * mn && cw >= P.min_side && ch >= P.min_side) keep |= rm;
    }
    store_bits_as_bytes(burr + base + (size_t)y * W + 32 * w, keep, W - 32 * w, 1u);
    n = __popc(keep);
  }
  if (px) {
    for (int o = 16; o; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if ((threadIdx.x & 31) == 0 && n) atomicAdd(px + blockIdx.z * px_stride, n);
  }
}

// ---- export in OpenCV numbering (ub_ccl) --------------------------------------
__global__ void k_mark_roots(const int* __restrict__ lab_all, int H, int W, const int* __restrict__ comp, size_t plane,
                             int* __restrict__ mark, int K) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  const size_t base = (size_t)blockIdx.z * H * W;
  const int p = y * W + x;
  if (lab_all[base + p] == p) mark[(size_t)blockIdx.z * K + comp[5 * plane + base + p]] = 1;
}
// inclusive prefix sum of mark[b][0..K) in place, one 1024-thread block per frame
__global__ void __launch_bounds__(1024) k_scan(int* __restrict__ mark, int K, int* __restrict__ n_labels) {
  __shared__ int s_w[32];
  __shared__ int s_carry;
  int* m = mark + (size_t)blockIdx.x * K;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (int i0 = 0; i0 < K; i0 += 1024) {
    const int i = i0 + threadIdx.x;
    int v = i < K ? m[i] : 0;
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
    if (lane == 31) s_w[warp] = v;
    __syncthreads();
    if (warp == 0) {
      int t = s_w[lane];
      for (int o = 1; o < 32; o <<= 1) { int u = __shfl_up_sync(0xffffffffu, t, o); if (lane >= o) t += u; }
      s_w[lane] = t;
    }
    __syncthreads();
    const int carry = s_carry;
    v += carry + (warp ? s_w[warp - 1] : 0);
    if (i < K) m[i] = v;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = v;
    __syncthreads();
  }
  if (threadIdx.x == 0) n_labels[blockIdx.x] = s_carry + 1;
}
__global__ void k_relabel(const int* __restrict__ lab_all, int H, int W, const int* __restrict__ comp, size_t plane,
                          const int* __restrict__ rank, int K, int* __restrict__ labels, int* __restrict__ stats, int max_labels,
                          const int* __restrict__ bgacc) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  const size_t base = (size_t)blockIdx.z * H * W;
  const int p = y * W + x;
  const int r = lab_all[base + p];
  int out = 0;
  if (r >= 0) {
    const int* c = comp + base + r;
    out = rank[(size_t)blockIdx.z * K + c[5 * plane]];
    if (r == p && stats && out < max_labels) {
      int* s = stats + ((size_t)blockIdx.z * max_labels + out) * 5;
      s[0] = c[plane]; s[1] = c[3 * plane]; s[2] = c[2 * plane] - c[plane] + 1; s[3] = c[4 * plane] - c[3 * plane] + 1; s[4] = c[0];
    }
  }
  if (labels) labels[base + p] = out;
  if (p == 0 && stats && max_labels > 0) {
    const int* a = bgacc + blockIdx.z * 8;
    int* s = stats + (size_t)blockIdx.z * max_labels * 5;
    s[0] = a[1]; s[1] = a[3]; s[2] = a[2] - a[1] + 1; s[3] = a[4] - a[3] + 1; s[4] = a[0];
  }
}
// ---- run-based export (ub_ccl's path): the first-appearance keys of the roots as a BIT plane, ranks by popcount ----
// markbits[b][key >> 5] |= 1 << (key & 31) for every root; background accumulators from the zero bits of every word
__global__ void __launch_bounds__(256) k_cclb_mark(const uint32_t* __restrict__ bits, const int* __restrict__ lab_all, int H, int W, int Wd,
                                                   const int* __restrict__ comp, size_t plane, uint32_t* __restrict__ markbits, int KW,
                                                   int* __restrict__ bgacc) {
  __shared__ int s_bg[5];                                       // background count / min x / max x / min y / max y of this block
  const int t = threadIdx.y * 32 + threadIdx.x;
  if (t < 5) s_bg[t] = (t == 1 || t == 3) ? INT_MAX : (t == 0 ? 0 : -1);
  __syncthreads();
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (w < Wd && y < H) {
    const uint32_t word = bits[((size_t)blockIdx.z * H + y) * Wd + w];
    const size_t base = (size_t)blockIdx.z * H * W;
    uint32_t c = word;
    while (c) {
      int a, len;
      c &= ~first_run(c, a, len);
      const int p = y * W + 32 * w + a;
      if (lab_all[base + p] == p) {
        const int key = comp[5 * plane + base + p];
        atomicOr(markbits + (size_t)blockIdx.z * KW + (key >> 5), 1u << (key & 31));
      }
    }
    const int rem = W - 32 * w;
    const uint32_t bg = ~word & (rem >= 32 ? 0xffffffffu : ((1u << rem) - 1u));
    if (bg) {
      atomicAdd(&s_bg[0], __popc(bg));
      atomicMin(&s_bg[1], 32 * w + __ffs(bg) - 1); atomicMax(&s_bg[2], 32 * w + 31 - __clz(bg));
      atomicMin(&s_bg[3], y); atomicMax(&s_bg[4], y);
    }
  }
  __syncthreads();
  if (t == 0 && s_bg[0]) {
    int* a = bgacc + blockIdx.z * 8;
    atomicAdd(a, s_bg[0]);
    atomicMin(a + 1, s_bg[1]); atomicMax(a + 2, s_bg[2]);
    atomicMin(a + 3, s_bg[3]); atomicMax(a + 4, s_bg[4]);
  }
}
// wordprefix[b][i] = number of marked keys in words [0, i); one block per frame
__global__ void __launch_bounds__(1024) k_scan_bits(const uint32_t* __restrict__ markbits, int KW, int* __restrict__ wordprefix,
                                                    int* __restrict__ n_labels) {
  __shared__ int s_w[32];
  const uint32_t* m = markbits + (size_t)blockIdx.x * KW;
  int* wp = wordprefix + (size_t)blockIdx.x * KW;
  const int per = (KW + 1023) / 1024, i0 = threadIdx.x * per, i1 = min(i0 + per, KW);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int sum = 0;
  for (int i = i0; i < i1; ++i) sum += __popc(m[i]);
  int v = sum;
  for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
  if (lane == 31) s_w[warp] = v;
  __syncthreads();
  if (warp == 0) {
    int t = s_w[lane];
    for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, t, o); if (lane >= o) t += u; }
    s_w[lane] = t;
  }
  __syncthreads();
  int run = v - sum + (warp ? s_w[warp - 1] : 0);               // exclusive prefix of this thread's chunk
  for (int i = i0; i < i1; ++i) { wp[i] = run; run += __popc(m[i]); }
  if (threadIdx.x == 1023) n_labels[blockIdx.x] = s_w[31] + 1;
}
// labels (OpenCV numbering) for the 32 pixels of every word + the stats row of every root
__global__ void __launch_bounds__(256) k_cclb_relabel(const uint32_t* __restrict__ bits, const int* __restrict__ lab_all, int H, int W, int Wd,
                                                      const int* __restrict__ comp, size_t plane, const uint32_t* __restrict__ markbits,
                                                      const int* __restrict__ wordprefix, int KW, int* __restrict__ labels,
                                                      int* __restrict__ stats, int max_labels, const int* __restrict__ bgacc) {
  // a thread resolves the labels of the 32 pixels of its word; the warp (32 words of one row) then stores them through
  // shared memory so that every store instruction writes 512 contiguous bytes instead of 32 scattered words
  __shared__ int s_lab[8][32 * 33];
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  const bool act = w < Wd && y < H;
  const size_t base = (size_t)blockIdx.z * H * W;
  int* srow = s_lab[threadIdx.y] + threadIdx.x * 33;
  if (act) {
    uint32_t c = bits[((size_t)blockIdx.z * H + y) * Wd + w];
    int pos = 0;
    while (c) {
      int a, len;
      c &= ~first_run(c, a, len);
      const int p = y * W + 32 * w + a, r = lab_all[base + p];
      const int key = comp[5 * plane + base + r];
      const int lbl = wordprefix[(size_t)blockIdx.z * KW + (key >> 5)] +
                      __popc(markbits[(size_t)blockIdx.z * KW + (key >> 5)] & ((1u << (key & 31)) - 1u)) + 1;
      if (labels) {
        for (; pos < a; ++pos) srow[pos] = 0;
        for (; pos < a + len; ++pos) srow[pos] = lbl;
      }
      if (r == p && stats && lbl < max_labels) {
        const int* k = comp + base + r;
        int* s = stats + ((size_t)blockIdx.z * max_labels + lbl) * 5;
        s[0] = k[plane]; s[1] = k[3 * plane]; s[2] = k[2 * plane] - k[plane] + 1; s[3] = k[4 * plane] - k[3 * plane] + 1; s[4] = k[0];
      }
    }
    if (labels) for (; pos < 32; ++pos) srow[pos] = 0;
    if (w == 0 && y == 0 && stats && max_labels > 0) {
      const int* a = bgacc + blockIdx.z * 8;
      int* s = stats + (size_t)blockIdx.z * max_labels * 5;
      s[0] = a[1]; s[1] = a[3]; s[2] = a[2] - a[1] + 1; s[3] = a[4] - a[3] + 1; s[4] = a[0];
    }
  }
  if (!labels || y >= H) return;                              // uniform per warp (a warp is one row)
  __syncwarp();
  int* orow = labels + base + (size_t)y * W + (size_t)blockIdx.x * 1024;      // first pixel of this warp's 32 words
  const int npx = min(1024, W - blockIdx.x * 1024);                           // pixels of the row this warp covers
  const int* sw = s_lab[threadIdx.y];
  if ((W & 3) == 0 && (reinterpret_cast<uintptr_t>(orow) & 15) == 0) {
    for (int q = threadIdx.x * 4; q < npx; q += 128) {                        // q: pixel offset, 4 pixels per lane and step
      const int* sp = sw + (q >> 5) * 33 + (q & 31);
      *reinterpret_cast<int4*>(orow + q) = make_int4(sp[0], sp[1], sp[2], sp[3]);
    }
  } else {
    for (int q = threadIdx.x; q < npx; q += 32) orow[q] = sw[(q >> 5) * 33 + (q & 31)];
  }
}
__global__ void k_bg_init(int* bgacc, int B) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B * 8) { int k = i & 7; bgacc[i] = (k == 1 || k == 3) ? INT_MAX : (k == 2 || k == 4) ? -1 : 0; }
}

}  // namespace s2

using namespace s2;

// cv2.getStructuringElement(MORPH_ELLIPSE,(k,k)) as row runs (oracle/cvnp.py: ellipse)
static bool make_footprint(int k, Footprint& fp) {
  if (k < 1 || k > 25) return false;
  const int r = k / 2, c = k / 2;
  const double inv_r2 = r ? 1.0 / ((double)r * r) : 0.0;
  fp.nrows = 0; fp.top = 0; fp.bottom = 0;
  for (int i = 0; i < k; ++i) {
    const int dy = i - r;
    if (abs(dy) > r) continue;
    const int dx = (int)lrint(c * sqrt((r * r - dy * dy) * inv_r2));
    const int j1 = c - dx > 0 ? c - dx : 0, j2 = c + dx + 1 < k ? c + dx + 1 : k;
    if (j2 <= j1) continue;
    fp.dy[fp.nrows] = (int8_t)(i - k / 2);           // anchor (k/2, k/2)
    fp.xa[fp.nrows] = (int8_t)(j1 - k / 2);
    fp.xb[fp.nrows] = (int8_t)(j2 - 1 - k / 2);
    if (-(i - k / 2) > fp.top) fp.top = -(i - k / 2);
    if ((i - k / 2) > fp.bottom) fp.bottom = i - k / 2;
    ++fp.nrows;
  }
  return true;
}

// the rows of a footprint ordered so that every run contains the one before it (k_morph_cols extends one accumulated word
// run by run); false when the runs are not nested (never for cv2's ellipses)
static bool sort_footprint(const Footprint& fp, Footprint& out) {
  out = fp;
  for (int i = 1; i < out.nrows; ++i)                          // insertion sort by run width
    for (int j = i; j > 0 && (out.xb[j] - out.xa[j]) < (out.xb[j - 1] - out.xa[j - 1]); --j) {
      std::swap(out.dy[j], out.dy[j - 1]); std::swap(out.xa[j], out.xa[j - 1]); std::swap(out.xb[j], out.xb[j - 1]);
    }
  for (int j = 1; j < out.nrows; ++j)
    if (out.xa[j] > out.xa[j - 1] || out.xb[j] < out.xb[j - 1]) return false;
  return out.top + out.bottom + 1 <= 32 && out.nrows >= 1;
}

// does the compile-time table FpT<K> equal the sorted run-time footprint of size K ?
template <int K>
static bool fp_table_matches(const Footprint& fs) {
  using F = FpT<K>;
  if (fs.nrows != F::N || fs.top != F::TOP || fs.bottom != F::BOT) return false;
  for (int j = 0; j < F::N; ++j)
    if (fs.dy[j] != F::dy(j) || fs.xa[j] != F::xa(j) || fs.xb[j] != F::xb(j)) return false;
  return true;
}
static bool fp_table_ok(int k, const Footprint& fs) {
  switch (k) {
    case 2: return fp_table_matches<2>(fs);
    case 3: return fp_table_matches<3>(fs);
    case 5: return fp_table_matches<5>(fs);
    case 8: return fp_table_matches<8>(fs);
    case 25: return fp_table_matches<25>(fs);
    default: return false;
  }
}

template <int K>
static void launch_morph_cols(int op, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const uint8_t* src, uint8_t* dst, int H, int W,
                              const Footprint& fs, int value, int nstrips, int nblk, int R, int RS) {
  switch (op) {
    case 0: k_morph_cols<0, K><<<grid, block, smem, st>>>(src, dst, H, W, fs, (uint32_t)value, nstrips, nblk, R, RS); break;
    case 1: k_morph_cols<1, K><<<grid, block, smem, st>>>(src, dst, H, W, fs, (uint32_t)value, nstrips, nblk, R, RS); break;
    case 2: k_morph_cols<2, K><<<grid, block, smem, st>>>(src, dst, H, W, fs, (uint32_t)value, nstrips, nblk, R, RS); break;
    default: k_morph_cols<3, K><<<grid, block, smem, st>>>(src, dst, H, W, fs, (uint32_t)value, nstrips, nblk, R, RS); break;
  }
}

static int morph_once(ub_ctx* ctx, bool dil, int k, const uint8_t* src, uint8_t* dst, int B, int H, int W, int value,
                      const uint8_t* notmask, cudaStream_t st) {
  Footprint fp;
  if (!make_footprint(k, fp)) return ub_fail(ctx, UB_ERR_ARG, "ellipse size %d unsupported (1..25)", k);
  dim3 grid(cdiv(W, MT_W), cdiv(H, MT_H), B);
  if (dil) k_morph<true><<<grid, 256, 0, st>>>(src, dst, H, W, fp, value, notmask);
  else k_morph<false><<<grid, 256, 0, st>>>(src, dst, H, W, fp, value, notmask);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

#define RUN(x) do { int r__ = (x); if (r__ != UB_OK) return r__; } while (0)

template <int HALO>
static void launch_morph_words(int op, dim3 grid, cudaStream_t st, const uint8_t* src, uint8_t* dst, int H, int W, const Footprint& fp, int value) {
  switch (op) {
    case 0: k_morph_words<HALO, 0><<<grid, 256, 0, st>>>(src, dst, H, W, fp, (uint32_t)value); break;
    case 1: k_morph_words<HALO, 1><<<grid, 256, 0, st>>>(src, dst, H, W, fp, (uint32_t)value); break;
    case 2: k_morph_words<HALO, 2><<<grid, 256, 0, st>>>(src, dst, H, W, fp, (uint32_t)value); break;
    default: k_morph_words<HALO, 3><<<grid, 256, 0, st>>>(src, dst, H, W, fp, (uint32_t)value); break;
  }
}

// op: 0 dilate, 1 erode, 2 close (dilate->erode), 3 open (erode->dilate); tmp only for the per-pixel fallback (UB_MORPH_WORDS=0)
int morph_op(ub_ctx* ctx, int op, int k, const uint8_t* src, uint8_t* dst, uint8_t* tmp, int B, int H, int W, int value, cudaStream_t st) {
  TimeScope ts(ctx, UB_T_STENCIL, st);
  static const int words_env = getenv("UB_MORPH_WORDS") ? atoi(getenv("UB_MORPH_WORDS")) : 2;   // 2: column-walking, 1: tiled words, 0: per pixel
  if (op < 0 || op > 3) return ub_fail(ctx, UB_ERR_ARG, "bad morph op %d", op);
  if (words_env) {
    Footprint fp;
    if (!make_footprint(k, fp)) return ub_fail(ctx, UB_ERR_ARG, "ellipse size %d unsupported (1..25)", k);
    Footprint fs;
    if (words_env >= 2 && sort_footprint(fp, fs)) {           // column-walking kernel (cols_kernels.cuh)
      const int nrings = op >= 2 ? 2 : 1, reach = nrings * (fp.top + fp.bottom);
      int RS = 4;
      while (RS < fp.top + fp.bottom + 1) RS *= 2;
      const int warps = (size_t)8 * nrings * RS * 128 <= 48 * 1024 ? 8 : 4;
      const int Wd = (W + 31) / 32, nstrips = cdiv(Wd, MC_VALID);
      int R = reach * 2 > 32 ? reach * 2 : 32;
      if (R > H) R = H;
      const int nblk = cdiv(H, R);
      const dim3 grid(cdiv(nstrips * nblk, warps), 1, B), block(32, warps);
      const size_t smem = (size_t)warps * nrings * RS * 128;
      static const int tab_env = getenv("UB_MORPH_TABLES") ? atoi(getenv("UB_MORPH_TABLES")) : 1;
      const int kt = tab_env && fp_table_ok(k, fs) ? k : 0;      // a compile-time footprint when the size is one of the reference's
      switch (kt) {
        case 2: launch_morph_cols<2>(op, grid, block, smem, st, src, dst, H, W, fs, value, nstrips, nblk, R, RS); break;
        case 3: launch_morph_cols<3>(op, grid, block, smem, st, src, dst, H, W, fs, value, nstrips, nblk, R, RS); break;
        case 5: launch_morph_cols<5>(op, grid, block, smem, st, src, dst, H, W, fs, value, nstrips, nblk, R, RS); break;
        case 8: launch_morph_cols<8>(op, grid, block, smem, st, src, dst, H, W, fs, value, nstrips, nblk, R, RS); break;
        case 25: launch_morph_cols<25>(op, grid, block, smem, st, src, dst, H, W, fs, value, nstrips, nblk, R, RS); break;
        default: launch_morph_cols<0>(op, grid, block, smem, st, src, dst, H, W, fs, value, nstrips, nblk, R, RS); break;
      }
      UB_CHECK_LAUNCH(ctx);
      return UB_OK;
    }
    const int halo = (op >= 2 ? 2 : 1) * (k / 2);          // rows the operator reaches above / below (anchor k/2)
    dim3 grid(cdiv(W, FM_W), cdiv(H, FM_TH), B);
    if (halo <= 4) launch_morph_words<4>(op, grid, st, src, dst, H, W, fp, value);
    else if (halo <= 12) launch_morph_words<12>(op, grid, st, src, dst, H, W, fp, value);
    else launch_morph_words<24>(op, grid, st, src, dst, H, W, fp, value);
    UB_CHECK_LAUNCH(ctx);
    return UB_OK;
  }
  switch (op) {
    case 0: return morph_once(ctx, true, k, src, dst, B, H, W, value, nullptr, st);
    case 1: return morph_once(ctx, false, k, src, dst, B, H, W, value, nullptr, st);
    case 2: RUN(morph_once(ctx, true, k, src, tmp, B, H, W, value, nullptr, st));
            return morph_once(ctx, false, k, tmp, dst, B, H, W, value, nullptr, st);
    default: RUN(morph_once(ctx, false, k, src, tmp, B, H, W, value, nullptr, st));
             return morph_once(ctx, true, k, tmp, dst, B, H, W, value, nullptr, st);
  }
}

static int ccl_label(ub_ctx* ctx, const uint8_t* src, int* lab, int B, int H, int W, int* comp, size_t plane, cudaStream_t st) {
  const int segs = (W + 31) / 32;
  dim3 gseg(cdiv(H * segs * 32, 256), 1, B), gpix(cdiv(W, 256), H, B);
  k_ccl_init<<<gseg, 256, 0, st>>>(src, lab, H, W);
  k_ccl_merge<<<gpix, 256, 0, st>>>(lab, H, W);
  k_ccl_flatten<<<gpix, 256, 0, st>>>(lab, H, W, comp, plane);
  UB_CHECK_LAUNCH_N(ctx, 3);
  return UB_OK;
}

// run-based labelling of a u8 image on packed rows: bits (+ strong) planes and lab entries at run starts
static int cclb_label(ub_ctx* ctx, const uint8_t* src, uint32_t* bits, uint32_t* strong, int* lab, int B, int H, int W,
                      int* comp, size_t plane, cudaStream_t st) {
  const int Wd = (W + 31) / 32;
  dim3 g(cdiv(Wd, 32), cdiv(H, 8), B);
  k_cclb_init<<<g, dim3(32, 8), 0, st>>>(src, bits, strong, lab, H, W, Wd);
  k_cclb_merge<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd);
  k_cclb_flatten<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, comp, plane);
  UB_CHECK_LAUNCH_N(ctx, 3);
  return UB_OK;
}

// The fused four-pixel kernels need word-aligned rows (edge_fused.cuh); UB_EDGE_FUSED=0 keeps the per-pixel kernels.
static bool edge_fused_ok(const void* a, const void* b, int W) {
  static const int env = getenv("UB_EDGE_FUSED") ? atoi(getenv("UB_EDGE_FUSED")) : 1;
  return env && W % 4 == 0 && ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 3) == 0;
}
static dim3 edge_fused_grid(int B, int H, int W) { return dim3(cdiv(W, EF_TW), cdiv(H, EF_TH), B); }

int gaussian5(ub_ctx* ctx, const uint8_t* src, uint8_t* dst, int B, int H, int W, cudaStream_t st) {
  TimeScope ts(ctx, UB_T_STENCIL, st);
  static const int cols_env = getenv("UB_BLUR_COLS") ? atoi(getenv("UB_BLUR_COLS")) : 1;
  if (cols_env && edge_fused_ok(src, dst, W)) {                // column-walking kernel (cols_kernels.cuh)
    const int nstrips = cdiv(W / 4, BC_VALID), spc = nstrips < 8 ? nstrips : 8, gx = cdiv(nstrips, spc);
    // rows per warp: a multiple of five (the register window's period) that balances the 4-row halo against whole waves
    // of CTAs (2048 threads per SM)
    const long slots = (long)ctx->sm_count * (2048 / (32 * spc));
    int R = 5; double best = 1e30;
    for (int r = 10; r <= 160; r += 5) {
      const long ctas = (long)B * gx * cdiv(H, r);
      const double cost = (double)((ctas + slots - 1) / slots) * (r + 4 + 6);    // + a fixed start-up term per CTA
      if (cost < best) { best = cost; R = r; }
    }
    k_blur5_cols<<<dim3(gx, cdiv(H, R), B), dim3(32, spc), 0, st>>>(src, dst, H, W, nstrips, R);
    UB_CHECK_LAUNCH(ctx);
    return UB_OK;
  }
  if (edge_fused_ok(src, dst, W)) {
    k_edge_fused<2><<<edge_fused_grid(B, H, W), 256, 0, st>>>(src, dst, nullptr, nullptr, nullptr, H, W, (W + 31) / 32, 0, 0);
    UB_CHECK_LAUNCH(ctx);
    return UB_OK;
  }
  k_gauss5<<<dim3(cdiv(W, GT_W), cdiv(H, GT_H), B), dim3(GT_W, 4), 0, st>>>(src, dst, H, W);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

// cv2.Canny: candidates (stencil) then hysteresis (union-find).  cand: u8 scratch, lab: i32 scratch.
// blur != 0: src is the gray plane and the 5x5 blur is done inside the fused kernel (GaussianBlur + Canny of the burr stage).
// phase: 0 = candidates + hysteresis, 1 = the candidate kernel only, 2 = the hysteresis only (ub_pipeline runs the per-frame flood
// CTAs after the network so that no long-running CTA holds an SM a persistent conv kernel is waiting for); the union-find
// path does all its work in phase 1
static int canny_fused(ub_ctx* ctx, const uint8_t* src, uint8_t* dst, int* lab, int B, int H, int W, int low, int high, int blur, cudaStream_t st,
                       int phase = 0) {
  const int Wd = (W + 31) / 32;
  const size_t nwords = (size_t)B * H * Wd;
  uint32_t* bits = reinterpret_cast<uint32_t*>(ctx->d_lab[1]);
  uint32_t* strong = bits + nwords;
  uint8_t* flag = reinterpret_cast<uint8_t*>(ctx->d_comp);
  // hysteresis as an in-shared-memory flood, one CTA per frame, when both bit planes of a frame fit (cols_kernels.cuh)
  static const int hyst_env = getenv("UB_HYST_FRAME") ? atoi(getenv("UB_HYST_FRAME")) : 1;
  const size_t hf_smem = (size_t)2 * (H + 2) * (Wd + 2) * sizeof(uint32_t);
  const bool hyst_frame = hyst_env && hf_smem <= 200 * 1024;
  if (hyst_frame) lab = nullptr;                              // no union-find entries wanted from the edge kernel
  if (phase == 2 && !hyst_frame) return UB_OK;
  if (phase != 2) {
    TimeScope ts(ctx, UB_T_STENCIL, st);
    if (blur) k_edge_fused<1><<<edge_fused_grid(B, H, W), 256, 0, st>>>(src, nullptr, bits, strong, lab, H, W, Wd, low, high);
    else k_edge_fused<0><<<edge_fused_grid(B, H, W), 256, 0, st>>>(src, nullptr, bits, strong, lab, H, W, Wd, low, high);
    UB_CHECK_LAUNCH(ctx);
  }
  if (hyst_frame && phase == 1) return UB_OK;
  if (hyst_frame) {
    TimeScope ts(ctx, UB_T_CCL, st);
    static bool attr_set[64] = {false};
    if (ctx->device >= 0 && ctx->device < 64 && !attr_set[ctx->device]) {
      UB_CUDA(ctx, cudaFuncSetAttribute(k_hyst_frame, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      attr_set[ctx->device] = true;
    }
    int nthr = H * Wd >= 4096 ? 1024 : (H * Wd >= 1024 ? 512 : 128);
    k_hyst_frame<<<B, nthr, hf_smem, st>>>(bits, strong, dst, H, W, Wd);
    UB_CHECK_LAUNCH(ctx);
    return UB_OK;
  }
  TimeScope ts(ctx, UB_T_CCL, st);
  dim3 g(cdiv(Wd, 32), cdiv(H, 8), B);
  k_cclb_merge<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd);
  k_cclb_flatten<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, nullptr, 0);
  UB_CUDA(ctx, cudaMemsetAsync(flag, 0, (size_t)B * H * W, st));
  k_cclb_hyst_mark<<<g, dim3(32, 8), 0, st>>>(bits, strong, lab, flag, H, W, Wd);
  k_cclb_hyst_final<<<g, dim3(32, 8), 0, st>>>(bits, lab, flag, dst, H, W, Wd);
  UB_CHECK_LAUNCH_N(ctx, 4);
  return UB_OK;
}

int canny(ub_ctx* ctx, const uint8_t* src, uint8_t* dst, uint8_t* cand, int* lab, int B, int H, int W, int low, int high, cudaStream_t st) {
  if (edge_fused_ok(src, src, W)) return canny_fused(ctx, src, dst, lab, B, H, W, low, high, 0, st);
  dim3 gpix(cdiv(W, 256), H, B);
  {
    TimeScope ts(ctx, UB_T_STENCIL, st);
    k_canny_nms<<<dim3(cdiv(W, GT_W), cdiv(H, GT_H), B), dim3(GT_W, 4), 0, st>>>(src, cand, H, W, low, high);
    UB_CHECK_LAUNCH(ctx);
  }
  TimeScope ts(ctx, UB_T_CCL, st);
  static const int packed_env = getenv("UB_CCL_PACKED") ? atoi(getenv("UB_CCL_PACKED")) : 1;
  if (packed_env) {
    const int Wd = (W + 31) / 32;
    const size_t nwords = (size_t)B * H * Wd;
    uint32_t* bits = reinterpret_cast<uint32_t*>(ctx->d_lab[1]);       // free during the burr stage
    uint32_t* strong = bits + nwords;
    uint8_t* flag = reinterpret_cast<uint8_t*>(ctx->d_comp);            // [B*H*W] root flags (accumulator scratch, unused until the final CCL)
    RUN(cclb_label(ctx, cand, bits, strong, lab, B, H, W, nullptr, 0, st));
    UB_CUDA(ctx, cudaMemsetAsync(flag, 0, (size_t)B * H * W, st));
    dim3 g(cdiv(Wd, 32), cdiv(H, 8), B);
    k_cclb_hyst_mark<<<g, dim3(32, 8), 0, st>>>(bits, strong, lab, flag, H, W, Wd);
    k_cclb_hyst_final<<<g, dim3(32, 8), 0, st>>>(bits, lab, flag, dst, H, W, Wd);
    UB_CHECK_LAUNCH_N(ctx, 2);
    return UB_OK;
  }
  RUN(ccl_label(ctx, cand, lab, B, H, W, nullptr, 0, st));
  UB_CUDA(ctx, cudaMemsetAsync(dst, 0, (size_t)B * H * W, st));
  k_hyst_mark<<<gpix, 256, 0, st>>>(cand, lab, dst, H, W);
  k_hyst_final<<<gpix, 256, 0, st>>>(lab, dst, H, W);
  UB_CHECK_LAUNCH_N(ctx, 2);
  return UB_OK;
}

// GaussianBlur((5,5),1.0) -> Canny(low, high); scratch: d_tmp[0] (blurred plane), d_tmp[2], d_lab[0..1], d_comp
int blur_canny(ub_ctx* ctx, const uint8_t* gray, uint8_t* dst, int B, int H, int W, int low, int high, cudaStream_t st, int phase) {
  if (edge_fused_ok(gray, gray, W)) return canny_fused(ctx, gray, dst, ctx->d_lab[0], B, H, W, low, high, 1, st, phase);
  if (phase == 2) return UB_OK;
  RUN(gaussian5(ctx, gray, ctx->d_tmp[0], B, H, W, st));
  return canny(ctx, ctx->d_tmp[0], dst, ctx->d_tmp[2], ctx->d_lab[0], B, H, W, low, high, st);
}

int enh_edges(ub_ctx* ctx, const uint8_t* gray, int B, int H, int W, int mode, int sobel_t, int lap_t,
              const uint8_t* base_edges, uint8_t* out, cudaStream_t st) {
  TimeScope ts(ctx, UB_T_STENCIL, st);
  dim3 gpix(cdiv(W, 256), H, B);
  int* fmax = ctx->d_counts + (size_t)ctx->max_batch * 8;          // [B] after the bg accumulators
  if (mode & 1) {
    UB_CUDA(ctx, cudaMemsetAsync(fmax, 0, sizeof(int) * B, st));
    k_sobel_max<<<gpix, 256, 0, st>>>(gray, H, W, fmax);
  }
  k_enh_edges<<<gpix, 256, 0, st>>>(gray, H, W, fmax, mode, sobel_t, lap_t, base_edges, out);
  UB_CHECK_LAUNCH_N(ctx, (mode & 1) ? 2 : 1);
  return UB_OK;
}

// ---- ub_sobel_edges_static: the Sobel-magnitude edge map without a context and without any allocation ------------
// Scratch of the two-kernel scheme (k_sobel_max -> k_enh_edges) is the per-frame maximum of gx^2+gy^2; here it lives in a
// file-scope device array that every GPU loading the library holds until exit (UB_STATIC_MAX_FRAMES ints = 16 KB).  In
// host code the array's name is only a stand-in: its device address is looked up once per device with
// cudaGetSymbolAddress while that device is current.  All calls share the array (calls that overlap on the GPU race).
static __device__ __align__(256) int g_static_sobel_max[UB_STATIC_MAX_FRAMES];

int sobel_edges_static(const uint8_t* gray, int B, int H, int W, int thresh, uint8_t* dst, cudaStream_t st) {
  static int* addr[64] = {nullptr};                     // device address of the scratch array, per device ordinal
  if (!gray || !dst) return ub_fail(nullptr, UB_ERR_ARG, "ub_sobel_edges_static: null pointer");
  if (B < 1 || H < 1 || W < 1) return ub_fail(nullptr, UB_ERR_ARG, "ub_sobel_edges_static: empty shape %dx%dx%d", B, H, W);
  if (B > UB_STATIC_MAX_FRAMES)
    return ub_fail(nullptr, UB_ERR_ARG, "ub_sobel_edges_static: %d frames exceed the static scratch (UB_STATIC_MAX_FRAMES = %d); "
                   "split the batch or use ub_sobel_edges", B, UB_STATIC_MAX_FRAMES);
  if (H > 65535 || B > 65535) return ub_fail(nullptr, UB_ERR_ARG, "ub_sobel_edges_static: H / B exceed the grid limits");
  int dev = -1;
  UB_CUDA(nullptr, cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return ub_fail(nullptr, UB_ERR_ARG, "ub_sobel_edges_static: device ordinal %d", dev);
  if (!addr[dev]) {                                     // set-up, once per device, with that device current
    int* p = nullptr;
    UB_CUDA(nullptr, cudaGetSymbolAddress((void**)&p, g_static_sobel_max));
    addr[dev] = p;
  }
  int* fmax = addr[dev];
  dim3 gpix(cdiv(W, 256), H, B);
  UB_CUDA(nullptr, cudaMemsetAsync(fmax, 0, sizeof(int) * B, st));      // the array keeps the last call's maxima
  k_sobel_max<<<gpix, 256, 0, st>>>(gray, H, W, fmax);
  k_enh_edges<<<gpix, 256, 0, st>>>(gray, H, W, fmax, 1, thresh, 0, nullptr, dst);
  UB_CUDA(nullptr, cudaGetLastError());
  return UB_OK;
}

// labels / stats in OpenCV numbering (ub_ccl).  Run-based path: labelling on packed rows, the roots' first-appearance keys as a
// bit plane, ranks by popcount (a 32x smaller scan than one int per key), labels written 32 pixels per thread.
// UB_CCL_PACKED=0 keeps the per-pixel kernels.
int ccl_export(ub_ctx* ctx, const uint8_t* src, int B, int H, int W, int* labels, int* stats, int max_labels, int* n_labels, cudaStream_t st) {
  TimeScope ts(ctx, UB_T_CCL, st);
  const size_t plane = (size_t)ctx->max_batch * ctx->frame_h * ctx->frame_w;
  int* lab = ctx->d_lab[0];
  int* bg = ctx->d_counts;
  const int K = ((H + 1) / 2) * ((W + 1) / 2) * 4;
  static const int packed_env = getenv("UB_CCL_PACKED") ? atoi(getenv("UB_CCL_PACKED")) : 1;
  if (packed_env) {
    const int Wd = (W + 31) / 32, KW = (K + 31) / 32;
    const size_t nwords = (size_t)B * H * Wd;
    uint32_t* bits = reinterpret_cast<uint32_t*>(ctx->d_lab[1]);
    uint32_t* markbits = bits + nwords;                                  // [B][KW]
    int* wordprefix = reinterpret_cast<int*>(markbits + (size_t)B * KW); // [B][KW]   (nwords + 2*B*KW <= ints of d_lab[1])
    dim3 g(cdiv(Wd, 32), cdiv(H, 8), B);
    RUN(cclb_label(ctx, src, bits, nullptr, lab, B, H, W, ctx->d_comp, plane, st));
    k_cclb_stats<true><<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane);
    k_bg_init<<<cdiv(B * 8, 256), 256, 0, st>>>(bg, B);
    UB_CUDA(ctx, cudaMemsetAsync(markbits, 0, sizeof(uint32_t) * (size_t)B * KW, st));
    k_cclb_mark<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane, markbits, KW, bg);
    k_scan_bits<<<B, 1024, 0, st>>>(markbits, KW, wordprefix, n_labels ? n_labels : bg + (size_t)ctx->max_batch * 8 + ctx->max_batch);
    if (stats) UB_CUDA(ctx, cudaMemsetAsync(stats, 0, sizeof(int) * (size_t)B * max_labels * 5, st));
    k_cclb_relabel<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane, markbits, wordprefix, KW, labels, stats, max_labels, bg);
    UB_CHECK_LAUNCH_N(ctx, 5);
    return UB_OK;
  }
  int* mark = ctx->d_lab[1];
  const int segs = (W + 31) / 32;
  dim3 gseg(cdiv(H * segs * 32, 256), 1, B), gpix(cdiv(W, 256), H, B);
  RUN(ccl_label(ctx, src, lab, B, H, W, ctx->d_comp, plane, st));
  k_bg_init<<<cdiv(B * 8, 256), 256, 0, st>>>(bg, B);
  k_ccl_stats<<<gseg, 256, 0, st>>>(lab, H, W, ctx->d_comp, plane, bg);
  UB_CUDA(ctx, cudaMemsetAsync(mark, 0, sizeof(int) * (size_t)B * K, st));
  k_mark_roots<<<gpix, 256, 0, st>>>(lab, H, W, ctx->d_comp, plane, mark, K);
  k_scan<<<B, 1024, 0, st>>>(mark, K, n_labels ? n_labels : bg + (size_t)ctx->max_batch * 8 + ctx->max_batch);
  if (stats) UB_CUDA(ctx, cudaMemsetAsync(stats, 0, sizeof(int) * (size_t)B * max_labels * 5, st));
  k_relabel<<<gpix, 256, 0, st>>>(lab, H, W, ctx->d_comp, plane, mark, K, labels, stats, max_labels, bg);
  UB_CHECK_LAUNCH_N(ctx, 5);
  return UB_OK;
}

// First half of the burr stage: the edge map, which depends on the gray frame only (not on the net's masks), so
// ub_pipeline runs it on a second stream beside the network.  Result in ctx->d_tmp[1].
// phase: 0 = everything, 1 = what can run beside the network without holding SMs for long (gray-only stencils),
// 2 = the rest (per-frame hysteresis flood, the enhanced variant's extra edge maps); see canny_fused
int burr_edges(ub_ctx* ctx, int variant, const uint8_t* gray, int B, int H, int W, const ub_burr_params& P, cudaStream_t st, int phase) {
  uint8_t* t1 = ctx->d_tmp[1];
  if (variant == UB_VARIANT_LAPLACIAN) {      // src/refactor/burr_detector.py:44-51: wrapped |Laplacian| > threshold, nothing else
    if (phase == 2) return UB_OK;
    return enh_edges(ctx, gray, B, H, W, 2, 0, P.canny_low, nullptr, t1, st);
  }
  // edges: blur -> Canny (-> | Sobel | Laplacian for the enhanced variant)        [t1]
  RUN(blur_canny(ctx, gray, t1, B, H, W, P.canny_low, P.canny_high, st, phase));
  if (variant == UB_VARIANT_ENHANCED && phase != 1) RUN(enh_edges(ctx, gray, B, H, W, 3, 50, 15, t1, t1, st));
  return UB_OK;
}

// The whole burr stage for a batch.
int burr_stage(ub_ctx* ctx, int variant, const uint8_t* gray, const uint8_t* mask, int B, int H, int W,
               const ub_burr_params& P, uint8_t* burr, int32_t* burr_px, int px_stride, cudaStream_t st) {
  RUN(burr_edges(ctx, variant, gray, B, H, W, P, st, 0));
  return burr_finish(ctx, mask, B, H, W, P, burr, burr_px, px_stride, st);
}

// Second half: detection band, candidates, CLOSE / OPEN, components, filter (edges in ctx->d_tmp[1]).
int burr_finish(ub_ctx* ctx, const uint8_t* mask, int B, int H, int W, const ub_burr_params& P, uint8_t* burr,
                int32_t* burr_px, int px_stride, cudaStream_t st) {
  uint8_t *t0 = ctx->d_tmp[0], *t1 = ctx->d_tmp[1], *t2 = ctx->d_tmp[2], *t3 = ctx->d_tmp[3];
  int* lab = ctx->d_lab[0];
  const size_t n = (size_t)B * H * W;
  const size_t plane = (size_t)ctx->max_batch * ctx->frame_h * ctx->frame_w;
  static const int fuse_env = getenv("UB_MORPH_FUSED") ? atoi(getenv("UB_MORPH_FUSED")) : 1;
  MorphChain mc;
  const bool fp_ok = make_footprint(P.band_k, mc.band) && make_footprint(P.close_k, mc.close) && make_footprint(P.open_k, mc.open);
  if (!fp_ok) return ub_fail(ctx, UB_ERR_ARG, "ellipse sizes %d/%d/%d unsupported (1..25)", P.band_k, P.close_k, P.open_k);
  static const int cols_env = getenv("UB_BURR_COLS") ? atoi(getenv("UB_BURR_COLS")) : 1;
  static const int packed_env0 = getenv("UB_CCL_PACKED") ? atoi(getenv("UB_CCL_PACKED")) : 1;
  MorphChain ms;
  if (cols_env && packed_env0 && sort_footprint(mc.band, ms.band) && sort_footprint(mc.close, ms.close) && sort_footprint(mc.open, ms.open) &&
      P.band_k / 2 + 2 * (P.close_k / 2) + 2 * (P.open_k / 2) <= 31) {
    // band, candidates, CLOSE, OPEN on one walk down 32-pixel-per-lane strips (cols_kernels.cuh); writes the packed candidate
    // rows and their run starts' union-find entries directly, so k_cclb_init and the u8 candidate plane drop out
    const int Wd = (W + 31) / 32, nstrips = cdiv(Wd, MC_VALID);
    uint32_t* bits = reinterpret_cast<uint32_t*>(ctx->d_lab[1]);
    {
      TimeScope ts(ctx, UB_T_STENCIL, st);
      int RS = 4;
      while (RS < mc.band.top + mc.band.bottom + 1 || RS < mc.close.top + mc.close.bottom + 1 || RS < mc.open.top + mc.open.bottom + 1) RS *= 2;
      const int warps = (size_t)8 * 6 * RS * 128 <= 48 * 1024 ? 8 : ((size_t)4 * 6 * RS * 128 <= 48 * 1024 ? 4 : 2);
      // rows per warp: every task also walks the 13 (standard) rows of reach above and below its rows, so few long tasks beat many
      // short ones as long as every SM has a dozen warps (batch 64, ncu: 4 rows 67.5 us, 8: 54.7, 12: 51.4, 16: 50.5; profiles/r03b)
      long want = ((long)B * H * nstrips + (long)ctx->sm_count * 16 - 1) / ((long)ctx->sm_count * 16);
      int R = want < 12 ? 12 : (want > 64 ? 64 : (int)want);
      static const int rows_env = getenv("UB_CHAIN_ROWS") ? atoi(getenv("UB_CHAIN_ROWS")) : 0;       // tuning override
      if (rows_env > 0) R = rows_env;
      if (R > H) R = H;
      const int nblk = cdiv(H, R);
      static const int tab_env = getenv("UB_MORPH_TABLES") ? atoi(getenv("UB_MORPH_TABLES")) : 1;
      const bool tabs = tab_env && fp_table_ok(P.band_k, ms.band) && fp_table_ok(P.close_k, ms.close) && fp_table_ok(P.open_k, ms.open);
      const dim3 grid(cdiv(nstrips * nblk, warps), 1, B), block(32, warps);
      const size_t smem = (size_t)warps * 6 * RS * 128;
      if (tabs && P.band_k == 8 && P.close_k == 3 && P.open_k == 2)          // infer_two_stage_burr.py:78-97
        k_burr_chain_cols<8, 3, 2><<<grid, block, smem, st>>>(mask, t1, nullptr, bits, lab, H, W, ms, nstrips, nblk, R, RS);
      else if (tabs && P.band_k == 25 && P.close_k == 5 && P.open_k == 3)    // infer_enhanced_burr.py:82-116
        k_burr_chain_cols<25, 5, 3><<<grid, block, smem, st>>>(mask, t1, nullptr, bits, lab, H, W, ms, nstrips, nblk, R, RS);
      else
        k_burr_chain_cols<0, 0, 0><<<grid, block, smem, st>>>(mask, t1, nullptr, bits, lab, H, W, ms, nstrips, nblk, R, RS);
      UB_CHECK_LAUNCH(ctx);
    }
    TimeScope ts(ctx, UB_T_CCL, st);
    dim3 g(cdiv(Wd, 32), cdiv(H, 8), B);
    k_cclb_merge<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd);
    k_cclb_flatten<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane);
    k_cclb_stats<false><<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane);
    k_cclb_filter<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane, P, burr, burr_px, px_stride);
    UB_CHECK_LAUNCH_N(ctx, 4);
    return UB_OK;
  }
  if (fuse_env && 2 * (P.open_k / 2) + 2 * (P.close_k / 2) + P.band_k / 2 <= FM_HALO) {
    // band, candidates, CLOSE, OPEN in one pass over bit-packed tiles                -> t2
    TimeScope ts(ctx, UB_T_STENCIL, st);
    const dim3 gm(cdiv(W, FM_W), cdiv(H, FM_TH), B);
    if (2 * (P.open_k / 2) + 2 * (P.close_k / 2) + P.band_k / 2 <= 8) k_burr_morph<8><<<gm, 256, 0, st>>>(mask, t1, t2, H, W, mc);
    else k_burr_morph<FM_HALO><<<gm, 256, 0, st>>>(mask, t1, t2, H, W, mc);
    UB_CHECK_LAUNCH(ctx);
  } else {
    {
      TimeScope ts(ctx, UB_T_STENCIL, st);
      // detection band = dilate(mask) & ~mask                                      [t0]
      RUN(morph_once(ctx, true, P.band_k, mask, t0, B, H, W, 1, mask, st));
      // candidates = edges & band                                                  [t2]
      k_and01<<<(unsigned)((n / 4 + 255) / 256 + 1), 256, 0, st>>>(t1, t0, t2, n);
      UB_CHECK_LAUNCH(ctx);
    }
    RUN(morph_op(ctx, 2, P.close_k, t2, t1, t3, B, H, W, 1, st));                   // CLOSE -> t1
    RUN(morph_op(ctx, 3, P.open_k, t1, t2, t3, B, H, W, 1, st));                    // OPEN  -> t2
  }
  // components + filter
  TimeScope ts(ctx, UB_T_CCL, st);
  const int segs = (W + 31) / 32;
  dim3 gseg(cdiv(H * segs * 32, 256), 1, B), gpix(cdiv(W, 256), H, B);
  static const int packed_env = getenv("UB_CCL_PACKED") ? atoi(getenv("UB_CCL_PACKED")) : 1;
  if (packed_env) {
    const int Wd = (W + 31) / 32;
    uint32_t* bits = reinterpret_cast<uint32_t*>(ctx->d_lab[1]);
    RUN(cclb_label(ctx, t2, bits, nullptr, lab, B, H, W, ctx->d_comp, plane, st));
    dim3 g(cdiv(Wd, 32), cdiv(H, 8), B);
    k_cclb_stats<false><<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane);
    k_cclb_filter<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane, P, burr, burr_px, px_stride);
    UB_CHECK_LAUNCH_N(ctx, 2);
    return UB_OK;
  }
  RUN(ccl_label(ctx, t2, lab, B, H, W, ctx->d_comp, plane, st));
  k_ccl_stats<<<gseg, 256, 0, st>>>(lab, H, W, ctx->d_comp, plane, nullptr);
  k_ccl_filter<<<gpix, 256, 0, st>>>(lab, H, W, ctx->d_comp, plane, P, burr, burr_px, px_stride);
  UB_CHECK_LAUNCH_N(ctx, 2);
  return UB_OK;
}

// =====================================================================================================
// compute_diameter_metrics (src/utils/geometry_enhanced.py:113-187): for the cable and the tape class keep
// the largest 8-connected component of area >= min_area (_largest_connected_component, :79-110), take every
// row's width max_x - min_x + 1 (:64-67), smooth the two width profiles with cv2.GaussianBlur((1,31), 0) on
// float32 (:70-76), and report the medians over the rows where both are > 0 (:150-171).
// Device form: run-based labelling (above) -> best component per frame by a 64-bit atomicMax on
// (area, OpenCV label order) -> one warp per row for the extent -> one block per frame for the float32
// smoothing (cv2's exact operation order, oracle/cvnp.py: gauss31_f32), compaction, bitonic sort, medians.
// =====================================================================================================
namespace s2 {

__device__ __forceinline__ uint32_t eq_bytes01(uint32_t w, uint32_t cls4) {          // byte j = (byte j of w == cls) ? 1 : 0
  const uint32_t v = w ^ cls4;                                                         // zero byte where equal
  return ((((v & 0x7f7f7f7fu) + 0x7f7f7f7fu) | v) & 0x80808080u) >> 7 ^ 0x01010101u;
}
// 16 pixels per thread (one 16-byte vector each way) when both planes are 16-byte aligned
__global__ void k_class_mask(const uint8_t* __restrict__ pred, uint8_t* __restrict__ out, size_t n, int cls) {
  const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 16;
  if (i >= n) return;
  const uint32_t c4 = 0x01010101u * (uint32_t)cls;
  if (i + 16 <= n && ((reinterpret_cast<uintptr_t>(pred) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {
    const uint4 v = *reinterpret_cast<const uint4*>(pred + i);
    *reinterpret_cast<uint4*>(out + i) = make_uint4(eq_bytes01(v.x, c4), eq_bytes01(v.y, c4), eq_bytes01(v.z, c4), eq_bytes01(v.w, c4));
  } else {
    for (size_t k = i; k < n && k < i + 16; ++k) out[k] = pred[k] == cls;
  }
}

// best[b] = max over components with area >= min_area of (area << 32 | INT_MAX - first-appearance key)
__global__ void __launch_bounds__(256) k_diam_select(const uint32_t* __restrict__ bits, const int* __restrict__ lab_all, int H, int W, int Wd,
                                                     const int* __restrict__ comp, size_t plane, int min_area,
                                                     unsigned long long* __restrict__ best) {
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (w >= Wd || y >= H) return;
  uint32_t c = bits[((size_t)blockIdx.z * H + y) * Wd + w];
  const size_t base = (size_t)blockIdx.z * H * W;
  while (c) {
    int a, len;
    c &= ~first_run(c, a, len);
    const int p = y * W + 32 * w + a;
    if (lab_all[base + p] != p) continue;                       // roots only
    const int area = comp[base + p];
    if (area >= min_area)
      atomicMax(best + blockIdx.z, ((unsigned long long)(uint32_t)area << 32) | (uint32_t)(INT_MAX - comp[5 * plane + base + p]));
  }
}

// one warp per (frame, row): extent of the best component in the row -> width (0 when absent)
__global__ void __launch_bounds__(256) k_diam_rows(const uint32_t* __restrict__ bits, const int* __restrict__ lab_all, int H, int W, int Wd,
                                                   const int* __restrict__ comp, size_t plane,
                                                   const unsigned long long* __restrict__ best, float* __restrict__ widths,
                                                   int frame_stride) {
  const int y = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (y >= H) return;
  const unsigned long long b = best ? best[blockIdx.z] : ~0ull;              // best == nullptr: every run of the plane counts
  const int b_area = (int)(b >> 32), b_key = INT_MAX - (int)(uint32_t)b;
  const size_t base = (size_t)blockIdx.z * H * W;
  int lo = INT_MAX, hi = -1;
  if (b)
    for (int w = lane; w < Wd; w += 32) {
      uint32_t c = bits[((size_t)blockIdx.z * H + y) * Wd + w];
      while (c) {
        int a, len;
        c &= ~first_run(c, a, len);
        const int x = 32 * w + a;
        bool take = best == nullptr;
        if (!take) { const int r = lab_all[base + y * W + x]; take = comp[base + r] == b_area && comp[5 * plane + base + r] == b_key; }
        if (take) { lo = min(lo, x); hi = max(hi, x + len - 1); }
      }
    }
  for (int o = 16; o; o >>= 1) { lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
  if (lane == 0) widths[(size_t)blockIdx.z * frame_stride + y] = hi >= 0 ? (float)(hi - lo + 1) : 0.f;
}

// float32(cv2.getGaussianKernel(31, 0)) -- sigma = 0.3*((31-1)*0.5-1)+0.8 = 5; centre first (tests pin the table against cv2)
__constant__ float c_gauss31[16] = {0x1.476facp-4f, 0x1.40f3dap-4f, 0x1.2e43p-4f, 0x1.117f6ap-4f, 0x1.db88f4p-5f, 0x1.8d334p-5f,
                                    0x1.3ec2b8p-5f, 0x1.eb8fdp-6f, 0x1.6c286ep-6f, 0x1.033262p-6f, 0x1.628266p-7f, 0x1.d1dbc8p-8f,
                                    0x1.2616cp-8f, 0x1.64bf72p-9f, 0x1.9fc9dp-10f, 0x1.d19932p-11f};

constexpr int DM_MAXH = 2048;
// one block per frame: smooth both profiles, valid rows, medians.  out[b] = {dc, dt (float bits), valid_rows, cable_area, tape_area, 0, 0, 0}
__global__ void __launch_bounds__(1024) k_diam_finish(const float* __restrict__ wc_all, const float* __restrict__ wt_all, int H,
                                                      const unsigned long long* __restrict__ best_c, const unsigned long long* __restrict__ best_t,
                                                      int min_valid_rows, int32_t* __restrict__ out) {
  __shared__ float s_c[DM_MAXH], s_t[DM_MAXH];
  __shared__ int s_n;
  const float* wc = wc_all + (size_t)blockIdx.x * H;
  const float* wt = wt_all + (size_t)blockIdx.x * H;
  if (threadIdx.x == 0) s_n = 0;
  float sc[DM_MAXH / 1024], st[DM_MAXH / 1024];
#pragma unroll
  for (int q = 0; q < DM_MAXH / 1024; ++q) {
    const int y = threadIdx.x + q * 1024;
    sc[q] = st[q] = 0.f;
    if (y < H) {
      // cv2's symmetric column filter: s = k0*c; s += kj*(p[+j] + p[-j]), every product and sum rounded to binary32, no FMA
      float a = __fmul_rn(c_gauss31[0], wc[y]), b = __fmul_rn(c_gauss31[0], wt[y]);
      for (int j = 1; j < 16; ++j) {
        const int yp = reflect101(y + j, H), ym = reflect101(y - j, H);
        a = __fadd_rn(a, __fmul_rn(c_gauss31[j], __fadd_rn(wc[yp], wc[ym])));
        b = __fadd_rn(b, __fmul_rn(c_gauss31[j], __fadd_rn(wt[yp], wt[ym])));
      }
      sc[q] = a; st[q] = b;
    }
  }
  __syncthreads();
  // compaction of the valid rows (order is irrelevant for a median), padding with +inf for the sort
  for (int i = threadIdx.x; i < DM_MAXH; i += 1024) s_c[i] = s_t[i] = __int_as_float(0x7f800000);
  __syncthreads();
#pragma unroll
  for (int q = 0; q < DM_MAXH / 1024; ++q) {
    const int y = threadIdx.x + q * 1024;
    if (y < H && sc[q] > 0.f && st[q] > 0.f) { const int k = atomicAdd(&s_n, 1); s_c[k] = sc[q]; s_t[k] = st[q]; }
  }
  __syncthreads();
  const int n = s_n;
  int np2 = 1;
  while (np2 < n) np2 <<= 1;
  for (int size = 2; size <= np2; size <<= 1)
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = threadIdx.x; i < np2; i += 1024) {
        const int j = i ^ stride;
        if (j > i) {
          const bool up = (i & size) == 0;
          float x = s_c[i], y2 = s_c[j];
          if ((x > y2) == up) { s_c[i] = y2; s_c[j] = x; }
          x = s_t[i]; y2 = s_t[j];
          if ((x > y2) == up) { s_t[i] = y2; s_t[j] = x; }
        }
      }
      __syncthreads();
    }
  if (threadIdx.x == 0) {
    float dc = 0.f, dt = 0.f;
    if (n >= min_valid_rows && n > 0) {       // np.median on float32: middle element, or float32 mean of the two middle ones
      dc = (n & 1) ? s_c[n >> 1] : __fmul_rn(__fadd_rn(s_c[(n >> 1) - 1], s_c[n >> 1]), 0.5f);
      dt = (n & 1) ? s_t[n >> 1] : __fmul_rn(__fadd_rn(s_t[(n >> 1) - 1], s_t[n >> 1]), 0.5f);
    }
    int32_t* o = out + (size_t)blockIdx.x * 8;
    o[0] = __float_as_int(dc); o[1] = __float_as_int(dt); o[2] = n;
    o[3] = (int)(best_c[blockIdx.x] >> 32); o[4] = (int)(best_t[blockIdx.x] >> 32); o[5] = o[6] = o[7] = 0;
  }
}

// cv2.GaussianBlur(widths.reshape(-1,1), (1,31), 0) for nvec width profiles of H rows each (same arithmetic as k_diam_finish)
__global__ void __launch_bounds__(256) k_smooth31(const float* __restrict__ in, float* __restrict__ out, int H, int nvec) {
  const int y = blockIdx.x * 256 + threadIdx.x, v = blockIdx.y;
  if (y >= H || v >= nvec) return;
  const float* w = in + (size_t)v * H;
  float a = __fmul_rn(c_gauss31[0], w[y]);
  for (int j = 1; j < 16; ++j) a = __fadd_rn(a, __fmul_rn(c_gauss31[j], __fadd_rn(w[reflect101(y + j, H)], w[reflect101(y - j, H)])));
  out[(size_t)v * H + y] = a;
}

// pack bits only (no labelling): the thickness profile takes every pixel of the class, not the largest component
__global__ void __launch_bounds__(256) k_pack_rows(const uint8_t* __restrict__ src, uint32_t* __restrict__ bits, int H, int W, int Wd) {
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (w >= Wd || y >= H) return;
  bits[((size_t)blockIdx.z * H + y) * Wd + w] = pack32(src + (size_t)blockIdx.z * H * W, H, W, y, 32 * w);
}

}  // namespace s2

// _compute_width_per_row(mask, smooth=True, kernel_size=31) (src/utils/geometry_enhanced.py:45-76) of the cable and the tape
// class masks, as compute_thickness_profile calls it (:208-212).  out: float [B][2][H] (cable, tape).
int width_profiles(ub_ctx* ctx, const uint8_t* pred, int B, int H, int W, int cable_cls, int tape_cls, float* out, cudaStream_t st) {
  if (W < 16 || H < 16) return ub_fail(ctx, UB_ERR_ARG, "ub_width_profiles: %dx%d smaller than 16x16", H, W);
  TimeScope ts(ctx, UB_T_CCL, st);
  const size_t n = (size_t)B * H * W;
  const int Wd = (W + 31) / 32;
  uint8_t* m = ctx->d_tmp[0];
  uint32_t* bits = reinterpret_cast<uint32_t*>(ctx->d_lab[1]);
  float* raw = reinterpret_cast<float*>(ctx->d_tmp[1]);              // [B][2][H]
  dim3 g(cdiv(Wd, 32), cdiv(H, 8), B);
  for (int k = 0; k < 2; ++k) {
    k_class_mask<<<(unsigned)((n / 16 + 255) / 256 + 1), 256, 0, st>>>(pred, m, n, k ? tape_cls : cable_cls);
    k_pack_rows<<<g, dim3(32, 8), 0, st>>>(m, bits, H, W, Wd);
    k_diam_rows<<<dim3(cdiv(H, 8), 1, B), 256, 0, st>>>(bits, nullptr, H, W, Wd, nullptr, 0, nullptr, raw + (size_t)k * H, 2 * H);
    UB_CHECK_LAUNCH_N(ctx, 3);
  }
  k_smooth31<<<dim3(cdiv(H, 256), 2 * B), 256, 0, st>>>(raw, out, H, 2 * B);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

// pred: class ids u8 [B,H,W]; out: int32 [B,8] (see k_diam_finish).  Scratch: d_tmp[0], d_lab[0..1], d_comp, d_counts.
int diameter_metrics(ub_ctx* ctx, const uint8_t* pred, int B, int H, int W, int cable_cls, int tape_cls, int min_area,
                     int min_valid_rows, int32_t* out, cudaStream_t st) {
  if (H > DM_MAXH) return ub_fail(ctx, UB_ERR_ARG, "ub_diameter_metrics: H = %d > %d", H, DM_MAXH);
  TimeScope ts(ctx, UB_T_CCL, st);
  const size_t n = (size_t)B * H * W, plane = (size_t)ctx->max_batch * ctx->frame_h * ctx->frame_w;
  const int Wd = (W + 31) / 32;
  uint8_t* m = ctx->d_tmp[0];
  int* lab = ctx->d_lab[0];
  uint32_t* bits = reinterpret_cast<uint32_t*>(ctx->d_lab[1]);
  // widths: 2 x [B,H] float in d_tmp[1] (B*H*W bytes), best[2][B] in d_tmp[2]
  if (W < 16 || H < 16) return ub_fail(ctx, UB_ERR_ARG, "ub_diameter_metrics: %dx%d smaller than 16x16", H, W);
  float* widths = reinterpret_cast<float*>(ctx->d_tmp[1]);
  unsigned long long* best = reinterpret_cast<unsigned long long*>(ctx->d_tmp[2]);
  UB_CUDA(ctx, cudaMemsetAsync(best, 0, sizeof(unsigned long long) * 2 * B, st));
  dim3 g(cdiv(Wd, 32), cdiv(H, 8), B);
  for (int k = 0; k < 2; ++k) {
    k_class_mask<<<(unsigned)((n / 16 + 255) / 256 + 1), 256, 0, st>>>(pred, m, n, k ? tape_cls : cable_cls);
    RUN(cclb_label(ctx, m, bits, nullptr, lab, B, H, W, ctx->d_comp, plane, st));
    k_cclb_stats<true><<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane);
    k_diam_select<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane, min_area, best + (size_t)k * B);
    k_diam_rows<<<dim3(cdiv(H, 8), 1, B), 256, 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane, best + (size_t)k * B,
                                                       widths + (size_t)k * B * H, H);
    UB_CHECK_LAUNCH_N(ctx, 4);
  }
  k_diam_finish<<<B, 1024, 0, st>>>(widths, widths + (size_t)B * H, H, best, best + B, min_valid_rows, out);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

// =====================================================================================================
// analyze_defects (src/utils/geometry_enhanced.py:236-330): tape coverage, tape / cable component counts, largest
// tape component, holes = CLOSE(tape, ELLIPSE 5x5) - tape with components of area >= hole_min_size counted and
// summed, and the pixel count of every class id (defect_areas).
// out[b] = { tape_px, tape_components, tape_largest_area, holes (>= min size), hole_px (of those), cable_components, 0, 0,
//            count[class 0..7] }   (int32 [B][16])
// =====================================================================================================
namespace s2 {

// per-frame summary of a labelled plane: acc[0] += components, acc[1] = max area, acc[2] += components with area >= min_area,
// acc[3] += their area, acc[4] += foreground pixels
__global__ void __launch_bounds__(256) k_comp_summary(const uint32_t* __restrict__ bits, const int* __restrict__ lab_all, int H, int W, int Wd,
                                                      const int* __restrict__ comp, int min_area, int* __restrict__ acc_all, int acc_stride) {
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (w >= Wd || y >= H) return;
  uint32_t c = bits[((size_t)blockIdx.z * H + y) * Wd + w];
  if (!c) return;
  const size_t base = (size_t)blockIdx.z * H * W;
  int* acc = acc_all + (size_t)blockIdx.z * acc_stride;
  atomicAdd(acc + 4, __popc(c));
  while (c) {
    int a, len;
    c &= ~first_run(c, a, len);
    const int p = y * W + 32 * w + a;
    if (lab_all[base + p] != p) continue;
    const int area = comp[base + p];
    atomicAdd(acc, 1);
    atomicMax(acc + 1, area);
    if (area >= min_area) { atomicAdd(acc + 2, 1); atomicAdd(acc + 3, area); }
  }
}

// holes = closed & ~tape as bytes {0,1};  class histogram (ids 0..7) with one shared-memory histogram per block
__global__ void __launch_bounds__(256) k_holes_hist(const uint8_t* __restrict__ pred, const uint8_t* __restrict__ closed, const uint8_t* __restrict__ tape,
                                                    uint8_t* __restrict__ holes, int HW, int* __restrict__ hist_all, int hist_stride) {
  // 16 pixels per thread; the class histogram is counted in registers (byte-compare masks), reduced per warp, then per block
  __shared__ int s_h[8];
  if (threadIdx.x < 8) s_h[threadIdx.x] = 0;
  __syncthreads();
  const size_t base = (size_t)blockIdx.y * HW;
  const int i = (blockIdx.x * 256 + threadIdx.x) * 16;
  int cnt[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (i < HW) {
    const uint8_t *pp = pred + base + i, *pc = closed + base + i, *pt = tape + base + i;
    uint8_t* ph = holes + base + i;
    if (i + 16 <= HW && ((reinterpret_cast<uintptr_t>(pp) | reinterpret_cast<uintptr_t>(pc) | reinterpret_cast<uintptr_t>(pt) |
                          reinterpret_cast<uintptr_t>(ph)) & 15) == 0) {
      const uint4 vp = *reinterpret_cast<const uint4*>(pp), vc = *reinterpret_cast<const uint4*>(pc), vt = *reinterpret_cast<const uint4*>(pt);
      const uint32_t wp[4] = {vp.x, vp.y, vp.z, vp.w}, wc[4] = {vc.x, vc.y, vc.z, vc.w}, wt[4] = {vt.x, vt.y, vt.z, vt.w};
      uint32_t h[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) h[k] = (eq_bytes01(wc[k], 0u) ^ 0x01010101u) & eq_bytes01(wt[k], 0u);       // closed != 0 && tape == 0
      *reinterpret_cast<uint4*>(ph) = make_uint4(h[0], h[1], h[2], h[3]);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const uint32_t c4 = 0x01010101u * (uint32_t)c;
        const uint32_t sum = eq_bytes01(wp[0], c4) + eq_bytes01(wp[1], c4) + eq_bytes01(wp[2], c4) + eq_bytes01(wp[3], c4);   // <= 4 per byte lane
        cnt[c] = (int)((sum * 0x01010101u) >> 24);
      }
    } else {
      for (int k = 0; k < 16 && i + k < HW; ++k) {
        ph[k] = (pc[k] != 0 && pt[k] == 0) ? 1 : 0;
        const int c = pp[k];
#pragma unroll
        for (int q = 0; q < 8; ++q) cnt[q] += c == q;
      }
    }
  }
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const int v = __reduce_add_sync(0xffffffffu, cnt[c]);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(&s_h[c], v);
  }
  __syncthreads();
  if (threadIdx.x < 8 && s_h[threadIdx.x]) atomicAdd(hist_all + (size_t)blockIdx.y * hist_stride + threadIdx.x, s_h[threadIdx.x]);
}

__global__ void k_defects_pack(const int* __restrict__ acc, int B, int32_t* __restrict__ out) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const int* a = acc + (size_t)b * 16;
  int32_t* o = out + (size_t)b * 16;
  o[0] = a[4]; o[1] = a[0]; o[2] = a[1];           // tape pixels, components, largest area
  o[3] = a[5 + 2]; o[4] = a[5 + 3];                // holes of area >= hole_min_size, their pixels
  o[5] = a[10];                                    // cable components
}

}  // namespace s2

int analyze_defects(ub_ctx* ctx, const uint8_t* pred, int B, int H, int W, int cable_cls, int tape_cls, int hole_min_size,
                    int32_t* out, cudaStream_t st) {
  TimeScope ts(ctx, UB_T_CCL, st);
  const size_t n = (size_t)B * H * W, plane = (size_t)ctx->max_batch * ctx->frame_h * ctx->frame_w;
  const int Wd = (W + 31) / 32;
  uint8_t *tape = ctx->d_tmp[0], *other = ctx->d_tmp[1], *closed = ctx->d_tmp[2], *tmp = ctx->d_tmp[3];
  int* lab = ctx->d_lab[0];
  uint32_t* bits = reinterpret_cast<uint32_t*>(ctx->d_lab[1]);
  int* acc = ctx->d_counts;                        // [B][16]: three 5-int summaries per frame (tape, holes, cable)
  dim3 g(cdiv(Wd, 32), cdiv(H, 8), B);
  const unsigned gm = (unsigned)((n / 4 + 255) / 256 + 1);
  UB_CUDA(ctx, cudaMemsetAsync(acc, 0, sizeof(int) * 16 * B, st));
  UB_CUDA(ctx, cudaMemsetAsync(out, 0, sizeof(int32_t) * 16 * B, st));
  // tape: components, largest, pixels
  k_class_mask<<<gm, 256, 0, st>>>(pred, tape, n, tape_cls);
  RUN(cclb_label(ctx, tape, bits, nullptr, lab, B, H, W, ctx->d_comp, plane, st));
  k_cclb_stats<false><<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane);
  k_comp_summary<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, 0, acc, 16);
  // holes = CLOSE(tape, 5x5) - tape, class histogram
  RUN(morph_op(ctx, 2, 5, tape, closed, tmp, B, H, W, 1, st));
  k_holes_hist<<<dim3(cdiv(H * W, 256 * 16), B), 256, 0, st>>>(pred, closed, tape, other, H * W, out + 8, 16);
  RUN(cclb_label(ctx, other, bits, nullptr, lab, B, H, W, ctx->d_comp, plane, st));
  k_cclb_stats<false><<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane);
  k_comp_summary<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, hole_min_size, acc + 5, 16);
  // cable: component count
  k_class_mask<<<gm, 256, 0, st>>>(pred, other, n, cable_cls);
  RUN(cclb_label(ctx, other, bits, nullptr, lab, B, H, W, ctx->d_comp, plane, st));
  k_cclb_stats<false><<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane);
  k_comp_summary<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, 0, acc + 10, 16);
  k_defects_pack<<<cdiv(B, 128), 128, 0, st>>>(acc, B, out);
  UB_CHECK_LAUNCH_N(ctx, 10);
  return UB_OK;
}

// =====================================================================================================
// src/refactor/postprocess.py: filter_cable_by_shape (:12-77) and constrain_tape_to_ring (:80-118).
// Both end in "keep one component": the first by a float64 score over (area, bounding box, centroid), the second by
// area; ties go to the smaller OpenCV label (the loop's strict '>' / np.argmax), i.e. the smaller first-appearance key.
// =====================================================================================================
namespace s2 {

// sums[root] += sum of x over the run: exact integers, like cv2's double accumulation of m10 (the host checks that
// H*W*W/2 fits 32 bits)
__global__ void __launch_bounds__(256) k_cclb_sumx(const uint32_t* __restrict__ bits, const int* __restrict__ lab_all, int H, int W, int Wd,
                                                   uint32_t* __restrict__ sums) {
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (w >= Wd || y >= H) return;
  uint32_t c = bits[((size_t)blockIdx.z * H + y) * Wd + w];
  const size_t base = (size_t)blockIdx.z * H * W;
  while (c) {
    int a, len;
    c &= ~first_run(c, a, len);
    const int x = 32 * w + a;
    atomicAdd(sums + base + lab_all[base + y * W + x], (uint32_t)(len * (2 * x + len - 1) / 2));
  }
}

// score of a root as filter_cable_by_shape computes it (:41-66), or a negative value when a gate rejects it
__device__ __forceinline__ double cable_score(const int* __restrict__ comp, size_t plane, size_t idx, const uint32_t* __restrict__ sums,
                                              int min_area, double min_aspect, double max_offset, int roi_width) {
  const int area = comp[idx];
  if (area < min_area) return -1.0;
  const int w = comp[2 * plane + idx] - comp[plane + idx] + 1, h = comp[4 * plane + idx] - comp[3 * plane + idx] + 1;
  const double aspect = __ddiv_rn((double)max(w, h), __dadd_rn((double)min(w, h), 1e-6));
  if (aspect < min_aspect) return -1.0;
  const double cx = __ddiv_rn((double)sums[idx], (double)area);                     // cv2 centroid x = m10 / m00
  const double off = __ddiv_rn(fabs(__dsub_rn(cx, __ddiv_rn((double)roi_width, 2.0))), (double)roi_width);
  if (off > max_offset) return -1.0;
  return __dmul_rn(__dmul_rn((double)area, aspect), __dsub_rn(1.0, off));            // area * aspect * (1.0 - center_offset)
}

// pass 0: best[b] = max score bits;  pass 1: key[b] = min first-appearance key among the roots with that score
template <int PASS>
__global__ void __launch_bounds__(256) k_cable_select(const uint32_t* __restrict__ bits, const int* __restrict__ lab_all, int H, int W, int Wd,
                                                      const int* __restrict__ comp, size_t plane, const uint32_t* __restrict__ sums,
                                                      int min_area, double min_aspect, double max_offset, int roi_width,
                                                      unsigned long long* __restrict__ best, int* __restrict__ key) {
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (w >= Wd || y >= H) return;
  uint32_t c = bits[((size_t)blockIdx.z * H + y) * Wd + w];
  const size_t base = (size_t)blockIdx.z * H * W;
  while (c) {
    int a, len;
    c &= ~first_run(c, a, len);
    const int p = y * W + 32 * w + a;
    if (lab_all[base + p] != p) continue;
    const double s = cable_score(comp, plane, base + p, sums, min_area, min_aspect, max_offset, roi_width);
    if (s < 0.0) continue;
    const unsigned long long sb = (unsigned long long)__double_as_longlong(s) + 1ull;       // non-negative doubles order as integers; +1: 0 = none
    if (PASS == 0) atomicMax(best + blockIdx.z, sb);
    else if (sb == best[blockIdx.z]) atomicMin(key + blockIdx.z, comp[5 * plane + base + p]);
  }
}

// out = 255 on the runs of the component whose key is key[b] (all zero when key[b] == INT_MAX)
__global__ void __launch_bounds__(256) k_keep_component(const uint32_t* __restrict__ bits, const int* __restrict__ lab_all, int H, int W, int Wd,
                                                        const int* __restrict__ comp, size_t plane, const int* __restrict__ key,
                                                        uint8_t* __restrict__ out) {
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (w >= Wd || y >= H) return;
  uint32_t c = bits[((size_t)blockIdx.z * H + y) * Wd + w], keep = 0;
  const size_t base = (size_t)blockIdx.z * H * W;
  const int k = key[blockIdx.z];
  while (c) {
    int a, len;
    const uint32_t rm = first_run(c, a, len);
    c &= ~rm;
    if (comp[5 * plane + base + lab_all[base + y * W + 32 * w + a]] == k) keep |= rm;
  }
  store_bits_as_bytes(out + base + (size_t)y * W + 32 * w, keep, W - 32 * w, 255u);
}

// key[b] from best[b] = (area << 32 | INT_MAX - key) of k_diam_select
__global__ void k_key_from_best(const unsigned long long* __restrict__ best, int* __restrict__ key, int B) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) key[b] = best[b] ? INT_MAX - (int)(uint32_t)best[b] : INT_MAX;
}

// ring = dilated & ~eroded, constrained = tape & ring as {0,1}
__global__ void k_ring_and(const uint8_t* __restrict__ tape, const uint8_t* __restrict__ dil, const uint8_t* __restrict__ ero,
                           uint8_t* __restrict__ out, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (tape[i] != 0 && dil[i] != 0 && ero[i] == 0) ? 1 : 0;
}

}  // namespace s2

// shared tail: label `m`, statistics with keys, then the caller-specific selection fills key[b]; finally write the kept component
static int label_with_keys(ub_ctx* ctx, const uint8_t* m, int B, int H, int W, cudaStream_t st) {
  const size_t plane = (size_t)ctx->max_batch * ctx->frame_h * ctx->frame_w;
  const int Wd = (W + 31) / 32;
  uint32_t* bits = reinterpret_cast<uint32_t*>(ctx->d_lab[1]);
  RUN(cclb_label(ctx, m, bits, nullptr, ctx->d_lab[0], B, H, W, ctx->d_comp, plane, st));
  k_cclb_stats<true><<<dim3(cdiv(Wd, 32), cdiv(H, 8), B), dim3(32, 8), 0, st>>>(bits, ctx->d_lab[0], H, W, Wd, ctx->d_comp, plane);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

int filter_cable_by_shape(ub_ctx* ctx, const uint8_t* mask, int B, int H, int W, int min_area, double min_aspect, double max_offset,
                          int roi_width, uint8_t* out, cudaStream_t st) {
  TimeScope ts(ctx, UB_T_CCL, st);
  const size_t plane = (size_t)ctx->max_batch * ctx->frame_h * ctx->frame_w, n = (size_t)B * H * W;
  const int Wd = (W + 31) / 32;
  uint32_t* bits = reinterpret_cast<uint32_t*>(ctx->d_lab[1]);
  int* lab = ctx->d_lab[0];
  if ((double)H * W * W / 2.0 >= 4294967296.0) return ub_fail(ctx, UB_ERR_ARG, "ub_filter_cable_by_shape: %dx%d too large for 32-bit x sums", H, W);
  if (!ctx->d_sums) {                                  // centroid scratch: one u32 per pixel position, allocated on first use
    void* p = nullptr;
    UB_CUDA(ctx, cudaMalloc(&p, sizeof(uint32_t) * plane));
    ctx->d_sums = static_cast<uint32_t*>(p);
  }
  uint32_t* sums = ctx->d_sums;
  unsigned long long* best = reinterpret_cast<unsigned long long*>(ctx->d_tmp[2]);
  int* key = reinterpret_cast<int*>(ctx->d_tmp[3]);
  RUN(label_with_keys(ctx, mask, B, H, W, st));
  UB_CUDA(ctx, cudaMemsetAsync(sums, 0, sizeof(uint32_t) * n, st));
  UB_CUDA(ctx, cudaMemsetAsync(best, 0, sizeof(unsigned long long) * B, st));
  UB_CUDA(ctx, cudaMemsetAsync(key, 0x7f, sizeof(int) * B, st));                     // 0x7f7f7f7f: larger than any key
  dim3 g(cdiv(Wd, 32), cdiv(H, 8), B);
  k_cclb_sumx<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, sums);
  k_cable_select<0><<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane, sums, min_area, min_aspect, max_offset, roi_width, best, key);
  k_cable_select<1><<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane, sums, min_area, min_aspect, max_offset, roi_width, best, key);
  k_keep_component<<<g, dim3(32, 8), 0, st>>>(bits, lab, H, W, Wd, ctx->d_comp, plane, key, out);
  UB_CHECK_LAUNCH_N(ctx, 4);
  return UB_OK;
}

int constrain_tape_to_ring(ub_ctx* ctx, const uint8_t* tape, const uint8_t* cable, int B, int H, int W, int dilate_k, int erode_k,
                           uint8_t* out, cudaStream_t st) {
  const size_t plane = (size_t)ctx->max_batch * ctx->frame_h * ctx->frame_w, n = (size_t)B * H * W;
  const int Wd = (W + 31) / 32;
  uint8_t *dil = ctx->d_tmp[0], *ero = ctx->d_tmp[1], *con = ctx->d_tmp[2];
  RUN(morph_op(ctx, 0, dilate_k, cable, dil, nullptr, B, H, W, 1, st));
  RUN(morph_op(ctx, 1, erode_k, cable, ero, nullptr, B, H, W, 1, st));
  TimeScope ts(ctx, UB_T_CCL, st);
  k_ring_and<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(tape, dil, ero, con, n);
  RUN(label_with_keys(ctx, con, B, H, W, st));
  uint32_t* bits = reinterpret_cast<uint32_t*>(ctx->d_lab[1]);
  unsigned long long* best = reinterpret_cast<unsigned long long*>(ctx->d_tmp[0]);     // dil is dead
  int* key = reinterpret_cast<int*>(ctx->d_tmp[1]);                                    // ero is dead
  UB_CUDA(ctx, cudaMemsetAsync(best, 0, sizeof(unsigned long long) * B, st));
  dim3 g(cdiv(Wd, 32), cdiv(H, 8), B);
  k_diam_select<<<g, dim3(32, 8), 0, st>>>(bits, ctx->d_lab[0], H, W, Wd, ctx->d_comp, plane, 1, best);
  k_key_from_best<<<cdiv(B, 128), 128, 0, st>>>(best, key, B);
  k_keep_component<<<g, dim3(32, 8), 0, st>>>(bits, ctx->d_lab[0], H, W, Wd, ctx->d_comp, plane, key, out);
  UB_CHECK_LAUNCH_N(ctx, 4);
  return UB_OK;
}
