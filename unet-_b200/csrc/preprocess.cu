// preprocess.cu -- uint8 frame -> net input / gray, frame normalisation, and
// logits -> class masks.  All integer arithmetic follows OpenCV 4.13 bit-exactly
// (restated and pinned in oracle/cvnp.py).
#include <map>
#include <math.h>
#include "common.cuh"

// ---------------------------------------------------------------------------------
// cv2.resize(INTER_LINEAR) coefficient tables (oracle/cvnp.py: linear_coeffs).
// Table layout per axis: int4 {i0, i1, c0, c1} per destination index.
// ---------------------------------------------------------------------------------
struct ResizeCache {
  std::map<uint64_t, int4*> dev;
  std::vector<std::vector<int4>*> host;
};

static void build_axis(int src, int dst, bool is_x, std::vector<int4>& t) {
  t.resize(dst);
  const double scale = 1.0 / ((double)dst / (double)src);     // cv::resize: scale = 1./inv_scale
  for (int d = 0; d < dst; ++d) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = (int)floorf(f);
    f -= (float)s;
    int i0, i1;
    if (is_x) {
      if (s < 0) { s = 0; f = 0.f; }
      if (s >= src - 1) { s = src - 1; f = 0.f; }
      i0 = s; i1 = s + 1 < src ? s + 1 : src - 1;
    } else {
      i0 = s < 0 ? 0 : (s > src - 1 ? src - 1 : s);
      int s1 = s + 1;
      i1 = s1 < 0 ? 0 : (s1 > src - 1 ? src - 1 : s1);
    }
    int c0 = (int)lrintf((1.f - f) * 2048.f);
    int c1 = (int)lrintf(f * 2048.f);
    t[d] = make_int4(i0, i1, c0, c1);
  }
}

static int get_table(ub_ctx* ctx, int src, int dst, bool is_x, const int4** out, cudaStream_t st) {
  if (!ctx->cache) ctx->cache = new ResizeCache();
  ResizeCache* rc = static_cast<ResizeCache*>(ctx->cache);
  uint64_t key = ((uint64_t)src << 33) | ((uint64_t)dst << 1) | (is_x ? 1 : 0);
  auto it = rc->dev.find(key);
  if (it == rc->dev.end()) {
    auto* h = new std::vector<int4>();
    build_axis(src, dst, is_x, *h);
    rc->host.push_back(h);
    int4* d = nullptr;
    UB_CUDA(ctx, cudaMalloc(&d, sizeof(int4) * dst));
    // blocking (once per shape): a later call on another stream must not read the table before it has landed
    UB_CUDA(ctx, cudaMemcpy(d, h->data(), sizeof(int4) * dst, cudaMemcpyHostToDevice));
    it = rc->dev.emplace(key, d).first;
  }
  *out = it->second;
  return UB_OK;
}

void resize_cache_free(ub_ctx* ctx) {
  if (!ctx->cache) return;
  ResizeCache* rc = static_cast<ResizeCache*>(ctx->cache);
  for (auto& kv : rc->dev) cudaFree(kv.second);
  for (auto* h : rc->host) delete h;
  delete rc;
  ctx->cache = nullptr;
}

namespace {

// One bilinear sample of channel c, cv2's 11-bit fixed point (two-pass, >>4, >>16, +2 >>2).
template <class Src>
__device__ __forceinline__ int bilerp(const Src& S, int4 ty, int4 tx, int c) {
  int h0 = S(ty.x, tx.x, c) * tx.z + S(ty.x, tx.y, c) * tx.w;
  int h1 = S(ty.y, tx.x, c) * tx.z + S(ty.y, tx.y, c) * tx.w;
  return (((ty.z * (h0 >> 4)) >> 16) + ((ty.w * (h1 >> 4)) >> 16) + 2) >> 2;
}

struct SrcPlain {
  const uint8_t* p; int w;
  __device__ __forceinline__ int operator()(int y, int x, int c) const { return p[((size_t)y * w + x) * 3 + c]; }
};
// view of cv2.rotate(src, ROTATE_90_COUNTERCLOCKWISE): R[y][x] = src[x][sw-1-y]
struct SrcRot {
  const uint8_t* p; int w;
  __device__ __forceinline__ int operator()(int y, int x, int c) const { return p[((size_t)x * w + (w - 1 - y)) * 3 + c]; }
};

// preprocess_image (infer_two_stage_burr.py:122-127): BGR->RGB, resize, /255.
// A block takes PRE_ROWS padded rows of one image (blockIdx.x = row group; the last group also zeroes the tile padding
// behind the plane).  u/255 -> fp16 is a 256-entry shared-memory table built once per block (the IEEE division is exact
// either way).  STAGED: the source rows the group's vertical taps touch are one contiguous byte range of the frame; it
// is copied to shared memory first with aligned 16-byte loads, all in flight together, so the tap reads are
// shared-memory bytes instead of dependent global byte loads.  !STAGED reads the taps from global memory (frames too
// large for the shared-memory window, or UB_PRE_STAGED=0).  The first version (one position per thread, flat index)
// was issue- and latency-bound: 204 instructions per position, DRAM at 25 % (profiles/r01_glue.md).
constexpr int PRE_ROWS = 8, PRE_T = 288;              // 2 * 288 covers the 514-position pitch of a 512-wide plane
static_assert(PRE_T >= 256 + PRE_ROWS, "threads 256.. load the row coefficients");
constexpr int PRE_SMEM_MAX = 160 * 1024;
// XI: the network input is written in the "x-im2col" form conv_block.cu consumes: channel k = kx*3 + c of position (y, x)
// holds colour c of pixel (y, x + kx - 1) (zero outside the image), k = 0..8 of 16, so the first conv needs no x-taps.
// The group's rows are collected in shared memory (8 B per position) and written out with their neighbours.
template <bool STAGED, bool NCHW, bool XI = false>
__global__ void __launch_bounds__(PRE_T) k_preprocess(const uint8_t* __restrict__ frames, int B, int H, int W, int nh, int nw,
                                                    const int4* __restrict__ ty, const int4* __restrict__ tx,
                                                    act_t* __restrict__ x_pp, float* __restrict__ x_nchw, PPLayout L,
                                                    int stage_bytes) {
  extern __shared__ __align__(16) uint8_t s_src[];
  [[maybe_unused]] uint2* s_xi = reinterpret_cast<uint2*>(s_src + stage_bytes);     // [PRE_ROWS][pitch] (r|g<<16, b)
  __shared__ act_t s_lut[256];
  __shared__ int4 s_ty[PRE_ROWS];                                // the group's row coefficients (read once per position below)
  if (threadIdx.x < 256) s_lut[threadIdx.x] = f2a(__fdiv_rn((float)threadIdx.x, 255.0f));
  if (threadIdx.x >= 256 && threadIdx.x < 256 + PRE_ROWS) {
    const int yy = blockIdx.x * PRE_ROWS + threadIdx.x - 256;
    if (yy >= 1 && yy <= nh) s_ty[threadIdx.x - 256] = ty[yy - 1];
  }
  const int b = blockIdx.y;
  const int W3 = W * 3;
  const uint8_t* img = frames + (size_t)b * H * W3;
  bool staged = false;
  int sbase = 0;                                                 // tap (row r, byte o) is s_src[sbase + r * W3 + o]
  if (STAGED) {
    const int ya = max((int)blockIdx.x * PRE_ROWS, 1), yb = min((int)blockIdx.x * PRE_ROWS + PRE_ROWS - 1, nh);   // interior rows
    if (ya <= yb) {
      const int r_lo = ty[ya - 1].x, r_hi = ty[yb - 1].y;        // the tables are monotone, .x <= .y
      const uint8_t* P = img + (size_t)r_lo * W3;
      const int n = (r_hi - r_lo + 1) * W3;
      const int a = (int)(reinterpret_cast<uintptr_t>(P) & 15);
      const uint8_t* lo = frames;
      const uint8_t* hi = frames + (size_t)B * H * W3;
      if (a + n <= stage_bytes) {                                // always, by the host's sizing; else the taps stay in global memory
        for (int i = threadIdx.x; i * 16 < a + n; i += PRE_T) {
          const uint8_t* g = P - a + i * 16;                     // 16-byte aligned address
          if (g >= lo && g + 16 <= hi) {
            *reinterpret_cast<uint4*>(s_src + i * 16) = *reinterpret_cast<const uint4*>(g);
          } else {                                               // first / last vector of the buffer: stay inside it
            for (int k = 0; k < 16; ++k) s_src[i * 16 + k] = (g + k >= lo && g + k < hi) ? g[k] : (uint8_t)0;
          }
        }
        staged = true;
        sbase = a - r_lo * W3;
      }
    }
  }
  __syncthreads();
  // A thread owns a column of the group's rows: column coefficients loaded once, and the horizontal pass of a source
  // row stays in registers for the next output row whenever cv2's two vertical taps share it (always, unless the
  // resize shrinks by more than 2).  ty[] is the same for the whole block, so every branch on rows is uniform.
  act_t* dstp = x_pp ? x_pp + (((size_t)b * 2) * L.Ppl + L.G) * 8 : nullptr;
  const int y0 = blockIdx.x * PRE_ROWS, y1 = min(y0 + PRE_ROWS, nh + 2);
  auto hpass = [&](int r, int o0, int o1, int wz, int ww, int (&h)[3]) {
    if (STAGED && staged) {
      const int base = sbase + r * W3;                           // 32-bit shared-memory index
#pragma unroll
      for (int c = 0; c < 3; ++c) h[c] = (s_src[base + o0 + c] * wz + s_src[base + o1 + c] * ww) >> 4;
    } else {
      const uint8_t* row = img + (size_t)r * W3;
#pragma unroll
      for (int c = 0; c < 3; ++c) h[c] = (row[o0 + c] * wz + row[o1 + c] * ww) >> 4;
    }
  };
  for (int xx = threadIdx.x; xx < L.pitch; xx += PRE_T) {
    const bool colin = xx >= 1 && xx <= nw;
    int4 cx = make_int4(0, 0, 0, 0);
    if (colin) cx = tx[xx - 1];
    const int o0 = cx.x * 3, o1 = cx.y * 3;
    int ra = -1, rb = -1, Ha[3] = {0, 0, 0}, Hb[3] = {0, 0, 0};    // source rows held and their horizontal passes (>> 4)
    for (int yy = y0; yy < y1; ++yy) {
      uint4 v = make_uint4(0, 0, 0, 0);
      if (colin && yy >= 1 && yy <= nh) {
        const int4 cy = s_ty[yy - y0];
        int n0[3], n1[3];
        if (cy.x == ra) { n0[0] = Ha[0]; n0[1] = Ha[1]; n0[2] = Ha[2]; }
        else if (cy.x == rb) { n0[0] = Hb[0]; n0[1] = Hb[1]; n0[2] = Hb[2]; }
        else hpass(cy.x, o0, o1, cx.z, cx.w, n0);
        if (cy.y == rb) { n1[0] = Hb[0]; n1[1] = Hb[1]; n1[2] = Hb[2]; }
        else if (cy.y == cy.x) { n1[0] = n0[0]; n1[1] = n0[1]; n1[2] = n0[2]; }
        else hpass(cy.y, o0, o1, cx.z, cx.w, n1);
        ra = cy.x; rb = cy.y;
        uint32_t h[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {                            // RGB order: channel c reads B,G,R[2-c]
          Ha[c] = n0[c]; Hb[c] = n1[c];
          const int u = (((cy.z * n0[2 - c]) >> 16) + ((cy.w * n1[2 - c]) >> 16) + 2) >> 2;   // cv2's 11-bit fixed point
          h[c] = (uint32_t)*reinterpret_cast<const uint16_t*>(&s_lut[u]);
          if (NCHW) x_nchw[(((size_t)b * 3 + c) * nh + (yy - 1)) * nw + (xx - 1)] = __fdiv_rn((float)u, 255.0f);
        }
        v.x = h[0] | (h[1] << 16);
        v.y = h[2];                                              // channels 3..7 are zero (+0.0 in either 16-bit format)
      }
      if (XI) s_xi[(yy - y0) * L.pitch + xx] = make_uint2(v.x, v.y);
      else if (dstp) *reinterpret_cast<uint4*>(dstp + (size_t)(yy * L.pitch + xx) * 8) = v;
    }
  }
  if constexpr (XI) {
    __syncthreads();
    const size_t chunk1 = (size_t)L.Ppl * 8;
    for (int i = threadIdx.x; i < (y1 - y0) * L.pitch; i += PRE_T) {
      const int r = i / L.pitch, xx = i - r * L.pitch;
      uint4 c0 = make_uint4(0, 0, 0, 0), c1 = make_uint4(0, 0, 0, 0);
      if (xx >= 1 && xx <= nw) {                                 // border positions stay zero
        const uint2 m = s_xi[i], l = s_xi[i - 1], rt = s_xi[i + 1];     // borders of the row hold zeros
        c0.x = l.x; c0.y = (l.y & 0xffffu) | (m.x << 16); c0.z = (m.x >> 16) | (m.y << 16); c0.w = rt.x;
        c1.x = rt.y & 0xffffu;
      }
      act_t* o = dstp + (size_t)((y0 + r) * L.pitch + xx) * 8;
      *reinterpret_cast<uint4*>(o) = c0;
      *reinterpret_cast<uint4*>(o + chunk1) = c1;
    }
  }
  if (dstp && blockIdx.x + 1 == gridDim.x)                        // the tile padding behind the plane
    for (int q = L.npos + threadIdx.x; q < L.ntiles * 128; q += PRE_T) {
      *reinterpret_cast<uint4*>(dstp + (size_t)q * 8) = make_uint4(0, 0, 0, 0);
      if (XI) *reinterpret_cast<uint4*>(dstp + (size_t)L.Ppl * 8 + (size_t)q * 8) = make_uint4(0, 0, 0, 0);
    }
}

// cv2.cvtColor(BGR2GRAY) (infer_two_stage_burr.py:317): (3735 B + 19235 G + 9798 R + 2^14) >> 15
// 16 pixels per thread: three 16-byte loads (48 B of BGR), one 16-byte store, when both planes are 16-byte aligned
__device__ __forceinline__ uint32_t gray4(uint32_t w0, uint32_t w1, uint32_t w2) {      // 12 bytes b g r b g r ... -> 4 gray bytes
  uint8_t px[12];
  *reinterpret_cast<uint32_t*>(px) = w0; *reinterpret_cast<uint32_t*>(px + 4) = w1; *reinterpret_cast<uint32_t*>(px + 8) = w2;
  uint32_t o = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t g = (3735u * px[3 * k] + 19235u * px[3 * k + 1] + 9798u * px[3 * k + 2] + (1u << 14)) >> 15;
    o |= g << (8 * k);
  }
  return o;
}
__global__ void k_gray(const uint8_t* __restrict__ frames, uint8_t* __restrict__ gray, size_t npix) {
  const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 16;
  if (i >= npix) return;
  if (i + 16 <= npix && ((reinterpret_cast<uintptr_t>(frames) | reinterpret_cast<uintptr_t>(gray)) & 15) == 0) {
    const uint4* p = reinterpret_cast<const uint4*>(frames + i * 3);
    const uint4 a = p[0], b = p[1], c = p[2];
    *reinterpret_cast<uint4*>(gray + i) = make_uint4(gray4(a.x, a.y, a.z), gray4(a.w, b.x, b.y), gray4(b.z, b.w, c.x), gray4(c.y, c.z, c.w));
  } else {
    for (size_t k = i; k < npix && k < i + 16; ++k)
      gray[k] = (uint8_t)((3735u * frames[k * 3] + 19235u * frames[k * 3 + 1] + 9798u * frames[k * 3 + 2] + (1u << 14)) >> 15);
  }
}

// One cv2.addWeighted step on a u8 sample: fma(a, alpha, b*beta) in binary32, round-half-even (oracle/cvnp.py: add_weighted_u8)
__device__ __forceinline__ float aw_step(float a, float alpha, float b, float beta) {
  return rintf(__fmaf_rn(a, alpha, __fmul_rn(b, beta)));
}
// The blend chain of visualize_two_stage (infer_two_stage_burr.py:134-154).  The four steps only depend on the sample,
// its channel and four flags (inside the ROI, cable, tape, burr), so a CTA first tabulates the whole chain --
// 16 flag sets x 3 channels x 256 values = 12 KB of shared memory, 48 chain evaluations per thread -- and then
// maps OV_ROWS image rows through it, four pixels (12 B of BGR + 3 mask words) per thread and step.
constexpr int OV_ROWS = 64;
__device__ __forceinline__ uint8_t ov_chain(float v, int ch, int flags) {
  v = aw_step(v, 0.7f, (flags & 1) ? v : 0.f, 0.3f);
  v = aw_step(v, 0.6f, ((flags & 2) && ch == 1) ? 255.f : 0.f, 0.4f);      // cable (0,255,0)
  v = aw_step(v, 0.6f, ((flags & 4) && ch == 0) ? 255.f : 0.f, 0.4f);      // tape  (255,0,0)
  v = aw_step(v, 0.5f, ((flags & 8) && ch != 1) ? 255.f : 0.f, 0.5f);      // burr  (255,0,255)
  return (uint8_t)fminf(fmaxf(v, 0.f), 255.f);
}
__global__ void __launch_bounds__(256) k_overlay(const uint8_t* __restrict__ frames, const uint8_t* __restrict__ mc,
                                                 const uint8_t* __restrict__ mt, const uint8_t* __restrict__ mb,
                                                 uint8_t* __restrict__ out, int H, int W, ub_roi roi) {
  __shared__ uint8_t lut[16 * 3 * 256];
  for (int t = 0; t < 48; ++t) lut[t * 256 + threadIdx.x] = ov_chain((float)threadIdx.x, t % 3, t / 3);
  __syncthreads();
  const int y0 = blockIdx.x * OV_ROWS, y1 = min(y0 + OV_ROWS, H);
  const size_t fbase = (size_t)blockIdx.y * H * W;
  const int wq = (W + 3) / 4;
  const bool vec = (W & 3) == 0;                       // rows (and frames) start on word boundaries
  for (int i = threadIdx.x; i < (y1 - y0) * wq; i += 256) {
    const int r = i / wq, x = 4 * (i - r * wq), y = y0 + r;
    const size_t pix = fbase + (size_t)y * W + x;
    const bool yin = y >= roi.y1 && y < roi.y2;
    if (vec) {
      const uint32_t* p = reinterpret_cast<const uint32_t*>(frames + pix * 3);
      const uint32_t w[3] = {p[0], p[1], p[2]};
      const uint32_t c4 = *reinterpret_cast<const uint32_t*>(mc + pix), t4 = *reinterpret_cast<const uint32_t*>(mt + pix),
                     b4 = *reinterpret_cast<const uint32_t*>(mb + pix);
      uint32_t o[3] = {0u, 0u, 0u};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int flags = (yin && x + k >= roi.x1 && x + k < roi.x2 ? 1 : 0) | (((c4 >> (8 * k)) & 0xffu) ? 2 : 0) |
                          (((t4 >> (8 * k)) & 0xffu) ? 4 : 0) | (((b4 >> (8 * k)) & 0xffu) ? 8 : 0);
        const uint8_t* l = lut + flags * 768;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
          const int bi = 3 * k + ch;
          o[bi >> 2] |= (uint32_t)l[ch * 256 + ((w[bi >> 2] >> (8 * (bi & 3))) & 0xffu)] << (8 * (bi & 3));
        }
      }
      uint32_t* q = reinterpret_cast<uint32_t*>(out + pix * 3);
      q[0] = o[0]; q[1] = o[1]; q[2] = o[2];
    } else {
      for (int k = 0; k < 4 && x + k < W; ++k) {
        const int flags = (yin && x + k >= roi.x1 && x + k < roi.x2 ? 1 : 0) | (mc[pix + k] ? 2 : 0) | (mt[pix + k] ? 4 : 0) |
                          (mb[pix + k] ? 8 : 0);
        for (int ch = 0; ch < 3; ++ch) out[(pix + k) * 3 + ch] = lut[flags * 768 + ch * 256 + frames[(pix + k) * 3 + ch]];
      }
    }
  }
}

// rotate + resize (infer_two_stage_burr.py:275-280)
template <class Src>
__global__ void k_resize3(Src S0, size_t src_stride, uint8_t* __restrict__ dst, int dh, int dw,
                          const int4* __restrict__ ty, const int4* __restrict__ tx) {
  int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
  if (x >= dw) return;
  Src S = S0; S.p += (size_t)b * src_stride;
  int4 cy = ty[y], cx = tx[x];
  uint8_t* o = dst + (((size_t)b * dh + y) * dw + x) * 3;
#pragma unroll
  for (int c = 0; c < 3; ++c) o[c] = (uint8_t)bilerp(S, cy, cx, c);
}

// argmax over classes (softmax is monotone; ties -> lowest index like np.argmax)
__global__ void k_argmax(const float* __restrict__ logits, int C, size_t hw, uint8_t* __restrict__ pred) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  int b = blockIdx.y;
  if (i >= hw) return;
  const float* p = logits + (size_t)b * C * hw + i;
  float bv = p[0]; int best = 0;
  for (int c = 1; c < C; ++c) { float v = p[c * hw]; if (v > bv) { bv = v; best = c; } }
  pred[(size_t)b * hw + i] = (uint8_t)best;
}

// (pred==1),(pred==2) -> NEAREST upscale -> ROI (infer_two_stage_burr.py:303-314) + pixel counts.
// A thread owns 4 consecutive columns (their source columns are computed once, in double as cv2
// does) and walks MK_ROWS rows; counts are reduced per block -> two atomics per block instead of
// two per warp on the same address (the per-pixel f64 version took 187 us per 64 frames).
constexpr int MK_ROWS = 8;
__global__ void __launch_bounds__(256) k_masks(const uint8_t* __restrict__ pred, int nh, int nw, int H, int W, double sy, double sx,
                                               ub_roi roi, uint8_t* __restrict__ cable, uint8_t* __restrict__ tape,
                                               int32_t* __restrict__ stats) {
  __shared__ int s_cnt[2];
  if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, yb = blockIdx.y * MK_ROWS, b = blockIdx.z;
  int nc = 0, nt = 0;
  if (x4 < W) {
    int sxi[4]; bool in_x[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int x = x4 + k;
      sxi[k] = min((int)floor(x * sx), nw - 1);
      in_x[k] = x < W && x >= roi.x1 && x < roi.x2;
    }
    const bool vec = x4 + 4 <= W && (W & 3) == 0;
    for (int y = yb; y < min(yb + MK_ROWS, H); ++y) {
      uint32_t vc = 0, vt = 0;
      if (y >= roi.y1 && y < roi.y2) {
        const uint8_t* row = pred + ((size_t)b * nh + min((int)floor(y * sy), nh - 1)) * nw;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (in_x[k]) { const int cls = row[sxi[k]]; vc |= (uint32_t)(cls == 1) << (8 * k); vt |= (uint32_t)(cls == 2) << (8 * k); }
      }
      nc += __popc(vc); nt += __popc(vt);
      const size_t o = ((size_t)b * H + y) * W + x4;
      if (vec) { *reinterpret_cast<uint32_t*>(cable + o) = vc; *reinterpret_cast<uint32_t*>(tape + o) = vt; }
      else for (int k = 0; k < 4 && x4 + k < W; ++k) { cable[o + k] = (uint8_t)(vc >> (8 * k)); tape[o + k] = (uint8_t)(vt >> (8 * k)); }
    }
  }
  if (stats) {
    for (int o = 16; o; o >>= 1) { nc += __shfl_xor_sync(0xffffffffu, nc, o); nt += __shfl_xor_sync(0xffffffffu, nt, o); }
    if ((threadIdx.x & 31) == 0) { if (nc) atomicAdd(&s_cnt[0], nc); if (nt) atomicAdd(&s_cnt[1], nt); }
    __syncthreads();
    if (threadIdx.x == 0) {
      if (s_cnt[0]) atomicAdd(stats + b * 4 + 1, s_cnt[0]);
      if (s_cnt[1]) atomicAdd(stats + b * 4 + 2, s_cnt[1]);
    }
  }
}

__global__ void k_stats_init(int32_t* stats, int B) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B * 4) stats[i] = (i & 3) == 0 ? (i >> 2) : 0;
}

}  // namespace

int preprocess_frames(ub_ctx* ctx, const uint8_t* frames, int B, int H, int W, int nh, int nw,
                      act_t* x_pp, float* x_nchw, uint8_t* gray, cudaStream_t st, bool xi) {
  if (x_pp || x_nchw) {
    const int4 *ty, *tx;
    int r = get_table(ctx, H, nh, false, &ty, st); if (r) return r;
    r = get_table(ctx, W, nw, true, &tx, st); if (r) return r;
    PPLayout L = pp_layout(nh, nw);
    TimeScope ts(ctx, UB_T_PREPOST, st);
    dim3 grid(cdiv(nh + 2, PRE_ROWS), B);
    // source rows under PRE_ROWS output rows: the taps of the first and last row are (PRE_ROWS-1)*H/nh apart, + the second tap
    const size_t stage = ((size_t)((PRE_ROWS - 1) * (size_t)H / nh + 3) * W * 3 + 32 + 15) & ~(size_t)15;
    static const int staged_env = getenv("UB_PRE_STAGED") ? atoi(getenv("UB_PRE_STAGED")) : 1;
    const size_t xi_bytes = xi ? (size_t)PRE_ROWS * L.pitch * sizeof(uint2) : 0;
    auto launch = [&](auto kern, size_t smem) {
      if (smem + xi_bytes > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, PRE_SMEM_MAX + 40 * 1024);
      kern<<<grid, PRE_T, smem + xi_bytes, st>>>(frames, B, H, W, nh, nw, ty, tx, x_pp, x_nchw, L, (int)smem);
    };
    if (xi) {
      if (!x_pp || x_nchw || xi_bytes > 40 * 1024) return ub_fail(ctx, UB_ERR_ARG, "preprocess: x-im2col output needs a PP destination and a plane at most 638 wide");
      if (staged_env && stage <= (size_t)PRE_SMEM_MAX) launch(k_preprocess<true, false, true>, stage);
      else launch(k_preprocess<false, false, true>, 0);
    } else if (staged_env && stage <= (size_t)PRE_SMEM_MAX) {
      if (x_nchw) launch(k_preprocess<true, true>, stage); else launch(k_preprocess<true, false>, stage);
    } else {
      if (x_nchw) launch(k_preprocess<false, true>, 0); else launch(k_preprocess<false, false>, 0);
    }
    UB_CHECK_LAUNCH(ctx);
  }
  if (gray) {
    size_t npix = (size_t)B * H * W;
    TimeScope ts(ctx, UB_T_PREPOST, st);
    k_gray<<<(unsigned)((npix / 16 + 255) / 256 + 1), 256, 0, st>>>(frames, gray, npix);
    UB_CHECK_LAUNCH(ctx);
  }
  return UB_OK;
}

int overlay_impl(ub_ctx* ctx, const uint8_t* frames, const uint8_t* mc, const uint8_t* mt, const uint8_t* mb, int B, int H, int W,
                 ub_roi roi, uint8_t* out, cudaStream_t st) {
  TimeScope ts(ctx, UB_T_PREPOST, st);
  k_overlay<<<dim3(cdiv(H, OV_ROWS), B), 256, 0, st>>>(frames, mc, mt, mb, out, H, W, roi);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

int normalize_frames_impl(ub_ctx* ctx, const uint8_t* src, int B, int sh, int sw, int rot,
                          uint8_t* dst, int dh, int dw, cudaStream_t st) {
  int rh = rot ? sw : sh, rw = rot ? sh : sw;
  const int4 *ty, *tx;
  int r = get_table(ctx, rh, dh, false, &ty, st); if (r) return r;
  r = get_table(ctx, rw, dw, true, &tx, st); if (r) return r;
  TimeScope ts(ctx, UB_T_PREPOST, st);
  dim3 grid(cdiv(dw, 128), dh, B);
  size_t stride = (size_t)sh * sw * 3;
  if (rot) k_resize3<SrcRot><<<grid, 128, 0, st>>>(SrcRot{src, sw}, stride, dst, dh, dw, ty, tx);
  else k_resize3<SrcPlain><<<grid, 128, 0, st>>>(SrcPlain{src, sw}, stride, dst, dh, dw, ty, tx);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

int argmax_logits(ub_ctx* ctx, const float* logits, int B, int C, int nh, int nw, uint8_t* pred, cudaStream_t st) {
  size_t hw = (size_t)nh * nw;
  TimeScope ts(ctx, UB_T_PREPOST, st);
  k_argmax<<<dim3((unsigned)((hw + 255) / 256), B), 256, 0, st>>>(logits, C, hw, pred);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

int masks_from_pred(ub_ctx* ctx, const uint8_t* pred, int B, int nh, int nw, int H, int W, ub_roi roi,
                    uint8_t* cable, uint8_t* tape, int32_t* stats, cudaStream_t st) {
  TimeScope ts(ctx, UB_T_PREPOST, st);
  if (stats) k_stats_init<<<cdiv(B * 4, 256), 256, 0, st>>>(stats, B);
  double sy = 1.0 / ((double)H / (double)nh), sx = 1.0 / ((double)W / (double)nw);
  k_masks<<<dim3(cdiv(cdiv(W, 4), 256), cdiv(H, MK_ROWS), B), 256, 0, st>>>(pred, nh, nw, H, W, sy, sx, roi, cable, tape, stats);
  UB_CHECK_LAUNCH_N(ctx, stats ? 2 : 1);
  return UB_OK;
}
