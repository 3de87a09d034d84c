// net_glue.cu -- layout conversion, 2x2 max-pool, bilinear x2 upsample in the PP layout,
// and the layer schedule of NestedUNet.forward (src/models/unetpp.py:93-135).
#include <stdlib.h>
#include "common.cuh"

namespace {

__device__ __forceinline__ size_t pp_off(int b, int chunk, int nchunks, int Ppl, int p) {
  return (((size_t)b * nchunks + chunk) * Ppl + p) * 8;   // in bf16 elements
}

// fp32 NCHW -> bf16 PP (zero border, zero channel padding)
__global__ void k_nchw_to_pp(const float* __restrict__ x, int B, int C, int H, int W,
                             act_t* __restrict__ out, int nchunks, PPLayout L) {
  int q = blockIdx.x * blockDim.x + threadIdx.x;
  int chunk = blockIdx.y, b = blockIdx.z;
  if (q >= L.ntiles * 128) return;
  int yy = q / L.pitch, xx = q - yy * L.pitch;
  uint4 v = make_uint4(0, 0, 0, 0);
  if (q < L.npos && yy >= 1 && yy <= H && xx >= 1 && xx <= W) {
    act_t h[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      int c = chunk * 8 + j;
      float f = (c < C) ? x[(((size_t)b * C + c) * H + (yy - 1)) * W + (xx - 1)] : 0.f;
      h[j] = f2a(f);
    }
    v = *reinterpret_cast<uint4*>(h);
  }
  *reinterpret_cast<uint4*>(out + pp_off(b, chunk, nchunks, L.Ppl, L.G + q)) = v;
}

// fp32 NCHW (3 channels) -> x-im2col PP: channel k = kx*3 + c of position (y,x) = x[c][y][x + kx - 1], zero outside
__global__ void k_nchw_to_xi(const float* __restrict__ x, int B, int H, int W, act_t* __restrict__ out, PPLayout L) {
  int q = blockIdx.x * blockDim.x + threadIdx.x;
  int chunk = blockIdx.y, b = blockIdx.z;
  if (q >= L.ntiles * 128) return;
  int yy = q / L.pitch, xx = q - yy * L.pitch;
  uint4 v = make_uint4(0, 0, 0, 0);
  if (q < L.npos && yy >= 1 && yy <= H && xx >= 1 && xx <= W) {
    act_t h[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = chunk * 8 + j, kx = k / 3, c = k - kx * 3, xs = xx - 1 + kx - 1;
      const float f = (k < 9 && xs >= 0 && xs < W) ? x[(((size_t)b * 3 + c) * H + (yy - 1)) * W + xs] : 0.f;
      h[j] = f2a(f);
    }
    v = *reinterpret_cast<uint4*>(h);
  }
  *reinterpret_cast<uint4*>(out + pp_off(b, chunk, 2, L.Ppl, L.G + q)) = v;
}

__global__ void k_pp_to_nchw(const act_t* __restrict__ in, int B, int C, int H, int W,
                             float* __restrict__ out, int nchunks, PPLayout L) {
  int x = blockIdx.x * blockDim.x + threadIdx.x;
  int y = blockIdx.y, bc = blockIdx.z;
  int b = bc / nchunks, chunk = bc % nchunks;
  if (x >= W) return;
  int p = L.G + (y + 1) * L.pitch + x + 1;
  uint4 v = *reinterpret_cast<const uint4*>(in + pp_off(b, chunk, nchunks, L.Ppl, p));
  const act_t* h = reinterpret_cast<const act_t*>(&v);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    int c = chunk * 8 + j;
    if (c < C) out[(((size_t)b * C + c) * H + y) * W + x] = a2f(h[j]);
  }
}

__device__ __forceinline__ uint4 max8(uint4 a, uint4 b) {
  uint4 r;
  const act2_t* pa = reinterpret_cast<const act2_t*>(&a);
  const act2_t* pb = reinterpret_cast<const act2_t*>(&b);
  act2_t* pr = reinterpret_cast<act2_t*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) pr[i] = __hmax2(pa[i], pb[i]);
  return r;
}

// MaxPool2d(2,2) (src/models/unetpp.py:75): in [H,W] -> out [H/2,W/2], both PP.
__global__ void k_maxpool2(const act_t* __restrict__ in, act_t* __restrict__ out,
                           int nchunks, PPLayout Li, PPLayout Lo) {
  int q = blockIdx.x * blockDim.x + threadIdx.x;
  int chunk = blockIdx.y, b = blockIdx.z;
  if (q >= Lo.ntiles * 128) return;
  int yy = q / Lo.pitch, xx = q - yy * Lo.pitch;
  uint4 v = make_uint4(0, 0, 0, 0);
  if (q < Lo.npos && yy >= 1 && yy <= Lo.H && xx >= 1 && xx <= Lo.W) {
    int iy = 2 * (yy - 1), ix = 2 * (xx - 1);
    const act_t* base = in + pp_off(b, chunk, nchunks, Li.Ppl, Li.G + (iy + 1) * Li.pitch + ix + 1);
    uint4 a = *reinterpret_cast<const uint4*>(base);
    uint4 c = *reinterpret_cast<const uint4*>(base + 8);
    uint4 d = *reinterpret_cast<const uint4*>(base + (size_t)Li.pitch * 8);
    uint4 e = *reinterpret_cast<const uint4*>(base + (size_t)Li.pitch * 8 + 8);
    v = max8(max8(a, c), max8(d, e));
  }
  *reinterpret_cast<uint4*>(out + pp_off(b, chunk, nchunks, Lo.Ppl, Lo.G + q)) = v;
}

// Upsample(scale 2, bilinear, align_corners=True) (src/models/unetpp.py:76):
// src = dst * (in-1)/(out-1) in fp32, as ATen's area_pixel_compute_source_index does.
// A thread owns one output column of one (image, chunk) plane and walks DOWN UP_R output rows:
// the column terms (x0, lx) are computed once, every input row is loaded and x-blended once
// (the source row index advances by 0 or 1 per output row, block-uniformly), and an output costs
// one y-blend.  The blend is packed 16-bit arithmetic, a + l*(c-a).  With one thread per output
// and fp32 math the kernel was issue-bound at 220 instructions per 16-byte output, 2.3x its HBM
// time (profiles/r01_glue.md).  No index clamps are needed: x0+1 / y0+1 can only leave the image
// with weight 0, where they read the zero border.  Edge threads also write the zero border.
constexpr int UP_R = 8;
__global__ void __launch_bounds__(128) k_upsample2(const act_t* __restrict__ in, act_t* __restrict__ out,
                                                   PPLayout Li, PPLayout Lo, float sy, float sx) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;            // output image column
  const int oy0 = blockIdx.y * UP_R;
  const uint4 zero = make_uint4(0, 0, 0, 0);
  act_t* oplane = out + ((size_t)blockIdx.z * Lo.Ppl + Lo.G) * 8;
  if (blockIdx.x == 0 && blockIdx.y == gridDim.y - 1)             // tail the flat conv tiles over-read
    for (int q = Lo.npos + threadIdx.x; q < Lo.ntiles * 128; q += blockDim.x) *reinterpret_cast<uint4*>(oplane + (size_t)q * 8) = zero;
  if (x >= Lo.W) return;
  const bool left = x == 0, right = x == Lo.W - 1;
  auto put = [&](act_t* p, uint4 v) {                             // p = this thread's position in a padded row
    *reinterpret_cast<uint4*>(p) = v;
    if (left) *reinterpret_cast<uint4*>(p - 8) = zero;
    if (right) *reinterpret_cast<uint4*>(p + 8) = zero;
  };
  if (blockIdx.y == 0) put(oplane + (size_t)(x + 1) * 8, zero);
  if (blockIdx.y == gridDim.y - 1) put(oplane + ((size_t)(Lo.H + 1) * Lo.pitch + x + 1) * 8, zero);

  const float fx = sx * (float)x;
  const int x0 = (int)fx;
  const float lx = fx - (float)x0;
  const act2_t lx2 = f2a2(lx, lx);
  const act_t* pin = in + ((size_t)blockIdx.z * Li.Ppl + Li.G + Li.pitch + 1 + x0) * 8;     // image row 0, column x0
  act_t* po = oplane + ((size_t)(oy0 + 1) * Lo.pitch + x + 1) * 8;
  const int in_row = Li.pitch * 8, out_row = Lo.pitch * 8;
  act2_t top[4], bot[4];
  auto blend_row = [&](int y, act2_t* r) {
    const uint4 a = *reinterpret_cast<const uint4*>(pin + (size_t)y * in_row);
    const uint4 c = *reinterpret_cast<const uint4*>(pin + (size_t)y * in_row + 8);
    const act2_t* ha = reinterpret_cast<const act2_t*>(&a);
    const act2_t* hc = reinterpret_cast<const act2_t*>(&c);
#pragma unroll
    for (int j = 0; j < 4; ++j) r[j] = __hfma2(lx2, __hsub2(hc[j], ha[j]), ha[j]);
  };
  int cur = -2;
#pragma unroll
  for (int r = 0; r < UP_R; ++r) {
    const int oy = oy0 + r;
    if (oy >= Lo.H) break;
    const float fy = sy * (float)oy;
    const int y0 = (int)fy;
    const float ly = fy - (float)y0;
    const act2_t ly2 = f2a2(ly, ly);
    if (y0 != cur) {
      if (y0 == cur + 1) {
#pragma unroll
        for (int j = 0; j < 4; ++j) top[j] = bot[j];
      } else {
        blend_row(y0, top);
      }
      blend_row(y0 + 1, bot);
      cur = y0;
    }
    uint4 v;
    act2_t* o2 = reinterpret_cast<act2_t*>(&v);
#pragma unroll
    for (int j = 0; j < 4; ++j) o2[j] = __hfma2(ly2, __hsub2(bot[j], top[j]), top[j]);
    put(po + (size_t)r * out_row, v);
  }
}

}  // namespace

int nchw_to_pp(ub_ctx* ctx, const float* x, int B, int C, int H, int W, act_t* out,
               int Cpad, cudaStream_t st) {
  PPLayout L = pp_layout(H, W);
  TimeScope ts(ctx, UB_T_GLUE, st);
  dim3 grid(cdiv(L.ntiles * 128, 256), Cpad / 8, B);
  k_nchw_to_pp<<<grid, 256, 0, st>>>(x, B, C, H, W, out, Cpad / 8, L);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

int nchw_to_xi(ub_ctx* ctx, const float* x, int B, int H, int W, act_t* out, cudaStream_t st) {
  PPLayout L = pp_layout(H, W);
  TimeScope ts(ctx, UB_T_GLUE, st);
  dim3 grid(cdiv(L.ntiles * 128, 256), 2, B);
  k_nchw_to_xi<<<grid, 256, 0, st>>>(x, B, H, W, out, L);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

int pp_to_nchw(ub_ctx* ctx, const act_t* in, int B, int C, int H, int W, float* out,
               cudaStream_t st) {
  PPLayout L = pp_layout(H, W);
  int nchunks = (C + 7) / 8;
  TimeScope ts(ctx, UB_T_GLUE, st);
  dim3 grid(cdiv(W, 128), H, B * nchunks);
  k_pp_to_nchw<<<grid, 128, 0, st>>>(in, B, C, H, W, out, nchunks, L);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

int maxpool2_pp(ub_ctx* ctx, const act_t* in, int B, int C, int H, int W,
                act_t* out, cudaStream_t st) {
  PPLayout Li = pp_layout(H, W), Lo = pp_layout(H / 2, W / 2);
  TimeScope ts(ctx, UB_T_GLUE, st);
  dim3 grid(cdiv(Lo.ntiles * 128, 256), C / 8, B);
  k_maxpool2<<<grid, 256, 0, st>>>(in, out, C / 8, Li, Lo);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

int upsample2_pp(ub_ctx* ctx, const act_t* in, int B, int C, int H, int W,
                 act_t* out, cudaStream_t st) {
  PPLayout Li = pp_layout(H, W), Lo = pp_layout(2 * H, 2 * W);
  float sy = (2 * H > 1) ? (float)(H - 1) / (float)(2 * H - 1) : 0.f;
  float sx = (2 * W > 1) ? (float)(W - 1) / (float)(2 * W - 1) : 0.f;
  TimeScope ts(ctx, UB_T_GLUE, st);
  if ((long long)B * (C / 8) > 65535) return ub_fail(ctx, UB_ERR_ARG, "upsample: %d planes exceed the grid limit", B * (C / 8));
  const int tx = Lo.W >= 128 ? 128 : ((Lo.W + 31) & ~31);
  dim3 grid(cdiv(Lo.W, tx), cdiv(Lo.H, UP_R), B * (C / 8));
  k_upsample2<<<grid, tx, 0, st>>>(in, out, Li, Lo, sy, sx);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

// Buffer indices into ub_ctx::act (see api.cu for sizes).
enum { A_XIN = 0, A_T0, A_X00, A_P0, A_T1, A_X10, A_P1, A_T2, A_X20, A_P2, A_T3, A_X30, A_P3,
       A_T4, A_X40, A_UP4, A_X31, A_UP3, A_X22, A_UP2, A_X13, A_UP1, A_X04 };

int run_conv_layer(ub_ctx* ctx, const ConvLayer& L, const ConvArgs& a, cudaStream_t st) {
  if (ctx->conv_impl == 1) return conv3x3_simt(ctx, L, a, st);
  if (!L.fold) return conv3x3_tc(ctx, L, a, st);
  return conv3x3_rows(ctx, L, a, st);
}

static int run_conv(ub_ctx* ctx, int li, const act_t* s0, const act_t* s1,
                    act_t* out, int B, int H, int W, cudaStream_t st,
                    float* logits = nullptr, uint8_t* pred = nullptr, bool head = false, act_t* pool_out = nullptr, bool up1 = false) {
  ConvArgs a{};
  a.src0 = s0; a.src1 = s1; a.out = out; a.B = B; a.H = H; a.W = W; a.pool_out = pool_out; a.up1 = up1;
  a.sat = (ctx->sat_check || ctx->sat_check_once) ? ctx->d_sat : nullptr;
  if (head) {
    a.head_w = ctx->head_w; a.head_b = ctx->head_b; a.num_classes = ctx->num_classes;
    a.logits = logits; a.pred = pred;
  }
  return run_conv_layer(ctx, ctx->conv[li], a, st);
}

#define RUN(x) do { int r__ = (x); if (r__ != UB_OK) return r__; } while (0)

// The fused-pool epilogues write the interior of the pooled plane only: its zero border and over-read tail are set here,
// once per shape (the P buffers have no other writer).
static int pool_border(ub_ctx* ctx, int pslot, act_t* pooled, int C, int H, int W, cudaStream_t st) {
  int* hw = ctx->pool_border_hw[pslot];
  if (hw[0] != H || hw[1] != W) {
    UB_CUDA(ctx, cudaMemsetAsync(pooled, 0, pp_bytes(ctx->max_batch, C, ctx->net_h >> (pslot + 1), ctx->net_w >> (pslot + 1)), st));
    hw[0] = H; hw[1] = W;
  }
  return UB_OK;
}

// conv2 of an encoder block followed by MaxPool2d(2,2) (unetpp.py:105-108).  Where the layer's kernel has the fused
// variant the pool is written from the conv epilogue (interior only: the pooled plane's zero border and over-read tail
// are set here, once per shape -- the P buffers have no other writer); otherwise the pool kernel runs.
static int run_conv_pool(ub_ctx* ctx, int li, int pslot, const act_t* src, act_t* out, act_t* pooled, int C,
                         int B, int H, int W, cudaStream_t st) {
  if (!conv_rows_can_pool(ctx, ctx->conv[li])) {
    RUN(run_conv(ctx, li, src, nullptr, out, B, H, W, st));
    return maxpool2_pp(ctx, out, B, C, H, W, pooled, st);
  }
  RUN(pool_border(ctx, pslot, pooled, C, H, W, st));
  return run_conv(ctx, li, src, nullptr, out, B, H, W, st, nullptr, nullptr, false, pooled);
}

// conv1 of a decoder block on cat([skip, up(below)]) (unetpp.py:111-116).  Where the layer's kernel has the fused
// variant, `below` goes in at half resolution and is interpolated on the way into shared memory; otherwise the
// upsampled tensor is materialised in `up` first.
static int run_conv_up(ub_ctx* ctx, int li, const act_t* skip, const act_t* below, act_t* up, act_t* out, int Cbelow,
                       int B, int H, int W, cudaStream_t st) {
  if (conv_rows_can_upsample(ctx, ctx->conv[li]) && !(H & 3))
    return run_conv(ctx, li, skip, below, out, B, H, W, st, nullptr, nullptr, false, nullptr, true);
  RUN(upsample2_pp(ctx, below, B, Cbelow, H / 2, W / 2, up, st));
  return run_conv(ctx, li, skip, up, out, B, H, W, st);
}

// The network input comes from exactly one of x_nchw / frames; with frames the preprocess kernel writes the PP input.
int net_forward_pp(ub_ctx* ctx, int B, int H, int W, const float* x_nchw, const uint8_t* frames, int fh, int fw,
                   float* logits, uint8_t* pred, cudaStream_t st) {
  act_t** A = ctx->act;
  const int h1 = H / 2, w1 = W / 2, h2 = H / 4, w2 = W / 4, h3 = H / 8, w3 = W / 8, h4 = H / 16, w4 = W / 16;
  // (Running the high-resolution levels a few frames at a time so that their intermediates stay inside the 126 MB
  //  L2 was tried and is slower: the small launches cost more than the residency gains, profiles/r01_conv.md.)
  // conv0_0 as one kernel where the shapes allow: its first conv takes the input in x-im2col form
  const bool sat_on = ctx->sat_check || ctx->sat_check_once;
  const bool fuse0 = conv_block_can_fuse(ctx, ctx->conv[0], ctx->conv[1], H, W);
  if (frames) RUN(preprocess_frames(ctx, frames, B, fh, fw, H, W, A[A_XIN], nullptr, nullptr, st, fuse0));
  else if (fuse0) RUN(nchw_to_xi(ctx, x_nchw, B, H, W, A[A_XIN], st));
  else RUN(nchw_to_pp(ctx, x_nchw, B, 3, H, W, A[A_XIN], 16, st));
  // encoder (unetpp.py:104-108)
  if (fuse0) {
    RUN(pool_border(ctx, 0, A[A_P0], 32, H, W, st));
    RUN(conv_block0(ctx, ctx->conv[0], ctx->conv[1], A[A_XIN], A[A_X00], A[A_P0], B, H, W, sat_on ? ctx->d_sat : nullptr, st));
  } else {
    RUN(run_conv(ctx, 0, A[A_XIN], nullptr, A[A_T0], B, H, W, st));
    RUN(run_conv_pool(ctx, 1, 0, A[A_T0], A[A_X00], A[A_P0], 32, B, H, W, st));
  }
  RUN(run_conv(ctx, 2, A[A_P0], nullptr, A[A_T1], B, h1, w1, st));
  RUN(run_conv_pool(ctx, 3, 1, A[A_T1], A[A_X10], A[A_P1], 64, B, h1, w1, st));
  RUN(run_conv(ctx, 4, A[A_P1], nullptr, A[A_T2], B, h2, w2, st));
  RUN(run_conv(ctx, 5, A[A_T2], nullptr, A[A_X20], B, h2, w2, st));
  RUN(maxpool2_pp(ctx, A[A_X20], B, 128, h2, w2, A[A_P2], st));
  RUN(run_conv(ctx, 6, A[A_P2], nullptr, A[A_T3], B, h3, w3, st));
  RUN(run_conv(ctx, 7, A[A_T3], nullptr, A[A_X30], B, h3, w3, st));
  RUN(maxpool2_pp(ctx, A[A_X30], B, 256, h3, w3, A[A_P3], st));
  RUN(run_conv(ctx, 8, A[A_P3], nullptr, A[A_T4], B, h4, w4, st));
  RUN(run_conv(ctx, 9, A[A_T4], nullptr, A[A_X40], B, h4, w4, st));
  // decoder (unetpp.py:111-116): cat([skip, up(below)]) is never materialised -- the conv reads its K range from
  // two sources
  RUN(run_conv_up(ctx, 10, A[A_X30], A[A_X40], A[A_UP4], A[A_T3], 512, B, h3, w3, st));
  RUN(run_conv(ctx, 11, A[A_T3], nullptr, A[A_X31], B, h3, w3, st));
  RUN(run_conv_up(ctx, 12, A[A_X20], A[A_X31], A[A_UP3], A[A_T2], 256, B, h2, w2, st));
  RUN(run_conv(ctx, 13, A[A_T2], nullptr, A[A_X22], B, h2, w2, st));
  RUN(run_conv_up(ctx, 14, A[A_X10], A[A_X22], A[A_UP2], A[A_T1], 128, B, h1, w1, st));
  RUN(run_conv(ctx, 15, A[A_T1], nullptr, A[A_X13], B, h1, w1, st));
  RUN(run_conv_up(ctx, 16, A[A_X00], A[A_X13], A[A_UP1], A[A_T0], 64, B, H, W, st));
  // conv0_4.conv2 with the 1x1 `final` head (unetpp.py:119) and argmax fused in the epilogue
  RUN(run_conv(ctx, 17, A[A_T0], nullptr, A[A_X04], B, H, W, st, logits, pred, true));
  ctx->sat_check_once = 0;
  return UB_OK;
}
