// enhance.cu -- `enhance_image` of the reference (infer_enhanced_burr.py:38-66) on the device:
//   cvtColor(BGR2LAB) :46 -> CLAHE(clip 3.0, 8x8 tiles) on L :50-51 -> cvtColor(LAB2BGR) :55 ->
//   fastNlMeansDenoisingColored(None, 10, 10, 7, 21) :58 -> filter2D 3x3 sharpen :61-64.
// Every step reproduces OpenCV 4.x's uint8 arithmetic bit for bit (oracle/cvlab.py and
// oracle/cvnp.py hold the CPU restatement this file is tested against):
//   * colour conversion: the fixed-point table algorithm of imgproc/color_lab.cpp
//     (RGB2Lab_b, Lab2RGBinteger).  The tables are rebuilt on the host at ub_create with the
//     same binary32 arithmetic OpenCV's softfloat performs, including its truncating cbrt.
//   * CLAHE: per-tile histogram, clip + redistribute, fp32 LUT scale, fp32 bilinear blend of the
//     four neighbouring tile LUTs (no FMA contraction -- OpenCV's build has none there).
//   * non-local means: integer SSD over the 7x7 template for each of the 21x21 search offsets,
//     weight = table[ssd >> 6] (fixed point, 19096 = 1.0), rounded integer division.
//     The denoiser works in LINEAR-gamma Lab (COLOR_LBGR2Lab / COLOR_Lab2LBGR), L and the
//     two-channel (a,b) plane separately.
// NLM is > 99 % of the work (1.6e8 template SSDs per 800x448 plane); the other kernels are
// plain streaming kernels.
#include <math.h>
#include <string.h>
#include "common.cuh"

namespace {

constexpr int LAB_SHIFT = 12, GAMMA_SHIFT = 3, LAB_SHIFT2 = LAB_SHIFT + GAMMA_SHIFT;
constexpr int CBRT_TAB = 256 * 3 / 2 * (1 << GAMMA_SHIFT);       // 3072
constexpr int LAB_BASE = 1 << 14;
constexpr int INV_GAMMA_TAB = 1 << 12;
constexpr int NLM_TWS = 7, NLM_SWS = 21, NLM_TR = NLM_TWS / 2, NLM_SR = NLM_SWS / 2, NLM_BORDER = NLM_TR + NLM_SR;
constexpr int NLM_MAX_TAB = 2048;

struct EnhTables {
  uint16_t gamma[256];          // sRGBGammaTab_b
  uint16_t cbrt[CBRT_TAB];      // LabCbrtTab_b
  uint16_t yf[512];             // LabToYF_b: (y, ify) per L
  uint8_t inv_gamma[INV_GAMMA_TAB];   // sRGBInvGammaTab_b
  int fwd[9], inv[9];           // fixed-point matrices, rows applied to / producing (B,G,R) memory order
  int nlm_n[2];                 // live entries of the weight tables (beyond: weight 0)
  int nlm_w[2][NLM_MAX_TAB];    // almost_dist2weight_ for 1 and 2 channels (h = 10)
};

// ---- host: tables ---------------------------------------------------------------------------
// cv::cbrt(softfloat): quartic rational approximation in binary64 on the mantissa scaled into
// [0.125,1); the binary64 fraction is truncated (not rounded) to 23 bits.
float soft_cbrt(float xf) {
  if (xf == 0.f) return 0.f;
  uint32_t ix; memcpy(&ix, &xf, 4); ix &= 0x7fffffffu;
  int ex = (int)(ix >> 23) - 127;
  int shx = ex % 3; shx -= shx >= 0 ? 3 : 0;
  const int ex3 = (ex - shx) / 3;
  uint32_t fi = (ix & ((1u << 23) - 1)) | ((uint32_t)(shx + 127) << 23);
  float frf; memcpy(&frf, &fi, 4);
  volatile double fr = frf;
  volatile double num = ((((45.2548339756803022511987494 * fr + 192.2798368355061050458134625) * fr + 119.1654824285581628956914143) * fr + 13.43250139086239872172837314) * fr + 0.1636161226585754240958355063);
  volatile double den = ((((14.80884093219134573786480845 * fr + 151.9714051044435648658557668) * fr + 168.5254414101568283957668343) * fr + 33.9905941350215598754191872) * fr + 1.0);
  double q = num / den;
  uint64_t qi; memcpy(&qi, &q, 8); qi &= ~((1ull << 29) - 1); memcpy(&q, &qi, 8);
  float r = (float)q;                                       // exact: 23 fraction bits left
  uint32_t ri; memcpy(&ri, &r, 4); ri += (uint32_t)(ex3 << 23); memcpy(&r, &ri, 4);
  return r;
}

int cv_round(double v) { return (int)lrint(v); }            // round half to even, as cvRound

void build_tables(EnhTables& T, float h_l, float h_ab) {
  for (int i = 0; i < 256; ++i) {
    volatile float x = (float)i / 255.0f;
    double xd = x;
    double g = xd <= 0.04045 ? xd / 12.92 : pow((xd + 0.055) / 1.055, 2.4);
    volatile float gf = (float)g;
    volatile float v = 2040.0f * gf;
    T.gamma[i] = (uint16_t)cv_round(v);
  }
  for (int i = 0; i < INV_GAMMA_TAB; ++i) {
    volatile float x = (1.0f / INV_GAMMA_TAB) * (float)i;
    double xd = x;
    double g = xd <= 0.0031308 ? xd * 12.92 : 1.055 * pow(xd, 1.0 / 2.4) - 0.055;
    volatile float gf = (float)g;
    volatile float v = 255.0f * gf;
    int r = cv_round(v);
    T.inv_gamma[i] = (uint8_t)(r < 0 ? 0 : r > 255 ? 255 : r);
  }
  {
    volatile float scale = 1.0f / (255.0f * 8.0f);
    volatile float lthresh = 216.0f / 24389.0f, lscale = 841.0f / 108.0f, lbias = 16.0f / 116.0f;
    for (int i = 0; i < CBRT_TAB; ++i) {
      volatile float x = scale * (float)i;
      float y;
      if (x < lthresh) y = (float)fma((double)x, (double)lscale, (double)lbias);   // softfloat mulAdd: one rounding
      else y = soft_cbrt(x);
      volatile float v = 32768.0f * y;
      T.cbrt[i] = (uint16_t)cv_round(v);
    }
  }
  for (int i = 0; i < 256; ++i) {
    int y, ify;
    if (i <= 20) {
      volatile float a = (float)(i * LAB_BASE * 20 * 9) / (float)(17 * 29 * 29 * 29);
      y = cv_round(a);
      volatile float t0 = 16.0f / 116.0f, t1 = (float)(i * 5) / (float)(3 * 17 * 29);
      volatile float t2 = t0 + t1;
      volatile float t3 = (float)LAB_BASE * t2;
      ify = cv_round(t3);
    } else {
      volatile float f0 = (float)(i * 100 * LAB_BASE) / (float)(255 * 116), f1 = (float)(16 * LAB_BASE) / 116.0f;
      volatile float fy = f0 + f1;
      ify = cv_round(fy);
      volatile float c2 = fy * fy;
      volatile float c3 = c2 * fy;
      volatile float c4 = c3 / (float)(LAB_BASE * LAB_BASE);
      y = cv_round(c4);
    }
    T.yf[2 * i] = (uint16_t)y; T.yf[2 * i + 1] = (uint16_t)ify;
  }
  static const double M[9] = {0.412453, 0.357580, 0.180423, 0.212671, 0.715160, 0.072169, 0.019334, 0.119193, 0.950227};
  static const double Mi[9] = {3.240479, -1.53715, -0.498535, -0.969256, 1.875991, 0.041556, 0.055648, -0.204043, 1.057311};
  static const double D65[3] = {0.950456, 1.0, 1.088754};
  for (int r = 0; r < 3; ++r) {
    T.fwd[r * 3 + 2] = cv_round(4096.0 * M[r * 3 + 0] / D65[r]);
    T.fwd[r * 3 + 1] = cv_round(4096.0 * M[r * 3 + 1] / D65[r]);
    T.fwd[r * 3 + 0] = cv_round(4096.0 * M[r * 3 + 2] / D65[r]);
    T.inv[r + 0] = cv_round(4096.0 * Mi[r + 6] * D65[r]);     // blue row
    T.inv[r + 3] = cv_round(4096.0 * Mi[r + 3] * D65[r]);
    T.inv[r + 6] = cv_round(4096.0 * Mi[r + 0] * D65[r]);     // red row
  }
  // NLM weights (FastNlMeansDenoisingInvoker ctor, DistSquared::calcWeight)
  const int fixed_point_mult = 2147483647 / (NLM_SWS * NLM_SWS * 255);            // 19096
  const double mult = 64.0 / (NLM_TWS * NLM_TWS);                                 // bin shift 6: 64 >= 49
  for (int c = 0; c < 2; ++c) {
    const int cn = c + 1;
    const float h = c ? h_ab : h_l;
    const double den = (double)(h * h * (float)cn);
    const int n_all = (int)(255.0 * 255.0 * cn / mult + 1);
    int live = 0;
    for (int i = 0; i < n_all && i < NLM_MAX_TAB; ++i) {
      double w = exp(-(i * mult) / den);
      int wi = cv_round(fixed_point_mult * w);
      if (wi < 0.001 * fixed_point_mult) wi = 0;
      T.nlm_w[c][i] = wi;
      if (wi) live = i + 1;
    }
    T.nlm_n[c] = live;
  }
}

}  // namespace

namespace {

// ---- colour conversion ----------------------------------------------------------------------
__device__ __forceinline__ int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }
__device__ __forceinline__ uint8_t sat_u8(int v) { return (uint8_t)min(max(v, 0), 255); }

// RGB2Lab_b::operator() (color_lab.cpp); p = 3 interleaved bytes in B,G,R order
__device__ __forceinline__ void bgr_to_lab(const EnhTables* __restrict__ T, bool srgb, int c0, int c1, int c2, int& L, int& a, int& b) {
  const int p0 = srgb ? T->gamma[c0] : c0 << GAMMA_SHIFT;
  const int p1 = srgb ? T->gamma[c1] : c1 << GAMMA_SHIFT;
  const int p2 = srgb ? T->gamma[c2] : c2 << GAMMA_SHIFT;
  const int fx = T->cbrt[descale(p0 * T->fwd[0] + p1 * T->fwd[1] + p2 * T->fwd[2], LAB_SHIFT)];
  const int fy = T->cbrt[descale(p0 * T->fwd[3] + p1 * T->fwd[4] + p2 * T->fwd[5], LAB_SHIFT)];
  const int fz = T->cbrt[descale(p0 * T->fwd[6] + p1 * T->fwd[7] + p2 * T->fwd[8], LAB_SHIFT)];
  constexpr int Lscale = (116 * 255 + 50) / 100;
  constexpr int Lshift = -((16 * 255 * (1 << LAB_SHIFT2) + 50) / 100);
  L = min(max(descale(Lscale * fy + Lshift, LAB_SHIFT2), 0), 255);
  a = min(max(descale(500 * (fx - fy) + 128 * (1 << LAB_SHIFT2), LAB_SHIFT2), 0), 255);
  b = min(max(descale(200 * (fy - fz) + 128 * (1 << LAB_SHIFT2), LAB_SHIFT2), 0), 255);
}

// abToXZ_b entry for fxz value i (Lab2RGBinteger tables, evaluated instead of stored: 144 KB)
__device__ __forceinline__ int ab_to_xz(int i) {
  return i <= 3390 ? i * 108 / 841 - LAB_BASE * 16 / 116 * 108 / 841 : i * i / LAB_BASE * i / LAB_BASE;
}

// Lab2RGBinteger::process
__device__ __forceinline__ void lab_to_bgr(const EnhTables* __restrict__ T, bool srgb, int L, int aa, int bb, int& o0, int& o1, int& o2) {
  const int y = T->yf[2 * L], ify = T->yf[2 * L + 1];
  const int adiv = ((5 * aa * 53687 + (1 << 7)) >> 13) - 128 * LAB_BASE / 500;
  const int bdiv = ((bb * 41943 + (1 << 4)) >> 9) - 128 * LAB_BASE / 200 + 1;
  const int x = ab_to_xz(ify + adiv), z = ab_to_xz(ify - bdiv);
  constexpr int shift = LAB_SHIFT + (14 - 12);
  int v[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    int t = descale(T->inv[r * 3] * x + T->inv[r * 3 + 1] * y + T->inv[r * 3 + 2] * z, shift);
    t = min(max(t, 0), INV_GAMMA_TAB - 1);
    v[r] = srgb ? (int)T->inv_gamma[t] : (255 * t) >> 12;       // linearInvGammaTab_b = trunc(255 * t / 4096)
  }
  o0 = v[0]; o1 = v[1]; o2 = v[2];
}

// code: 0 BGR2LAB, 1 LBGR2Lab, 2 LAB2BGR, 3 Lab2LBGR.  4 pixels (12 bytes) per thread.
__global__ void __launch_bounds__(256) k_cvt_lab(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, size_t npix,
                                                 int code, const EnhTables* __restrict__ T) {
  const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t p0 = q * 4;
  if (p0 >= npix) return;
  uint8_t in[12], out[12];
  const bool full = p0 + 4 <= npix;
  if (full) {
    const uint32_t* s = reinterpret_cast<const uint32_t*>(src + p0 * 3);
    uint32_t w0 = s[0], w1 = s[1], w2 = s[2];
    memcpy(in, &w0, 4); memcpy(in + 4, &w1, 4); memcpy(in + 8, &w2, 4);
  } else {
    for (size_t i = 0; i < (npix - p0) * 3; ++i) in[i] = src[p0 * 3 + i];
  }
  const int n = full ? 4 : (int)(npix - p0);
  for (int i = 0; i < n; ++i) {
    int r0, r1, r2;
    if (code < 2) bgr_to_lab(T, code == 0, in[3 * i], in[3 * i + 1], in[3 * i + 2], r0, r1, r2);
    else lab_to_bgr(T, code == 2, in[3 * i], in[3 * i + 1], in[3 * i + 2], r0, r1, r2);
    out[3 * i] = (uint8_t)r0; out[3 * i + 1] = (uint8_t)r1; out[3 * i + 2] = (uint8_t)r2;
  }
  if (full) {
    uint32_t w0, w1, w2;
    memcpy(&w0, out, 4); memcpy(&w1, out + 4, 4); memcpy(&w2, out + 8, 4);
    uint32_t* d = reinterpret_cast<uint32_t*>(dst + p0 * 3);
    d[0] = w0; d[1] = w1; d[2] = w2;
  } else {
    for (int i = 0; i < n * 3; ++i) dst[p0 * 3 + i] = out[i];
  }
}

// ---- CLAHE ----------------------------------------------------------------------------------
// One CTA per (tile, frame): histogram of channel `ch` of an interleaved image, clip,
// redistribute, prefix sum, LUT (CLAHE_CalcLut_Body, imgproc/clahe.cpp).
__global__ void __launch_bounds__(256) k_clahe_lut(const uint8_t* __restrict__ src, int H, int W, int pstride, int ch,
                                                   int tiles_x, int tiles_y, int clip, float lut_scale, uint8_t* __restrict__ luts) {
  __shared__ int hist[256];
  __shared__ int scan[256];
  __shared__ int s_excess;
  const int t = threadIdx.x;
  const int tile = blockIdx.x, b = blockIdx.y;
  const int tw = W / tiles_x, th = H / tiles_y;
  const int x0 = (tile % tiles_x) * tw, y0 = (tile / tiles_x) * th;
  hist[t] = 0;
  if (t == 0) s_excess = 0;
  __syncthreads();
  const uint8_t* img = src + (size_t)b * H * W * pstride + ch;
  for (int i = t; i < tw * th; i += 256) {
    const int y = i / tw, x = i - y * tw;
    atomicAdd(&hist[img[((size_t)(y0 + y) * W + x0 + x) * pstride]], 1);
  }
  __syncthreads();
  int h = hist[t];
  if (h > clip) { atomicAdd(&s_excess, h - clip); h = clip; }
  __syncthreads();
  const int excess = s_excess;
  const int batch = excess / 256;
  int resid = excess - batch * 256;
  h += batch;
  if (resid) {
    const int step = max(256 / resid, 1);
    if (t % step == 0 && t / step < resid) h += 1;
  }
  scan[t] = h;
  __syncthreads();
  for (int off = 1; off < 256; off <<= 1) {
    const int v = t >= off ? scan[t - off] : 0;
    __syncthreads();
    scan[t] += v;
    __syncthreads();
  }
  const float f = __fmul_rn((float)scan[t], lut_scale);
  luts[((size_t)b * tiles_x * tiles_y + tile) * 256 + t] = sat_u8(__float2int_rn(f));
}

// Bilinear blend of the four neighbouring tile LUTs (CLAHE_Interpolation_Body); in place on
// channel `ch` of the interleaved image.
__global__ void __launch_bounds__(256) k_clahe_apply(uint8_t* __restrict__ img, int H, int W, int pstride, int ch,
                                                     int tiles_x, int tiles_y, const uint8_t* __restrict__ luts) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
  if (x >= W) return;
  const float inv_tw = __fdiv_rn(1.0f, (float)(W / tiles_x)), inv_th = __fdiv_rn(1.0f, (float)(H / tiles_y));
  const float txf = __fsub_rn(__fmul_rn((float)x, inv_tw), 0.5f);
  const float tyf = __fsub_rn(__fmul_rn((float)y, inv_th), 0.5f);
  int tx1 = (int)floorf(txf), ty1 = (int)floorf(tyf);
  const float xa = __fsub_rn(txf, (float)tx1), ya = __fsub_rn(tyf, (float)ty1);
  const float xa1 = __fsub_rn(1.0f, xa), ya1 = __fsub_rn(1.0f, ya);
  const int tx2 = min(tx1 + 1, tiles_x - 1), ty2 = min(ty1 + 1, tiles_y - 1);
  tx1 = max(tx1, 0); ty1 = max(ty1, 0);
  uint8_t* p = img + (((size_t)b * H + y) * W + x) * pstride + ch;
  const int v = *p;
  const uint8_t* L = luts + (size_t)b * tiles_x * tiles_y * 256 + v;
  const float l11 = L[(ty1 * tiles_x + tx1) * 256], l12 = L[(ty1 * tiles_x + tx2) * 256];
  const float l21 = L[(ty2 * tiles_x + tx1) * 256], l22 = L[(ty2 * tiles_x + tx2) * 256];
  const float top = __fadd_rn(__fmul_rn(l11, xa1), __fmul_rn(l12, xa));
  const float bot = __fadd_rn(__fmul_rn(l21, xa1), __fmul_rn(l22, xa));
  const float res = __fadd_rn(__fmul_rn(top, ya1), __fmul_rn(bot, ya));
  *p = sat_u8(__float2int_rn(res));
}

// ---- 3x3 sharpen (filter2D, BORDER_REFLECT_101, saturate) -------------------------------------
__global__ void __launch_bounds__(256) k_sharpen3(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int H, int W) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;   // i over W*3 bytes of a row
  if (i >= W * 3) return;
  const int x = i / 3, c = i - x * 3;
  const uint8_t* img = src + (size_t)b * H * W * 3;
  const int ym = y == 0 ? (H > 1 ? 1 : 0) : y - 1, yp = y == H - 1 ? (H > 1 ? H - 2 : 0) : y + 1;
  const int xm = x == 0 ? (W > 1 ? 1 : 0) : x - 1, xp = x == W - 1 ? (W > 1 ? W - 2 : 0) : x + 1;
  const int ys[3] = {ym, y, yp}, xs[3] = {xm, x, xp};
  int acc = 0;
#pragma unroll
  for (int dy = 0; dy < 3; ++dy)
#pragma unroll
    for (int dx = 0; dx < 3; ++dx) acc -= img[((size_t)ys[dy] * W + xs[dx]) * 3 + c];
  acc += 10 * img[((size_t)y * W + x) * 3 + c];
  dst[(((size_t)b * H + y) * W + x) * 3 + c] = sat_u8(acc);
}

}  // namespace

namespace {

// ---- non-local means ------------------------------------------------------------------------
// FastNlMeansDenoisingInvoker<uchar | Vec2b, int, unsigned, DistSquared, int>::operator().
// CTA = 8 warps side by side; a warp owns NLM_WCOLS = 26 output columns x RT rows (RT = 20: see nlm()).  The source
// patch (tile + 13-pixel REFLECT_101 border) sits in shared memory.  For each of the 441 search
// offsets a lane walks DOWN its column keeping the vertical 7-sum of squared differences in
// registers (ring of 7), the horizontal 7-sum is four shuffles (lanes 26..31 only feed their
// neighbours), and the per-pixel accumulators (sum w*p, sum w) stay in registers across all
// offsets -- shared memory is read three times per (pixel, offset), never written.
constexpr int NLM_WARPS = 8, NLM_WCOLS = 26, NLM_TCOLS = NLM_WARPS * NLM_WCOLS;   // 208 columns per CTA
constexpr int NLM_PW = NLM_TCOLS + 2 * NLM_BORDER + 6;                             // 240: lanes 26..31 read up to +5 beyond

__device__ __forceinline__ int reflect101(int p, int n) {
  if (n == 1) return 0;
  while (p < 0 || p >= n) p = p < 0 ? -p : 2 * n - 2 - p;
  return p;
}

template <int CN> struct NlmPix;
template <> struct NlmPix<1> {
  typedef uint8_t T;
  static __device__ __forceinline__ int dist(T a, T b) { const int d = (int)a - (int)b; return d * d; }
};
template <> struct NlmPix<2> {
  typedef uint16_t T;
  static __device__ __forceinline__ int dist(T a, T b) {
    const int d0 = (int)(a & 255) - (int)(b & 255), d1 = (int)(a >> 8) - (int)(b >> 8);
    return d0 * d0 + d1 * d1;
  }
};

template <int CN, int RT>
__global__ void __launch_bounds__(NLM_WARPS * 32) k_nlm(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int H, int W,
                                                        int pstride, int ch, const int* __restrict__ wtab, int ntab) {
  typedef typename NlmPix<CN>::T T;
  constexpr int PH = RT + 2 * NLM_BORDER;
  extern __shared__ __align__(16) uint8_t nlm_smem[];
  int* s_tab = reinterpret_cast<int*>(nlm_smem);
  T* patch = reinterpret_cast<T*>(nlm_smem + NLM_MAX_TAB * 4);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int x0 = blockIdx.x * NLM_TCOLS, y0 = blockIdx.y * RT, b = blockIdx.z;
  const uint8_t* img = src + (size_t)b * H * W * pstride + ch;

  for (int i = threadIdx.x; i < ntab; i += blockDim.x) s_tab[i] = wtab[i];
  for (int i = threadIdx.x; i < PH * NLM_PW; i += blockDim.x) {
    const int r = i / NLM_PW, c = i - r * NLM_PW;
    const int gy = reflect101(y0 - NLM_BORDER + r, H), gx = reflect101(x0 - NLM_BORDER + c, W);
    const uint8_t* p = img + ((size_t)gy * W + gx) * pstride;
    if (CN == 1) patch[i] = p[0];
    else patch[i] = (T)(p[0] | (p[1] << 8));
  }
  __syncthreads();

  unsigned est[CN][RT], wsum[RT];
#pragma unroll
  for (int r = 0; r < RT; ++r) {
    wsum[r] = 0;
#pragma unroll
    for (int c = 0; c < CN; ++c) est[c][r] = 0;
  }
  // lane l computes the squared differences of patch column col_d = SR + warp*26 + l (= output column - TR);
  // after the horizontal sum lane l (< 26) holds the template SSD of output column warp*26 + l.
  const int col_d = NLM_SR + warp * NLM_WCOLS + lane;
  for (int o = 0; o < NLM_SWS * NLM_SWS; ++o) {
    const int dy = o / NLM_SWS - NLM_SR, dx = o - (o / NLM_SWS) * NLM_SWS - NLM_SR;
    const T* pc = patch + NLM_SR * NLM_PW + col_d;                  // d2 row j = -TR  -> patch row BORDER - TR = SR
    const T* ps = patch + (NLM_SR + dy) * NLM_PW + col_d + dx;
    int ring[NLM_TWS];
    int V = 0;
#pragma unroll
    for (int j = 0; j < RT + 2 * NLM_TR; ++j) {
      const int d = NlmPix<CN>::dist(pc[j * NLM_PW], ps[j * NLM_PW]);
      if (j >= NLM_TWS) V -= ring[j % NLM_TWS];
      V += d;
      ring[j % NLM_TWS] = d;
      if (j >= 2 * NLM_TR) {
        const int r = j - 2 * NLM_TR;                               // output row of this step
        const int a2 = V + __shfl_down_sync(0xffffffffu, V, 1);
        const int a4 = a2 + __shfl_down_sync(0xffffffffu, a2, 2);
        const int ssd = a4 + __shfl_down_sync(0xffffffffu, a2, 4) + __shfl_down_sync(0xffffffffu, V, 6);
        const int idx = ssd >> 6;
        if (idx < ntab) {
          const unsigned w = (unsigned)s_tab[idx];
          const T p = ps[(r + NLM_TR) * NLM_PW + NLM_TR];           // shifted pixel of (row r, column + TR)
          wsum[r] += w;
          if (CN == 1) est[0][r] += w * (unsigned)p;
          else { est[0][r] += w * (unsigned)(p & 255); est[CN - 1][r] += w * (unsigned)(p >> 8); }
        }
      }
    }
  }
  const int x = x0 + warp * NLM_WCOLS + lane;
  if (lane < NLM_WCOLS && x < W) {
#pragma unroll
    for (int r = 0; r < RT; ++r) {
      const int y = y0 + r;
      if (y < H) {
        uint8_t* q = dst + (((size_t)b * H + y) * W + x) * pstride + ch;
#pragma unroll
        for (int c = 0; c < CN; ++c) q[c] = (uint8_t)min((est[c][r] + wsum[r] / 2) / wsum[r], 255u);
      }
    }
  }
}

template <int CN, int RT>
int launch_nlm(ub_ctx* ctx, const uint8_t* src, uint8_t* dst, int B, int H, int W, int pstride, int ch, const EnhTables* T, cudaStream_t st) {
  typedef typename NlmPix<CN>::T PT;
  const size_t smem = NLM_MAX_TAB * 4 + (size_t)(RT + 2 * NLM_BORDER) * NLM_PW * sizeof(PT);
  static bool attr_done = false;
  if (!attr_done) {
    UB_CUDA(ctx, cudaFuncSetAttribute(k_nlm<CN, RT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_done = true;
  }
  EnhTables* host = reinterpret_cast<EnhTables*>(ctx->enh_host);
  dim3 grid(cdiv(W, NLM_TCOLS), cdiv(H, RT), B);
  k_nlm<CN, RT><<<grid, NLM_WARPS * 32, smem, st>>>(src, dst, H, W, pstride, ch, T->nlm_w[CN - 1], host->nlm_n[CN - 1]);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

}  // namespace

// ---- host entry points (called from api.cu) ---------------------------------------------------
int enhance_init(ub_ctx* ctx) {
  EnhTables* T = new EnhTables;
  build_tables(*T, 10.0f, 10.0f);
  if (T->nlm_n[0] >= NLM_MAX_TAB || T->nlm_n[1] >= NLM_MAX_TAB) { delete T; return ub_fail(ctx, UB_ERR_ARG, "NLM weight table too long"); }
  ctx->enh_host = T;
  UB_CUDA(ctx, cudaMalloc(&ctx->d_enh, sizeof(EnhTables)));
  UB_CUDA(ctx, cudaMemcpy(ctx->d_enh, T, sizeof(EnhTables), cudaMemcpyHostToDevice));
  return UB_OK;
}

void enhance_free(ub_ctx* ctx) {
  delete reinterpret_cast<EnhTables*>(ctx->enh_host);
  ctx->enh_host = nullptr;
  if (ctx->d_enh) cudaFree(ctx->d_enh);
  ctx->d_enh = nullptr;
}

int cvt_lab(ub_ctx* ctx, int code, const uint8_t* src, uint8_t* dst, int B, int H, int W, cudaStream_t st) {
  if (code < 0 || code > 3) return ub_fail(ctx, UB_ERR_ARG, "colour code %d (0 BGR2LAB, 1 LBGR2Lab, 2 LAB2BGR, 3 Lab2LBGR)", code);
  TimeScope ts(ctx, UB_T_ENHANCE, st);
  const size_t npix = (size_t)B * H * W;
  k_cvt_lab<<<(unsigned)((npix / 4 + 256) / 256), 256, 0, st>>>(src, dst, npix, code, reinterpret_cast<const EnhTables*>(ctx->d_enh));
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

// CLAHE on channel `ch` of an image with `pstride` interleaved channels; src may equal dst.
int clahe(ub_ctx* ctx, const uint8_t* src, uint8_t* dst, int B, int H, int W, int pstride, int ch, double clip_limit,
          int tiles_x, int tiles_y, cudaStream_t st) {
  if (tiles_x < 1 || tiles_y < 1 || tiles_x * tiles_y > 64 || W % tiles_x || H % tiles_y)
    return ub_fail(ctx, UB_ERR_ARG, "CLAHE: %dx%d does not divide into %dx%d tiles (<= 64 tiles; OpenCV would pad)", W, H, tiles_x, tiles_y);
  // the per-tile LUTs live at the start of d_comp (24 bytes per frame pixel of the context)
  if ((size_t)B * tiles_x * tiles_y * 256 > (size_t)24 * ctx->max_batch * ctx->frame_h * ctx->frame_w)
    return ub_fail(ctx, UB_ERR_ARG, "CLAHE: %d images x %d tiles need more LUT scratch than a %dx%d context holds", B,
                   tiles_x * tiles_y, ctx->frame_w, ctx->frame_h);
  TimeScope ts(ctx, UB_T_ENHANCE, st);
  const int area = (W / tiles_x) * (H / tiles_y);
  int clip = (int)(clip_limit * area / 256);
  if (clip < 1) clip = 1;
  const float lut_scale = 255.0f / (float)area;
  uint8_t* luts = reinterpret_cast<uint8_t*>(ctx->d_comp);              // B * 64 * 256 bytes of scratch
  if (src != dst) UB_CUDA(ctx, cudaMemcpyAsync(dst, src, (size_t)B * H * W * pstride, cudaMemcpyDeviceToDevice, st));
  k_clahe_lut<<<dim3(tiles_x * tiles_y, B), 256, 0, st>>>(src, H, W, pstride, ch, tiles_x, tiles_y, clip, lut_scale, luts);
  k_clahe_apply<<<dim3(cdiv(W, 256), H, B), 256, 0, st>>>(dst, H, W, pstride, ch, tiles_x, tiles_y, luts);
  UB_CHECK_LAUNCH_N(ctx, 2);
  return UB_OK;
}

// fastNlMeansDenoising(h = 10, 7, 21) on `cn` (1 or 2) channels starting at `ch` of an interleaved image
int nlm(ub_ctx* ctx, const uint8_t* src, uint8_t* dst, int B, int H, int W, int pstride, int ch, int cn, cudaStream_t st) {
  if (cn < 1 || cn > 2 || ch < 0 || ch + cn > pstride) return ub_fail(ctx, UB_ERR_ARG, "NLM: %d channels at %d of %d", cn, ch, pstride);
  if (src == dst) return ub_fail(ctx, UB_ERR_ARG, "NLM cannot run in place");
  TimeScope ts(ctx, UB_T_ENHANCE, st);
  const EnhTables* T = reinterpret_cast<const EnhTables*>(ctx->d_enh);
  // rows per warp walk (RT): the accumulators of RT rows live in registers, so RT sets the register count and with it the warps per
  // SM that hide the SHFL / LDS latency of the walk.  One channel: 98 registers / 16 warps per SM at RT = 32, 64 / 28 at RT = 20
  // (ncu: issue slots 46 % -> 59 % busy, short-scoreboard stalls either way).  Two channels: 126 registers at RT = 32 AND at 24, yet
  // 24 rows take half the time -- the walk is unrolled over RT + 6 steps inside the loop over the 441 offsets, and the 38-step body
  // is what hurts there.  Per 800x448 frame, batch 16, L / ab channels (profiles/r03k_nlm_rows_sweep.log): 0.309 / 0.457 ms at
  // RT = 32, 0.201 / 0.248 at 28, 0.211 / 0.232 at 24, 0.201 / 0.215 at 20 (the default), 0.204 / 0.250 at 16, 0.207 / 0.264 at 8.
  static const int rt_env = getenv("UB_NLM_RT") ? atoi(getenv("UB_NLM_RT")) : 20;
#define UB_NLM_CASE(R_)                                                                         \
  if (rt_env == R_)                                                                             \
    return cn == 1 ? launch_nlm<1, R_>(ctx, src, dst, B, H, W, pstride, ch, T, st)              \
                   : launch_nlm<2, R_>(ctx, src, dst, B, H, W, pstride, ch, T, st);
  UB_NLM_CASE(8) UB_NLM_CASE(16) UB_NLM_CASE(24) UB_NLM_CASE(28) UB_NLM_CASE(32)
#undef UB_NLM_CASE
  return cn == 1 ? launch_nlm<1, 20>(ctx, src, dst, B, H, W, pstride, ch, T, st)
                 : launch_nlm<2, 20>(ctx, src, dst, B, H, W, pstride, ch, T, st);
}

int sharpen3(ub_ctx* ctx, const uint8_t* src, uint8_t* dst, int B, int H, int W, cudaStream_t st) {
  if (src == dst) return ub_fail(ctx, UB_ERR_ARG, "sharpen cannot run in place");
  TimeScope ts(ctx, UB_T_ENHANCE, st);
  k_sharpen3<<<dim3(cdiv(W * 3, 256), H, B), 256, 0, st>>>(src, dst, H, W);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

// enhance_image for a batch of BGR frames; scratch: two 3-channel images (ctx->d_lab[0], d_lab[1])
int enhance_frames(ub_ctx* ctx, const uint8_t* frames, uint8_t* out, int B, int H, int W, cudaStream_t st) {
  uint8_t* A = reinterpret_cast<uint8_t*>(ctx->d_lab[0]);
  uint8_t* Bf = reinterpret_cast<uint8_t*>(ctx->d_lab[1]);
  int r;
  if ((r = cvt_lab(ctx, 0, frames, A, B, H, W, st))) return r;                    // :46  BGR -> Lab
  if ((r = clahe(ctx, A, A, B, H, W, 3, 0, 3.0, 8, 8, st))) return r;            // :50-51 CLAHE on L, in place
  if ((r = cvt_lab(ctx, 2, A, A, B, H, W, st))) return r;                         // :55  Lab -> BGR
  if ((r = cvt_lab(ctx, 1, A, A, B, H, W, st))) return r;                         // :58  linear BGR -> Lab (inside cv2)
  if ((r = nlm(ctx, A, Bf, B, H, W, 3, 0, 1, st))) return r;                      //      L, h = 10
  if ((r = nlm(ctx, A, Bf, B, H, W, 3, 1, 2, st))) return r;                      //      (a,b), hColor = 10
  if ((r = cvt_lab(ctx, 3, Bf, Bf, B, H, W, st))) return r;                       //      Lab -> linear BGR
  return sharpen3(ctx, Bf, out, B, H, W, st);                                     // :61-64
}
