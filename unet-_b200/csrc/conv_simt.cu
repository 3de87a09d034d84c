// conv_simt.cu -- plain CUDA-core 3x3 conv in the PP layout.  NOT the product path: it
// exists so the tcgen05 kernel (conv_tc.cu) can be checked kernel-against-kernel on the
// device, and is selected only through ub_debug_set_conv_impl(ctx, 1) (a test hook; never a fallback).
#include "common.cuh"

namespace {

struct SimtParams {
  const act_t* src0; const act_t* src1; act_t* out;
  const act_t* w; const float* bias;
  int B, cin0, cin1, cout, kc, bn, fold;
  PPLayout L;
  const float* head_w; const float* head_b; int ncls; float* logits; uint8_t* pred;
};

__global__ void __launch_bounds__(128) k_conv_simt(SimtParams P) {
  const PPLayout& L = P.L;
  int q = blockIdx.x * blockDim.x + threadIdx.x;          // position within the padded image
  int b = blockIdx.z;
  if (q >= L.ntiles * 128) return;
  int p = L.G + q;
  int yy = q / L.pitch, xx = q - yy * L.pitch;
  bool interior = (q < L.npos) && yy >= 1 && yy <= L.H && xx >= 1 && xx <= L.W;
  const bool head = P.head_w != nullptr;
  int nco_chunks = P.cout / 8;
  int co_begin = head ? 0 : blockIdx.y, co_end = head ? nco_chunks : blockIdx.y + 1;
  int kb0 = P.cin0 / P.kc, kbn = (P.cin0 + P.cin1) / P.kc, cpk = P.kc / 8;
  int nch0 = P.cin0 / 8, nch1 = P.cin1 / 8;
  float lg[8];
#pragma unroll
  for (int c = 0; c < 8; ++c) lg[c] = (head && c < P.ncls) ? P.head_b[c] : 0.f;

  for (int coc = co_begin; coc < co_end; ++coc) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (interior) {
      int nt = (coc * 8) / P.bn, r0 = (coc * 8) % P.bn;
      for (int kb = 0; kb < kbn; ++kb) {
        const act_t* plane0;
        if (kb < kb0) plane0 = P.src0 + ((size_t)b * nch0 + (size_t)kb * cpk) * L.Ppl * 8;
        else          plane0 = P.src1 + ((size_t)b * nch1 + (size_t)(kb - kb0) * cpk) * L.Ppl * 8;
        for (int tap = 0; tap < 9; ++tap) {
          int off = (tap / 3 - 1) * L.pitch + (tap % 3 - 1);
          for (int cc = 0; cc < cpk; ++cc) {
            uint4 iv = *reinterpret_cast<const uint4*>(plane0 + ((size_t)cc * L.Ppl + p + off) * 8);
            const act_t* ih = reinterpret_cast<const act_t*>(&iv);
            float fi[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) fi[j] = a2f(ih[j]);
            const act_t* wp =
                P.fold ? P.w + ((((size_t)kb * 3 + tap / 3) * cpk + cc) * (3 * P.cout) + (tap % 3) * P.cout + coc * 8) * 8
                       : P.w + (((((size_t)nt * kbn + kb) * 9 + tap) * cpk + cc) * P.bn + r0) * 8;
#pragma unroll
            for (int r = 0; r < 8; ++r) {
              uint4 wv = *reinterpret_cast<const uint4*>(wp + r * 8);
              const act_t* wh = reinterpret_cast<const act_t*>(&wv);
#pragma unroll
              for (int j = 0; j < 8; ++j) acc[r] = fmaf(fi[j], a2f(wh[j]), acc[r]);
            }
          }
        }
      }
    }
    float v[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) v[r] = interior ? fminf(fmaxf(acc[r] + P.bias[coc * 8 + r], 0.f), UB_ACT_MAX) : 0.f;
    if (head) {
#pragma unroll
      for (int c = 0; c < 8; ++c)
        if (c < P.ncls)
#pragma unroll
          for (int r = 0; r < 8; ++r) lg[c] = fmaf(P.head_w[c * P.cout + coc * 8 + r], v[r], lg[c]);
    } else {
      act_t h[8];
#pragma unroll
      for (int r = 0; r < 8; ++r) h[r] = f2a(v[r]);
      *reinterpret_cast<uint4*>(P.out + (((size_t)b * nco_chunks + coc) * L.Ppl + p) * 8) =
          *reinterpret_cast<uint4*>(h);
    }
  }
  if (head && interior) {
    size_t hw = (size_t)L.H * L.W, pix = (size_t)(yy - 1) * L.W + (xx - 1);
    int best = 0; float bv = lg[0];
#pragma unroll
    for (int c = 0; c < 8; ++c)
      if (c < P.ncls) {
        if (P.logits) P.logits[((size_t)b * P.ncls + c) * hw + pix] = lg[c];
        if (c > 0 && lg[c] > bv) { bv = lg[c]; best = c; }
      }
    if (P.pred) P.pred[(size_t)b * hw + pix] = (uint8_t)best;
  }
}

}  // namespace

int conv3x3_simt(ub_ctx* ctx, const ConvLayer& Lr, const ConvArgs& a, cudaStream_t st) {
  SimtParams P{};
  P.src0 = a.src0; P.src1 = a.src1; P.out = a.out; P.w = Lr.w; P.bias = Lr.bias;
  P.B = a.B; P.cin0 = Lr.cin0; P.cin1 = Lr.cin1; P.cout = Lr.cout; P.kc = Lr.kc; P.bn = Lr.bn; P.fold = Lr.fold;
  P.L = pp_layout(a.H, a.W);
  P.head_w = a.head_w; P.head_b = a.head_b; P.ncls = a.num_classes; P.logits = a.logits; P.pred = a.pred;
  if (a.head_w && Lr.cout != 32) return ub_fail(ctx, UB_ERR_ARG, "fused head needs cout == 32");
  TimeScope ts(ctx, UB_T_CONV, st);
  dim3 grid(cdiv(P.L.ntiles * 128, 128), a.head_w ? 1 : Lr.cout / 8, a.B);
  k_conv_simt<<<grid, 128, 0, st>>>(P);
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}
