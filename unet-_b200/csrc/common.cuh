// common.cuh -- shared declarations of libunetb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

// ---------------------------------------------------------------------------------
// 16-bit activation / weight type of the network path.
// Default: IEEE fp16 operands, fp32 accumulation.  Measured on B200 with random-init
// weights (DESIGN.md "numerics"): bf16 operands give 1.1-1.4e-2 max-relative logit error
// and 99.1-99.5 % argmax agreement -- outside BASELINE.json's tolerance (1e-2, 99.9 %) --
// while fp16 operands (3 more mantissa bits, same tcgen05 kind::f16 rate, same bytes) give
// 1.6e-3 and >= 99.9 %.  The epilogue saturates at the fp16 maximum.  -DUB_ACT_BF16
// rebuilds the library with bf16 operands for comparison.
// ---------------------------------------------------------------------------------
#ifdef UB_ACT_BF16
typedef __nv_bfloat16 act_t;
typedef __nv_bfloat162 act2_t;
#define UB_MMA_FMT 1u                 // tcgen05 instruction descriptor a/b format: BF16
#define UB_ACT_MAX 3.0e38f
__host__ __device__ __forceinline__ act_t f2a(float f) { return __float2bfloat16_rn(f); }
__host__ __device__ __forceinline__ float a2f(act_t h) { return __bfloat162float(h); }
__device__ __forceinline__ act2_t f2a2(float a, float b) { return __floats2bfloat162_rn(a, b); }
#else
typedef __half act_t;
typedef __half2 act2_t;
#define UB_MMA_FMT 0u                 // F16
#define UB_ACT_MAX 65504.0f
__host__ __device__ __forceinline__ act_t f2a(float f) { return __float2half_rn(f); }
__host__ __device__ __forceinline__ float a2f(act_t h) { return __half2float(h); }
__device__ __forceinline__ act2_t f2a2(float a, float b) { return __floats2half2_rn(a, b); }
#endif
#include <stdio.h>
#include <string>
#include <vector>
#include "../../include/unetb200.h"

// ---------------------------------------------------------------------------------
// Activation layout ("pixel-planar, zero-padded", PP)
//
// A feature map [B, C, H, W] is stored in bf16 as  [B][C/8][Ppl][8]:
// for every image and every group of 8 channels ("chunk") one plane of Ppl positions,
// 16 bytes per position.  Position of pixel (y,x):  p = G + (y+1)*(W+2) + (x+1), i.e.
// the plane is the image with a one-pixel zero border, flattened row-major, preceded by
// a guard of G >= W+3 positions and followed by a guard of >= 128 + W+3 positions.
//
// Why: a 3x3 tap (ky,kx) of ANY output position p is the input position
// p + (ky-1)*(W+2) + (kx-1) -- a constant offset, borders included.  A conv is then nine
// shifted GEMMs over a flat position axis: a tile of 128 consecutive positions is a
// UMMA M=128 operand whose rows are 16 B apart (no-swizzle K-major core matrices), the
// tap shift is a plain address offset, and every global->shared move is a contiguous
// 1-D bulk copy (no tensor map, no im2col).  Outputs computed at border positions are
// garbage and are overwritten with zeros by the epilogue, which keeps the zero border.
// ---------------------------------------------------------------------------------
struct PPLayout {
  int H, W;        // image size
  int pitch;       // W + 2
  int G;           // leading guard (positions), multiple of 8
  int npos;        // (H+2)*(W+2)
  int ntiles;      // ceil(npos / 128)
  int Ppl;         // positions per plane, multiple of 8
};

static inline PPLayout pp_layout(int H, int W) {
  PPLayout L;
  L.H = H; L.W = W; L.pitch = W + 2;
  L.G = ((W + 3) + 7) & ~7;
  L.npos = (H + 2) * (W + 2);
  L.ntiles = (L.npos + 127) / 128;
  L.Ppl = ((L.G + L.npos + 400 + (W + 3)) + 7) & ~7;   // covers the over-reads of both tilings
  return L;
}
static inline size_t pp_bytes(int B, int C, int H, int W) {
  PPLayout L = pp_layout(H, W);
  return (size_t)B * (C / 8) * L.Ppl * 16;
}

// ---------------------------------------------------------------------------------
// Conv layer description (weights already folded / repacked on the device)
// ---------------------------------------------------------------------------------
struct ConvLayer {
  int cin0, cin1;      // channels of source 0 (skip) and source 1 (upsampled); cin1 may be 0
  int cout;
  int kc;              // channels per k-block (16 / 32 / 64); divides cin0 and cin1
  int bn;              // N tile (<= 256); divides cout
  int fold = 0;        // 1: kx-folded weights [kblock][ky][kc/8][kx*cout+co][8] (conv_rows.cu, cout <= 64)
  // weights bf16 [cout/bn][kblocks][9 taps][kc/8][bn][8], bias fp32 [cout]
  act_t* w = nullptr;
  // fold layers, CTA-pair kernels: the same weights split by halves of N' = 3*cout,
  // [half][kblock][ky][kc/8][N'/2][8] (each CTA of a pair holds one half, contiguous)
  act_t* w2 = nullptr;
  // layer 0 only: x-taps folded into K for the x-im2col input, [ky][2 chunks][32][8] with k = kx*3 + c (conv_block.cu)
  act_t* wx = nullptr;
  float* bias = nullptr;
  float bias_host[64] = {0};   // host mirror of the first 64 biases (conv_rows passes them as kernel parameters)
};

// ---------------------------------------------------------------------------------
// Context
// ---------------------------------------------------------------------------------
struct Timing {
  bool on = false;
  struct Rec { cudaEvent_t a, b; int cls; };
  std::vector<Rec> recs;
  std::vector<cudaEvent_t> pool;
  float ms[UB_T_NCLASS] = {0};
  int64_t launches[UB_T_NCLASS] = {0};
};

struct ub_ctx {
  int device = 0, sm_count = 148;
  int max_batch = 0, frame_h = 0, frame_w = 0, net_h = 0, net_w = 0, num_classes = 3;
  int conv_impl = 0;                 // 0 tcgen05, 1 simt, 2 tcgen05 with aligned A segments (flat kernel only)
  int64_t launches = 0;              // kernels launched so far
  std::string err;
  bool weights_set = false;
  ConvLayer conv[18];
  float* head_w = nullptr;           // [num_classes][32] fp32
  float* head_b = nullptr;           // [num_classes]
  float head_host[8 * 33] = {0};     // host mirror: [num_classes][32] weights then [num_classes] biases
  // activation arena (PP layout, sized for max_batch at net_h x net_w)
  act_t* act[24] = {nullptr};
  int pool_border_hw[2][2] = {{0, 0}, {0, 0}};   // (H, W) for which the zero borders of A_P0 / A_P1 were last set (fused pool)
  // pre/post + stage-2 scratch
  uint8_t* d_frames = nullptr;       // [B,fh,fw,3]   (ub_pipeline_host staging)
  uint8_t* d_gray = nullptr;         // [B,fh,fw]
  uint8_t* d_pred = nullptr;         // [B,nh,nw]
  uint8_t* d_mask[3] = {nullptr};    // cable, tape, burr [B,fh,fw]
  uint8_t* d_tmp[4] = {nullptr};     // u8 planes [B,fh,fw]
  int32_t* d_lab[2] = {nullptr};     // i32 planes [B,fh,fw]
  int32_t* d_comp = nullptr;         // per-pixel component accumulators
  uint32_t* d_sums = nullptr;        // [B*H*W] per-root x sums (ub_filter_cable_by_shape; allocated on first use)
  int32_t* d_counts = nullptr;       // small per-frame counters
  int32_t* d_stats = nullptr;        // [B,4]
  float* d_logits = nullptr;         // [B,C,nh,nw] fp32 (module path / debug)
  int* d_rs_tab = nullptr;           // resize coefficient tables
  void* cache = nullptr;             // host-side cache of resize tables (ResizeCache*)
  void* enh_host = nullptr;          // enhance.cu lookup tables (EnhTables*), host copy
  void* d_enh = nullptr;             // ... and device copy
  cudaStream_t own_stream = nullptr;
  // ub_pipeline_submit / _wait: two slots so the H2D copy of batch i+1 (copy_stream) overlaps the compute of batch i
  uint8_t* d_frames2 = nullptr;        // frames of slot 1 (slot 0 uses d_frames), allocated at the first submit
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_h2d[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr};
  int slot_batch[2] = {0, 0};
  uint8_t* d_src[2] = {nullptr, nullptr};   // ub_pipeline_submit_src: raw source frames per slot (allocated on first use)
  size_t src_bytes = 0;
  int fuse_block0 = 1;                  // conv0_0 (both convs + pool) as one kernel (conv_block.cu; UB_FUSE_BLOCK0 / ub_set_option "fuse_block0")
  int overlap = 1;                      // gray + edge map on the aux stream beside the network (UB_OVERLAP / ub_set_option)
  int sat_check = 0;                    // conv epilogues count outputs that hit the 16-bit maximum (ub_set_option "check_saturation")
  int sat_check_once = 0;               // ... for the next forward only (armed by ub_set_weights)
  unsigned long long* d_sat = nullptr;  // the counter
  cudaStream_t aux_stream = nullptr;   // ub_pipeline: gray + edge map run here beside the network
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  void* graphs = nullptr;              // ub_pipeline_graph: captured steps (StepGraphs*, api.cu)
  Timing tm;
};

extern thread_local std::string g_last_error;
int ub_fail(ub_ctx* ctx, int code, const char* fmt, ...);

#define UB_CUDA(ctx, call)                                                              \
  do {                                                                                  \
    cudaError_t e__ = (call);                                                           \
    if (e__ != cudaSuccess)                                                             \
      return ub_fail(ctx, UB_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call,     \
                     cudaGetErrorString(e__));                                          \
  } while (0)

// after a group of n kernel launches: count them and surface launch errors
#define UB_CHECK_LAUNCH_N(ctx, n) do { (ctx)->launches += (n); UB_CUDA(ctx, cudaGetLastError()); } while (0)
#define UB_CHECK_LAUNCH(ctx) UB_CHECK_LAUNCH_N(ctx, 1)

// RAII bracket that times one kernel class with CUDA events when timing is enabled.
struct TimeScope {
  ub_ctx* c; cudaStream_t s; int idx = -1;
  TimeScope(ub_ctx* ctx, int cls, cudaStream_t st);
  ~TimeScope();
};

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// ---- entry points implemented across translation units ---------------------------
// conv_simt.cu / conv_tc.cu : out = relu(conv3x3(cat(src0,src1)) + bias), PP layout in/out.
// If head_w != nullptr (cout must be 32) the 1x1 head is fused: logits fp32 NCHW and/or
// pred u8 are written instead of `out`.
struct ConvArgs {
  const act_t* src0; const act_t* src1; act_t* out;
  int B, H, W;
  const float* head_w; const float* head_b; int num_classes;
  float* logits; uint8_t* pred;
  unsigned long long* sat;     // non-null: count outputs clamped to the largest finite 16-bit value
  act_t* pool_out;             // non-null: also write MaxPool2d(2,2) of the output into this PP plane (conv_rows only)
  bool up1;                    // src1 is [B, cin1, H/2, W/2]: bilinear x2 upsample fused into the operand path (conv_rows only)
};
int conv3x3_simt(ub_ctx* ctx, const ConvLayer& L, const ConvArgs& a, cudaStream_t st);
int conv3x3_tc(ub_ctx* ctx, const ConvLayer& L, const ConvArgs& a, cudaStream_t st);
int conv_tc_init(ub_ctx* ctx);
int conv3x3_rows(ub_ctx* ctx, const ConvLayer& L, const ConvArgs& a, cudaStream_t st);
int conv_rows_init(ub_ctx* ctx);
bool conv_rows_can_pool(const ub_ctx* ctx, const ConvLayer& L);   // the layer's kernel has a fused-pool variant
bool conv_rows_can_upsample(const ub_ctx* ctx, const ConvLayer& L);   // ... a fused-upsample variant (ConvArgs::up1)
// conv_block.cu: ConvBlock conv0_0 (conv1 -> conv2 -> 2x2 max-pool) in one kernel, intermediate kept in shared memory
int conv_block_init(ub_ctx* ctx);
bool conv_block_can_fuse(const ub_ctx* ctx, const ConvLayer& l1, const ConvLayer& l2, int H, int W);
int conv_block0(ub_ctx* ctx, const ConvLayer& l1, const ConvLayer& l2, const act_t* src, act_t* out, act_t* pool_out,
                int B, int H, int W, unsigned long long* sat, cudaStream_t st);
int run_conv_layer(ub_ctx* ctx, const ConvLayer& L, const ConvArgs& a, cudaStream_t st);   // picks the kernel

// activation buffers (ub_ctx::act): channels and pyramid level (spatial = net >> level); order = enum in net_glue.cu
static const int kActC[23] = {16, 32, 32, 32, 64, 64, 64, 128, 128, 128, 256, 256, 256, 512, 512, 512, 256, 256, 128, 128, 64, 64, 0};
static const int kActL[23] = {0, 0, 0, 1, 1, 1, 2, 2, 2, 3, 3, 3, 4, 4, 4, 3, 3, 2, 2, 1, 1, 0, 0};

// net_glue.cu
int nchw_to_pp(ub_ctx*, const float* x, int B, int C, int H, int W, act_t* out, int Cpad, cudaStream_t);
int nchw_to_xi(ub_ctx*, const float* x, int B, int H, int W, act_t* out, cudaStream_t);   // 3-channel input -> x-im2col PP (16 channels)
int pp_to_nchw(ub_ctx*, const act_t* in, int B, int C, int H, int W, float* out, cudaStream_t);
int maxpool2_pp(ub_ctx*, const act_t* in, int B, int C, int H, int W, act_t* out, cudaStream_t);
int upsample2_pp(ub_ctx*, const act_t* in, int B, int C, int H, int W, act_t* out, cudaStream_t);
// Whole forward.  The network input comes from exactly one of: x_nchw (fp32 NCHW [B,3,H,W]) or frames (u8 BGR
// [B,fh,fw,3], preprocess fused per sub-batch); writes logits (fp32 NCHW, nullable) and/or pred (u8, nullable).
int net_forward_pp(ub_ctx*, int B, int H, int W, const float* x_nchw, const uint8_t* frames, int fh, int fw,
                   float* logits, uint8_t* pred, cudaStream_t);

// preprocess.cu
int preprocess_frames(ub_ctx*, const uint8_t* frames, int B, int H, int W, int nh, int nw,
                      act_t* x_pp, float* x_nchw, uint8_t* gray, cudaStream_t, bool xi = false);   // xi: x-im2col network input (conv_block.cu)
int masks_from_pred(ub_ctx*, const uint8_t* pred, int B, int nh, int nw, int H, int W, ub_roi roi,
                    uint8_t* cable, uint8_t* tape, int32_t* stats, cudaStream_t);
int argmax_logits(ub_ctx*, const float* logits, int B, int C, int nh, int nw, uint8_t* pred, cudaStream_t);

// enhance.cu (enhance_image: Lab conversions, CLAHE, non-local means, sharpen); images u8 interleaved
int enhance_init(ub_ctx*);
void enhance_free(ub_ctx*);
int cvt_lab(ub_ctx*, int code, const uint8_t* src, uint8_t* dst, int B, int H, int W, cudaStream_t);
int clahe(ub_ctx*, const uint8_t* src, uint8_t* dst, int B, int H, int W, int pstride, int ch, double clip_limit,
          int tiles_x, int tiles_y, cudaStream_t);
int nlm(ub_ctx*, const uint8_t* src, uint8_t* dst, int B, int H, int W, int pstride, int ch, int cn, cudaStream_t);
int sharpen3(ub_ctx*, const uint8_t* src, uint8_t* dst, int B, int H, int W, cudaStream_t);
int enhance_frames(ub_ctx*, const uint8_t* frames, uint8_t* out, int B, int H, int W, cudaStream_t);

// stage2.cu
int burr_edges(ub_ctx*, int variant, const uint8_t* gray, int B, int H, int W, const ub_burr_params& p, cudaStream_t, int phase = 0);
int burr_finish(ub_ctx*, const uint8_t* mask, int B, int H, int W, const ub_burr_params& p, uint8_t* burr,
                int32_t* burr_px, int px_stride, cudaStream_t);
int burr_stage(ub_ctx*, int variant, const uint8_t* gray, const uint8_t* mask, int B, int H, int W,
               const ub_burr_params& p, uint8_t* burr, int32_t* burr_px, int px_stride, cudaStream_t);
