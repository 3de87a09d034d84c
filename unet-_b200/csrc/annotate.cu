// annotate.cu -- the vector half of visualize_two_stage (infer_two_stage_burr.py:156-168) on the device:
//   cv2.findContours(mask, RETR_EXTERNAL, CHAIN_APPROX_SIMPLE) + cv2.drawContours(img, contours, -1, colour, 2)
// as two kernels, with no contour lists and no border following, and a generic "paint where the layer is set" kernel for
// what the host rasterises (the ROI rectangle and the Hershey text).
//
// Why a local rule is enough.  Suzuki border following (what findContours runs) walks the cracks between a component and
// the 4-connected background region outside it, counter-clockwise: at a border pixel p it scans p's eight neighbours in
// chain-code order from the one after the previous border pixel and steps to the first foreground one.  The zeros it
// scans form an arc that starts with / contains an orthogonal neighbour of p lying in that background region.  So the set
// of steps of all OUTER borders is: for every foreground pixel p and every orthogonal neighbour z = p + D[d] (d even) that
// is EXTERIOR background (4-connected to the image frame: RETR_EXTERNAL drops holes and whatever lies inside them) --
// if p + D[d+1] is foreground the walk steps there, else if p + D[d+2] is foreground it steps there, else it stays at p
// and the next crack (p, d+2) decides.  drawContours(thickness 2) paints, for every segment between consecutive contour
// points, a filled polygon and a disc at its end; the union over a straight run of unit steps equals the painted long
// segment (all eight directions, any length), so the image is the union over all steps of the unit step's stamp plus the
// disc stamp on every border pixel.  Stamps are bit masks taken from cv2.line (tools/make_stamps.py;
// tests/test_host_logic.py checks them against the installed OpenCV).  Checked against cv2 on random, blob, line and
// speckle masks including ones that touch the image border (tests/test_gpu_annotate.py, tools/contour_proto.py).
// Thickness 3 (the burr contours): same rule with 7x7 stamps.  OpenCV's thick-line rasteriser is not translation invariant
// when an end point lies within one pixel of the image border (its polygon clipping then paints EXTRA pixels, never
// fewer): for a frame with a border pixel of the mask in that two-pixel band the kernel raises redo[frame], and the caller
// redraws that frame's contours with cv2 over the device's (a superset in the same colour) -- exact in every case.
#include "common.cuh"
#include "flood.cuh"

namespace ann {

// 32 pixels of row y starting at x -> one word, bit i = pixel x+i non-zero; 0 outside the image
__device__ __forceinline__ uint32_t pack32(const uint8_t* __restrict__ img, int H, int W, int y, int x) {
  if (y < 0 || y >= H || x >= W || x + 32 <= 0) return 0u;
  const uint8_t* p = img + (size_t)y * W + x;
  uint32_t word = 0;
  if (x >= 0 && x + 32 <= W && ((reinterpret_cast<uintptr_t>(p) & 15) == 0)) {
    const uint4 a = *reinterpret_cast<const uint4*>(p), b = *reinterpret_cast<const uint4*>(p + 16);
    const uint32_t v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t t = (((v[k] & 0x7f7f7f7fu) + 0x7f7f7f7fu) | v[k]) & 0x80808080u;      // bit 7 of byte j: byte j non-zero
      word = ((t * 0x00204081u) >> 28) * (1u << (4 * k)) + word;
    }
  } else {
    for (int i = 0; i < 32; ++i)
      if (x + i >= 0 && x + i < W && p[i]) word |= 1u << i;
  }
  return word;
}

// Exterior background of one frame: flood of the background (4-connectivity) from the image frame, both bit planes of
// the frame in shared memory with a border of one word / one row that is exterior by definition.  ext [B][H][Wd] words,
// bit = 1: background pixel connected to the outside (bits beyond W in the last word are set).
__global__ void __launch_bounds__(1024) k_exterior_frame(const uint8_t* __restrict__ mask, uint32_t* __restrict__ ext, int H, int W, int Wd) {
  extern __shared__ uint32_t ex_smem[];
  const int Wp = Wd + 2, NP = (H + 2) * Wp;
  uint32_t* C = ex_smem;                       // background (may be flooded)
  volatile uint32_t* S = ex_smem + NP;         // flooded so far
  const int tid = threadIdx.x, nthr = blockDim.x;
  for (int i = tid; i < NP; i += nthr) { C[i] = 0u; S[i] = 0xffffffffu; }        // the frame around the image is exterior
  __syncthreads();
  const uint8_t* img = mask + (size_t)blockIdx.x * H * W;
  for (int i = tid; i < H * Wd; i += nthr) {
    const int r = i / Wd, w = i - r * Wd, idx = (r + 1) * Wp + w + 1;
    const int rem = W - 32 * w;
    const uint32_t outside = rem >= 32 ? 0u : (0xffffffffu << rem);              // columns >= W of the last word
    C[idx] = ~pack32(img, H, W, r, 32 * w) | outside;
    S[idx] = outside;
  }
  __syncthreads();
  fl::flood<false>(C, S, H, Wd);
  uint32_t* out = ext + (size_t)blockIdx.x * H * Wd;
  for (int i = tid; i < H * Wd; i += nthr) {
    const int r = i / Wd, w = i - r * Wd;
    out[i] = S[(r + 1) * Wp + w + 1];
  }
}

// Stamps of cv2.line(img, p, p + D[d], colour, T) around p for T = 2 (5x5, bit (dy+2)*5 + dx+2) and T = 3 (7x7, bit
// (dy+3)*7 + dx+3); entry 8: the single point (the disc drawn at every contour vertex).  Chain codes D: 0 E, 1 NE, 2 N,
// 3 NW, 4 W, 5 SW, 6 S, 7 SE (y down).  Printed by tools/make_stamps.py; tests/test_host_logic.py compares them with the
// installed OpenCV through ub_contour_stamp.
#define UB_STAMPS2 {0x67980ull, 0x67bccull, 0x239c4ull, 0x33de6ull, 0x33cc0ull, 0x67bcc0ull, 0x473880ull, 0xcf7980ull, 0x23880ull}
#define UB_STAMPS3 {0xc3cfcf0c00ull, 0x41c7df1c10ull, 0x41c7cf8e08ull, 0x41c7c7c704ull, 0x61e7e78600ull, 0x1071f7c70400ull, \
                    0x20e3e7c70400ull, 0x41c7c7c70400ull, 0x41c7c70400ull}
__constant__ unsigned long long c_stamp[2][9] = {UB_STAMPS2, UB_STAMPS3};
static const unsigned long long h_stamp[2][9] = {UB_STAMPS2, UB_STAMPS3};

// One thread per 32-pixel word of the mask: border pixels (foreground with an exterior orthogonal neighbour) are few, each
// paints the stamps of its steps.  Three 34-bit windows (rows y-1, y, y+1; bit i+1 = pixel 32w + i) of the foreground and
// of the exterior plane give every neighbour of every pixel of the word by shifts.  Outside the image: background, exterior.
template <int T>       // thickness 2 or 3
__global__ void __launch_bounds__(256) k_contour_paint(const uint8_t* __restrict__ mask, const uint32_t* __restrict__ ext, uint8_t* __restrict__ frames,
                                                       int H, int W, int Wd, uchar3 colour, int* __restrict__ redo) {
  constexpr int R = T, SIDE = 2 * T + 1;
  const int w = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
  if (w >= Wd || y >= H) return;
  const uint8_t* img = mask + (size_t)blockIdx.z * H * W;
  const uint32_t* ex = ext + (size_t)blockIdx.z * H * Wd;
  const uint32_t fg = pack32(img, H, W, y, 32 * w);
  if (!fg) return;
  unsigned long long F[3], E[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int yy = y - 1 + k;
    if (yy < 0 || yy >= H) { F[k] = 0ull; E[k] = ~0ull; continue; }
    const uint32_t fl = w > 0 ? pack32(img, H, W, yy, 32 * (w - 1)) : 0u, fm = k == 1 ? fg : pack32(img, H, W, yy, 32 * w),
                   fr = w + 1 < Wd ? pack32(img, H, W, yy, 32 * (w + 1)) : 0u;
    const uint32_t el = w > 0 ? ex[yy * Wd + w - 1] : 0xffffffffu, em = ex[yy * Wd + w], er = w + 1 < Wd ? ex[yy * Wd + w + 1] : 0xffffffffu;
    F[k] = (unsigned long long)(fl >> 31) | ((unsigned long long)fm << 1) | ((unsigned long long)(fr & 1u) << 33);
    E[k] = (unsigned long long)(el >> 31) | ((unsigned long long)em << 1) | ((unsigned long long)(er & 1u) << 33);
  }
  // exterior orthogonal neighbour: E, N, W, S of pixel i are window bits (row 1, i+2), (row 0, i+1), (row 1, i), (row 2, i+1)
  uint32_t cand = fg & (uint32_t)((E[1] >> 2) | (E[0] >> 1) | E[1] | (E[2] >> 1));
  uint8_t* frame = frames + (size_t)blockIdx.z * H * W * 3;
  while (cand) {
    const int i = __ffs(cand) - 1;
    cand &= cand - 1;
    // neighbours in chain-code order: 0 E, 1 NE, 2 N, 3 NW, 4 W, 5 SW, 6 S, 7 SE
    const uint32_t r0 = (uint32_t)(F[0] >> i) & 7u, r1 = (uint32_t)(F[1] >> i) & 7u, r2 = (uint32_t)(F[2] >> i) & 7u;   // bit 0: x-1, 1: x, 2: x+1
    const uint32_t nb = ((r1 >> 2) & 1u) | (((r0 >> 2) & 1u) << 1) | (((r0 >> 1) & 1u) << 2) | ((r0 & 1u) << 3) | ((r1 & 1u) << 4) |
                        ((r2 & 1u) << 5) | (((r2 >> 1) & 1u) << 6) | (((r2 >> 2) & 1u) << 7);
    const uint32_t eo = ((uint32_t)(E[1] >> (i + 2)) & 1u) | (((uint32_t)(E[0] >> (i + 1)) & 1u) << 2) | (((uint32_t)(E[1] >> i) & 1u) << 4) |
                        (((uint32_t)(E[2] >> (i + 1)) & 1u) << 6);
    unsigned long long stamp = c_stamp[T - 2][8];
#pragma unroll
    for (int d = 0; d < 8; d += 2)
      if ((eo >> d) & 1u) {
        if ((nb >> (d + 1)) & 1u) stamp |= c_stamp[T - 2][d + 1];
        else if ((nb >> ((d + 2) & 7)) & 1u) stamp |= c_stamp[T - 2][(d + 2) & 7];
      }
    const int x = 32 * w + i;
    if (T == 3 && redo && (x <= 1 || y <= 1 || x >= W - 2 || y >= H - 2)) redo[blockIdx.z] = 1;
    while (stamp) {
      const int b = __ffsll((long long)stamp) - 1;
      stamp &= stamp - 1;
      const int yy = y + b / SIDE - R, xx = x + b % SIDE - R;
      if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
        uint8_t* p = frame + ((size_t)yy * W + xx) * 3;
        p[0] = colour.x; p[1] = colour.y; p[2] = colour.z;
      }
    }
  }
}

// frames[q] = palette[layer[q] - 1] where layer[q] != 0.  layer: one plane shared by the batch (stride 0) or one per frame.
__global__ void __launch_bounds__(256) k_paint_layer(uint8_t* __restrict__ frames, const uint8_t* __restrict__ layer, size_t layer_stride,
                                                     size_t npix, uchar3 c0, uchar3 c1, uchar3 c2) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npix) return;
  const uint8_t v = layer[(size_t)blockIdx.y * layer_stride + i];
  if (!v) return;
  const uchar3 c = v == 1 ? c0 : (v == 2 ? c1 : c2);
  uint8_t* p = frames + ((size_t)blockIdx.y * npix + i) * 3;
  p[0] = c.x; p[1] = c.y; p[2] = c.z;
}

}  // namespace ann

int draw_contours_impl(ub_ctx* ctx, uint8_t* frames, const uint8_t* mask, int B, int H, int W, int b, int g, int r, int thickness,
                       int* redo, cudaStream_t st) {
  if (thickness != 2 && thickness != 3) return ub_fail(ctx, UB_ERR_ARG, "ub_draw_contours: thickness %d (2 or 3)", thickness);
  if (thickness == 3 && !redo) return ub_fail(ctx, UB_ERR_ARG, "ub_draw_contours: thickness 3 needs the redo flags");
  const int Wd = (W + 31) / 32;
  const size_t smem = (size_t)2 * (H + 2) * (Wd + 2) * sizeof(uint32_t);
  if (smem > 200 * 1024) return ub_fail(ctx, UB_ERR_ARG, "ub_draw_contours: a %dx%d frame does not fit the per-frame flood (%zu B of shared memory)", H, W, smem);
  if ((size_t)B * H * Wd * sizeof(uint32_t) > (size_t)ctx->max_batch * ctx->frame_h * ctx->frame_w * sizeof(int))
    return ub_fail(ctx, UB_ERR_ARG, "ub_draw_contours: scratch too small");
  TimeScope ts(ctx, UB_T_PREPOST, st);
  static bool attr_set[64] = {false};
  if (ctx->device >= 0 && ctx->device < 64 && !attr_set[ctx->device]) {
    UB_CUDA(ctx, cudaFuncSetAttribute(ann::k_exterior_frame, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set[ctx->device] = true;
  }
  uint32_t* ext = reinterpret_cast<uint32_t*>(ctx->d_lab[1]);
  const int nthr = H * Wd >= 4096 ? 1024 : (H * Wd >= 1024 ? 512 : 128);
  ann::k_exterior_frame<<<B, nthr, smem, st>>>(mask, ext, H, W, Wd);
  const uchar3 col = make_uchar3((unsigned char)b, (unsigned char)g, (unsigned char)r);
  const dim3 grid(cdiv(Wd, 32), cdiv(H, 8), B), block(32, 8);
  if (redo) UB_CUDA(ctx, cudaMemsetAsync(redo, 0, sizeof(int) * B, st));
  if (thickness == 2) ann::k_contour_paint<2><<<grid, block, 0, st>>>(mask, ext, frames, H, W, Wd, col, redo);
  else ann::k_contour_paint<3><<<grid, block, 0, st>>>(mask, ext, frames, H, W, Wd, col, redo);
  UB_CHECK_LAUNCH_N(ctx, 2);
  return UB_OK;
}

int paint_layer_impl(ub_ctx* ctx, uint8_t* frames, const uint8_t* layer, int per_frame, int B, int H, int W, const uint8_t* bgr9, cudaStream_t st) {
  TimeScope ts(ctx, UB_T_PREPOST, st);
  const size_t npix = (size_t)H * W;
  ann::k_paint_layer<<<dim3((unsigned)((npix + 255) / 256), B), 256, 0, st>>>(frames, layer, per_frame ? npix : 0, npix,
      make_uchar3(bgr9[0], bgr9[1], bgr9[2]), make_uchar3(bgr9[3], bgr9[4], bgr9[5]), make_uchar3(bgr9[6], bgr9[7], bgr9[8]));
  UB_CHECK_LAUNCH(ctx);
  return UB_OK;
}

unsigned long long contour_stamp(int thickness, int code) {
  return (thickness == 2 || thickness == 3) && code >= 0 && code < 9 ? ann::h_stamp[thickness - 2][code] : 0ull;
}
