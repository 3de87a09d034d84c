// api.cu -- the C ABI of libunetb200.so (include/unetb200.h): context, weight folding /
// repacking, and thin wrappers that validate arguments and enqueue the kernels.
#include <math.h>
#include <stdarg.h>
#include <string.h>
#include <map>
#include <vector>
#include "common.cuh"

thread_local std::string g_last_error;

int ub_fail(ub_ctx* ctx, int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_last_error = buf;
  if (ctx) ctx->err = buf;
  return code;
}

// ---------------------------------------------------------------- timing
TimeScope::TimeScope(ub_ctx* ctx, int cls, cudaStream_t st) : c(ctx), s(st) {
  if (!c->tm.on) return;
  auto get = [&]() {
    cudaEvent_t e;
    if (!c->tm.pool.empty()) { e = c->tm.pool.back(); c->tm.pool.pop_back(); }
    else cudaEventCreate(&e);
    return e;
  };
  Timing::Rec r{get(), get(), cls};
  cudaEventRecord(r.a, s);
  c->tm.recs.push_back(r);
  idx = (int)c->tm.recs.size() - 1;
}
TimeScope::~TimeScope() {
  if (idx >= 0) cudaEventRecord(c->tm.recs[idx].b, s);
}

// forward declarations of functions defined in other translation units
void resize_cache_free(ub_ctx* ctx);
int normalize_frames_impl(ub_ctx*, const uint8_t*, int, int, int, int, uint8_t*, int, int, cudaStream_t);
int overlay_impl(ub_ctx*, const uint8_t*, const uint8_t*, const uint8_t*, const uint8_t*, int, int, int, ub_roi, uint8_t*, cudaStream_t);
int morph_op(ub_ctx*, int op, int k, const uint8_t* src, uint8_t* dst, uint8_t* tmp, int B, int H, int W, int value, cudaStream_t);
int gaussian5(ub_ctx*, const uint8_t*, uint8_t*, int, int, int, cudaStream_t);
int canny(ub_ctx*, const uint8_t*, uint8_t*, uint8_t*, int*, int, int, int, int, int, cudaStream_t);
int blur_canny(ub_ctx*, const uint8_t*, uint8_t*, int, int, int, int, int, cudaStream_t, int phase = 0);
int draw_contours_impl(ub_ctx*, uint8_t*, const uint8_t*, int, int, int, int, int, int, int, int*, cudaStream_t);
int paint_layer_impl(ub_ctx*, uint8_t*, const uint8_t*, int, int, int, int, const uint8_t*, cudaStream_t);
unsigned long long contour_stamp(int, int);
int diameter_metrics(ub_ctx*, const uint8_t*, int, int, int, int, int, int, int, int32_t*, cudaStream_t);
int analyze_defects(ub_ctx*, const uint8_t*, int, int, int, int, int, int, int32_t*, cudaStream_t);
int width_profiles(ub_ctx*, const uint8_t*, int, int, int, int, int, float*, cudaStream_t);
int filter_cable_by_shape(ub_ctx*, const uint8_t*, int, int, int, int, double, double, int, uint8_t*, cudaStream_t);
int constrain_tape_to_ring(ub_ctx*, const uint8_t*, const uint8_t*, int, int, int, int, int, uint8_t*, cudaStream_t);
int enh_edges(ub_ctx*, const uint8_t*, int, int, int, int, int, int, const uint8_t*, uint8_t*, cudaStream_t);
int ccl_export(ub_ctx*, const uint8_t*, int, int, int, int*, int*, int, int*, cudaStream_t);
int sobel_edges_static(const uint8_t*, int, int, int, int, uint8_t*, cudaStream_t);

// ---------------------------------------------------------------- layer tables
namespace {
// keeps the caller's current device across an entry point that has to switch to the context's
struct StepKey {
  const void* frames; void* stats; void* masks; int B, variant, overlap, has_roi, has_params; ub_roi roi; ub_burr_params params;
};
struct StepGraph { StepKey key; cudaGraphExec_t exec = nullptr; int calls = 0; int64_t launches = 0; };
struct StepGraphs { std::vector<StepGraph> v; };
constexpr size_t kMaxStepGraphs = 8;

void step_graphs_free(ub_ctx* ctx) {
  if (!ctx->graphs) return;
  StepGraphs* g = static_cast<StepGraphs*>(ctx->graphs);
  for (auto& e : g->v) if (e.exec) cudaGraphExecDestroy(e.exec);
  delete g;
  ctx->graphs = nullptr;
}

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; if (prev != dev) cudaSetDevice(dev); else prev = -1; }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
struct LayerDef { const char* block; int idx; int cin0, cin1, cout, kc, bn; int real_cin0; int fold; };
// cin0 of the first layer is the 3 input channels zero-padded to 16 (one UMMA K step)
const LayerDef kLayers[18] = {
    {"conv0_0", 1, 16, 0, 32, 16, 32, 3, 1},     {"conv0_0", 2, 32, 0, 32, 32, 32, 32, 1},
    {"conv1_0", 1, 32, 0, 64, 32, 64, 32, 1},    {"conv1_0", 2, 64, 0, 64, 64, 64, 64, 1},
    {"conv2_0", 1, 64, 0, 128, 16, 128, 64, 0},  {"conv2_0", 2, 128, 0, 128, 16, 128, 128, 0},
    {"conv3_0", 1, 128, 0, 256, 16, 256, 128, 0}, {"conv3_0", 2, 256, 0, 256, 16, 256, 256, 0},
    {"conv4_0", 1, 256, 0, 512, 16, 256, 256, 0}, {"conv4_0", 2, 512, 0, 512, 16, 256, 512, 0},
    {"conv3_1", 1, 256, 512, 256, 16, 256, 256, 0}, {"conv3_1", 2, 256, 0, 256, 16, 256, 256, 0},
    {"conv2_2", 1, 128, 256, 128, 16, 128, 128, 0}, {"conv2_2", 2, 128, 0, 128, 16, 128, 128, 0},
    {"conv1_3", 1, 64, 128, 64, 32, 64, 64, 1},  {"conv1_3", 2, 64, 0, 64, 64, 64, 64, 1},
    {"conv0_4", 1, 32, 64, 32, 32, 32, 32, 1},   {"conv0_4", 2, 32, 0, 32, 32, 32, 32, 1},
};
cudaStream_t as_stream(void* s) { return static_cast<cudaStream_t>(s); }

template <class T>
int dmalloc(ub_ctx* ctx, T** p, size_t n) {
  cudaError_t e = cudaMalloc((void**)p, n * sizeof(T));
  if (e != cudaSuccess) return ub_fail(ctx, UB_ERR_NOMEM, "cudaMalloc(%zu bytes): %s", n * sizeof(T), cudaGetErrorString(e));
  e = cudaMemset(*p, 0, n * sizeof(T));
  if (e != cudaSuccess) return ub_fail(ctx, UB_ERR_CUDA, "cudaMemset: %s", cudaGetErrorString(e));
  return UB_OK;
}
}  // namespace

#define RUN(x) do { int r__ = (x); if (r__ != UB_OK) return r__; } while (0)
#define NEED_CTX(ctx) if (!(ctx)) return ub_fail(nullptr, UB_ERR_ARG, "null context")

extern "C" {

int ub_version(void) { return 100; }

const char* ub_last_error(ub_ctx* ctx) { return ctx ? ctx->err.c_str() : g_last_error.c_str(); }

int ub_create(ub_ctx** out, int device, int max_batch, int frame_h, int frame_w, int net_h, int net_w, int num_classes) {
  if (!out) return ub_fail(nullptr, UB_ERR_ARG, "out is null");
  *out = nullptr;
  if (max_batch < 1 || frame_h < 1 || frame_w < 1 || net_h < 16 || net_w < 16 || (net_h & 15) || (net_w & 15) ||
      num_classes < 1 || num_classes > 8)
    return ub_fail(nullptr, UB_ERR_ARG, "bad sizes: batch %d frame %dx%d net %dx%d classes %d (net dims must be multiples of 16)",
                   max_batch, frame_h, frame_w, net_h, net_w, num_classes);
  cudaDeviceProp prop;
  cudaError_t e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess) return ub_fail(nullptr, UB_ERR_CUDA, "cudaGetDeviceProperties(%d): %s", device, cudaGetErrorString(e));
  if (prop.major != 10)
    return ub_fail(nullptr, UB_ERR_ARCH, "device %d is sm_%d%d; libunetb200 is built for sm_100a only (no fallback path)",
                   device, prop.major, prop.minor);
  DeviceGuard guard(device);
  if ((e = cudaGetLastError()) != cudaSuccess) return ub_fail(nullptr, UB_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
  ub_ctx* c = new ub_ctx();
  if (const char* ov = getenv("UB_OVERLAP")) c->overlap = atoi(ov);
  if (const char* fb = getenv("UB_FUSE_BLOCK0")) c->fuse_block0 = atoi(fb);
  c->device = device; c->sm_count = prop.multiProcessorCount;
  c->max_batch = max_batch; c->frame_h = frame_h; c->frame_w = frame_w; c->net_h = net_h; c->net_w = net_w;
  c->num_classes = num_classes;
  if (const char* impl = getenv("UB_CONV_IMPL")) {        // 2 = the aligned-A variant of the tcgen05 kernel; the CUDA-core comparator (1) cannot
    if (atoi(impl) == 2) c->conv_impl = 2;                 // be selected from the environment, only per context by ub_debug_set_conv_impl
  }
  int r = UB_OK;
  auto fail = [&](int code) { ub_destroy(c); return code; };
  for (int i = 0; i < 22 && r == UB_OK; ++i) {
    size_t bytes = pp_bytes(max_batch, kActC[i], net_h >> kActL[i], net_w >> kActL[i]);
    r = dmalloc(c, &c->act[i], bytes / sizeof(act_t));
  }
  if (r) return fail(r);
  const size_t fpix = (size_t)max_batch * frame_h * frame_w, npix = (size_t)max_batch * net_h * net_w;
  const size_t ipl = (size_t)max_batch * (frame_h + 1) * (frame_w + 1);
  if ((r = dmalloc(c, &c->d_frames, fpix * 3)) || (r = dmalloc(c, &c->d_gray, fpix)) || (r = dmalloc(c, &c->d_pred, npix)) ||
      (r = dmalloc(c, &c->d_lab[0], ipl)) || (r = dmalloc(c, &c->d_lab[1], ipl)) || (r = dmalloc(c, &c->d_comp, fpix * 6)) ||
      (r = dmalloc(c, &c->d_counts, (size_t)max_batch * 16)) || (r = dmalloc(c, &c->d_stats, (size_t)max_batch * 8)) ||
      (r = dmalloc(c, &c->d_logits, npix * num_classes)) || (r = dmalloc(c, &c->d_sat, 1)))
    return fail(r);
  for (int i = 0; i < 3 && !r; ++i) r = dmalloc(c, &c->d_mask[i], fpix);
  for (int i = 0; i < 4 && !r; ++i) r = dmalloc(c, &c->d_tmp[i], fpix);
  if (r) return fail(r);
  if ((r = conv_tc_init(c)) != UB_OK || (r = conv_rows_init(c)) != UB_OK || (r = conv_block_init(c)) != UB_OK || (r = enhance_init(c)) != UB_OK) return fail(r);
  e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->aux_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
  for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
    e = cudaEventCreateWithFlags(&c->ev_h2d[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_done[i], cudaEventDisableTiming);
  }
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming);
  if (e != cudaSuccess) { ub_fail(nullptr, UB_ERR_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(e)); return fail(UB_ERR_CUDA); }
  *out = c;
  return UB_OK;
}

void ub_destroy(ub_ctx* c) {
  if (!c) return;
  DeviceGuard guard(c->device);
  cudaDeviceSynchronize();
  step_graphs_free(c);
  cudaFree(c->d_src[0]); cudaFree(c->d_src[1]); cudaFree(c->d_sat);
  for (auto& p : c->act) cudaFree(p);
  for (auto& L : c->conv) { cudaFree(L.w); cudaFree(L.w2); cudaFree(L.wx); cudaFree(L.bias); }
  cudaFree(c->head_w); cudaFree(c->head_b);
  cudaFree(c->d_frames); cudaFree(c->d_gray); cudaFree(c->d_pred);
  for (auto& p : c->d_mask) cudaFree(p);
  for (auto& p : c->d_tmp) cudaFree(p);
  for (auto& p : c->d_lab) cudaFree(p);
  cudaFree(c->d_sums); cudaFree(c->d_comp); cudaFree(c->d_counts); cudaFree(c->d_stats); cudaFree(c->d_logits);
  resize_cache_free(c);
  enhance_free(c);
  for (auto& r : c->tm.recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
  for (auto& e : c->tm.pool) cudaEventDestroy(e);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  if (c->aux_stream) cudaStreamDestroy(c->aux_stream);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  for (int i = 0; i < 2; ++i) { if (c->ev_h2d[i]) cudaEventDestroy(c->ev_h2d[i]); if (c->ev_done[i]) cudaEventDestroy(c->ev_done[i]); }
  cudaFree(c->d_frames2);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->ev_join) cudaEventDestroy(c->ev_join);
  delete c;
}

int ub_debug_set_conv_impl(ub_ctx* ctx, int impl) {
  NEED_CTX(ctx);
  if (impl < 0 || impl > 2) return ub_fail(ctx, UB_ERR_ARG, "conv impl %d (0 tcgen05, 1 simt, 2 tcgen05 aligned-A)", impl);
  ctx->conv_impl = impl;
  step_graphs_free(ctx);                 // captured steps hold the old kernel choice
  return UB_OK;
}

// Fold eval-mode BN into the conv, repack to bf16 [cout/bn][kblock][tap][kc/8][bn][8].
int ub_set_weights(ub_ctx* ctx, const ub_named_tensor* t, int n) {
  NEED_CTX(ctx);
  if (!t || n <= 0) return ub_fail(ctx, UB_ERR_ARG, "no tensors");
  DeviceGuard guard(ctx->device);
  cudaDeviceSynchronize();
  step_graphs_free(ctx);                 // captured steps hold the old weight pointers
  std::map<std::string, const ub_named_tensor*> m;
  for (int i = 0; i < n; ++i) if (t[i].name) m[t[i].name] = &t[i];
  auto get = [&](const std::string& name, int64_t numel, const float** out) -> int {
    auto it = m.find(name);
    if (it == m.end()) return ub_fail(ctx, UB_ERR_ARG, "missing key \"%s\" in state_dict", name.c_str());
    if (it->second->numel != numel)
      return ub_fail(ctx, UB_ERR_ARG, "size mismatch for %s: got %lld elements, expected %lld", name.c_str(),
                     (long long)it->second->numel, (long long)numel);
    *out = static_cast<const float*>(it->second->data_host);
    return UB_OK;
  };
  for (int li = 0; li < 18; ++li) {
    const LayerDef& d = kLayers[li];
    const std::string cv = std::string(d.block) + ".conv" + std::to_string(d.idx);
    const std::string bn = std::string(d.block) + ".bn" + std::to_string(d.idx);
    const int cin_real = d.real_cin0 + d.cin1, cin_pad = d.cin0 + d.cin1;
    const float *w, *b, *g, *be, *mu, *var;
    RUN(get(cv + ".weight", (int64_t)d.cout * cin_real * 9, &w));
    RUN(get(cv + ".bias", d.cout, &b));
    RUN(get(bn + ".weight", d.cout, &g));
    RUN(get(bn + ".bias", d.cout, &be));
    RUN(get(bn + ".running_mean", d.cout, &mu));
    RUN(get(bn + ".running_var", d.cout, &var));
    const int kbn = cin_pad / d.kc, cpk = d.kc / 8, nnt = d.cout / d.bn;
    // wp2: the CTA-pair layout (conv_rows: halves of N' = 3*cout; conv_tc with bn >= 64: halves of every n-tile, per stage)
    const bool tc_pair = !d.fold && d.bn >= 64;
    const int kys = d.bn >= 256 ? 1 : 3;               // kernel rows per pipeline stage of k_conv_tc (Cfg::KYS)
    std::vector<act_t> wp((size_t)d.cout * cin_pad * 9), wp2(d.fold || tc_pair ? wp.size() : 0);
    std::vector<float> bias(d.cout);
    for (int co = 0; co < d.cout; ++co) {
      const double scale = (double)g[co] / sqrt((double)var[co] + 1e-5);
      bias[co] = (float)(((double)b[co] - (double)mu[co]) * scale + (double)be[co]);
      const int nt = co / d.bn, r = co % d.bn;
      for (int ci = 0; ci < cin_pad; ++ci) {
        // padded channel -> real channel of the torch weight (only layer 0 has padding)
        int cr = -1;
        if (ci < d.cin0) { if (ci < d.real_cin0) cr = ci; }
        else cr = d.real_cin0 + (ci - d.cin0);
        const int kb = ci / d.kc, cc = (ci % d.kc) / 8, j = ci % 8;
        for (int tap = 0; tap < 9; ++tap) {
          const float v = cr < 0 ? 0.f : (float)((double)w[((size_t)co * cin_real + cr) * 9 + tap] * scale);
          const size_t o = d.fold ? (((((size_t)kb * 3 + tap / 3) * cpk + cc) * (3 * d.cout) + (tap % 3) * d.cout + co) * 8) + j
                                  : ((((((size_t)nt * kbn + kb) * 9 + tap) * cpk + cc) * d.bn + r) * 8) + j;
          wp[o] = f2a(v);
          if (d.fold) {
            const int nph = 3 * d.cout / 2, n = (tap % 3) * d.cout + co;
            wp2[((((((size_t)(n / nph) * kbn + kb) * 3 + tap / 3) * cpk + cc) * nph) + n % nph) * 8 + j] = wp[o];
          } else if (tc_pair) {
            // [n-tile][k-block][stage][rank][row of the stage][kx][chunk][bn/2][8]
            const int ky = tap / 3, kx = tap % 3, bh = d.bn / 2, nst = 3 / kys;
            const size_t stg = ((size_t)nt * kbn + kb) * nst + ky / kys;
            wp2[((((((stg * 2 + r / bh) * kys + ky % kys) * 3 + kx) * cpk + cc) * bh) + r % bh) * 8 + j] = wp[o];
          }
        }
      }
    }
    (void)nnt;
    ConvLayer& L = ctx->conv[li];
    L.cin0 = d.cin0; L.cin1 = d.cin1; L.cout = d.cout; L.kc = d.kc; L.bn = d.bn; L.fold = d.fold;
    if (!L.w) RUN(dmalloc(ctx, &L.w, wp.size()));
    if (!L.bias) RUN(dmalloc(ctx, &L.bias, bias.size()));
    if (li == 0) {                 // x-taps folded into K (k = kx*3 + c) for the x-im2col input of conv_block.cu: [ky][2][32][8]
      std::vector<act_t> wx((size_t)3 * 16 * d.cout, f2a(0.f));
      for (int co = 0; co < d.cout; ++co) {
        const double scale = (double)g[co] / sqrt((double)var[co] + 1e-5);
        for (int ky = 0; ky < 3; ++ky)
          for (int kx = 0; kx < 3; ++kx)
            for (int c = 0; c < d.real_cin0; ++c) {
              const int k = kx * 3 + c;
              wx[(((size_t)ky * 2 + k / 8) * d.cout + co) * 8 + k % 8] = f2a((float)((double)w[((size_t)co * cin_real + c) * 9 + ky * 3 + kx] * scale));
            }
      }
      if (!L.wx) RUN(dmalloc(ctx, &L.wx, wx.size()));
      UB_CUDA(ctx, cudaMemcpy(L.wx, wx.data(), wx.size() * sizeof(act_t), cudaMemcpyHostToDevice));
    }
    if (d.fold || tc_pair) {
      if (!L.w2) RUN(dmalloc(ctx, &L.w2, wp2.size()));
      UB_CUDA(ctx, cudaMemcpy(L.w2, wp2.data(), wp2.size() * sizeof(act_t), cudaMemcpyHostToDevice));
    }
    UB_CUDA(ctx, cudaMemcpy(L.w, wp.data(), wp.size() * sizeof(act_t), cudaMemcpyHostToDevice));
    UB_CUDA(ctx, cudaMemcpy(L.bias, bias.data(), bias.size() * sizeof(float), cudaMemcpyHostToDevice));
    for (int i = 0; i < 64 && i < d.cout; ++i) L.bias_host[i] = bias[i];
  }
  const float *hw, *hb;
  RUN(get("final.weight", (int64_t)ctx->num_classes * 32, &hw));
  RUN(get("final.bias", ctx->num_classes, &hb));
  if (!ctx->head_w) RUN(dmalloc(ctx, &ctx->head_w, (size_t)ctx->num_classes * 32));
  if (!ctx->head_b) RUN(dmalloc(ctx, &ctx->head_b, (size_t)ctx->num_classes));
  UB_CUDA(ctx, cudaMemcpy(ctx->head_w, hw, sizeof(float) * ctx->num_classes * 32, cudaMemcpyHostToDevice));
  UB_CUDA(ctx, cudaMemcpy(ctx->head_b, hb, sizeof(float) * ctx->num_classes, cudaMemcpyHostToDevice));
  memcpy(ctx->head_host, hw, sizeof(float) * ctx->num_classes * 32);
  memcpy(ctx->head_host + ctx->num_classes * 32, hb, sizeof(float) * ctx->num_classes);
  ctx->weights_set = true;
  ctx->sat_check_once = 1;            // the first forward with these weights counts saturated outputs
  return UB_OK;
}

int ub_set_option(ub_ctx* ctx, const char* name, int value) {
  NEED_CTX(ctx);
  if (!name) return ub_fail(ctx, UB_ERR_ARG, "null option name");
  if (!strcmp(name, "overlap")) ctx->overlap = value != 0;
  else if (!strcmp(name, "fuse_block0")) { ctx->fuse_block0 = value != 0; step_graphs_free(ctx); }
  else if (!strcmp(name, "check_saturation")) { ctx->sat_check = value != 0; if (!value) ctx->sat_check_once = 0; }
  else return ub_fail(ctx, UB_ERR_ARG, "unknown option \"%s\"", name);
  return UB_OK;
}

int ub_get_stat(ub_ctx* ctx, const char* name, int64_t* value, int reset) {
  NEED_CTX(ctx);
  if (!name || !value) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  if (strcmp(name, "saturated")) return ub_fail(ctx, UB_ERR_ARG, "unknown statistic \"%s\"", name);
  DeviceGuard guard(ctx->device);
  unsigned long long v = 0;
  UB_CUDA(ctx, cudaDeviceSynchronize());
  UB_CUDA(ctx, cudaMemcpy(&v, ctx->d_sat, sizeof v, cudaMemcpyDeviceToHost));
  if (reset) UB_CUDA(ctx, cudaMemset(ctx->d_sat, 0, sizeof v));
  *value = (int64_t)v;
  return UB_OK;
}

static int check_net_shape(ub_ctx* ctx, int B, int H, int W) {
  if (B < 1 || B > ctx->max_batch) return ub_fail(ctx, UB_ERR_ARG, "batch %d outside [1,%d]", B, ctx->max_batch);
  if (H < 16 || W < 16 || (H & 15) || (W & 15)) return ub_fail(ctx, UB_ERR_ARG, "net input %dx%d must be multiples of 16", H, W);
  // buffers are sized for net_h x net_w; any shape whose padded planes are not larger fits
  for (int l = 0; l < 5; ++l)
    if (pp_layout(H >> l, W >> l).Ppl > pp_layout(ctx->net_h >> l, ctx->net_w >> l).Ppl)
      return ub_fail(ctx, UB_ERR_ARG, "net input %dx%d exceeds the context's %dx%d", H, W, ctx->net_h, ctx->net_w);
  if (!ctx->weights_set) return ub_fail(ctx, UB_ERR_STATE, "ub_set_weights has not been called");
  return UB_OK;
}

int ub_forward(ub_ctx* ctx, const float* x, int B, int H, int W, float* logits, void* stream) {
  NEED_CTX(ctx);
  if (!x || !logits) return ub_fail(ctx, UB_ERR_ARG, "null tensor");
  RUN(check_net_shape(ctx, B, H, W));
  cudaStream_t st = as_stream(stream);
  return net_forward_pp(ctx, B, H, W, x, nullptr, 0, 0, logits, nullptr, st);
}

int ub_preprocess(ub_ctx* ctx, const uint8_t* frames, int B, int H, int W, float* x_nchw, uint8_t* gray, void* stream) {
  NEED_CTX(ctx);
  if (!frames || B < 1 || H < 2 || W < 2) return ub_fail(ctx, UB_ERR_ARG, "bad frames argument");
  return preprocess_frames(ctx, frames, B, H, W, ctx->net_h, ctx->net_w, nullptr, x_nchw, gray, as_stream(stream));
}

int ub_normalize_frames(ub_ctx* ctx, const uint8_t* src, int B, int sh, int sw, int rot, uint8_t* dst, int dh, int dw, void* stream) {
  NEED_CTX(ctx);
  if (!src || !dst || B < 1 || sh < 2 || sw < 2 || dh < 1 || dw < 1) return ub_fail(ctx, UB_ERR_ARG, "bad argument");
  return normalize_frames_impl(ctx, src, B, sh, sw, rot, dst, dh, dw, as_stream(stream));
}

int ub_masks_from_logits(ub_ctx* ctx, const float* logits, int B, int nh, int nw, int H, int W, const ub_roi* roi,
                         uint8_t* pred, uint8_t* cable, uint8_t* tape, void* stream) {
  NEED_CTX(ctx);
  if (!logits || !roi || !cable || !tape) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  if (B > ctx->max_batch || (size_t)nh * nw > (size_t)ctx->net_h * ctx->net_w) return ub_fail(ctx, UB_ERR_ARG, "shape exceeds context");
  cudaStream_t st = as_stream(stream);
  uint8_t* p = pred ? pred : ctx->d_pred;
  RUN(argmax_logits(ctx, logits, B, ctx->num_classes, nh, nw, p, st));
  return masks_from_pred(ctx, p, B, nh, nw, H, W, *roi, cable, tape, nullptr, st);
}

static int check_frame_shape(ub_ctx* ctx, int B, int H, int W) {
  if (B < 1 || B > ctx->max_batch || H < 1 || W < 1 || (size_t)H * W > (size_t)ctx->frame_h * ctx->frame_w)
    return ub_fail(ctx, UB_ERR_ARG, "frames %dx%dx%d exceed the context (%d x %dx%d)", B, H, W, ctx->max_batch, ctx->frame_h, ctx->frame_w);
  return UB_OK;
}

int ub_overlay(ub_ctx* ctx, const uint8_t* frames, const uint8_t* mc, const uint8_t* mt, const uint8_t* mb, int B, int H, int W,
               const ub_roi* roi, uint8_t* out, void* stream) {
  NEED_CTX(ctx);
  if (!frames || !mc || !mt || !mb || !out || !roi || B < 1 || H < 1 || W < 1) return ub_fail(ctx, UB_ERR_ARG, "ub_overlay: null pointer or empty shape");
  return overlay_impl(ctx, frames, mc, mt, mb, B, H, W, *roi, out, as_stream(stream));
}

int ub_draw_contours(ub_ctx* ctx, uint8_t* frames, const uint8_t* mask, int B, int H, int W, int b, int g, int r, int thickness,
                     int32_t* redo, void* stream) {
  NEED_CTX(ctx);
  if (!frames || !mask) return ub_fail(ctx, UB_ERR_ARG, "ub_draw_contours: null pointer");
  RUN(check_frame_shape(ctx, B, H, W));
  return draw_contours_impl(ctx, frames, mask, B, H, W, b, g, r, thickness, redo, as_stream(stream));
}

int ub_paint_layer(ub_ctx* ctx, uint8_t* frames, const uint8_t* layer, int per_frame, int B, int H, int W, const uint8_t* bgr9, void* stream) {
  NEED_CTX(ctx);
  if (!frames || !layer || !bgr9) return ub_fail(ctx, UB_ERR_ARG, "ub_paint_layer: null pointer");
  RUN(check_frame_shape(ctx, B, H, W));
  return paint_layer_impl(ctx, frames, layer, per_frame, B, H, W, bgr9, as_stream(stream));
}

unsigned long long ub_contour_stamp(int thickness, int code) { return contour_stamp(thickness, code); }

int ub_diameter_metrics(ub_ctx* ctx, const uint8_t* pred, int B, int H, int W, int cable_cls, int tape_cls, int min_area,
                        int min_valid_rows, int32_t* out, void* stream) {
  NEED_CTX(ctx);
  if (!pred || !out) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  RUN(check_frame_shape(ctx, B, H, W));
  return diameter_metrics(ctx, pred, B, H, W, cable_cls, tape_cls, min_area, min_valid_rows, out, as_stream(stream));
}

int ub_analyze_defects(ub_ctx* ctx, const uint8_t* pred, int B, int H, int W, int cable_cls, int tape_cls, int hole_min_size,
                       int32_t* out, void* stream) {
  NEED_CTX(ctx);
  if (!pred || !out) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  RUN(check_frame_shape(ctx, B, H, W));
  return analyze_defects(ctx, pred, B, H, W, cable_cls, tape_cls, hole_min_size, out, as_stream(stream));
}

int ub_width_profiles(ub_ctx* ctx, const uint8_t* pred, int B, int H, int W, int cable_cls, int tape_cls, float* out, void* stream) {
  NEED_CTX(ctx);
  if (!pred || !out) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  RUN(check_frame_shape(ctx, B, H, W));
  return width_profiles(ctx, pred, B, H, W, cable_cls, tape_cls, out, as_stream(stream));
}

int ub_filter_cable_by_shape(ub_ctx* ctx, const uint8_t* mask, int B, int H, int W, int min_area, double min_aspect,
                             double max_center_offset, int roi_width, uint8_t* out, void* stream) {
  NEED_CTX(ctx);
  if (!mask || !out || roi_width < 1) return ub_fail(ctx, UB_ERR_ARG, "null argument or roi_width < 1");
  RUN(check_frame_shape(ctx, B, H, W));
  return filter_cable_by_shape(ctx, mask, B, H, W, min_area, min_aspect, max_center_offset, roi_width, out, as_stream(stream));
}

int ub_constrain_tape_to_ring(ub_ctx* ctx, const uint8_t* tape, const uint8_t* cable, int B, int H, int W, int dilate_k, int erode_k,
                              uint8_t* out, void* stream) {
  NEED_CTX(ctx);
  if (!tape || !cable || !out) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  RUN(check_frame_shape(ctx, B, H, W));
  return constrain_tape_to_ring(ctx, tape, cable, B, H, W, dilate_k, erode_k, out, as_stream(stream));
}

int ub_gaussian5(ub_ctx* ctx, const uint8_t* src, int B, int H, int W, uint8_t* dst, void* stream) {
  NEED_CTX(ctx);
  RUN(check_frame_shape(ctx, B, H, W));
  return gaussian5(ctx, src, dst, B, H, W, as_stream(stream));
}

int ub_canny(ub_ctx* ctx, const uint8_t* src, int B, int H, int W, int low, int high, uint8_t* dst, void* stream) {
  NEED_CTX(ctx);
  RUN(check_frame_shape(ctx, B, H, W));
  return canny(ctx, src, dst, ctx->d_tmp[3], ctx->d_lab[0], B, H, W, low, high, as_stream(stream));
}

int ub_blur_canny(ub_ctx* ctx, const uint8_t* gray, int B, int H, int W, int low, int high, uint8_t* dst, void* stream) {
  NEED_CTX(ctx);
  RUN(check_frame_shape(ctx, B, H, W));
  return blur_canny(ctx, gray, dst, B, H, W, low, high, as_stream(stream));
}

int ub_morph(ub_ctx* ctx, int op, int k, const uint8_t* src, int B, int H, int W, int value, uint8_t* dst, void* stream) {
  NEED_CTX(ctx);
  RUN(check_frame_shape(ctx, B, H, W));
  return morph_op(ctx, op, k, src, dst, ctx->d_tmp[3], B, H, W, value, as_stream(stream));
}

int ub_ccl(ub_ctx* ctx, const uint8_t* src, int B, int H, int W, int32_t* labels, int32_t* stats, int max_labels,
           int32_t* n_labels, void* stream) {
  NEED_CTX(ctx);
  RUN(check_frame_shape(ctx, B, H, W));
  return ccl_export(ctx, src, B, H, W, labels, stats, max_labels, n_labels, as_stream(stream));
}

int ub_sobel_edges(ub_ctx* ctx, const uint8_t* gray, int B, int H, int W, int thresh, uint8_t* dst, void* stream) {
  NEED_CTX(ctx);
  RUN(check_frame_shape(ctx, B, H, W));
  return enh_edges(ctx, gray, B, H, W, 1, thresh, 0, nullptr, dst, as_stream(stream));
}
int ub_sobel_edges_static(const uint8_t* gray, int B, int H, int W, int thresh, uint8_t* dst, void* stream) {
  return sobel_edges_static(gray, B, H, W, thresh, dst, as_stream(stream));
}
int ub_laplacian_edges(ub_ctx* ctx, const uint8_t* gray, int B, int H, int W, int thresh, uint8_t* dst, void* stream) {
  NEED_CTX(ctx);
  RUN(check_frame_shape(ctx, B, H, W));
  return enh_edges(ctx, gray, B, H, W, 2, 0, thresh, nullptr, dst, as_stream(stream));
}

static ub_burr_params default_params(int variant) {
  ub_burr_params p;
  if (variant == UB_VARIANT_ENHANCED) p = {50, 500, 6, 5, 30, 100, 25, 5, 3};
  else if (variant == UB_VARIANT_LAPLACIAN) p = {20, 500, 1 << 20, 0, 30, 0, 21, 1, 1};       // BurrConfig defaults, src/refactor/config.py:98-103
  else p = {30, 800, 5, 4, 50, 150, 8, 3, 2};
  return p;
}

int ub_burr(ub_ctx* ctx, int variant, const uint8_t* gray, const uint8_t* mask, int B, int H, int W,
            const ub_burr_params* params, uint8_t* burr, int32_t* burr_px, void* stream) {
  NEED_CTX(ctx);
  if (!gray || !mask || !burr) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  if (variant < UB_VARIANT_STANDARD || variant > UB_VARIANT_LAPLACIAN) return ub_fail(ctx, UB_ERR_ARG, "variant %d", variant);
  RUN(check_frame_shape(ctx, B, H, W));
  cudaStream_t st = as_stream(stream);
  ub_burr_params p = params ? *params : default_params(variant);
  if (burr_px) UB_CUDA(ctx, cudaMemsetAsync(burr_px, 0, sizeof(int32_t) * B, st));
  return burr_stage(ctx, variant, gray, mask, B, H, W, p, burr, burr_px, 1, st);
}

static int check_enh_shape(ub_ctx* ctx, int B, int H, int W) {
  RUN(check_frame_shape(ctx, B, H, W));
  if (H < 14 || W < 14) return ub_fail(ctx, UB_ERR_ARG, "image %dx%d smaller than the 13-pixel NLM border", H, W);
  return UB_OK;
}

int ub_cvt_lab(ub_ctx* ctx, int code, const uint8_t* src, int B, int H, int W, uint8_t* dst, void* stream) {
  NEED_CTX(ctx);
  if (!src || !dst) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  RUN(check_frame_shape(ctx, B, H, W));
  return cvt_lab(ctx, code, src, dst, B, H, W, as_stream(stream));
}

int ub_clahe(ub_ctx* ctx, const uint8_t* src, int B, int H, int W, double clip_limit, int tiles_x, int tiles_y, uint8_t* dst, void* stream) {
  NEED_CTX(ctx);
  if (!src || !dst) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  RUN(check_frame_shape(ctx, B, H, W));
  return clahe(ctx, src, dst, B, H, W, 1, 0, clip_limit, tiles_x, tiles_y, as_stream(stream));
}

int ub_nlm(ub_ctx* ctx, const uint8_t* src, int B, int H, int W, int cn, uint8_t* dst, void* stream) {
  NEED_CTX(ctx);
  if (!src || !dst) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  RUN(check_enh_shape(ctx, B, H, W));
  return nlm(ctx, src, dst, B, H, W, cn, 0, cn, as_stream(stream));
}

int ub_sharpen3(ub_ctx* ctx, const uint8_t* src, int B, int H, int W, uint8_t* dst, void* stream) {
  NEED_CTX(ctx);
  if (!src || !dst) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  RUN(check_frame_shape(ctx, B, H, W));
  return sharpen3(ctx, src, dst, B, H, W, as_stream(stream));
}

int ub_enhance(ub_ctx* ctx, const uint8_t* frames, int B, int H, int W, uint8_t* dst, void* stream) {
  NEED_CTX(ctx);
  if (!frames || !dst) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  RUN(check_enh_shape(ctx, B, H, W));
  return enhance_frames(ctx, frames, dst, B, H, W, as_stream(stream));
}

int ub_pipeline(ub_ctx* ctx, const uint8_t* frames, int B, int variant, const ub_roi* roi, const ub_burr_params* params,
                int32_t* frame_stats, uint8_t* masks, void* stream) {
  NEED_CTX(ctx);
  if (!frames || !frame_stats) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  const int H = ctx->frame_h, W = ctx->frame_w, nh = ctx->net_h, nw = ctx->net_w;
  RUN(check_net_shape(ctx, B, nh, nw));
  cudaStream_t st = as_stream(stream);
  const bool do_enhance = (variant & UB_FLAG_ENHANCE_IMAGE) != 0;
  variant &= ~UB_FLAG_ENHANCE_IMAGE;
  if (variant != UB_VARIANT_STANDARD && variant != UB_VARIANT_ENHANCED) return ub_fail(ctx, UB_ERR_ARG, "variant %d", variant);
  if (do_enhance) {
    // the enhanced frames feed BOTH the net and the gray conversion (infer_enhanced_burr.py:279-280);
    // they live in the CCL scratch, which stage 2 only touches after the net has consumed them
    uint8_t* enh = reinterpret_cast<uint8_t*>(ctx->d_comp) + (size_t)ctx->max_batch * H * W * 4;
    RUN(check_enh_shape(ctx, B, H, W));
    RUN(enhance_frames(ctx, frames, enh, B, H, W, st));
    frames = enh;
  }
  ub_roi r;
  if (roi) r = *roi;
  else if (variant == UB_VARIANT_ENHANCED) r = {200, 0, 600, 448};                     // VERTICAL_ROI, infer_enhanced_burr.py:23-28
  else r = {(int)(140 * ((double)W / 512)), 0, (int)(270 * ((double)W / 512)), (int)(512 * ((double)H / 512))};   // :37-47
  ub_burr_params p = params ? *params : default_params(variant);
  const size_t fpix = (size_t)H * W;
  uint8_t *cable = ctx->d_mask[0], *tape = ctx->d_mask[1], *burr = ctx->d_mask[2];
  // The gray frame and the edge map depend on the frames only: they run on a second stream beside the network
  // (fork / join with events; the integer kernels fill the SMs the tensor-core kernels leave idle at their tails).
  const int overlap_env = ctx->overlap;
  cudaStream_t sa = overlap_env ? ctx->aux_stream : st;
  if (overlap_env) {
    UB_CUDA(ctx, cudaEventRecord(ctx->ev_fork, st));
    UB_CUDA(ctx, cudaStreamWaitEvent(sa, ctx->ev_fork, 0));
  }
  RUN(preprocess_frames(ctx, frames, B, H, W, nh, nw, nullptr, nullptr, ctx->d_gray, sa));          // gray only
  // hyst_late: the short-CTA stencils run beside the network, the per-frame hysteresis flood (one long-running CTA per frame) only
  // once the last conv kernel has been issued -- beside k_masks -- so that it never holds an SM a persistent conv kernel needs
  static const int late_env = getenv("UB_HYST_LATE") ? atoi(getenv("UB_HYST_LATE")) : 1;
  const bool late = overlap_env && late_env;
  RUN(burr_edges(ctx, variant, ctx->d_gray, B, H, W, p, sa, late ? 1 : 0));
  if (overlap_env && !late) UB_CUDA(ctx, cudaEventRecord(ctx->ev_join, sa));
  RUN(net_forward_pp(ctx, B, nh, nw, nullptr, frames, H, W, nullptr, ctx->d_pred, st));               // preprocess fused per sub-batch
  if (late) {
    UB_CUDA(ctx, cudaEventRecord(ctx->ev_fork, st));
    UB_CUDA(ctx, cudaStreamWaitEvent(sa, ctx->ev_fork, 0));
    RUN(burr_edges(ctx, variant, ctx->d_gray, B, H, W, p, sa, 2));
    UB_CUDA(ctx, cudaEventRecord(ctx->ev_join, sa));
  }
  RUN(masks_from_pred(ctx, ctx->d_pred, B, nh, nw, H, W, r, cable, tape, frame_stats, st));
  if (overlap_env) UB_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_join, 0));
  RUN(burr_finish(ctx, cable, B, H, W, p, burr, frame_stats + 3, 4, st));
  if (masks) {
    for (int k = 0; k < 3; ++k)
      UB_CUDA(ctx, cudaMemcpy2DAsync(masks + k * fpix, 3 * fpix, ctx->d_mask[k], fpix, fpix, B, cudaMemcpyDeviceToDevice, st));
  }
  return UB_OK;
}

// ub_pipeline as a replayed CUDA graph.  A step is ~35 launches on two streams; at batch 64 the host issues them well
// ahead of the device, at batch 1-4 (the reference loop is batch 1, infer_two_stage_burr.py:292-299) issuing them IS the
// step time.  Graphs are keyed on everything that is baked into the launches (pointers, batch, variant, ROI, parameters,
// the overlap switch).  The first call with a key runs the step eagerly (lazy allocations, one-time table uploads and the
// armed saturation check happen there), the second captures and instantiates, later calls only launch.

int ub_pipeline_graph(ub_ctx* ctx, const uint8_t* frames, int B, int variant, const ub_roi* roi, const ub_burr_params* params,
                      int32_t* frame_stats, uint8_t* masks, void* stream) {
  NEED_CTX(ctx);
  cudaStream_t st = as_stream(stream);
  // the legacy default stream cannot be captured, and event-timed runs want their brackets: both run the plain step
  if (st == nullptr || st == cudaStreamLegacy || ctx->tm.on || ctx->sat_check || ctx->sat_check_once)
    return ub_pipeline(ctx, frames, B, variant, roi, params, frame_stats, masks, stream);
  if (!ctx->graphs) ctx->graphs = new StepGraphs();
  StepGraphs* G = static_cast<StepGraphs*>(ctx->graphs);
  StepKey k;
  memset(&k, 0, sizeof k);
  k.frames = frames; k.stats = frame_stats; k.masks = masks; k.B = B; k.variant = variant; k.overlap = ctx->overlap;
  k.has_roi = roi != nullptr; k.has_params = params != nullptr;
  if (roi) k.roi = *roi;
  if (params) memcpy(&k.params, params, sizeof *params);
  StepGraph* e = nullptr;
  for (auto& c : G->v) if (!memcmp(&c.key, &k, sizeof k)) { e = &c; break; }
  if (!e) {
    if (G->v.size() >= kMaxStepGraphs) {                       // drop the oldest
      if (G->v.front().exec) cudaGraphExecDestroy(G->v.front().exec);
      G->v.erase(G->v.begin());
    }
    G->v.push_back(StepGraph());
    e = &G->v.back();
    e->key = k;
  }
  if (e->calls++ == 0) return ub_pipeline(ctx, frames, B, variant, roi, params, frame_stats, masks, stream);
  if (!e->exec) {
    DeviceGuard guard(ctx->device);
    const int64_t l0 = ctx->launches;
    UB_CUDA(ctx, cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    const int r = ub_pipeline(ctx, frames, B, variant, roi, params, frame_stats, masks, stream);
    cudaGraph_t graph = nullptr;
    const cudaError_t ce = cudaStreamEndCapture(st, &graph);
    e->launches = ctx->launches - l0;
    ctx->launches = l0;                                        // nothing has run yet
    if (r != UB_OK) { if (graph) cudaGraphDestroy(graph); e->calls = 0; return r; }
    if (ce != cudaSuccess) { e->calls = 0; return ub_fail(ctx, UB_ERR_CUDA, "cudaStreamEndCapture: %s", cudaGetErrorString(ce)); }
    const cudaError_t ie = cudaGraphInstantiate(&e->exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ie != cudaSuccess) { e->exec = nullptr; e->calls = 0; return ub_fail(ctx, UB_ERR_CUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(ie)); }
  }
  UB_CUDA(ctx, cudaGraphLaunch(e->exec, st));
  ctx->launches += e->launches;
  return UB_OK;
}

int ub_pipeline_host(ub_ctx* ctx, const uint8_t* frames_host, int B, int variant, const ub_roi* roi,
                     const ub_burr_params* params, int32_t* stats_host, uint8_t* masks_host) {
  NEED_CTX(ctx);
  if (!frames_host || !stats_host) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  if (B < 1 || B > ctx->max_batch) return ub_fail(ctx, UB_ERR_ARG, "batch %d outside [1,%d]", B, ctx->max_batch);
  cudaStream_t st = ctx->own_stream;
  const size_t fbytes = (size_t)ctx->frame_h * ctx->frame_w * 3;
  UB_CUDA(ctx, cudaMemcpyAsync(ctx->d_frames, frames_host, fbytes * B, cudaMemcpyHostToDevice, st));
  RUN(ub_pipeline(ctx, ctx->d_frames, B, variant, roi, params, ctx->d_stats, nullptr, st));
  UB_CUDA(ctx, cudaMemcpyAsync(stats_host, ctx->d_stats, sizeof(int32_t) * 4 * B, cudaMemcpyDeviceToHost, st));
  if (masks_host) {
    const size_t fpix = (size_t)ctx->frame_h * ctx->frame_w;
    for (int k = 0; k < 3; ++k)
      UB_CUDA(ctx, cudaMemcpy2DAsync(masks_host + k * fpix, 3 * fpix, ctx->d_mask[k], fpix, fpix, B, cudaMemcpyDeviceToHost, st));
  }
  UB_CUDA(ctx, cudaStreamSynchronize(st));
  return UB_OK;
}

// Streaming form of ub_pipeline_host.  submit(slot) enqueues: H2D copy of the frames on the copy stream (as soon as the
// slot's previous batch has been consumed), the whole step on the compute stream, D2H copy of the statistics; wait(slot)
// blocks until that batch's statistics are in stats_host.  With two slots in flight (submit i+1, then wait i) the copy of
// the next batch overlaps the compute of the current one -- how a video loop feeds the device.
int ub_pipeline_submit(ub_ctx* ctx, const uint8_t* frames_host, int B, int variant, const ub_roi* roi,
                       const ub_burr_params* params, int32_t* stats_host, int slot) {
  NEED_CTX(ctx);
  if (!frames_host || !stats_host) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  if (slot < 0 || slot > 1) return ub_fail(ctx, UB_ERR_ARG, "slot %d (0 or 1)", slot);
  if (B < 1 || B > ctx->max_batch) return ub_fail(ctx, UB_ERR_ARG, "batch %d outside [1,%d]", B, ctx->max_batch);
  if (ctx->slot_batch[slot]) return ub_fail(ctx, UB_ERR_STATE, "slot %d is still in flight: call ub_pipeline_wait first", slot);
  const size_t fbytes = (size_t)ctx->frame_h * ctx->frame_w * 3;
  if (slot == 1 && !ctx->d_frames2) {
    cudaError_t e = cudaMalloc(&ctx->d_frames2, fbytes * ctx->max_batch);
    if (e != cudaSuccess) return ub_fail(ctx, UB_ERR_NOMEM, "cudaMalloc(%zu bytes): %s", fbytes * ctx->max_batch, cudaGetErrorString(e));
  }
  uint8_t* d_fr = slot ? ctx->d_frames2 : ctx->d_frames;
  int32_t* d_st = ctx->d_stats + (size_t)slot * ctx->max_batch * 4;
  // the slot's frame buffer was last read by its previous batch, whose completion ub_pipeline_wait has observed
  UB_CUDA(ctx, cudaMemcpyAsync(d_fr, frames_host, fbytes * B, cudaMemcpyHostToDevice, ctx->copy_stream));
  UB_CUDA(ctx, cudaEventRecord(ctx->ev_h2d[slot], ctx->copy_stream));
  cudaStream_t st = ctx->own_stream;
  UB_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_h2d[slot], 0));
  RUN(ub_pipeline(ctx, d_fr, B, variant, roi, params, d_st, nullptr, st));
  UB_CUDA(ctx, cudaMemcpyAsync(stats_host, d_st, sizeof(int32_t) * 4 * B, cudaMemcpyDeviceToHost, st));
  UB_CUDA(ctx, cudaEventRecord(ctx->ev_done[slot], st));
  ctx->slot_batch[slot] = B;
  return UB_OK;
}

int ub_pipeline_submit_src(ub_ctx* ctx, const uint8_t* src_host, int B, int src_h, int src_w, int rotate, int variant,
                           const ub_roi* roi, const ub_burr_params* params, int32_t* stats_host, int slot) {
  NEED_CTX(ctx);
  if (!src_host || !stats_host) return ub_fail(ctx, UB_ERR_ARG, "null argument");
  if (slot < 0 || slot > 1) return ub_fail(ctx, UB_ERR_ARG, "slot %d (0 or 1)", slot);
  if (B < 1 || B > ctx->max_batch || src_h < 2 || src_w < 2) return ub_fail(ctx, UB_ERR_ARG, "batch %d / source %dx%d", B, src_h, src_w);
  if (ctx->slot_batch[slot]) return ub_fail(ctx, UB_ERR_STATE, "slot %d is still in flight: call ub_pipeline_wait first", slot);
  const size_t fbytes = (size_t)ctx->frame_h * ctx->frame_w * 3, sbytes = (size_t)src_h * src_w * 3;
  if (sbytes * ctx->max_batch > ctx->src_bytes) {              // (re)allocate both source slots for this source size
    UB_CUDA(ctx, cudaDeviceSynchronize());
    for (int i = 0; i < 2; ++i) { cudaFree(ctx->d_src[i]); ctx->d_src[i] = nullptr; }
    ctx->src_bytes = 0;
    for (int i = 0; i < 2; ++i) {
      cudaError_t e = cudaMalloc(&ctx->d_src[i], sbytes * ctx->max_batch);
      if (e != cudaSuccess) return ub_fail(ctx, UB_ERR_NOMEM, "cudaMalloc(%zu bytes): %s", sbytes * ctx->max_batch, cudaGetErrorString(e));
    }
    ctx->src_bytes = sbytes * ctx->max_batch;
  }
  if (slot == 1 && !ctx->d_frames2) {
    cudaError_t e = cudaMalloc(&ctx->d_frames2, fbytes * ctx->max_batch);
    if (e != cudaSuccess) return ub_fail(ctx, UB_ERR_NOMEM, "cudaMalloc(%zu bytes): %s", fbytes * ctx->max_batch, cudaGetErrorString(e));
  }
  uint8_t* d_fr = slot ? ctx->d_frames2 : ctx->d_frames;
  int32_t* d_st = ctx->d_stats + (size_t)slot * ctx->max_batch * 4;
  UB_CUDA(ctx, cudaMemcpyAsync(ctx->d_src[slot], src_host, sbytes * B, cudaMemcpyHostToDevice, ctx->copy_stream));
  UB_CUDA(ctx, cudaEventRecord(ctx->ev_h2d[slot], ctx->copy_stream));
  cudaStream_t st = ctx->own_stream;
  UB_CUDA(ctx, cudaStreamWaitEvent(st, ctx->ev_h2d[slot], 0));
  RUN(normalize_frames_impl(ctx, ctx->d_src[slot], B, src_h, src_w, rotate, d_fr, ctx->frame_h, ctx->frame_w, st));
  RUN(ub_pipeline(ctx, d_fr, B, variant, roi, params, d_st, nullptr, st));
  UB_CUDA(ctx, cudaMemcpyAsync(stats_host, d_st, sizeof(int32_t) * 4 * B, cudaMemcpyDeviceToHost, st));
  UB_CUDA(ctx, cudaEventRecord(ctx->ev_done[slot], st));
  ctx->slot_batch[slot] = B;
  return UB_OK;
}

int ub_pipeline_wait(ub_ctx* ctx, int slot) {
  NEED_CTX(ctx);
  if (slot < 0 || slot > 1) return ub_fail(ctx, UB_ERR_ARG, "slot %d (0 or 1)", slot);
  if (!ctx->slot_batch[slot]) return ub_fail(ctx, UB_ERR_STATE, "slot %d has nothing in flight", slot);
  UB_CUDA(ctx, cudaEventSynchronize(ctx->ev_done[slot]));
  ctx->slot_batch[slot] = 0;
  return UB_OK;
}

int ub_debug_conv(ub_ctx* ctx, int layer, const float* x0, const float* x1, int B, int H, int W, float* out, void* stream) {
  NEED_CTX(ctx);
  if (layer < 0 || layer > 17 || !x0 || !out) return ub_fail(ctx, UB_ERR_ARG, "bad argument");
  if (!ctx->weights_set) return ub_fail(ctx, UB_ERR_STATE, "ub_set_weights has not been called");
  const LayerDef& d = kLayers[layer];
  const int lvl = layer < 10 ? layer / 2 : (layer < 12 ? 3 : layer < 14 ? 2 : layer < 16 ? 1 : 0);
  if (B > ctx->max_batch || pp_layout(H, W).Ppl > pp_layout(ctx->net_h >> lvl, ctx->net_w >> lvl).Ppl)
    return ub_fail(ctx, UB_ERR_ARG, "shape exceeds the context at pyramid level %d", lvl);
  if ((d.cin1 != 0) != (x1 != nullptr)) return ub_fail(ctx, UB_ERR_ARG, "layer %d takes %d sources", layer, d.cin1 ? 2 : 1);
  cudaStream_t st = as_stream(stream);
  // scratch: buffers of the same level that are large enough: in0 -> x/skip slot, in1 -> up slot, out -> t slot
  static const int in0[5] = {2, 5, 8, 11, 14}, in1[5] = {21, 19, 17, 15, 15}, outb[5] = {1, 4, 7, 10, 13};
  act_t* s0 = layer == 0 ? ctx->act[0] : ctx->act[in0[lvl]];
  RUN(nchw_to_pp(ctx, x0, B, d.real_cin0, H, W, s0, d.cin0, st));
  if (x1) RUN(nchw_to_pp(ctx, x1, B, d.cin1, H, W, ctx->act[in1[lvl]], d.cin1, st));
  ConvArgs a{};
  a.src0 = s0; a.src1 = x1 ? ctx->act[in1[lvl]] : nullptr; a.out = ctx->act[outb[lvl]]; a.B = B; a.H = H; a.W = W;
  a.sat = ctx->sat_check ? ctx->d_sat : nullptr;
  RUN(run_conv_layer(ctx, ctx->conv[layer], a, st));
  return pp_to_nchw(ctx, ctx->act[outb[lvl]], B, d.cout, H, W, out, st);
}

// conv1 of a decoder block from the half-resolution `below` tensor: fused = 1 runs the layer's fused-upsample kernel
// (error if it has none), fused = 0 materialises Upsample(x2) first -- the two must agree bit for bit.
int ub_debug_conv_up(ub_ctx* ctx, int layer, const float* x0, const float* x1_low, int B, int H, int W, int fused, float* out,
                     void* stream) {
  NEED_CTX(ctx);
  if (layer < 10 || layer > 16 || (layer & 1) || !x0 || !x1_low || !out || (H & 1) || (W & 1)) return ub_fail(ctx, UB_ERR_ARG, "bad argument");
  if (!ctx->weights_set) return ub_fail(ctx, UB_ERR_STATE, "ub_set_weights has not been called");
  const LayerDef& d = kLayers[layer];
  const int lvl = layer < 12 ? 3 : layer < 14 ? 2 : layer < 16 ? 1 : 0;
  if (B > ctx->max_batch || pp_layout(H, W).Ppl > pp_layout(ctx->net_h >> lvl, ctx->net_w >> lvl).Ppl)
    return ub_fail(ctx, UB_ERR_ARG, "shape exceeds the context at pyramid level %d", lvl);
  cudaStream_t st = as_stream(stream);
  static const int in0[4] = {2, 5, 8, 11}, upb[4] = {21, 19, 17, 15}, outb[4] = {1, 4, 7, 10}, lowb[4] = {4, 7, 10, 13};
  act_t *s0 = ctx->act[in0[lvl]], *low = ctx->act[lowb[lvl]], *up = ctx->act[upb[lvl]], *o = ctx->act[outb[lvl]];
  RUN(nchw_to_pp(ctx, x0, B, d.cin0, H, W, s0, d.cin0, st));
  RUN(nchw_to_pp(ctx, x1_low, B, d.cin1, H / 2, W / 2, low, d.cin1, st));
  ConvArgs a{};
  a.src0 = s0; a.out = o; a.B = B; a.H = H; a.W = W;
  if (fused) {
    if (!conv_rows_can_upsample(ctx, ctx->conv[layer]) || (H & 3)) return ub_fail(ctx, UB_ERR_ARG, "layer %d has no fused-upsample kernel", layer);
    a.src1 = low; a.up1 = true;
  } else {
    RUN(upsample2_pp(ctx, low, B, d.cin1, H / 2, W / 2, up, st));
    a.src1 = up;
  }
  RUN(run_conv_layer(ctx, ctx->conv[layer], a, st));
  return pp_to_nchw(ctx, o, B, d.cout, H, W, out, st);
}

int ub_debug_activation(ub_ctx* ctx, int index, int B, int C, int H, int W, float* out, void* stream) {
  NEED_CTX(ctx);
  if (index < 0 || index > 21 || !out || C > kActC[index]) return ub_fail(ctx, UB_ERR_ARG, "bad argument");
  return pp_to_nchw(ctx, ctx->act[index], B, C, H, W, out, as_stream(stream));
}

int64_t ub_launch_count(ub_ctx* ctx) { return ctx ? ctx->launches : 0; }

int ub_timing_enable(ub_ctx* ctx, int on) {
  NEED_CTX(ctx);
  ctx->tm.on = on != 0;
  return UB_OK;
}

int ub_timing_read(ub_ctx* ctx, float* ms, int64_t* launches, int reset) {
  NEED_CTX(ctx);
  UB_CUDA(ctx, cudaDeviceSynchronize());
  for (auto& r : ctx->tm.recs) {
    float t = 0.f;
    if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess) { ctx->tm.ms[r.cls] += t; ctx->tm.launches[r.cls] += 1; }
    ctx->tm.pool.push_back(r.a); ctx->tm.pool.push_back(r.b);
  }
  ctx->tm.recs.clear();
  for (int i = 0; i < UB_T_NCLASS; ++i) {
    if (ms) ms[i] = ctx->tm.ms[i];
    if (launches) launches[i] = ctx->tm.launches[i];
    if (reset) { ctx->tm.ms[i] = 0; ctx->tm.launches[i] = 0; }
  }
  return UB_OK;
}

}  // extern "C"
