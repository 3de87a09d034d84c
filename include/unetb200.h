/* unetb200.h -- C ABI of libunetb200.so: the B200 (sm_100a) implementation of the
 * per-frame inspection hot path of Chenxu1103/UNET-.
 *
 * The reference has no FFI / plugin layer of its own (it is pure Python calling
 * torch and OpenCV), so each entry point below names the reference statements it
 * replaces (file:line under /root/reference).  The Python host layer in unet-_b200/
 * binds these with ctypes and re-creates the reference's own contracts on top
 * (NestedUNet module, preprocess_image, detect_burrs_on_cable, ...).
 *
 * Conventions
 *   - every function returns 0 on success or a negative ub_status; no C++ exception
 *     crosses the ABI; ub_last_error() gives the message of the last failure.
 *   - pointers suffixed _dev are device pointers borrowed for the duration of the
 *     enqueued work; pointers suffixed _host are host pointers.
 *   - `stream` is a cudaStream_t passed as void*; work is enqueued asynchronously on
 *     it unless a function says otherwise.  One ctx per (thread, GPU); not thread-safe.
 *   - there is no CPU fallback: ub_create fails on anything that is not sm_100.
 */
#ifndef UNETB200_H
#define UNETB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ub_ctx ub_ctx;

typedef enum ub_status {
  UB_OK = 0,
  UB_ERR_ARG = -1,       /* bad argument / shape */
  UB_ERR_CUDA = -2,      /* CUDA runtime error (message in ub_last_error) */
  UB_ERR_ARCH = -3,      /* device is not sm_100 */
  UB_ERR_STATE = -4,     /* e.g. weights not set */
  UB_ERR_NOMEM = -5
} ub_status;

/* One entry of a state_dict (src/models/unetpp.py:40-91; 134 entries for the 3-class
 * model).  `data_host` is fp32 (int64 for *.num_batches_tracked, which is ignored). */
typedef struct ub_named_tensor {
  const char* name;
  const void* data_host;
  int64_t numel;
} ub_named_tensor;

typedef struct ub_roi { int32_t x1, y1, x2, y2; } ub_roi;   /* infer_two_stage_burr.py:37-47 */

/* Component filter of infer_two_stage_burr.py:103-117 / infer_enhanced_burr.py:121-136. */
typedef struct ub_burr_params {
  int32_t min_area, max_area;   /* keep iff min_area <= area <= max_area            */
  int32_t max_aspect;           /* and max(w,h) <= max_aspect*min(w,h)  (5 or 6)     */
  int32_t min_side;             /* and w >= min_side and h >= min_side  (4 or 5)     */
  int32_t canny_low, canny_high;/* 50/150 standard, 30/100 enhanced                  */
  int32_t band_k, close_k, open_k; /* ellipse sizes: 8/3/2 standard, 25/5/3 enhanced */
} ub_burr_params;

enum { UB_VARIANT_STANDARD = 0,   /* detect_burrs_on_cable, infer_two_stage_burr.py:50-119 */
       UB_VARIANT_ENHANCED = 1,   /* detect_burrs_enhanced, infer_enhanced_burr.py:69-138  */
       /* get_burr_mask_rulebased, src/refactor/burr_detector.py:11-66 (ub_burr only): edges = wrapped |Laplacian| >
        * canny_low (the field holds BurrConfig.laplacian_threshold; canny_high unused), band_k = 2*band_out+1,
        * close_k = open_k = 1 (no morphology), area filter only (max_aspect large, min_side 0); output {0,1} */
       UB_VARIANT_LAPLACIAN = 2,
       /* OR-ed into `variant` of ub_pipeline / ub_pipeline_host: run enhance_image on every frame
        * before the net and the gray conversion (--enhance, infer_enhanced_burr.py:279-280) */
       UB_FLAG_ENHANCE_IMAGE = 0x100 };

/* ---- lifetime ------------------------------------------------------------------- */
/* Allocates all scratch for batches of up to max_batch frames of frame_h x frame_w and
 * a net input of net_h x net_w (multiples of 16).  num_classes <= 8. */
int ub_create(ub_ctx** out, int device, int max_batch, int frame_h, int frame_w,
              int net_h, int net_w, int num_classes);
void ub_destroy(ub_ctx* ctx);
const char* ub_last_error(ub_ctx* ctx);          /* ctx may be NULL */
int ub_version(void);

/* ---- stage 1: the network -------------------------------------------------------- */
/* Replaces model.load_state_dict(...) + .eval() (infer_two_stage_burr.py:214-217):
 * folds eval-mode BatchNorm (eps 1e-5) into each conv, repacks to bf16 K-blocked
 * layout, uploads.  Synchronous. */
int ub_set_weights(ub_ctx* ctx, const ub_named_tensor* tensors, int n);

/* Replaces NestedUNet.forward in eval mode (src/models/unetpp.py:93-135):
 * x fp32 NCHW [B,3,H,W] -> logits fp32 NCHW [B,num_classes,H,W]; H,W multiples of 16,
 * H*W <= net_h*net_w, B <= max_batch. */
int ub_forward(ub_ctx* ctx, const float* x_nchw_dev, int B, int H, int W,
               float* logits_nchw_dev, void* stream);

/* Debug only (like ub_debug_conv): which conv kernel this context runs.  0 = the tcgen05 tensor-core kernels (the product path,
 * always the default), 1 = a plain CUDA-core kernel that exists as the comparator of the kernel-vs-kernel parity tests
 * (tests/test_gpu_net.py), 2 = the tcgen05 kernel with aligned A segments.  Nothing in the library selects 1 on its own: it
 * is not a fallback, and it cannot be switched on from the environment. */
int ub_debug_set_conv_impl(ub_ctx* ctx, int impl);

/* ---- pre / post processing around the net ---------------------------------------- */
/* Replaces preprocess_image (infer_two_stage_burr.py:122-127) and the BGR2GRAY of :317
 * for a batch: frames u8 BGR [B,H,W,3] -> x fp32 NCHW [B,3,net_h,net_w] (RGB, /255;
 * nullable) and gray u8 [B,H,W] (nullable). */
int ub_preprocess(ub_ctx* ctx, const uint8_t* frames_bgr_dev, int B, int H, int W,
                  float* x_nchw_dev, uint8_t* gray_dev, void* stream);

/* Replaces cv2.rotate(ROTATE_90_COUNTERCLOCKWISE) + cv2.resize(INTER_LINEAR)
 * (infer_two_stage_burr.py:275-280): src u8 [B,sh,sw,3] -> dst u8 [B,dh,dw,3]. */
int ub_normalize_frames(ub_ctx* ctx, const uint8_t* src_dev, int B, int sh, int sw,
                        int rotate_ccw, uint8_t* dst_dev, int dh, int dw, void* stream);

/* Replaces softmax/argmax/class masks/NEAREST resize/ROI (infer_two_stage_burr.py:299-314):
 * logits fp32 NCHW [B,C,nh,nw] -> pred u8 [B,nh,nw] (nullable), mask_cable / mask_tape
 * u8 [B,H,W] in {0,1}. */
int ub_masks_from_logits(ub_ctx* ctx, const float* logits_dev, int B, int nh, int nw,
                         int H, int W, const ub_roi* roi, uint8_t* pred_dev,
                         uint8_t* mask_cable_dev, uint8_t* mask_tape_dev, void* stream);

/* Visualisation blend of visualize_two_stage (infer_two_stage_burr.py:134-154; infer_enhanced_burr.py:157-175):
 * the four cv2.addWeighted steps -- dim outside the ROI (0.7/0.3), tint cable (0,255,0) and tape (255,0,0) at
 * 0.6/0.4 and burr (255,0,255) at 0.5/0.5 -- in one pass, bit-exact with OpenCV's binary32 fma + round-half-even.
 * frames / out u8 [B,H,W,3] BGR, masks u8 [B,H,W] (non-zero = set).  Boxes, contours and text stay with the caller. */
int ub_overlay(ub_ctx* ctx, const uint8_t* frames_bgr_dev, const uint8_t* mask_cable_dev, const uint8_t* mask_tape_dev,
               const uint8_t* mask_burr_dev, int B, int H, int W, const ub_roi* roi, uint8_t* out_bgr_dev, void* stream);

/* The contour half of visualize_two_stage (infer_two_stage_burr.py:160-166; infer_enhanced_burr.py:181-187):
 * cv2.findContours(mask, RETR_EXTERNAL, CHAIN_APPROX_SIMPLE) + cv2.drawContours(img, contours, -1, (b,g,r), thickness) in
 * place on frames u8 [B,H,W,3] BGR, mask u8 [B,H,W] (non-zero = set); thickness 2 (cable, tape) or 3 (burr).  The exterior
 * background is flooded per frame in shared memory, every foreground pixel next to it paints the stamps of its border
 * steps (csrc/annotate.cu).  Thickness 2: byte-identical to OpenCV's result.  Thickness 3: byte-identical unless a border
 * pixel of the mask lies within one pixel of the image border, where OpenCV's polygon clipping paints extra pixels; for
 * those frames redo[frame] is set to 1 (0 otherwise) and the caller redraws that frame's contours with cv2 on top -- a
 * superset in the same colour, so the result is exact either way.  redo int32 [B] on the device: required for thickness
 * 3, nullable for 2.  (H+2) x (W/32+2) x 8 bytes must fit 200 KB of shared memory (800 x 448, 1024 x 576 ...). */
int ub_draw_contours(ub_ctx* ctx, uint8_t* frames_bgr_dev, const uint8_t* mask_dev, int B, int H, int W, int b, int g, int r,
                     int thickness, int32_t* redo_dev, void* stream);

/* Paints what the host rasterised -- the ROI box and label (infer_two_stage_burr.py:156-158), Hershey text -- over frames
 * u8 [B,H,W,3] on the device: frames[q] = colour[layer[q] - 1] where layer[q] is 1, 2 or 3.  layer u8 [H,W] shared by the
 * batch (per_frame = 0) or [B,H,W]; bgr9: three BGR triples in HOST memory. */
int ub_paint_layer(ub_ctx* ctx, uint8_t* frames_bgr_dev, const uint8_t* layer_dev, int per_frame, int B, int H, int W,
                   const uint8_t* bgr9_host, void* stream);

/* The stamp ub_draw_contours paints for a unit border step in chain-code direction `code` (0..7: E, NE, N, NW, W, SW, S, SE)
 * or for a single point (code 8): thickness 2 -> 5x5, bit (dy+2)*5 + dx+2; thickness 3 -> 7x7, bit (dy+3)*7 + dx+3; 0 for
 * anything else.  Lets the CPU tests compare the tables with cv2.line of the installed OpenCV.  No context, no device. */
unsigned long long ub_contour_stamp(int thickness, int code);

/* compute_diameter_metrics (src/utils/geometry_enhanced.py:113-187): pred u8 [B,H,W] class ids -> out int32 [B][8] =
 * { dc_px, dt_px (binary32 bit patterns: medians of the Gaussian-smoothed row widths of the largest cable / tape
 * component over the rows where both are > 0; 0 when valid_rows < min_valid_rows), valid_rows, cable_area, tape_area
 * (pixels of the kept components: coverage numerators), 0, 0, 0 }.  min_area: the reference passes 50 (:143-144).
 * 16 <= H <= 2048, W >= 16.  The mm values and the differences are scalar host arithmetic (ops.compute_diameter_metrics). */
int ub_diameter_metrics(ub_ctx* ctx, const uint8_t* pred_dev, int B, int H, int W, int cable_cls, int tape_cls,
                        int min_area, int min_valid_rows, int32_t* out_dev, void* stream);

/* analyze_defects (src/utils/geometry_enhanced.py:236-330): pred u8 [B,H,W] class ids (< 8) -> out int32 [B][16] =
 * { tape_px, tape_components, tape_largest_area, holes, hole_px, cable_components, 0, 0, pixel count of class 0..7 } where
 * holes are the 8-connected components of CLOSE(tape, ELLIPSE 5x5) - tape with area >= hole_min_size.  The ratios
 * (coverage, hole ratio, largest-area ratio) are scalar host arithmetic (ops.analyze_defects). */
int ub_analyze_defects(ub_ctx* ctx, const uint8_t* pred_dev, int B, int H, int W, int cable_cls, int tape_cls,
                       int hole_min_size, int32_t* out_dev, void* stream);

/* _compute_width_per_row(mask, smooth=True, kernel_size=31) (src/utils/geometry_enhanced.py:45-76) for the cable and the
 * tape class of pred u8 [B,H,W], as compute_thickness_profile uses it (:208-212): out float [B][2][H] = per-row extent
 * max_x - min_x + 1 of every pixel of the class (0 for empty rows) after cv2's float32 GaussianBlur((1,31), 0). */
int ub_width_profiles(ub_ctx* ctx, const uint8_t* pred_dev, int B, int H, int W, int cable_cls, int tape_cls,
                      float* out_dev, void* stream);

/* filter_cable_by_shape (src/refactor/postprocess.py:12-77): of the 8-connected components of mask (non-zero = set) with
 * area >= min_area, max(w,h)/(min(w,h)+1e-6) >= min_aspect and |centroid_x - roi_width/2| / roi_width <= max_center_offset,
 * keep the one with the largest area*aspect*(1-offset) (float64, first label on ties) -> out u8 {0,255}; all zero if none. */
int ub_filter_cable_by_shape(ub_ctx* ctx, const uint8_t* mask_dev, int B, int H, int W, int min_area, double min_aspect,
                             double max_center_offset, int roi_width, uint8_t* out_dev, void* stream);
/* constrain_tape_to_ring (src/refactor/postprocess.py:80-118): ring = dilate(cable, ELLIPSE(dilate_k)) - erode(cable,
 * ELLIPSE(erode_k)); the largest component of tape & ring -> out u8 {0,255}.  Binary masks ({0,1} or {0,255}). */
int ub_constrain_tape_to_ring(ub_ctx* ctx, const uint8_t* tape_dev, const uint8_t* cable_dev, int B, int H, int W,
                              int dilate_k, int erode_k, uint8_t* out_dev, void* stream);

/* ---- stage 2: individual operators (each = one cv2 call of the reference) --------- */
int ub_gaussian5(ub_ctx* ctx, const uint8_t* src_dev, int B, int H, int W, uint8_t* dst_dev, void* stream);      /* :85  */
int ub_canny(ub_ctx* ctx, const uint8_t* src_dev, int B, int H, int W, int low, int high, uint8_t* dst_dev, void* stream); /* :86 */
/* GaussianBlur((5,5),1.0) followed by Canny(low, high) in one fused pass (:85-86; infer_enhanced_burr.py:89-90). */
int ub_blur_canny(ub_ctx* ctx, const uint8_t* gray_dev, int B, int H, int W, int low, int high, uint8_t* dst_dev, void* stream);
/* op: 0 dilate, 1 erode, 2 close, 3 open with cv2 MORPH_ELLIPSE(k,k), k <= 25, on binary images:
 * any non-zero input is foreground, the output is {0, value}. */
int ub_morph(ub_ctx* ctx, int op, int k, const uint8_t* src_dev, int B, int H, int W, int value, uint8_t* dst_dev, void* stream); /* :78-80,:92-97 */
/* cv2.connectedComponentsWithStats(conn 8) (:100): labels i32 [B,H,W] in OpenCV order,
 * stats i32 [B,max_labels,5] (left,top,width,height,area; row 0 = background),
 * n_labels i32 [B] (includes background).  Components beyond max_labels are labelled
 * but their stats rows are dropped. */
int ub_ccl(ub_ctx* ctx, const uint8_t* src_dev, int B, int H, int W, int32_t* labels_dev,
           int32_t* stats_dev, int max_labels, int32_t* n_labels_dev, void* stream);
/* Sobel-magnitude / Laplacian edge maps of infer_enhanced_burr.py:93-102 -> {0,255}. */
int ub_sobel_edges(ub_ctx* ctx, const uint8_t* gray_dev, int B, int H, int W, int thresh, uint8_t* dst_dev, void* stream);
int ub_laplacian_edges(ub_ctx* ctx, const uint8_t* gray_dev, int B, int H, int W, int thresh, uint8_t* dst_dev, void* stream);

/* ub_sobel_edges without a context and without any allocation, for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  This entry point has no counterpart in the reference.
 * gray u8 [B,H,W] and dst u8 [B,H,W] are device pointers on the
 * CURRENT device; the result equals ub_sobel_edges' bit for bit.  The kernel pair's scratch (one int per frame: the
 * maximum squared gradient) is a static device array of UB_STATIC_MAX_FRAMES entries (16 KB per GPU, held until the
 * library is unloaded); B > UB_STATIC_MAX_FRAMES returns UB_ERR_ARG (message via ub_last_error(NULL)).  The call
 * performs no cudaMalloc / cudaFree; the array is cleared on `stream` before use.  All calls share the array: calls that
 * overlap on the GPU (different streams, no ordering) race -- serialise them. */
#define UB_STATIC_MAX_FRAMES 4096
int ub_sobel_edges_static(const uint8_t* gray_dev, int B, int H, int W, int thresh, uint8_t* dst_dev, void* stream);

/* Whole detect_burrs_on_cable / detect_burrs_enhanced for a batch:
 * gray u8 [B,H,W], mask_cable u8 [B,H,W] in {0,1} -> burr_mask u8 [B,H,W] in {0,1},
 * burr_px i32 [B] (nullable). */
int ub_burr(ub_ctx* ctx, int variant, const uint8_t* gray_dev, const uint8_t* mask_cable_dev,
            int B, int H, int W, const ub_burr_params* params, uint8_t* burr_mask_dev,
            int32_t* burr_px_dev, void* stream);

/* ---- enhance_image (infer_enhanced_burr.py:38-66; --enhance only) -------------------- */
/* The individual cv2 calls, bit-exact for uint8.  Images are u8 [B,H,W,3] on the device.
 * code: 0 COLOR_BGR2LAB (:46), 1 COLOR_LBGR2Lab, 2 COLOR_LAB2BGR (:55), 3 COLOR_Lab2LBGR
 * (1 and 3 are what fastNlMeansDenoisingColored uses internally).  src may equal dst. */
int ub_cvt_lab(ub_ctx* ctx, int code, const uint8_t* src_dev, int B, int H, int W, uint8_t* dst_dev, void* stream);
/* cv2.createCLAHE(clip_limit, (tiles_x,tiles_y)).apply on a single-channel u8 [B,H,W] (:50-51);
 * W % tiles_x == 0, H % tiles_y == 0, tiles_x*tiles_y <= 64 (800x448 with 8x8 qualifies; OpenCV
 * pads other sizes, this entry point rejects them).  src may equal dst. */
int ub_clahe(ub_ctx* ctx, const uint8_t* src_dev, int B, int H, int W, double clip_limit, int tiles_x, int tiles_y,
             uint8_t* dst_dev, void* stream);
/* cv2.fastNlMeansDenoising(src, None, 10, 7, 21) on u8 [B,H,W,cn], cn = 1 or 2; H,W >= 14. */
int ub_nlm(ub_ctx* ctx, const uint8_t* src_dev, int B, int H, int W, int cn, uint8_t* dst_dev, void* stream);
/* cv2.filter2D(src, -1, [[-1,-1,-1],[-1,9,-1],[-1,-1,-1]]) on u8 [B,H,W,3] (:61-64). */
int ub_sharpen3(ub_ctx* ctx, const uint8_t* src_dev, int B, int H, int W, uint8_t* dst_dev, void* stream);
/* Whole enhance_image for a batch: frames u8 BGR [B,H,W,3] -> u8 BGR [B,H,W,3]
 * (BGR2LAB, CLAHE(3.0, 8x8) on L, LAB2BGR, fastNlMeansDenoisingColored(10,10,7,21), sharpen). */
int ub_enhance(ub_ctx* ctx, const uint8_t* frames_bgr_dev, int B, int H, int W, uint8_t* dst_dev, void* stream);

/* ---- the whole per-frame step (body of the loop, infer_two_stage_burr.py:286-343) -- */
/* frames u8 BGR [B,frame_h,frame_w,3] -> frame_stats i32 [B,4] = (index in batch,
 * cable_px, tape_px, burr_px); masks (nullable) u8 [B,3,frame_h,frame_w] = cable, tape,
 * burr.  roi == NULL -> map_roi_to_original of the frame size. */
int ub_pipeline(ub_ctx* ctx, const uint8_t* frames_bgr_dev, int B, int variant,
                const ub_roi* roi, const ub_burr_params* params,
                int32_t* frame_stats_dev, uint8_t* masks_dev, void* stream);

/* ub_pipeline replayed as a CUDA graph, for small batches where issuing the ~35 launches of a step from the host is the
 * step time (the reference loop is batch 1, infer_two_stage_burr.py:292-299).  Same arguments and results as ub_pipeline.
 * Graphs are cached per (pointers, B, variant, roi, params): the first call with a combination runs the plain step, the
 * second captures it, later calls launch the captured graph.  `stream` must not be the legacy default stream (that, and
 * runs with event timing or the saturation check on, take the plain path).  Pointers must stay valid while reused. */
int ub_pipeline_graph(ub_ctx* ctx, const uint8_t* frames_bgr_dev, int B, int variant,
                      const ub_roi* roi, const ub_burr_params* params,
                      int32_t* frame_stats_dev, uint8_t* masks_dev, void* stream);

/* Same, from HOST buffers (pinned for full speed): copies frames host->device, runs the
 * step, copies the stats (and masks if non-NULL) back, and waits for completion. */
int ub_pipeline_host(ub_ctx* ctx, const uint8_t* frames_bgr_host, int B, int variant,
                     const ub_roi* roi, const ub_burr_params* params,
                     int32_t* frame_stats_host, uint8_t* masks_host);

/* Streaming form for a loop over batches (the while-loop of infer_two_stage_burr.py:268-361 with the frames
 * arriving from the host): ub_pipeline_submit enqueues the host->device copy, the step and the copy of the
 * statistics into stats_host for `slot` (0 or 1) and returns; ub_pipeline_wait blocks until that slot's
 * statistics have arrived.  Submitting batch i+1 before waiting for batch i overlaps its copy with the
 * compute of batch i.  frames_host / stats_host must stay valid (and should be pinned) until the wait. */
int ub_pipeline_submit(ub_ctx* ctx, const uint8_t* frames_bgr_host, int B, int variant,
                       const ub_roi* roi, const ub_burr_params* params, int32_t* frame_stats_host, int slot);
int ub_pipeline_wait(ub_ctx* ctx, int slot);

/* ub_pipeline_submit for a loop whose frames arrive at the SOURCE resolution (infer_two_stage_burr.py:270-280,
 * infer_enhanced_burr.py:273-276): src u8 [B,src_h,src_w,3] on the host is copied to the device on the copy stream,
 * rotated 90 degrees counter-clockwise (rotate != 0) and resized to the context's frame size on the device, then the step
 * runs as in ub_pipeline_submit.  Wait with ub_pipeline_wait(slot). */
int ub_pipeline_submit_src(ub_ctx* ctx, const uint8_t* src_bgr_host, int B, int src_h, int src_w, int rotate, int variant,
                           const ub_roi* roi, const ub_burr_params* params, int32_t* frame_stats_host, int slot);

/* ---- run-time switches and counters ------------------------------------------------- */
/* name = "overlap" (1: gray / edge map on a second stream beside the network, default), "fuse_block0" (1: the first
 * encoder block -- both convs and the pool, unetpp.py:23-26,105 -- runs as one kernel, default; 0: three launches, for A/B
 * and parity checks), "check_saturation" (1: every conv
 * epilogue counts the outputs it clamps to the largest finite 16-bit value; 0: off.  ub_set_weights arms the check for the
 * first forward after it, so a checkpoint loaded through infer_two_stage_burr.py:214-217 whose activations overflow the
 * operand type is reported once at no steady-state cost). */
int ub_set_option(ub_ctx* ctx, const char* name, int value);
/* name = "saturated": outputs clamped so far while the check was on (synchronises the device); reset != 0 zeroes it. */
int ub_get_stat(ub_ctx* ctx, const char* name, int64_t* value, int reset);

/* ---- test hooks (layer-level parity checks; not used by the product path) ------------ */
/* Runs conv layer `layer` (0..17 in network order; BN folded, ReLU) on NCHW fp32 inputs:
 * x0 [B,cin0,H,W] (layer 0: 3 channels), x1 [B,cin1,H,W] or NULL -> out [B,cout,H,W]. */
int ub_debug_conv(ub_ctx* ctx, int layer, const float* x0_dev, const float* x1_dev, int B, int H, int W,
                  float* out_dev, void* stream);
/* conv1 of a decoder block (layer 10, 12, 14 or 16) on cat([x0, Upsample(x2, bilinear, align_corners)(x1_low)])
 * (src/models/unetpp.py:76,111-116): x0 [B,cin0,H,W], x1_low [B,cin1,H/2,W/2] -> out [B,cout,H,W].  fused != 0
 * interpolates inside the conv kernel (layers 14 and 16; UB_ERR_ARG for the others), fused == 0 materialises the
 * upsampled tensor first; the two results are bit-identical. */
int ub_debug_conv_up(ub_ctx* ctx, int layer, const float* x0_dev, const float* x1_low_dev, int B, int H, int W,
                     int fused, float* out_dev, void* stream);
/* Copies activation buffer `index` of the last forward (order: x_in,t0,x0_0,p0,t1,x1_0,p1,t2,
 * x2_0,p2,t3,x3_0,p3,t4,x4_0,up4,x3_1,up3,x2_2,up2,x1_3,up1) to NCHW fp32 [B,C,H,W]. */
int ub_debug_activation(ub_ctx* ctx, int index, int B, int C, int H, int W, float* out_dev, void* stream);

/* ---- instrumentation -------------------------------------------------------------- */
/* When enabled, every kernel class is bracketed by CUDA events on the launch stream;
 * ub_timing_read sums them (synchronises).  classes: 0 conv3x3, 1 net glue (pool,
 * upsample, layout), 2 pre/post, 3 stage-2 stencil/morphology, 4 CCL/hysteresis, 5 enhance_image. */
enum { UB_T_CONV = 0, UB_T_GLUE = 1, UB_T_PREPOST = 2, UB_T_STENCIL = 3, UB_T_CCL = 4, UB_T_ENHANCE = 5, UB_T_NCLASS = 6 };
int ub_timing_enable(ub_ctx* ctx, int on);
int ub_timing_read(ub_ctx* ctx, float* ms_per_class, int64_t* scopes_per_class, int reset);
/* Number of kernels this context has launched so far. */
int64_t ub_launch_count(ub_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* UNETB200_H */
