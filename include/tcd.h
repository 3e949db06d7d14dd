/* libtcd -- C ABI of the B200-native per-frame detection hot path.
 *
 * The reference (filthyrake/telescope_cam_detection) is 100 % Python and has no FFI; its
 * boundary for this path is the Python class surface.  The package
 * telescope_cam_detection_b200 re-creates those classes and marshals into the entry points
 * below through ctypes.  Every entry point names the reference code it replaces.
 *
 * Conventions: extern "C", opaque handles, plain pointers and sizes, int status
 * (0 = TCD_OK), a thread-local last-error string, caller-provided CUDA stream (void*,
 * NULL = the handle's own stream).  No hidden allocation after create.  There is no CPU
 * fallback: every compute entry fails with TCD_ERR_CUDA when no sm_100 device is present.
 */
#ifndef TCD_H_
#define TCD_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  TCD_OK = 0,
  TCD_ERR_INVALID = 1,
  TCD_ERR_CUDA = 2,
  TCD_ERR_NOMEM = 3,
  TCD_ERR_UNSUPPORTED = 4
} tcd_status;

const char* tcd_last_error(void);
int tcd_version(void);
/* number of kernels this library has launched in this process (bench.py gpu_launches) */
uint64_t tcd_launch_count(void);
/* bytes of camera frames this library has copied host->device in this process (bench.py e2e.h2d_bytes_per_step);
 * detect/submit upload only the source rows the resize taps with non-zero weight (see tcd_detector_submit) */
uint64_t tcd_h2d_byte_count(void);

/* ------------------------------------------------------------------ preprocess (K1+K2)
 * Replaces YOLOXDetector.preprocess, src/yolox_detector.py:186-220 (HWC u8 -> float,
 * bilinear stretch, align_corners=False, no antialias) and, with other settings,
 * SpeciesClassifier.preprocess src/species_classifier.py:298-352.                      */
typedef struct {
  int32_t out_h, out_w;   /* network input size (640, 640)                                */
  int32_t mode;           /* 0 = stretch (what the reference code does), 1 = letterbox    */
  int32_t swap_rb;        /* 0 = keep BGR (reference), 1 = BGR->RGB                       */
  float scale;            /* multiplied in before mean/std (1.0 reference; 1/255 typical) */
  float mean[3];          /* subtracted after scale, per OUTPUT channel                   */
  float inv_std[3];       /* multiplied after the subtraction                             */
  float pad_value;        /* letterbox fill (114)                                         */
} tcd_preproc_cfg;

/* frames[b]: HWC u8 with row stride row_stride[b] bytes, on the device.
 * out_layout 0: f32 NCHW [B,3,out_h,out_w]                (parity layout, = reference)
 * out_layout 1: bf16 space-to-depth NHWC [B,out_h/2,out_w/2,16] (Focus stem input:
 *               channels = TL,BL,TR,BR patches x 3, 4 zero pad channels)                 */
tcd_status tcd_preprocess(const tcd_preproc_cfg* cfg, const uint8_t* const* frames,
                          const int32_t* h, const int32_t* w, const int64_t* row_stride, int32_t B,
                          int32_t out_layout, void* out_dev, void* stream);

/* ------------------------------------------------------------------ conv engine (K3)
 * One BaseConv / Conv2d of the detector: implicit GEMM on tcgen05 with TMA-fed operands,
 * folded-BN bias + optional SiLU + optional residual in the epilogue.  Exposed for the
 * per-layer parity tests; the detector below owns the full layer plan.
 * x: bf16 NHWC [N,H,W,x_ctot] (device), reads channels [x_coff, x_coff+cin).
 * w: f32 [cout,cin,k,k] and bias f32 [cout] on the HOST (packed per call: test entry).
 * y: bf16 NHWC [N,Ho,Wo,y_ctot] (device), writes channels [y_coff, y_coff+cout).
 * resid (optional, device): bf16 NHWC [N,Ho,Wo,r_ctot], channels [r_coff, r_coff+cout).  */
tcd_status tcd_conv2d_nhwc_bf16(const void* x, int32_t N, int32_t H, int32_t W, int32_t x_ctot,
                                int32_t x_coff, int32_t cin, const float* w, const float* bias,
                                int32_t cout, int32_t k, int32_t stride, int32_t silu,
                                const void* resid, int32_t r_ctot, int32_t r_coff, void* y,
                                int32_t y_ctot, int32_t y_coff, void* stream);

/* The same with `groups` groups: cout / groups output channels per group, group g reading the cin input channels
 * at x_coff + g * cin; w is [cout][cin][k][k].  (The second 3x3 convs of the YOLOX head's cls and reg branches run
 * as ONE such launch over the [cls | reg] buffer.)                                         */
tcd_status tcd_conv2d_nhwc_bf16_grouped(const void* x, int32_t N, int32_t H, int32_t W, int32_t x_ctot,
                                        int32_t x_coff, int32_t cin, const float* w, const float* bias,
                                        int32_t cout, int32_t k, int32_t stride, int32_t silu,
                                        const void* resid, int32_t r_ctot, int32_t r_coff, void* y,
                                        int32_t y_ctot, int32_t y_coff, int32_t groups, void* stream);

/* ------------------------------------------------------------------ detector (A1-A5)
 * Replaces YOLOXDetector.load_model / detect / detect_batch, src/yolox_detector.py:84-184,
 * 284-406 together with upstream YOLOX-S forward and yolox.utils.postprocess.            */
typedef struct tcd_detector tcd_detector;

typedef struct {
  int32_t input_h, input_w;   /* multiples of 32; 640x640 for BASELINE                    */
  int32_t max_batch;          /* frames per detect_batch call the arena is sized for      */
  int32_t num_classes;        /* 80                                                       */
  int32_t max_out;            /* rows per frame in the result table                       */
  int32_t max_src_h, max_src_w; /* expected camera frame: initial size of the host-frame staging
                                   buffers; a larger frame grows its lane's buffer on arrival      */
  tcd_preproc_cfg pre;        /* out_h/out_w ignored (taken from input_h/input_w)         */
} tcd_detector_cfg;

/* packed_weights: blob written by telescope_cam_detection_b200.weights.pack_yolox()
 * (BN already folded, fp32, canonical layer order; layout documented in DESIGN.md).      */
tcd_status tcd_create_detector(const tcd_detector_cfg* cfg, const void* packed_weights,
                               size_t packed_bytes, int32_t device, tcd_detector** out);
void tcd_destroy_detector(tcd_detector* det);

/* conf / nms thresholds of yolox.utils.postprocess (src/yolox_detector.py:309-314);
 * cheap, lock-free for the caller, never re-captures a CUDA graph.                       */
tcd_status tcd_detector_set_thresholds(tcd_detector* det, float conf_thre, float nms_thre);

/* Per-camera class filter, evaluated on the device on the rows that survive NMS, in score order -- the
 * wildlife mask of YOLOXDetector._format_model_output_to_detections (src/yolox_detector.py:258,
 * src/coco_constants.py:23-29) followed by InferenceEngine._post_process_detections
 * (src/inference_engine_yolox.py:505-534): ensure_valid_bbox on the box un-scaled to camera pixels
 * (ratio = orig / network size, fp32 product like the reference's np.float32 * float), per-class confidence
 * (the override, else the global threshold), global minimum area, per-class min / max area, stop at max_det.
 * All comparisons in double, like the Python floats they replace: the rows kept are exactly the dicts the
 * reference's loop keeps.  Tables are per detector; table 0 is "no filter" and cannot be set.            */
#define TCD_MAX_CLASSES 128
#define TCD_MAX_FILTER_TABLES 64
typedef struct {
  double cls_conf[TCD_MAX_CLASSES];      /* confidence < cls_conf[c]  -> dropped                          */
  double cls_min_area[TCD_MAX_CLASSES];  /* area < cls_min_area[c]    -> dropped (-inf: no constraint)    */
  double cls_max_area[TCD_MAX_CLASSES];  /* area > cls_max_area[c]    -> dropped (+inf: no constraint)    */
  double min_area;                       /* global min_box_area; <= 0: off                                */
  uint64_t class_mask[2];                /* bit c set: class c allowed (wildlife_only: {0,14,15,16,21})   */
  int32_t max_det;                       /* keep the first max_det survivors                              */
  int32_t reserved;
} tcd_class_filter;
tcd_status tcd_detector_set_class_filter(tcd_detector* det, int32_t table, const tcd_class_filter* f);
/* Parity tap: the filter of `table` applied on the device to caller-supplied rows (HOST f32 [n,7], network pixels,
 * score order) of a frame of orig_h x orig_w; keep_host[i] = 1 for the rows that survive (max_det included).     */
tcd_status tcd_detector_filter_rows(tcd_detector* det, int32_t table, const float* rows_host, int32_t n,
                                    int32_t orig_h, int32_t orig_w, int32_t* keep_host);

/* frames[b]: HWC u8 BGR.  frames_on_device = 0: host pointers (pinned or pageable), the
 * H2D copy is part of the call; 1: device pointers.  out_dets: HOST [B,max_out,7] f32 rows
 * [x1,y1,x2,y2,obj_conf,class_conf,class_id] in network-input pixel units, score-descending
 * exactly like postprocess(); out_counts: HOST [B] int32 (rows kept, clamped to max_out).
 * Synchronous on return (results are in host memory).                                    */
tcd_status tcd_detect_batch(tcd_detector* det, const uint8_t* const* frames, const int32_t* h,
                            const int32_t* w, const int64_t* row_stride, int32_t B,
                            int32_t frames_on_device, float* out_dets, int32_t* out_counts);

/* Split-phase variant used by the pipelined coordinator / bench: enqueue as pipeline slot
 * `slot` (0 .. tcd_detector_num_slots()-1), then wait.  A slot is (engine, lane): engine
 * slot / n_lanes owns the stream, the activation arena and the CUDA graphs, lane slot % n_lanes owns
 * the frame staging and the result table of ONE in-flight batch (n_engines = 4, env TCD_ENGINES 1..8;
 * n_lanes = 4, env TCD_LANES 1..4).  Dynamic batching: a submit uploads and preprocesses its frames at
 * once, the network + decode/NMS launch of an engine is deferred until its lanes are full or one of
 * them is waited on, and then runs ONCE over all the frames its lanes hold (up to n_lanes x max_batch)
 * -- under load the per-launch latencies are paid per group of batches, a lone batch runs at its own
 * size.  Frames of a lane upload on the engine's copy stream while the engine computes.              */
tcd_status tcd_detect_batch_async(tcd_detector* det, int32_t slot, const uint8_t* const* frames,
                                  const int32_t* h, const int32_t* w, const int64_t* row_stride,
                                  int32_t B, int32_t frames_on_device);
/* The same with a class-filter table per frame (filter_table[b] in [0, TCD_MAX_FILTER_TABLES); NULL = all 0):
 * frames of different cameras, each with its own thresholds, share one batch as they do in
 * SharedInferenceCoordinator._process_batch (src/shared_inference_coordinator.py:226-288).               */
tcd_status tcd_detector_submit(tcd_detector* det, int32_t slot, const uint8_t* const* frames,
                               const int32_t* h, const int32_t* w, const int64_t* row_stride,
                               int32_t B, int32_t frames_on_device, const int32_t* filter_table);
tcd_status tcd_detect_batch_wait(tcd_detector* det, int32_t slot, float* out_dets,
                                 int32_t* out_counts);

/* Parity taps: run preprocess+network only and copy the decoded head output
 * [B,n_anchors,5+num_classes] f32 (what upstream YOLOX.forward returns in eval) to HOST. */
tcd_status tcd_forward(tcd_detector* det, const uint8_t* const* frames, const int32_t* h,
                       const int32_t* w, const int64_t* row_stride, int32_t B,
                       int32_t frames_on_device, float* out_pred_host);
/* postprocess only, on a caller-supplied decoded prediction (HOST f32 [B,A,5+C]).        */
tcd_status tcd_postprocess(tcd_detector* det, const float* pred_host, int32_t B, float* out_dets,
                           int32_t* out_counts);
int32_t tcd_detector_num_anchors(const tcd_detector* det);
/* per-phase device times of the last synchronous detect_batch, ms:
 * [0]=h2d [1]=preprocess [2]=network [3]=decode+nms [4]=d2h                              */
tcd_status tcd_detector_last_timing(const tcd_detector* det, float out_ms[5]);
/* CUDA stream (cudaStream_t) of the engine that pipeline slot `slot` runs on -- lets a caller record
 * its own events around async calls (bench.py times with events on THIS stream).           */
void* tcd_detector_stream(tcd_detector* det, int32_t slot);
int32_t tcd_detector_num_slots(const tcd_detector* det);
/* Device copies of a slot's result table ([B,max_out,7] f32 rows, score-descending) and row counts ([B] i32), valid from
 * tcd_detect_batch_wait(slot) until the next submit on that slot.  Lets a multi-GPU host forward the table to the
 * serving rank (NCCL gather) without a host round trip; no counterpart in the single-GPU reference.                  */
tcd_status tcd_detector_result_device(tcd_detector* det, int32_t slot, float** out_dets, int32_t** out_counts);
/* Measurement aid: in-kernel timeline (clock64 / globaltimer stamps of the first and last CTA)
 * of conv op `idx` of the batch-B layer plan; see conv_umma.cuh for the slot meaning (2x16 milestones + 4x16 per-tile stamps of CTA 0).   */
tcd_status tcd_detector_trace_op(tcd_detector* det, int32_t B, int32_t idx, int32_t reps,
                                 uint64_t out128[128], int32_t* grid, int32_t* smem_bytes);
/* Measurement aid: one eager (graph-less) pass per iteration with a CUDA-event pair around every
 * kernel.  Ops in launch order: 0 = preprocess, 1..n_layers = layer plan (tcgen05 convs,
 * upsample, SPP pool), then decode_filter, sort, nms_mask, nms_reduce.  op_ms[i] = mean ms.  */
tcd_status tcd_detector_profile(tcd_detector* det, const uint8_t* const* frames, const int32_t* h,
                                const int32_t* w, const int64_t* row_stride, int32_t B,
                                int32_t frames_on_device, int32_t iters, float* op_ms, int32_t cap,
                                int32_t* n_ops);
const char* tcd_detector_op_name(tcd_detector* det, int32_t B, int32_t idx);
int32_t tcd_detector_op_kind(tcd_detector* det, int32_t B, int32_t idx); /* 0 conv, 1 other */
double tcd_detector_op_flops(tcd_detector* det, int32_t B, int32_t idx); /* 2*MACs, convs only */
/* debug tap: copy activation buffer `name` (bf16 NHWC) of the last forward to HOST as f32 */
tcd_status tcd_detector_read_activation(tcd_detector* det, const char* name, float* out_host,
                                        size_t out_elems, int32_t* shape4);

/* ------------------------------------------------------------------ motion filter (K8+K9)
 * Replaces cv2.createBackgroundSubtractorMOG2(...).apply + threshold(200) + GaussianBlur
 * (21x21) + threshold(25) + countNonZero(roi) in MotionFilter.has_motion_in_bbox,
 * src/motion_filter.py:56-60,68-127.                                                     */
typedef struct tcd_mog2 tcd_mog2;
tcd_status tcd_mog2_create(int32_t h, int32_t w, int32_t history, float var_threshold,
                           int32_t detect_shadows, int32_t device, tcd_mog2** out);
void tcd_mog2_destroy(tcd_mog2* m);
tcd_status tcd_mog2_reset(tcd_mog2* m);
/* One background-model update.  frame: HWC u8 BGR (host or device).  `repeats` >= 1 applies
 * the same frame that many times (the reference calls apply() once per detection,
 * src/motion_filter.py:88,161-162); the mask of the LAST apply is kept on the device.
 * fg_mask_host (optional): raw MOG2 mask {0,127,255}.                                     */
tcd_status tcd_mog2_apply(tcd_mog2* m, const uint8_t* frame, int64_t row_stride,
                          int32_t frame_on_device, int32_t repeats, uint8_t* fg_mask_host);
/* Mask clean-up on the last mask: (mask > shadow_thr) -> blur ksize x ksize (OpenCV u8
 * fixed-point, REFLECT_101) -> (> blur_thr), then countNonZero inside each clamped integer
 * ROI (x1,y1,x2,y2).  ksize == 0 skips the blur (motion_blur_size <= 0).
 * clean_mask_host (optional): the cleaned 0/255 mask.                                    */
tcd_status tcd_mog2_roi_counts(tcd_mog2* m, int32_t shadow_thr, int32_t ksize, int32_t blur_thr,
                               const int32_t* rois, int32_t n_rois, int32_t* counts_host,
                               uint8_t* clean_mask_host);

/* device time (CUDA events on the handle's stream) of the last call's kernels, ms:
 * [0] one MOG2 apply, [1] mask clean-up (threshold + blur + threshold) + ROI counts       */
tcd_status tcd_mog2_last_ms(const tcd_mog2* m, float out_ms[2]);

/* ------------------------------------------------------------------ empty-frame gate (K10)
 * Replaces EmptyFrameFilter.has_motion, src/empty_frame_filter.py:51-98.                 */
typedef struct tcd_empty_filter tcd_empty_filter;
tcd_status tcd_empty_filter_create(int32_t h, int32_t w, int32_t blur_size, int32_t device,
                                   tcd_empty_filter** out);
void tcd_empty_filter_destroy(tcd_empty_filter* f);
tcd_status tcd_empty_filter_reset(tcd_empty_filter* f);
/* frame: HWC u8 BGR (channels = 3) or an already-gray plane (channels = 1, the reference
 * accepts both, :65-68).  changed_pixels_host: count of |cur-prev| > threshold (strict), or -1
 * on the first frame after create/reset.  gray_blur_host (optional): the blurred gray plane.   */
tcd_status tcd_empty_filter_step(tcd_empty_filter* f, const uint8_t* frame, int64_t row_stride,
                                 int32_t channels, int32_t frame_on_device, int32_t threshold,
                                 int32_t* changed_pixels_host, uint8_t* gray_blur_host);

float tcd_empty_filter_last_ms(const tcd_empty_filter* f);  /* fused gray+blur+absdiff+count kernel, ms */

/* ------------------------------------------------------------------ stream bank (K8-K10, batched)
 * The per-camera filters of BASELINE configs[2] (64 streams, 8 per GPU) for all streams of one GPU at once:
 * frames are uploaded ONCE and shared by the empty-frame gate, the background model and the detector
 * (tcd_stream_bank_frame -> tcd_detector_submit(..., frames_on_device = 1)); the gate is ONE kernel launch
 * over the streams, MotionFilter.filter_detections ONE background-model pass + ONE clean-up launch per round of
 * up to 16 detections per stream, with no host synchronisation in between.  Same arithmetic as tcd_mog2_* /
 * tcd_empty_filter_* (bit-exact vs cv2; src/empty_frame_filter.py:51-98, src/motion_filter.py:68-176).      */
typedef struct tcd_stream_bank tcd_stream_bank;
tcd_status tcd_stream_bank_create(int32_t n_streams, int32_t h, int32_t w, int32_t history, float var_threshold,
                                  int32_t detect_shadows, int32_t blur_size, int32_t device, tcd_stream_bank** out);
void tcd_stream_bank_destroy(tcd_stream_bank* b);
/* what: 1 = background model (MotionFilter.reset_background), 2 = gate reference plane (EmptyFrameFilter.reset) */
tcd_status tcd_stream_bank_reset(tcd_stream_bank* b, int32_t stream, int32_t what);
/* frames_host[s] (HWC u8 BGR, row stride strides[s]; strides NULL = tight) or NULL when stream s has no frame */
tcd_status tcd_stream_bank_upload(tcd_stream_bank* b, const uint8_t* const* frames_host, const int64_t* strides);
const uint8_t* tcd_stream_bank_frame(tcd_stream_bank* b, int32_t stream);   /* device frame, row stride w*3 */
void* tcd_stream_bank_cuda_stream(tcd_stream_bank* b);
/* changed_host[s]: pixels with |blur(gray) - previous| > threshold; -1 first frame; -2 no frame this step */
tcd_status tcd_stream_bank_gate(tcd_stream_bank* b, int32_t threshold, int32_t* changed_host);
/* boxes of stream s = rois[offsets[s] .. offsets[s+1]) (clamped integer x1,y1,x2,y2); counts_host[i] = non-zero
 * pixels of the cleaned mask inside box i after its own model update (the reference updates once per detection) */
tcd_status tcd_stream_bank_motion(tcd_stream_bank* b, const int32_t* rois, const int32_t* offsets, int32_t shadow_thr,
                                  int32_t ksize, int32_t blur_thr, int32_t* counts_host);
tcd_status tcd_stream_bank_last_ms(const tcd_stream_bank* b, float out_ms[3]);

/* ------------------------------------------------------------------ frame ring (ingest)
 * Replaces the copy chain of the capture loop (src/stream_capture_gpu_ffmpeg.py:253-283: stdout.read -> bytes ->
 * np.frombuffer -> torch.from_numpy(..).to(device) from pageable memory -> two .clone()s): raw bgr24 frames are read
 * straight into a PINNED slot, uploaded once (asynchronously, on the ring's own copy stream) into the matching DEVICE
 * slot, and consumers use the device slot itself (tcd_detector_submit(..., frames_on_device = 1),
 * tcd_stream_bank, tcd_motion_apply ...).  A slot is rewritten n_slots frames later: size the ring for the deepest
 * consumer queue.  One ring per camera, one producer thread per ring; slots are plain indices the caller cycles.  */
typedef struct tcd_frame_ring tcd_frame_ring;
tcd_status tcd_frame_ring_create(int32_t n_slots, int64_t frame_bytes, int32_t device, tcd_frame_ring** out);
void tcd_frame_ring_destroy(tcd_frame_ring* r);
uint8_t* tcd_frame_ring_host(tcd_frame_ring* r, int32_t slot);          /* pinned host slot */
const uint8_t* tcd_frame_ring_device(tcd_frame_ring* r, int32_t slot);  /* device slot */
/* blocking read()s from fd until the slot holds one frame; *got < frame_bytes = end of stream.  Waits for the slot's
 * previous upload first.  (ctypes calls it without the GIL: a capture thread costs the interpreter nothing.) */
tcd_status tcd_frame_ring_read_fd(tcd_frame_ring* r, int32_t slot, int32_t fd, int64_t* got);
tcd_status tcd_frame_ring_publish(tcd_frame_ring* r, int32_t slot);     /* start the slot's upload */
/* order `stream` (a cudaStream_t) after the slot's upload; NULL = wait on the host */
tcd_status tcd_frame_ring_wait(tcd_frame_ring* r, int32_t slot, void* stream);

/* ------------------------------------------------------------------ crop enhancement
 * Replaces ImageEnhancer.enhance_clahe_bilateral (src/image_enhancement.py:146-184): cv2.cvtColor BGR2LAB ->
 * cv2.createCLAHE(clipLimit, tileGridSize).apply(L) -> cv2.cvtColor LAB2BGR -> cv2.bilateralFilter(d, sigmaColor,
 * sigmaSpace); single-channel crops skip the colour conversions.  One call enhances a BATCH of crops in three kernel
 * launches.  Colour conversions and CLAHE equal the library's 8-bit results bit for bit; the bilateral filter equals
 * oracle/enhance.py bit for bit and the library within its own path-to-path spread. */
typedef struct tcd_enhance_cfg {
  float clip_limit;          /* CLAHE clipLimit (reference default 2.0) */
  int32_t grid_x, grid_y;    /* CLAHE tileGridSize (8, 8) */
  int32_t d;                 /* bilateral diameter (9); 3..15 */
  float sigma_color, sigma_space; /* 75, 75 */
} tcd_enhance_cfg;
typedef struct tcd_enhancer tcd_enhancer;
tcd_status tcd_enhancer_create(const tcd_enhance_cfg* cfg, int32_t device, tcd_enhancer** out);
void tcd_enhancer_destroy(tcd_enhancer* e);
/* imgs[i] / outs[i]: HWC u8 with `channels` (1 or 3) channels, h[i] x w[i], row strides in bytes (NULL or 0 = tight);
 * host or device pointers as the two flags say (device crops may be views into a frame).  Returns when the outputs
 * are complete. */
tcd_status tcd_enhance_clahe_bilateral(tcd_enhancer* e, const uint8_t* const* imgs, const int32_t* h, const int32_t* w,
                                       const int64_t* stride, int32_t n, int32_t channels, int32_t imgs_on_device,
                                       uint8_t* const* outs, const int64_t* out_stride, int32_t outs_on_device);
tcd_status tcd_enhancer_stream(tcd_enhancer* e, void** stream);
/* No counterpart in the reference.  tcd_enhance_clahe_bilateral for ONE
 * crop for callers that cannot allocate: no handle, and no cudaMalloc / cudaFree / allocating library call inside --
 * all device scratch (descriptor, colour tables, tile LUTs, the intermediate image) lives in a static 1 MiB device array
 * inside libtcd, resident on every GPU that loads the library until exit.  img_dev / out_dev: DEVICE pointers, HWC u8,
 * `channels` 1 or 3, row strides in bytes (0 = tight).  Limits: h * w * channels <= tcd_enhance_noalloc_max_image_bytes()
 * (786 432 = 512 x 512 x 3) and grid_x * grid_y <= 64; larger requests fail with TCD_ERR_INVALID and a message that names
 * the limit.  Work is enqueued on `stream` (a cudaStream_t, NULL = default) and the call returns once it is complete.
 * Every call shares the one scratch array: calls that overlap on the GPU (other host threads / streams) race and must
 * be serialised by the caller.  Results equal tcd_enhance_clahe_bilateral's bit for bit. */
tcd_status tcd_enhance_clahe_bilateral_noalloc(const tcd_enhance_cfg* cfg, int32_t device, const uint8_t* img_dev, int32_t h,
                                               int32_t w, int64_t stride, int32_t channels, uint8_t* out_dev,
                                               int64_t out_stride, void* stream);
int64_t tcd_enhance_noalloc_max_image_bytes(void);
/* the 8-bit colour-conversion tables the kernels use (host code, no GPU needed): gamma[256], cbrt[3072], yf[512],
 * inv[4096], forward / inverse matrix coefficients cf[9] / ci[9] -- exported so that they can be checked without a GPU */
tcd_status tcd_enhance_tables(uint16_t* gamma, uint16_t* cbrt, uint16_t* yf, uint8_t* inv, int32_t* cf, int32_t* ci);

/* ------------------------------------------------------------------ JPEG encode
 * Replaces cv2.imencode('.jpg', frame, [cv2.IMWRITE_JPEG_QUALITY, q]) on the result path (src/web_server.py:923: one
 * encode per camera per MJPEG tick; src/snapshot_saver.py:160 ring-buffer compression, :358-404 cv2.imwrite) for a
 * batch of BGR u8 frames, host or device.  Baseline sequential JPEG, YCbCr 4:2:0, Annex K Huffman tables, IJG quality
 * scale; the files are byte-identical to the reference's encoder (libjpeg-turbo inside cv2). */
typedef struct tcd_jpeg_encoder tcd_jpeg_encoder;
tcd_status tcd_jpeg_encoder_create(int32_t max_h, int32_t max_w, int32_t max_batch, int32_t device, tcd_jpeg_encoder** out);
void tcd_jpeg_encoder_destroy(tcd_jpeg_encoder* e);
/* out[i]: HOST buffer of out_cap[i] bytes (tcd_jpeg_max_bytes never fails); out_size[i] = file length */
tcd_status tcd_jpeg_encode(tcd_jpeg_encoder* e, const uint8_t* const* frames, const int32_t* h, const int32_t* w,
                           const int64_t* stride, int32_t n, int32_t frames_on_device, int32_t quality,
                           uint8_t* const* out, const int64_t* out_cap, int64_t* out_size);
int64_t tcd_jpeg_max_bytes(int32_t h, int32_t w);

/* ------------------------------------------------------------------ Stage-2 crops (K13)
 * Replaces the per-detection slice frame[y1:y2, x1:x2] (src/two_stage_pipeline_yolox.py:289)
 * + SpeciesClassifier.preprocess (src/species_classifier.py:311-350) for ALL crops of a
 * frame in one launch.  boxes: HOST int32 [N,4] (x1,y1,x2,y2) already padded + clamped.
 * out_layout 0: f32 NCHW [N,3,S,S]; 1: bf16 NHWC [N,S,S,4].  out is a DEVICE pointer.     */
tcd_status tcd_crop_resize(const uint8_t* frame_dev, int32_t h, int32_t w, int64_t row_stride,
                           const int32_t* boxes_host, int32_t n, int32_t S,
                           const tcd_preproc_cfg* cfg, int32_t out_layout, void* out_dev,
                           void* stream);

/* Stage-2 crops straight into the classifier body: the same gather as tcd_crop_resize
 * (src/two_stage_pipeline_yolox.py:289 + src/species_classifier.py:311-350), written as the bf16 PATCH ROWS the
 * patch-embedding GEMM of a vision transformer reads (what timm's PatchEmbed Conv2d computes inside
 * `self.model(input_tensor)`, src/species_classifier.py:386-388): rows_dev[(i*G + py)*G + px][c*patch*patch + ky*patch
 * + kx], G = S / patch, k_pad >= 3*patch*patch columns per row (the caller keeps the padding columns zero).
 * boxes_host: int32 [n,4], copied into boxes_scratch_dev (device, >= 16*n bytes, caller-owned: nothing is allocated).
 * Asynchronous on `stream`. */
tcd_status tcd_crop_patches(const uint8_t* frame_dev, int32_t h, int32_t w, int64_t row_stride,
                            const int32_t* boxes_host, int32_t* boxes_scratch_dev, int32_t n, int32_t S, int32_t patch,
                            int32_t k_pad, const tcd_preproc_cfg* cfg, void* rows_dev, void* stream);

/* the same patch rows from an already preprocessed batch (x_dev: f32 NCHW [n,3,S,S], what SpeciesClassifier.preprocess
 * returns, src/species_classifier.py:298-352); asynchronous on `stream` */
tcd_status tcd_patchify_f32(const float* x_dev, int32_t n, int32_t S, int32_t patch, int32_t k_pad, void* rows_dev,
                            void* stream);

/* ------------------------------------------------------------------ layer programs (RT-DETR, Stage-2 classifier)
 * A program is an ordered list of kernels over device buffers it owns, replayed as one CUDA
 * graph.  The host (telescope_cam_detection_b200/rtdetr_engine.py) lays RT-DETRv2-R50 out as
 * such a list -- the replacement of RTDETRDetector.load_model / self.model(img, orig_size),
 * src/rtdetr_detector.py:60-204,257,340 -- from tcgen05 convs / linears plus the transformer
 * kernels of csrc/tfm_kernels.cuh; eva02_engine.py does the same for the Stage-2 classifier body
 * (timm.create_model(...) / self.model(input_tensor), src/species_classifier.py:258-262,386-388).
 * Nothing is allocated after the last add_*.                                                  */
typedef struct tcd_program tcd_program;
typedef struct {
  char name[64];
  const void* x; int32_t N, H, W, x_ctot, x_coff, cin; int64_t x_nstride;   /* bf16 NHWC input view  */
  const float* w_host; const float* bias_host;        /* f32 [cout][cin][k][k], [cout] (BN folded)     */
  int32_t cout, k, stride, act;                       /* act: 0 none 1 SiLU 2 ReLU 3 GELU; 4 gated     */
  /* act 4 (token linears on CTA pairs only): 32-column slabs alternate gate / value, the output holds      */
  /* silu(gate) * value and is cout / 2 columns wide (the fused fc1_g || fc1_x of a SwiGLU MLP)            */
  void* y; int32_t y_ctot, y_coff; int64_t y_nstride; /* bf16 output view (NULL when out_f32 is set)   */
  const void* resid; int32_t r_ctot, r_coff, resid_pre; int64_t r_nstride;
  float* out_f32; int32_t out_ld, out_col, rows_per_img, row_off;  /* f32 row output (logits etc.)    */
  /* LayerNorm applied by the CONSUMER (token linears on CTA pairs only): x is the UN-normalised tensor, ln_stats holds   */
  /* (mean, 1/std) per row (TCD_OP_ROW_STATS) and ln_colsum_host the sums over cin of the (bf16-rounded) weight rows:     */
  /* y = rstd * (x W^T - mean * colsum) + bias -- the normalised tensor is never written.  NULL: plain linear.            */
  const float* ln_stats; const float* ln_colsum_host;
  /* ln_parts > 0: ln_stats holds ln_parts partial (sum, sum of squares) pairs per row instead -- written by the PRODUCER of  */
  /* x (stats_out of the launches that wrote it, or TCD_OP_ROW_STATS) -- over ln_count columns; eps = ln_eps.               */
  int32_t ln_parts, ln_count; float ln_eps;
  /* stats_out (CTA-pair GEMM only): every 256-column tile (128 for act 4) of this launch also writes the (sum, sum of      */
  /* squares) of its output values per row to stats_out[row * stats_parts + stats_slot0 + tile]: LayerNorm statistics for   */
  /* the next consumer without another pass over the tensor.                                                               */
  float* stats_out; int32_t stats_parts, stats_slot0;
} tcd_conv_desc;
enum {
  TCD_OP_LAYERNORM = 1, TCD_OP_ADD_ROWS, TCD_OP_MASK_ROWS, TCD_OP_MHA, TCD_OP_MSDA, TCD_OP_ROWMAX,
  TCD_OP_TOPK, TCD_OP_GATHER, TCD_OP_REF_INIT, TCD_OP_REF_UPDATE, TCD_OP_DETR_POST, TCD_OP_MAXPOOL,
  TCD_OP_AVGPOOL, TCD_OP_UPSAMPLE,
  /* Stage-2 classifier body (EVA02 vision transformer) */
  TCD_OP_VIT_ASSEMBLE, TCD_OP_MHA_ROPE, TCD_OP_SWIGLU_LN, TCD_OP_SOFTMAX, TCD_OP_TAKE, TCD_OP_HEAD_GEMV, TCD_OP_ROW_STATS
};
typedef struct {
  char name[64];
  int32_t kind;
  void* p[8];      /* device pointers, meaning per kind (see csrc/program.cu) */
  int32_t i[12];
  float f[4];
} tcd_op_desc;
tcd_status tcd_program_create(int32_t device, tcd_program** out);
void tcd_program_destroy(tcd_program* prog);
tcd_status tcd_program_alloc(tcd_program* prog, size_t bytes, void** out_dev);  /* zero-filled */
tcd_status tcd_program_upload(tcd_program* prog, const void* host, size_t bytes, void** out_dev);
tcd_status tcd_program_add_conv(tcd_program* prog, const tcd_conv_desc* d);
tcd_status tcd_program_add_op(tcd_program* prog, const tcd_op_desc* d);
/* enqueue one replay on the program's stream (graph != 0: captured CUDA graph); asynchronous */
tcd_status tcd_program_run(tcd_program* prog, int32_t graph);
tcd_status tcd_program_sync(tcd_program* prog);
void* tcd_program_stream(tcd_program* prog);
int32_t tcd_program_num_ops(const tcd_program* prog);
const char* tcd_program_op_name(const tcd_program* prog, int32_t idx);
double tcd_program_op_flops(const tcd_program* prog, int32_t idx);   /* 2*MACs, convs / linears only */
/* eager pass with a CUDA-event pair around every op; op_ms[i] = mean ms over iters          */
tcd_status tcd_program_profile(tcd_program* prog, int32_t iters, float* op_ms, int32_t cap);
/* PIL-style antialiased resize + BGR->RGB + /255 (RTDETRDetector.preprocess,
 * src/rtdetr_detector.py:206-236; T.Resize on a PIL image, :176-180) of n HWC u8 frames (host or
 * device) into bf16 NHWC [n,out_h,out_w,16] at out_dev; out_u8_host (optional): the resized u8 RGB
 * image [n,out_h,out_w,3] for bit-exact parity against PIL.                                   */
tcd_status tcd_pil_resize(const uint8_t* const* frames, const int32_t* h, const int32_t* w,
                          const int64_t* row_stride, int32_t n, int32_t frames_on_device, int32_t out_h,
                          int32_t out_w, void* out_dev, uint8_t* out_u8_host, void* stream);

/* ------------------------------------------------------------------ small utilities    */
tcd_status tcd_malloc_device(size_t bytes, void** out);
tcd_status tcd_free_device(void* p);
tcd_status tcd_malloc_pinned(size_t bytes, void** out);
tcd_status tcd_free_pinned(void* p);
tcd_status tcd_memcpy_h2d(void* dst, const void* src, size_t bytes);
tcd_status tcd_memcpy_d2h(void* dst, const void* src, size_t bytes);
tcd_status tcd_device_synchronize(void);

#ifdef __cplusplus
}
#endif
#endif /* TCD_H_ */
