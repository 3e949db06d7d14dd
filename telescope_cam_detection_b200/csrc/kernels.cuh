// Argument structs and host launchers shared between the kernel TUs and the C-ABI glue.
#pragma once
#include "tcd_common.cuh"

namespace tcd {

// ---- preprocess (preproc.cu)
struct PreFrame {
  const uint8_t* ptr;
  long long stride;
  int h, w;
  int content_h, content_w;  // letterbox content size (== out size in stretch mode)
  float scale_y, scale_x;
  // vertical ratio an odd integer k (1080p -> 640 letterbox: 3): output row d taps source row k*d + (k-1)/2 with
  // weight exactly 1 and its neighbour with weight exactly 0 -- only every k-th row is read (and uploaded)
  int row_step, row_phase;   // 1, 0 in the general case
};
int preproc_launch(const tcd_preproc_cfg* cfg, const PreFrame* frames_host, const PreFrame* frames_dev,
                   int B, int layout, void* out_dev, cudaStream_t stream);
void preproc_describe(const tcd_preproc_cfg* cfg, const uint8_t* ptr, int h, int w, long long stride,
                      PreFrame* f);

// ---- decode / filter / NMS (postproc.cu)
constexpr int kPostMaxWords = 132;      // cap <= 8448 candidates per frame -> 132 mask words per row
constexpr int kMaxFilterTables = 64;    // per-camera class-filter tables of one detector (0 = pass-through)
struct FrameMeta {                      // per frame of a batch: which filter table, and the un-scaling of its boxes
  int table;
  float ratio_w, ratio_h;               // orig / network size, as np.float32
  int pad;
};
struct PostBuffers {
  int B, A, C, cap, max_out;
  const float* pred;      // [B,A,5+C] raw head output or decoded prediction
  float* decoded;         // optional [B,A,5+C]
  const float* thr;       // device [2] = {conf, nms}
  int* cand_count;        // [B]
  float* cand;            // [B,cap,8]
  int* cand_anchor;       // [B,cap]
  float* srec;            // [B,cap,8] sorted rows
  float4* sbox;           // [B,cap] sorted offset boxes
  int* n_sorted;          // [B]
  unsigned long long* mask;  // [B,cap,cap/64]
  float* out_dets;        // [B,max_out,7]
  int* out_counts;        // [B]
  int n_levels;
  int lvl_start[4], lvl_w[4], lvl_stride[4];
  const int* maybe_count;         // [B] rows the prediction convs wrote (objectness could still reach conf); nullptr:
  const int* maybe_idx;           // [B,cap] their anchor indices                                   every row is there
  const FrameMeta* meta;          // device [B] or nullptr (no class filter)
  const tcd_class_filter* tables; // device [kMaxFilterTables]
};
// evs (optional): 8 events, recorded as a pair around each of the 4 kernels (profiling pass)
int postproc_launch(const PostBuffers& pb, int raw, cudaStream_t stream, cudaEvent_t* evs = nullptr);
int class_filter_rows_launch(const tcd_class_filter* tables, int table, const float* rows_dev, int n, float rw, float rh,
                             int* keep_dev, cudaStream_t stream);

// ---- motion (motion.cu)
constexpr int kMogModes = 5;
constexpr int kMogMaxRepeats = 16;
struct Mog2Params {
  int npix, w, h;
  long long stride;
  float Tb, TB, Tg, varInit, varMin, varMax, CT, tau;
  int detect_shadows;
  int repeats;
  float alpha[kMogMaxRepeats];
};
// SoA over pixels: weight[m][npix], var[m][npix], mean[m][c][npix] (f32), nmodes[npix] (u8)
struct Mog2State {
  float* weight;
  float* var;
  float* mean;
  uint8_t* nmodes;
};
constexpr int kBlurMaxK = 31;
struct BlurArgs {
  int w, h, ksize;         // ksize odd, <= 31
  long long in_stride;
  int src_mode;            // 0: u8 plane binarised (p > pre_thr ? 255 : 0); 1: BGR -> gray; 2: u8 plane as is
  int pre_thr;
  int post_mode;           // 0: write (v > post_thr ? 255 : 0); 1: absdiff vs prev, count, prev = v
  int post_thr;
  int first;               // post_mode 1: no previous frame yet (just store)
  uint16_t k[kBlurMaxK];   // 8.8 fixed-point taps, sum 256
};
// One stream of a bank, as the batched kernels see it (device array, rewritten per launch)
struct BankStream {
  const uint8_t* frame;        // this step's frame (device, HWC u8 BGR)
  long long stride;
  Mog2State st;
  uint8_t* fg;                 // [kMogMaxRepeats][npix] raw MOG2 masks, one plane per model update of the launch
  uint8_t* prev;               // gate: previous blurred gray plane
  unsigned int* counter;       // gate: changed-pixel counter
  unsigned int* roi_counts;    // [kMogMaxRepeats] counters of this launch's boxes
  int active, first;           // gate: stream has a frame this step / has no previous plane yet
  int repeats;                 // motion: model updates (= boxes) of this launch, 0 = none
  float alpha[kMogMaxRepeats];
  int roi[kMogMaxRepeats][4];  // clamped integer boxes (x1,y1,x2,y2)
};
int bank_gate_launch(const BankStream* streams_dev, int n_streams, const BlurArgs& a, cudaStream_t stream);
int bank_mog2_launch(const BankStream* streams_dev, int n_streams, const Mog2Params& p, cudaStream_t stream);
int bank_roi_launch(const BankStream* streams_dev, int n_streams, const BlurArgs& a, cudaStream_t stream);
int mog2_launch(const uint8_t* frame, const Mog2State& st, const Mog2Params& p, uint8_t* fgmask,
                cudaStream_t stream);
int blur_launch(const uint8_t* src, const BlurArgs& a, uint8_t* dst, uint8_t* prev,
                unsigned int* counter, cudaStream_t stream);
int threshold_launch(const uint8_t* src, int n, int thr, uint8_t* dst, cudaStream_t stream);
int roi_count_launch(const uint8_t* mask, int w, int h, const int* rois_dev, int n,
                     unsigned int* counts_dev, cudaStream_t stream);

// ---- stage-2 crops (crop_resize.cu)
int crop_resize_launch(const uint8_t* frame, int h, int w, long long stride, const int* boxes_dev,
                       int n, int S, const tcd_preproc_cfg* cfg, int layout, void* out,
                       cudaStream_t stream);

// ---- crop enhancement (enhance.cu)
struct EnhTables {            // device pointers
  const uint16_t* gamma;      // [256]  sRGB gamma x 8
  const uint16_t* cbrt;       // [3072] cube root, 1.15 fixed point
  const uint16_t* yf;         // [256][2] L -> (y, fy), 14-bit
  const uint8_t* inv;         // [4096] inverse gamma
};
struct EnhImg {
  const uint8_t* src;         // crop, HWC u8 (cn channels)
  long long src_stride;
  uint8_t* mid;               // after CLAHE (tight rows)
  uint8_t* dst;
  long long dst_stride;
  uint8_t* luts;              // [gy*gx][256]
  int h, w, cn, th, tw, clip;
  float lut_scale, inv_tw, inv_th;
};
void enhance_host_tables(uint16_t* gamma, uint16_t* cbrt, uint16_t* yf, uint8_t* inv, int32_t* cf, int32_t* ci);
int enhance_set_constants(const int32_t* cf, const int32_t* ci);
int enhance_set_window(const int8_t* dy, const int8_t* dx, const float* w, int maxk);
int enhance_launch(const EnhImg* imgs_dev, int n, int cn, int gx, int gy, int max_h, int max_w, const EnhTables& t,
                   const float* color_w_dev, int radius, int maxk, cudaStream_t stream);

// ---- JPEG encode (jpeg.cu)
struct JpegImg {
  const uint8_t* src;          // BGR u8 HWC
  long long stride;
  int h, w, mcus_x, mcus_y, blocks_x, blocks_y;   // blocks_*: real luma blocks, ceil(w/8) x ceil(h/8)
  const uint16_t* qtab;        // [2][64] natural order: luma, chroma
  int16_t* coef;               // [mcus][6][64] zigzag order
  unsigned int* bit_cnt;       // [mcus*6]
  unsigned long long* bit_off; // [mcus*6]
  unsigned int* words;         // zeroed bit stream (big-endian inside each word)
  unsigned long long* meta;    // [0] total bits, [1] file bytes, [2] overflow flag
  uint8_t* out;                // file buffer: header (written by the host side) + scan + EOI
  long long out_cap;
  int header_len;
};
int jpeg_set_tables(const uint8_t* zigzag_pos, const uint16_t* code, const uint8_t* size);
int jpeg_launch(const JpegImg* imgs_dev, int n, int max_mcus_x, int max_mcus_y, cudaStream_t stream);

int require_sm100();

}  // namespace tcd
