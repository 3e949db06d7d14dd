// C ABI glue, part 3: MOG2 motion filter and empty-frame gate handles.
#include <algorithm>
#include <atomic>
#include <math.h>
#include <string.h>
#include <vector>

#include "kernels.cuh"

using namespace tcd;

namespace tcd {
extern std::atomic<uint64_t> g_h2d_bytes;

// OpenCV's bit-exact u8 GaussianBlur uses an 8.8 fixed-point 1-D kernel built by
// getGaussianKernelBitExact + getGaussianKernelFixedPoint_ED (imgproc/src/smooth.dispatch.cpp,
// OpenCV 4.x; dependency of the reference, not vendored): sigma = 0.3*((n-1)*0.5-1)+0.8 when
// sigma<=0; small fixed tables for n<=7; g normalised in double; then left-to-right error
// diffusion: v = round_half_even(g[i]*256 + err), err carries, centre tap = 256 - 2*sum.
// For n = 21 this yields [0,2,2,4,6,11,15,20,25,28,30,...] as recovered in SURVEY App. C.
int gaussian_kernel_q8(int n, uint16_t* out) {
  if (n < 1 || n > kBlurMaxK || (n & 1) == 0) return TCD_ERR_INVALID;
  std::vector<double> g(n);
  static const double small[4][7] = {{1.0},
                                     {0.25, 0.5, 0.25},
                                     {0.0625, 0.25, 0.375, 0.25, 0.0625},
                                     {0.03125, 0.109375, 0.21875, 0.28125, 0.21875, 0.109375, 0.03125}};
  if (n <= 7) {
    for (int i = 0; i < n; ++i) g[i] = small[n / 2][i];
  } else {
    const double sigma = ((n - 1) * 0.5 - 1) * 0.3 + 0.8;
    const double scale2x = -0.5 / (sigma * sigma);
    double sum = 0;
    for (int i = 0; i < n; ++i) {
      const double x = i - (n - 1) * 0.5;
      g[i] = exp(scale2x * x * x);
      sum += g[i];
    }
    for (int i = 0; i < n; ++i) g[i] /= sum;
  }
  const int n2 = n / 2;
  double err = 0;
  long long sum = 0;
  for (int i = 0; i < n2; ++i) {
    const double adj = g[i] * 256.0 + err;
    const long long v0 = (long long)nearbyint(adj);  // round half to even (default FP mode)
    err = adj - (double)v0;
    out[i] = out[n - 1 - i] = (uint16_t)v0;
    sum += v0;
  }
  out[n2] = (uint16_t)(256 - 2 * sum);
  return TCD_OK;
}

}  // namespace tcd

struct tcd_mog2 {
  int h, w, history, device, detect_shadows;
  float var_threshold;
  long long nframes = 0;
  Mog2State st{};
  uint8_t* frame_dev = nullptr;   // staging for host frames
  uint8_t* fg = nullptr;          // raw MOG2 mask
  uint8_t* clean = nullptr;       // cleaned mask
  int* rois_dev = nullptr;
  unsigned int* counts_dev = nullptr;
  int roi_cap = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};  // apply begin/end, clean-up begin/end
  float last_ms[2] = {0.f, 0.f};                             // per apply; blur+threshold+roi count
};

struct tcd_empty_filter {
  int h, w, blur_size, device;
  bool have_prev = false;
  uint8_t* frame_dev = nullptr;
  uint8_t* prev = nullptr;
  unsigned int* counter = nullptr;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[2] = {nullptr, nullptr};
  float last_ms = 0.f;
  uint16_t k[kBlurMaxK];
};

extern "C" {

tcd_status tcd_mog2_create(int32_t h, int32_t w, int32_t history, float var_threshold,
                           int32_t detect_shadows, int32_t device, tcd_mog2** out) {
  TCD_REQUIRE(out && h > 0 && w > 0 && history > 0, "mog2_create: bad argument");
  TCD_CUDA_OK(cudaSetDevice(device));
  int rc = require_sm100();
  if (rc) return (tcd_status)rc;
  tcd_mog2* m = new tcd_mog2();
  m->h = h; m->w = w; m->history = history; m->device = device;
  m->detect_shadows = detect_shadows; m->var_threshold = var_threshold;
  const size_t np = (size_t)h * w;
  TCD_CUDA_OK(cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking));
  for (auto& e : m->ev) TCD_CUDA_OK(cudaEventCreate(&e));
  TCD_CUDA_OK(cudaMalloc((void**)&m->st.weight, np * kMogModes * 4));
  TCD_CUDA_OK(cudaMalloc((void**)&m->st.var, np * kMogModes * 4));
  TCD_CUDA_OK(cudaMalloc((void**)&m->st.mean, np * kMogModes * 3 * 4));
  TCD_CUDA_OK(cudaMalloc((void**)&m->st.nmodes, np));
  TCD_CUDA_OK(cudaMalloc((void**)&m->frame_dev, np * 3));
  TCD_CUDA_OK(cudaMalloc((void**)&m->fg, np));
  TCD_CUDA_OK(cudaMalloc((void**)&m->clean, np));
  m->roi_cap = 1024;
  TCD_CUDA_OK(cudaMalloc((void**)&m->rois_dev, sizeof(int) * 4 * m->roi_cap));
  TCD_CUDA_OK(cudaMalloc((void**)&m->counts_dev, sizeof(unsigned int) * m->roi_cap));
  *out = m;
  return tcd_mog2_reset(m);
}

void tcd_mog2_destroy(tcd_mog2* m) {
  if (!m) return;
  cudaSetDevice(m->device);
  cudaStreamSynchronize(m->stream);
  void* ptrs[] = {m->st.weight, m->st.var, m->st.mean, m->st.nmodes, m->frame_dev, m->fg, m->clean,
                  m->rois_dev, m->counts_dev};
  for (void* p : ptrs) if (p) cudaFree(p);
  for (auto& e : m->ev) if (e) cudaEventDestroy(e);
  cudaStreamDestroy(m->stream);
  delete m;
}

tcd_status tcd_mog2_reset(tcd_mog2* m) {
  TCD_REQUIRE(m, "mog2_reset: null handle");
  TCD_CUDA_OK(cudaSetDevice(m->device));
  const size_t np = (size_t)m->h * m->w;
  m->nframes = 0;
  TCD_CUDA_OK(cudaMemsetAsync(m->st.nmodes, 0, np, m->stream));
  TCD_CUDA_OK(cudaMemsetAsync(m->fg, 0, np, m->stream));
  TCD_CUDA_OK(cudaStreamSynchronize(m->stream));
  return TCD_OK;
}

tcd_status tcd_mog2_apply(tcd_mog2* m, const uint8_t* frame, int64_t row_stride,
                          int32_t frame_on_device, int32_t repeats, uint8_t* fg_mask_host) {
  TCD_REQUIRE(m && frame && repeats >= 1, "mog2_apply: bad argument");
  TCD_CUDA_OK(cudaSetDevice(m->device));
  const size_t tight = (size_t)m->w * 3;
  if (row_stride <= 0) row_stride = (int64_t)tight;
  const uint8_t* src = frame;
  long long stride = row_stride;
  if (!frame_on_device) {
    TCD_CUDA_OK(cudaMemcpy2DAsync(m->frame_dev, tight, frame, (size_t)row_stride, tight, (size_t)m->h,
                                  cudaMemcpyHostToDevice, m->stream));
    src = m->frame_dev;
    stride = (long long)tight;
  }
  Mog2Params p;
  p.npix = m->h * m->w; p.w = m->w; p.h = m->h; p.stride = stride;
  // OpenCV defaults (probed on cv2 4.13)
  p.Tb = m->var_threshold; p.TB = 0.9f; p.Tg = 9.0f; p.varInit = 15.0f; p.varMin = 4.0f;
  p.varMax = 75.0f; p.CT = 0.05f; p.tau = 0.5f;
  p.detect_shadows = m->detect_shadows;
  int left = repeats;
  cudaEventRecord(m->ev[0], m->stream);
  while (left > 0) {
    const int n = left < kMogMaxRepeats ? left : kMogMaxRepeats;
    p.repeats = n;
    for (int i = 0; i < n; ++i) {
      ++m->nframes;
      const long long den = 2 * m->nframes < m->history ? 2 * m->nframes : m->history;
      p.alpha[i] = (float)(1.0 / (double)den);
    }
    int rc = mog2_launch(src, m->st, p, m->fg, m->stream);
    if (rc) return (tcd_status)rc;
    left -= n;
  }
  cudaEventRecord(m->ev[1], m->stream);
  if (fg_mask_host)
    TCD_CUDA_OK(cudaMemcpyAsync(fg_mask_host, m->fg, (size_t)m->h * m->w, cudaMemcpyDeviceToHost, m->stream));
  TCD_CUDA_OK(cudaStreamSynchronize(m->stream));
  float t = 0.f;
  if (cudaEventElapsedTime(&t, m->ev[0], m->ev[1]) == cudaSuccess) m->last_ms[0] = t / (float)repeats;
  return TCD_OK;
}

tcd_status tcd_mog2_roi_counts(tcd_mog2* m, int32_t shadow_thr, int32_t ksize, int32_t blur_thr,
                               const int32_t* rois, int32_t n_rois, int32_t* counts_host,
                               uint8_t* clean_mask_host) {
  TCD_REQUIRE(m && n_rois >= 0 && (n_rois == 0 || (rois && counts_host)), "mog2_roi_counts: bad argument");
  TCD_REQUIRE(n_rois <= m->roi_cap, "mog2_roi_counts: %d ROIs exceed capacity %d", n_rois, m->roi_cap);
  TCD_CUDA_OK(cudaSetDevice(m->device));
  const int np = m->h * m->w;
  int rc;
  cudaEventRecord(m->ev[2], m->stream);
  if (ksize > 1) {
    BlurArgs a;
    memset(&a, 0, sizeof(a));
    a.w = m->w; a.h = m->h; a.ksize = ksize; a.in_stride = m->w;
    a.src_mode = 0; a.pre_thr = shadow_thr; a.post_mode = 0; a.post_thr = blur_thr;
    rc = gaussian_kernel_q8(ksize, a.k);
    TCD_REQUIRE(rc == TCD_OK, "mog2_roi_counts: blur size %d unsupported (odd, <= %d)", ksize, kBlurMaxK);
    rc = blur_launch(m->fg, a, m->clean, nullptr, nullptr, m->stream);
  } else {
    rc = threshold_launch(m->fg, np, shadow_thr, m->clean, m->stream);
  }
  if (rc) return (tcd_status)rc;
  if (n_rois > 0) {
    for (int i = 0; i < n_rois; ++i) {
      const int32_t* r = rois + 4 * i;
      TCD_REQUIRE(r[0] >= 0 && r[1] >= 0 && r[2] <= m->w && r[3] <= m->h, "mog2_roi_counts: ROI %d not clamped", i);
    }
    TCD_CUDA_OK(cudaMemcpyAsync(m->rois_dev, rois, sizeof(int) * 4 * n_rois, cudaMemcpyHostToDevice, m->stream));
    TCD_CUDA_OK(cudaMemsetAsync(m->counts_dev, 0, sizeof(unsigned int) * n_rois, m->stream));
    rc = roi_count_launch(m->clean, m->w, m->h, m->rois_dev, n_rois, m->counts_dev, m->stream);
    if (rc) return (tcd_status)rc;
    cudaEventRecord(m->ev[3], m->stream);
    TCD_CUDA_OK(cudaMemcpyAsync(counts_host, m->counts_dev, sizeof(int) * n_rois, cudaMemcpyDeviceToHost, m->stream));
  }
  if (clean_mask_host)
    TCD_CUDA_OK(cudaMemcpyAsync(clean_mask_host, m->clean, (size_t)np, cudaMemcpyDeviceToHost, m->stream));
  TCD_CUDA_OK(cudaStreamSynchronize(m->stream));
  float t = 0.f;
  if (n_rois > 0 && cudaEventElapsedTime(&t, m->ev[2], m->ev[3]) == cudaSuccess) m->last_ms[1] = t;
  return TCD_OK;
}

/* device time (CUDA events on the handle's stream) of the kernels of the last call, ms:
 * [0] one MOG2 apply, [1] mask clean-up (threshold + blur + threshold) + ROI counts        */
tcd_status tcd_mog2_last_ms(const tcd_mog2* m, float out_ms[2]) {
  TCD_REQUIRE(m && out_ms, "mog2_last_ms: null argument");
  out_ms[0] = m->last_ms[0];
  out_ms[1] = m->last_ms[1];
  return TCD_OK;
}

tcd_status tcd_empty_filter_create(int32_t h, int32_t w, int32_t blur_size, int32_t device,
                                   tcd_empty_filter** out) {
  TCD_REQUIRE(out && h > 0 && w > 0, "empty_filter_create: bad argument");
  TCD_CUDA_OK(cudaSetDevice(device));
  int rc = require_sm100();
  if (rc) return (tcd_status)rc;
  tcd_empty_filter* f = new tcd_empty_filter();
  f->h = h; f->w = w; f->blur_size = blur_size; f->device = device;
  rc = gaussian_kernel_q8(blur_size, f->k);
  if (rc) {
    set_last_error("empty_filter_create: blur size %d unsupported (odd, <= %d)", blur_size, kBlurMaxK);
    delete f;
    return TCD_ERR_INVALID;
  }
  const size_t np = (size_t)h * w;
  TCD_CUDA_OK(cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking));
  for (auto& e : f->ev) TCD_CUDA_OK(cudaEventCreate(&e));
  TCD_CUDA_OK(cudaMalloc((void**)&f->frame_dev, np * 3));
  TCD_CUDA_OK(cudaMalloc((void**)&f->prev, np));
  TCD_CUDA_OK(cudaMalloc((void**)&f->counter, sizeof(unsigned int)));
  *out = f;
  return TCD_OK;
}

void tcd_empty_filter_destroy(tcd_empty_filter* f) {
  if (!f) return;
  cudaSetDevice(f->device);
  cudaStreamSynchronize(f->stream);
  cudaFree(f->frame_dev);
  cudaFree(f->prev);
  cudaFree(f->counter);
  for (auto& e : f->ev) if (e) cudaEventDestroy(e);
  cudaStreamDestroy(f->stream);
  delete f;
}

tcd_status tcd_empty_filter_reset(tcd_empty_filter* f) {
  TCD_REQUIRE(f, "empty_filter_reset: null handle");
  f->have_prev = false;
  return TCD_OK;
}

tcd_status tcd_empty_filter_step(tcd_empty_filter* f, const uint8_t* frame, int64_t row_stride,
                                 int32_t channels, int32_t frame_on_device, int32_t threshold,
                                 int32_t* changed_pixels_host, uint8_t* gray_blur_host) {
  TCD_REQUIRE(f && frame && changed_pixels_host && (channels == 1 || channels == 3),
              "empty_filter_step: bad argument");
  TCD_CUDA_OK(cudaSetDevice(f->device));
  const size_t tight = (size_t)f->w * channels;
  if (row_stride <= 0) row_stride = (int64_t)tight;
  const uint8_t* src = frame;
  long long stride = row_stride;
  if (!frame_on_device) {
    TCD_CUDA_OK(cudaMemcpy2DAsync(f->frame_dev, tight, frame, (size_t)row_stride, tight, (size_t)f->h,
                                  cudaMemcpyHostToDevice, f->stream));
    src = f->frame_dev;
    stride = (long long)tight;
  }
  BlurArgs a;
  memset(&a, 0, sizeof(a));
  a.w = f->w; a.h = f->h; a.ksize = f->blur_size; a.in_stride = stride;
  a.src_mode = channels == 3 ? 1 : 2; a.post_mode = 1; a.post_thr = threshold; a.first = f->have_prev ? 0 : 1;
  memcpy(a.k, f->k, sizeof(a.k));
  TCD_CUDA_OK(cudaMemsetAsync(f->counter, 0, sizeof(unsigned int), f->stream));
  cudaEventRecord(f->ev[0], f->stream);
  int rc = blur_launch(src, a, nullptr, f->prev, f->counter, f->stream);
  if (rc) return (tcd_status)rc;
  cudaEventRecord(f->ev[1], f->stream);
  unsigned int cnt = 0;
  TCD_CUDA_OK(cudaMemcpyAsync(&cnt, f->counter, sizeof(cnt), cudaMemcpyDeviceToHost, f->stream));
  if (gray_blur_host)
    TCD_CUDA_OK(cudaMemcpyAsync(gray_blur_host, f->prev, (size_t)f->h * f->w, cudaMemcpyDeviceToHost, f->stream));
  TCD_CUDA_OK(cudaStreamSynchronize(f->stream));
  *changed_pixels_host = f->have_prev ? (int32_t)cnt : -1;
  f->have_prev = true;
  float t = 0.f;
  if (cudaEventElapsedTime(&t, f->ev[0], f->ev[1]) == cudaSuccess) f->last_ms = t;
  return TCD_OK;
}

float tcd_empty_filter_last_ms(const tcd_empty_filter* f) { return f ? f->last_ms : 0.f; }

}  // extern "C"

// ------------------------------------------------------------------ stream bank: the per-camera filters, batched
// BASELINE configs[2] (SURVEY 8(d) cfg 3): the streams of one GPU share ONE launch per kernel per step instead of
// one synchronous C call per stream (and per detection): frames are uploaded once and shared by the gate, the
// background model and the detector's preprocess (tcd_stream_bank_frame -> frames_on_device = 1).
struct tcd_stream_bank {
  int n = 0, h = 0, w = 0, history = 0, device = 0, detect_shadows = 1, blur_size = 21;
  float var_threshold = 16.f;
  size_t npix = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t copy = nullptr;       // uploads: the next step's frames arrive while this step is being filtered
  cudaEvent_t ev_up[2] = {nullptr, nullptr};
  std::vector<Mog2State> st;
  std::vector<long long> nframes;
  std::vector<char> have_prev, has_frame, has_frame_buf[2];
  uint8_t* frames = nullptr;         // current step's frames: one of frame_buf[2], each [n][h*w*3]
  uint8_t* frame_buf[2] = {nullptr, nullptr};
  int cur = 0, pending = -1;         // buffer in use / buffer an upload has filled but gate() has not adopted yet
  uint8_t* fg = nullptr;             // [n][kMogMaxRepeats][npix]
  uint8_t* prev = nullptr;           // [n][npix]
  unsigned int* gate_counts = nullptr;   // [n]
  unsigned int* roi_counts = nullptr;    // [roi_cap]
  int roi_cap = 0;
  BankStream* desc_dev = nullptr;    // [kDescRing][n]
  BankStream* desc_host = nullptr;   // pinned mirror
  int ring_pos = 0;
  unsigned int* counts_pinned = nullptr;   // [max(n, roi_cap)]
  cudaEvent_t ev[6] = {};
  float last_ms[3] = {0.f, 0.f, 0.f};
  uint16_t k[kBlurMaxK];
};
static constexpr int kDescRing = 64;

extern "C" {

tcd_status tcd_stream_bank_create(int32_t n_streams, int32_t h, int32_t w, int32_t history, float var_threshold,
                                  int32_t detect_shadows, int32_t blur_size, int32_t device, tcd_stream_bank** out) {
  TCD_REQUIRE(out && n_streams >= 1 && n_streams <= 64 && h > 0 && w > 0 && history > 0, "stream_bank_create: bad argument");
  TCD_REQUIRE(blur_size == 21, "stream_bank_create: the batched kernels are the 21-tap case (got %d)", blur_size);
  TCD_CUDA_OK(cudaSetDevice(device));
  int rc = require_sm100();
  if (rc) return (tcd_status)rc;
  tcd_stream_bank* b = new tcd_stream_bank();
  b->n = n_streams; b->h = h; b->w = w; b->history = history; b->device = device;
  b->detect_shadows = detect_shadows; b->var_threshold = var_threshold; b->blur_size = blur_size;
  b->npix = (size_t)h * w;
  gaussian_kernel_q8(blur_size, b->k);
  b->st.resize(n_streams);
  b->nframes.assign(n_streams, 0);
  b->have_prev.assign(n_streams, 0);
  b->has_frame.assign(n_streams, 0);
  const size_t np = b->npix;
  TCD_CUDA_OK(cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking));
  TCD_CUDA_OK(cudaStreamCreateWithFlags(&b->copy, cudaStreamNonBlocking));
  for (auto& e : b->ev) TCD_CUDA_OK(cudaEventCreate(&e));
  for (auto& e : b->ev_up) TCD_CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  for (int k = 0; k < 2; ++k) b->has_frame_buf[k].assign(n_streams, 0);
  for (int i = 0; i < n_streams; ++i) {
    TCD_CUDA_OK(cudaMalloc((void**)&b->st[i].weight, np * kMogModes * 4));
    TCD_CUDA_OK(cudaMalloc((void**)&b->st[i].var, np * kMogModes * 4));
    TCD_CUDA_OK(cudaMalloc((void**)&b->st[i].mean, np * kMogModes * 3 * 4));
    TCD_CUDA_OK(cudaMalloc((void**)&b->st[i].nmodes, np));
    TCD_CUDA_OK(cudaMemset(b->st[i].nmodes, 0, np));
  }
  for (int k = 0; k < 2; ++k) TCD_CUDA_OK(cudaMalloc((void**)&b->frame_buf[k], np * 3 * n_streams));
  b->frames = b->frame_buf[0];
  TCD_CUDA_OK(cudaMalloc((void**)&b->fg, np * kMogMaxRepeats * n_streams));
  TCD_CUDA_OK(cudaMemset(b->fg, 0, np * kMogMaxRepeats * n_streams));
  TCD_CUDA_OK(cudaMalloc((void**)&b->prev, np * n_streams));
  TCD_CUDA_OK(cudaMalloc((void**)&b->gate_counts, sizeof(unsigned int) * n_streams));
  b->roi_cap = 4096;
  TCD_CUDA_OK(cudaMalloc((void**)&b->roi_counts, sizeof(unsigned int) * b->roi_cap));
  TCD_CUDA_OK(cudaMalloc((void**)&b->desc_dev, sizeof(BankStream) * kDescRing * n_streams));
  TCD_CUDA_OK(cudaMallocHost((void**)&b->desc_host, sizeof(BankStream) * kDescRing * n_streams));
  TCD_CUDA_OK(cudaMallocHost((void**)&b->counts_pinned, sizeof(unsigned int) * (size_t)std::max(n_streams, b->roi_cap)));
  *out = b;
  return TCD_OK;
}

void tcd_stream_bank_destroy(tcd_stream_bank* b) {
  if (!b) return;
  cudaSetDevice(b->device);
  cudaStreamSynchronize(b->stream);
  for (auto& s : b->st) {
    cudaFree(s.weight); cudaFree(s.var); cudaFree(s.mean); cudaFree(s.nmodes);
  }
  cudaStreamSynchronize(b->copy);
  void* ptrs[] = {b->frame_buf[0], b->frame_buf[1], b->fg, b->prev, b->gate_counts, b->roi_counts, b->desc_dev};
  for (void* p : ptrs) if (p) cudaFree(p);
  if (b->desc_host) cudaFreeHost(b->desc_host);
  if (b->counts_pinned) cudaFreeHost(b->counts_pinned);
  for (auto& e : b->ev) if (e) cudaEventDestroy(e);
  for (auto& e : b->ev_up) if (e) cudaEventDestroy(e);
  cudaStreamDestroy(b->copy);
  cudaStreamDestroy(b->stream);
  delete b;
}

tcd_status tcd_stream_bank_reset(tcd_stream_bank* b, int32_t stream, int32_t what) {
  TCD_REQUIRE(b && stream >= 0 && stream < b->n, "stream_bank_reset: bad argument");
  TCD_CUDA_OK(cudaSetDevice(b->device));
  if (what & 1) {   // background model (MotionFilter.reset_background)
    b->nframes[stream] = 0;
    TCD_CUDA_OK(cudaMemsetAsync(b->st[stream].nmodes, 0, b->npix, b->stream));
  }
  if (what & 2) b->have_prev[stream] = 0;   // gate (EmptyFrameFilter.reset)
  return TCD_OK;
}

/* frames_host[s]: the NEXT step's frame of stream s (HWC u8 BGR, row stride strides[s] or tight when strides is
 * NULL), NULL = the stream has no frame in that step.  One copy per frame on the bank's copy stream into the buffer
 * that is not in use: asynchronous, and it may be issued while the current step is still being filtered (the two
 * frame buffers alternate; tcd_stream_bank_gate adopts the uploaded one). */
tcd_status tcd_stream_bank_upload(tcd_stream_bank* b, const uint8_t* const* frames_host, const int64_t* strides) {
  TCD_REQUIRE(b && frames_host, "stream_bank_upload: null argument");
  TCD_REQUIRE(b->pending < 0, "stream_bank_upload: the previous upload has not been adopted by a gate call yet");
  TCD_CUDA_OK(cudaSetDevice(b->device));
  const size_t tight = (size_t)b->w * 3;
  const int k = b->cur ^ 1;
  for (int s = 0; s < b->n; ++s) {
    b->has_frame_buf[k][s] = frames_host[s] != nullptr;
    if (!frames_host[s]) continue;
    const size_t st = strides && strides[s] > 0 ? (size_t)strides[s] : tight;
    uint8_t* dst = b->frame_buf[k] + (size_t)s * b->npix * 3;
    if (st == tight) TCD_CUDA_OK(cudaMemcpyAsync(dst, frames_host[s], tight * b->h, cudaMemcpyHostToDevice, b->copy));
    else TCD_CUDA_OK(cudaMemcpy2DAsync(dst, tight, frames_host[s], st, tight, (size_t)b->h, cudaMemcpyHostToDevice, b->copy));
    g_h2d_bytes.fetch_add((uint64_t)tight * b->h, std::memory_order_relaxed);
  }
  TCD_CUDA_OK(cudaEventRecord(b->ev_up[k], b->copy));
  b->pending = k;
  return TCD_OK;
}

const uint8_t* tcd_stream_bank_frame(tcd_stream_bank* b, int32_t stream) {
  if (!b || stream < 0 || stream >= b->n) return nullptr;
  return b->frames + (size_t)stream * b->npix * 3;
}
void* tcd_stream_bank_cuda_stream(tcd_stream_bank* b) { return b ? (void*)b->stream : nullptr; }

static BankStream* bank_next_desc(tcd_stream_bank* b, BankStream** host) {
  if (b->ring_pos == kDescRing) {   // every slot of the descriptor ring is in flight: drain once
    cudaStreamSynchronize(b->stream);
    b->ring_pos = 0;
  }
  *host = b->desc_host + (size_t)b->ring_pos * b->n;
  return b->desc_dev + (size_t)(b->ring_pos++) * b->n;
}

/* EmptyFrameFilter.has_motion of every stream that got a frame in the last upload, ONE launch:
 * changed_host[s] = changed-pixel count, -1 = first frame of the stream, -2 = no frame this step. */
tcd_status tcd_stream_bank_gate(tcd_stream_bank* b, int32_t threshold, int32_t* changed_host) {
  TCD_REQUIRE(b && changed_host, "stream_bank_gate: null argument");
  TCD_CUDA_OK(cudaSetDevice(b->device));
  if (b->pending >= 0) {   // adopt the uploaded frames: they become "this step" for the gate, the detector and the model
    b->cur = b->pending;
    b->pending = -1;
    b->frames = b->frame_buf[b->cur];
    b->has_frame = b->has_frame_buf[b->cur];
    TCD_CUDA_OK(cudaStreamWaitEvent(b->stream, b->ev_up[b->cur], 0));
  }
  BankStream* host = nullptr;
  BankStream* dev = bank_next_desc(b, &host);
  for (int s = 0; s < b->n; ++s) {
    BankStream& d = host[s];
    memset(&d, 0, sizeof(d));
    d.frame = b->frames + (size_t)s * b->npix * 3;
    d.stride = (long long)b->w * 3;
    d.prev = b->prev + (size_t)s * b->npix;
    d.counter = b->gate_counts + s;
    d.active = b->has_frame[s] ? 1 : 0;
    d.first = b->have_prev[s] ? 0 : 1;
  }
  TCD_CUDA_OK(cudaMemcpyAsync(dev, host, sizeof(BankStream) * b->n, cudaMemcpyHostToDevice, b->stream));
  TCD_CUDA_OK(cudaMemsetAsync(b->gate_counts, 0, sizeof(unsigned int) * b->n, b->stream));
  BlurArgs a;
  memset(&a, 0, sizeof(a));
  a.w = b->w; a.h = b->h; a.ksize = b->blur_size; a.in_stride = (long long)b->w * 3;
  a.src_mode = 1; a.post_mode = 1; a.post_thr = threshold;
  memcpy(a.k, b->k, sizeof(a.k));
  cudaEventRecord(b->ev[1], b->stream);
  int rc = bank_gate_launch(dev, b->n, a, b->stream);
  if (rc) return (tcd_status)rc;
  cudaEventRecord(b->ev[2], b->stream);
  TCD_CUDA_OK(cudaMemcpyAsync(b->counts_pinned, b->gate_counts, sizeof(unsigned int) * b->n, cudaMemcpyDeviceToHost, b->stream));
  TCD_CUDA_OK(cudaStreamSynchronize(b->stream));
  b->ring_pos = 0;
  for (int s = 0; s < b->n; ++s) {
    if (!b->has_frame[s]) { changed_host[s] = -2; continue; }
    changed_host[s] = b->have_prev[s] ? (int32_t)b->counts_pinned[s] : -1;
    b->have_prev[s] = 1;
  }
  float t = 0.f;
  if (cudaEventElapsedTime(&t, b->ev[1], b->ev[2]) == cudaSuccess) b->last_ms[0] = t;
  return TCD_OK;
}

/* MotionFilter.filter_detections of every stream, batched.  Boxes of stream s: rois[offsets[s] .. offsets[s+1])
 * (clamped integer x1,y1,x2,y2; a degenerate box still costs its model update, like the reference's early return
 * after apply()).  Per box, in order: one background-model update with the stream's frame, (mask > shadow_thr),
 * 21x21 blur, (> blur_thr), count inside the box -> counts_host.  Rounds of up to 16 boxes per stream share one
 * model pass (state read and written once); nothing synchronises with the host until every count is back. */
tcd_status tcd_stream_bank_motion(tcd_stream_bank* b, const int32_t* rois, const int32_t* offsets, int32_t shadow_thr,
                                  int32_t ksize, int32_t blur_thr, int32_t* counts_host) {
  TCD_REQUIRE(b && offsets && (offsets[b->n] == 0 || (rois && counts_host)), "stream_bank_motion: null argument");
  TCD_REQUIRE(ksize == 21, "stream_bank_motion: the batched clean-up is the 21-tap case (got %d)", ksize);
  TCD_CUDA_OK(cudaSetDevice(b->device));
  const int total = offsets[b->n];
  TCD_REQUIRE(total <= b->roi_cap, "stream_bank_motion: %d boxes exceed capacity %d", total, b->roi_cap);
  if (total == 0) return TCD_OK;
  int max_n = 0;
  for (int s = 0; s < b->n; ++s) {
    const int n_s = offsets[s + 1] - offsets[s];
    TCD_REQUIRE(n_s >= 0 && (n_s == 0 || b->has_frame[s]), "stream_bank_motion: stream %d has boxes but no frame", s);
    max_n = std::max(max_n, n_s);
    for (int i = offsets[s]; i < offsets[s + 1]; ++i) {
      const int32_t* r = rois + 4 * i;
      TCD_REQUIRE(r[0] >= 0 && r[1] >= 0 && r[2] <= b->w && r[3] <= b->h, "stream_bank_motion: box %d not clamped", i);
    }
  }
  TCD_CUDA_OK(cudaMemsetAsync(b->roi_counts, 0, sizeof(unsigned int) * total, b->stream));
  Mog2Params p;
  memset(&p, 0, sizeof(p));
  p.npix = (int)b->npix; p.w = b->w; p.h = b->h; p.stride = (long long)b->w * 3;
  p.Tb = b->var_threshold; p.TB = 0.9f; p.Tg = 9.0f; p.varInit = 15.0f; p.varMin = 4.0f;
  p.varMax = 75.0f; p.CT = 0.05f; p.tau = 0.5f;
  p.detect_shadows = b->detect_shadows;
  BlurArgs a;
  memset(&a, 0, sizeof(a));
  a.w = b->w; a.h = b->h; a.ksize = ksize; a.in_stride = b->w;
  a.src_mode = 0; a.pre_thr = shadow_thr; a.post_mode = 2; a.post_thr = blur_thr;
  memcpy(a.k, b->k, sizeof(a.k));
  cudaEventRecord(b->ev[3], b->stream);
  int rounds = 0;
  for (int c0 = 0; c0 < max_n; c0 += kMogMaxRepeats, ++rounds) {
    BankStream* host = nullptr;
    BankStream* dev = bank_next_desc(b, &host);
    for (int s = 0; s < b->n; ++s) {
      BankStream& d = host[s];
      memset(&d, 0, sizeof(d));
      const int n_s = offsets[s + 1] - offsets[s];
      const int reps = std::max(0, std::min(kMogMaxRepeats, n_s - c0));
      d.frame = b->frames + (size_t)s * b->npix * 3;
      d.stride = (long long)b->w * 3;
      d.st = b->st[s];
      d.fg = b->fg + (size_t)s * kMogMaxRepeats * b->npix;
      d.roi_counts = b->roi_counts + offsets[s] + c0;
      d.repeats = reps;
      for (int i = 0; i < reps; ++i) {
        ++b->nframes[s];
        const long long den = 2 * b->nframes[s] < b->history ? 2 * b->nframes[s] : b->history;
        d.alpha[i] = (float)(1.0 / (double)den);
        memcpy(d.roi[i], rois + 4 * (offsets[s] + c0 + i), 4 * sizeof(int));
      }
    }
    TCD_CUDA_OK(cudaMemcpyAsync(dev, host, sizeof(BankStream) * b->n, cudaMemcpyHostToDevice, b->stream));
    int rc = bank_mog2_launch(dev, b->n, p, b->stream);
    if (rc == TCD_OK) rc = bank_roi_launch(dev, b->n, a, b->stream);
    if (rc) return (tcd_status)rc;
  }
  cudaEventRecord(b->ev[4], b->stream);
  TCD_CUDA_OK(cudaMemcpyAsync(b->counts_pinned, b->roi_counts, sizeof(unsigned int) * total, cudaMemcpyDeviceToHost, b->stream));
  TCD_CUDA_OK(cudaStreamSynchronize(b->stream));
  b->ring_pos = 0;
  for (int i = 0; i < total; ++i) counts_host[i] = (int32_t)b->counts_pinned[i];
  float t = 0.f;
  if (cudaEventElapsedTime(&t, b->ev[3], b->ev[4]) == cudaSuccess) { b->last_ms[1] = t; b->last_ms[2] = (float)rounds; }
  return TCD_OK;
}

/* [0] gate kernel ms of the last gate call, [1] model updates + clean-up + counts of the last motion call, [2] its rounds */
tcd_status tcd_stream_bank_last_ms(const tcd_stream_bank* b, float out_ms[3]) {
  TCD_REQUIRE(b && out_ms, "stream_bank_last_ms: null argument");
  for (int i = 0; i < 3; ++i) out_ms[i] = b->last_ms[i];
  return TCD_OK;
}

}  // extern "C"
