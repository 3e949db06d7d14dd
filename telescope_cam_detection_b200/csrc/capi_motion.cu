// C ABI glue, part 3: MOG2 motion filter and empty-frame gate handles.
#include <math.h>
#include <string.h>
#include <vector>

#include "kernels.cuh"

using namespace tcd;

namespace tcd {

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
