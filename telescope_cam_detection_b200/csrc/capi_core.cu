// C ABI glue, part 1: error plumbing, utilities, stand-alone kernel entries
// (tcd_preprocess, tcd_conv2d_nhwc_bf16, tcd_crop_resize).  Declarations: include/tcd.h.
#include <atomic>
#include <map>
#include <mutex>
#include <stdarg.h>
#include <string.h>
#include <vector>

#include "conv_umma.cuh"
#include "kernels.cuh"
#include "tfm_kernels.cuh"

namespace tcd {

std::atomic<uint64_t> g_launch_count{0};
std::atomic<uint64_t> g_h2d_bytes{0};
static thread_local char g_err[1024] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int ensure_dyn_smem(const void* fn, size_t bytes) {
  if (bytes <= 48 * 1024) return TCD_OK;
  static std::mutex mu;
  static std::map<std::pair<int, const void*>, size_t> high;
  int dev = 0;
  TCD_CUDA_OK(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lk(mu);
  size_t& h = high[std::make_pair(dev, fn)];
  if (bytes <= h) return TCD_OK;
  TCD_CUDA_OK(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  h = bytes;
  return TCD_OK;
}

// Refuse to run anywhere but on a Blackwell data-centre part: no fallback path exists.
int require_sm100() {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    set_last_error("no CUDA device: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  int major = 0;
  e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  if (e != cudaSuccess || major != 10) {
    set_last_error("libtcd needs an sm_100 GPU (found compute capability major %d): %s", major,
                   e == cudaSuccess ? "unsupported device" : cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return TCD_OK;
}

}  // namespace tcd

using namespace tcd;

extern "C" {

const char* tcd_last_error(void) { return g_err; }
int tcd_version(void) { return 100; }
uint64_t tcd_launch_count(void) { return g_launch_count.load(); }
uint64_t tcd_h2d_byte_count(void) { return g_h2d_bytes.load(); }

tcd_status tcd_malloc_device(size_t bytes, void** out) {
  TCD_CUDA_OK(cudaMalloc(out, bytes));
  return TCD_OK;
}
tcd_status tcd_free_device(void* p) {
  TCD_CUDA_OK(cudaFree(p));
  return TCD_OK;
}
tcd_status tcd_malloc_pinned(size_t bytes, void** out) {
  TCD_CUDA_OK(cudaMallocHost(out, bytes));
  return TCD_OK;
}
tcd_status tcd_free_pinned(void* p) {
  TCD_CUDA_OK(cudaFreeHost(p));
  return TCD_OK;
}
tcd_status tcd_memcpy_h2d(void* dst, const void* src, size_t bytes) {
  TCD_CUDA_OK(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
  return TCD_OK;
}
tcd_status tcd_memcpy_d2h(void* dst, const void* src, size_t bytes) {
  TCD_CUDA_OK(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost));
  return TCD_OK;
}
tcd_status tcd_device_synchronize(void) {
  TCD_CUDA_OK(cudaDeviceSynchronize());
  return TCD_OK;
}

tcd_status tcd_preprocess(const tcd_preproc_cfg* cfg, const uint8_t* const* frames, const int32_t* h,
                          const int32_t* w, const int64_t* row_stride, int32_t B, int32_t out_layout,
                          void* out_dev, void* stream) {
  TCD_REQUIRE(cfg && frames && h && w && out_dev && B > 0, "tcd_preprocess: null argument");
  TCD_REQUIRE(cfg->out_h % 2 == 0 && cfg->out_w % 2 == 0, "tcd_preprocess: output size must be even");
  int rc = require_sm100();
  if (rc) return (tcd_status)rc;
  std::vector<PreFrame> pf(B);
  for (int b = 0; b < B; ++b)
    preproc_describe(cfg, frames[b], h[b], w[b], row_stride ? row_stride[b] : (int64_t)w[b] * 3, &pf[b]);
  PreFrame* dev = nullptr;
  TCD_CUDA_OK(cudaMalloc((void**)&dev, sizeof(PreFrame) * B));
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemcpyAsync(dev, pf.data(), sizeof(PreFrame) * B, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) {
    rc = preproc_launch(cfg, pf.data(), dev, B, out_layout, out_dev, st);
    e = cudaStreamSynchronize(st);
  }
  cudaFree(dev);
  if (e != cudaSuccess) {
    set_last_error("tcd_preprocess: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return (tcd_status)rc;
}

tcd_status tcd_conv2d_nhwc_bf16(const void* x, int32_t N, int32_t H, int32_t W, int32_t x_ctot,
                                int32_t x_coff, int32_t cin, const float* w, const float* bias,
                                int32_t cout, int32_t k, int32_t stride, int32_t silu,
                                const void* resid, int32_t r_ctot, int32_t r_coff, void* y,
                                int32_t y_ctot, int32_t y_coff, void* stream) {
  return tcd_conv2d_nhwc_bf16_grouped(x, N, H, W, x_ctot, x_coff, cin, w, bias, cout, k, stride, silu, resid, r_ctot,
                                      r_coff, y, y_ctot, y_coff, 1, stream);
}

tcd_status tcd_conv2d_nhwc_bf16_grouped(const void* x, int32_t N, int32_t H, int32_t W, int32_t x_ctot,
                                        int32_t x_coff, int32_t cin, const float* w, const float* bias,
                                        int32_t cout, int32_t k, int32_t stride, int32_t silu,
                                        const void* resid, int32_t r_ctot, int32_t r_coff, void* y,
                                        int32_t y_ctot, int32_t y_coff, int32_t groups, void* stream) {
  TCD_REQUIRE(x && w && bias && y && groups >= 1, "tcd_conv2d: null argument");
  int rc = require_sm100();
  if (rc) return (tcd_status)rc;
  const int cin_pad = (cin + 15) / 16 * 16;
  std::vector<uint16_t> hw((size_t)k * k * cout * cin_pad);
  conv_pack_weights(w, cout, cin, k, cout, cin_pad, hw.data());
  void* dw = nullptr;
  float* db = nullptr;
  TCD_CUDA_OK(cudaMalloc(&dw, hw.size() * 2));
  TCD_CUDA_OK(cudaMalloc((void**)&db, sizeof(float) * cout));
  TCD_CUDA_OK(cudaMemcpy(dw, hw.data(), hw.size() * 2, cudaMemcpyHostToDevice));
  TCD_CUDA_OK(cudaMemcpy(db, bias, sizeof(float) * cout, cudaMemcpyHostToDevice));
  ConvDesc d;
  d.x = x; d.N = N; d.H = H; d.W = W; d.x_ctot = x_ctot; d.x_coff = x_coff; d.cin = cin_pad;
  d.w = dw; d.bias = db; d.cout = cout; d.cout_pad = cout; d.k = k; d.stride = stride; d.silu = silu;
  d.y = y; d.y_ctot = y_ctot; d.y_coff = y_coff;
  d.resid = resid; d.r_ctot = r_ctot; d.r_coff = r_coff;
  d.groups = groups;
  ConvPlan plan;
  rc = conv_plan_build(d, &plan);
  if (rc == TCD_OK) rc = conv_plan_launch(plan, (cudaStream_t)stream);
  cudaError_t e = cudaStreamSynchronize((cudaStream_t)stream);
  cudaFree(dw);
  cudaFree(db);
  if (rc == TCD_OK && e != cudaSuccess) {
    set_last_error("tcd_conv2d: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return (tcd_status)rc;
}

tcd_status tcd_crop_resize(const uint8_t* frame_dev, int32_t h, int32_t w, int64_t row_stride,
                           const int32_t* boxes_host, int32_t n, int32_t S,
                           const tcd_preproc_cfg* cfg, int32_t out_layout, void* out_dev,
                           void* stream) {
  TCD_REQUIRE(frame_dev && cfg && (n == 0 || (boxes_host && out_dev)), "tcd_crop_resize: null argument");
  if (n == 0) return TCD_OK;
  int rc = require_sm100();
  if (rc) return (tcd_status)rc;
  for (int i = 0; i < n; ++i) {
    const int32_t* b = boxes_host + 4 * i;
    TCD_REQUIRE(b[0] >= 0 && b[1] >= 0 && b[2] <= w && b[3] <= h, "tcd_crop_resize: box %d outside frame", i);
  }
  int* dev = nullptr;
  TCD_CUDA_OK(cudaMalloc((void**)&dev, sizeof(int) * 4 * n));
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemcpyAsync(dev, boxes_host, sizeof(int) * 4 * n, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) {
    rc = crop_resize_launch(frame_dev, h, w, row_stride, dev, n, S, cfg, out_layout, out_dev, st);
    e = cudaStreamSynchronize(st);
  }
  cudaFree(dev);
  if (e != cudaSuccess) {
    set_last_error("tcd_crop_resize: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return (tcd_status)rc;
}

tcd_status tcd_crop_patches(const uint8_t* frame_dev, int32_t h, int32_t w, int64_t row_stride,
                            const int32_t* boxes_host, int32_t* boxes_scratch_dev, int32_t n, int32_t S, int32_t patch,
                            int32_t k_pad, const tcd_preproc_cfg* cfg, void* rows_dev, void* stream) {
  TCD_REQUIRE(frame_dev && cfg && (n == 0 || (boxes_host && boxes_scratch_dev && rows_dev)), "tcd_crop_patches: null argument");
  if (n == 0) return TCD_OK;
  int rc = require_sm100();
  if (rc) return (tcd_status)rc;
  for (int i = 0; i < n; ++i) {
    const int32_t* b = boxes_host + 4 * i;
    TCD_REQUIRE(b[0] >= 0 && b[1] >= 0 && b[2] <= w && b[3] <= h, "tcd_crop_patches: box %d outside frame", i);
  }
  cudaStream_t st = (cudaStream_t)stream;
  TCD_CUDA_OK(cudaMemcpyAsync(boxes_scratch_dev, boxes_host, sizeof(int) * 4 * n, cudaMemcpyHostToDevice, st));
  return (tcd_status)crop_patch_launch(frame_dev, row_stride, boxes_scratch_dev, n, S, patch, k_pad, cfg, rows_dev, st);
}

tcd_status tcd_patchify_f32(const float* x_dev, int32_t n, int32_t S, int32_t patch, int32_t k_pad, void* rows_dev,
                            void* stream) {
  TCD_REQUIRE(x_dev && rows_dev && n >= 1, "tcd_patchify_f32: bad argument");
  int rc = require_sm100();
  if (rc) return (tcd_status)rc;
  return (tcd_status)patchify_f32_launch(x_dev, rows_dev, n, S, patch, k_pad, (cudaStream_t)stream);
}

}  // extern "C"
