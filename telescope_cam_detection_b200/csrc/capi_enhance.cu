// C ABI of the crop enhancer (include/tcd.h, "crop enhancement"): a handle owns the device tables, the bilateral
// window and a workspace that grows with the largest batch seen.
#include <cmath>
#include <mutex>
#include <vector>

#include "kernels.cuh"
#include "tcd_common.cuh"

using namespace tcd;

struct tcd_enhancer {
  int device = 0;
  tcd_enhance_cfg cfg{};
  cudaStream_t stream = nullptr;
  uint8_t* tables = nullptr;      // gamma | cbrt | yf | inv
  EnhTables t{};
  float* color_w[2] = {nullptr, nullptr};   // [256] (1 channel), [768] (3 channels)
  int radius = 0, maxk = 0;
  std::vector<int8_t> dy, dx;
  std::vector<float> sw;
  int32_t cf[9], ci[9];
  // workspace
  uint8_t* ws = nullptr;
  size_t ws_bytes = 0;
  EnhImg* desc_dev = nullptr;
  int desc_cap = 0;
  std::mutex mu;
};

extern "C" {

tcd_status tcd_enhance_tables(uint16_t* gamma, uint16_t* cbrt, uint16_t* yf, uint8_t* inv, int32_t* cf, int32_t* ci) {
  TCD_REQUIRE(gamma && cbrt && yf && inv && cf && ci, "enhance_tables: null argument");
  enhance_host_tables(gamma, cbrt, yf, inv, cf, ci);
  return TCD_OK;
}

tcd_status tcd_enhancer_create(const tcd_enhance_cfg* cfg, int32_t device, tcd_enhancer** out) {
  TCD_REQUIRE(cfg && out, "enhancer_create: null argument");
  TCD_REQUIRE(cfg->grid_x >= 1 && cfg->grid_y >= 1 && cfg->grid_x * cfg->grid_y <= 4096, "enhancer_create: tile grid %dx%d",
              cfg->grid_x, cfg->grid_y);
  TCD_REQUIRE(cfg->d >= 3 && cfg->d <= 15, "enhancer_create: bilateral diameter %d (3..15; the sigma-derived default of d <= 0 "
              "is not built)", cfg->d);
  TCD_REQUIRE(cfg->sigma_color > 0 && cfg->sigma_space > 0, "enhancer_create: sigmas must be positive");
  TCD_CUDA_OK(cudaSetDevice(device));
  if (int rc = require_sm100()) return (tcd_status)rc;
  tcd_enhancer* e = new tcd_enhancer();
  e->device = device;
  e->cfg = *cfg;
  std::vector<uint16_t> gamma(256), cbrt(3072), yf(512);
  std::vector<uint8_t> inv(4096);
  enhance_host_tables(gamma.data(), cbrt.data(), yf.data(), inv.data(), e->cf, e->ci);
  const size_t tb = 512 + 6144 + 1024 + 4096;
  cudaError_t ce = cudaMalloc((void**)&e->tables, tb);
  if (ce == cudaSuccess) ce = cudaMemcpy(e->tables, gamma.data(), 512, cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMemcpy(e->tables + 512, cbrt.data(), 6144, cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMemcpy(e->tables + 6656, yf.data(), 1024, cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMemcpy(e->tables + 7680, inv.data(), 4096, cudaMemcpyHostToDevice);
  e->t.gamma = (const uint16_t*)e->tables;
  e->t.cbrt = (const uint16_t*)(e->tables + 512);
  e->t.yf = (const uint16_t*)(e->tables + 6656);
  e->t.inv = e->tables + 7680;
  // bilateral tables, as the library builds them: double exp, narrowed to float
  e->radius = cfg->d / 2;
  const double gcc = -0.5 / ((double)cfg->sigma_color * cfg->sigma_color), gsc = -0.5 / ((double)cfg->sigma_space * cfg->sigma_space);
  for (int i = -e->radius; i <= e->radius; ++i)
    for (int j = -e->radius; j <= e->radius; ++j) {
      const double r = std::sqrt((double)i * i + (double)j * j);
      if (r > e->radius || (i == 0 && j == 0)) continue;
      e->sw.push_back((float)std::exp(r * r * gsc));
      e->dy.push_back((int8_t)i);
      e->dx.push_back((int8_t)j);
    }
  e->maxk = (int)e->sw.size();
  for (int cn = 1; cn <= 3 && ce == cudaSuccess; cn += 2) {
    std::vector<float> cw(256 * cn);
    for (int i = 0; i < 256 * cn; ++i) cw[i] = (float)std::exp((double)i * i * gcc);
    float*& dst = e->color_w[cn == 3];
    ce = cudaMalloc((void**)&dst, cw.size() * sizeof(float));
    if (ce == cudaSuccess) ce = cudaMemcpy(dst, cw.data(), cw.size() * sizeof(float), cudaMemcpyHostToDevice);
  }
  if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
  if (ce != cudaSuccess) {
    set_last_error("enhancer_create: %s", cudaGetErrorString(ce));
    delete e;
    return TCD_ERR_CUDA;
  }
  *out = e;
  return TCD_OK;
}

void tcd_enhancer_destroy(tcd_enhancer* e) {
  if (!e) return;
  cudaSetDevice(e->device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  cudaFree(e->tables);
  cudaFree(e->color_w[0]);
  cudaFree(e->color_w[1]);
  cudaFree(e->ws);
  cudaFree(e->desc_dev);
  if (e->stream) cudaStreamDestroy(e->stream);
  delete e;
}

/* n crops of `channels` (1 or 3) channels each.  imgs[i] / outs[i]: HWC u8, host or device as the two flags say
 * (device crops may be views into a frame: pointer + row stride).  Host crops are staged through the workspace; the
 * call returns when the outputs are complete (host outputs) or enqueued on the enhancer's stream and synchronised
 * (device outputs), so the caller may use them at once. */
tcd_status tcd_enhance_clahe_bilateral(tcd_enhancer* e, const uint8_t* const* imgs, const int32_t* h, const int32_t* w,
                                       const int64_t* stride, int32_t n, int32_t channels, int32_t imgs_on_device,
                                       uint8_t* const* outs, const int64_t* out_stride, int32_t outs_on_device) {
  TCD_REQUIRE(e && imgs && h && w && outs && n >= 1, "enhance: null argument");
  TCD_REQUIRE(channels == 1 || channels == 3, "enhance: %d channels (1 or 3)", channels);
  std::lock_guard<std::mutex> lock(e->mu);
  TCD_CUDA_OK(cudaSetDevice(e->device));
  const int gx = e->cfg.grid_x, gy = e->cfg.grid_y, cn = channels;
  // workspace layout per crop: [staged source (host input)] [mid] [staged output (host output)] [luts]
  std::vector<EnhImg> d(n);
  std::vector<size_t> off_src(n), off_mid(n), off_out(n), off_lut(n);
  size_t need = 0;
  int max_h = 0, max_w = 0;
  auto take = [&](size_t bytes) { const size_t o = need; need += (bytes + 255) & ~(size_t)255; return o; };
  for (int i = 0; i < n; ++i) {
    TCD_REQUIRE(h[i] >= 2 && w[i] >= 2 && h[i] <= 8192 && w[i] <= 8192, "enhance: crop %d is %dx%d", i, h[i], w[i]);
    TCD_REQUIRE(e->radius < h[i] && e->radius < w[i], "enhance: crop %d (%dx%d) is smaller than the bilateral radius %d", i, h[i],
                w[i], e->radius);
    const size_t img_bytes = (size_t)h[i] * w[i] * cn;
    off_src[i] = imgs_on_device ? 0 : take(img_bytes);
    off_mid[i] = take(img_bytes);
    off_out[i] = outs_on_device ? 0 : take(img_bytes);
    off_lut[i] = take((size_t)gx * gy * 256);
    max_h = std::max(max_h, (int)h[i]);
    max_w = std::max(max_w, (int)w[i]);
  }
  if (need > e->ws_bytes) {
    TCD_CUDA_OK(cudaStreamSynchronize(e->stream));
    cudaFree(e->ws);
    e->ws = nullptr;
    e->ws_bytes = 0;
    TCD_CUDA_OK(cudaMalloc((void**)&e->ws, need));
    e->ws_bytes = need;
  }
  if (n > e->desc_cap) {
    TCD_CUDA_OK(cudaStreamSynchronize(e->stream));
    cudaFree(e->desc_dev);
    e->desc_dev = nullptr;
    e->desc_cap = 0;
    TCD_CUDA_OK(cudaMalloc((void**)&e->desc_dev, sizeof(EnhImg) * (size_t)n));
    e->desc_cap = n;
  }
  for (int i = 0; i < n; ++i) {
    const size_t tight = (size_t)w[i] * cn;
    const size_t st = stride && stride[i] > 0 ? (size_t)stride[i] : tight;
    EnhImg& m = d[i];
    if (imgs_on_device) {
      m.src = imgs[i];
      m.src_stride = (long long)st;
    } else {
      TCD_CUDA_OK(cudaMemcpy2DAsync(e->ws + off_src[i], tight, imgs[i], st, tight, (size_t)h[i], cudaMemcpyHostToDevice, e->stream));
      m.src = e->ws + off_src[i];
      m.src_stride = (long long)tight;
    }
    m.mid = e->ws + off_mid[i];
    const size_t ost = out_stride && out_stride[i] > 0 ? (size_t)out_stride[i] : tight;
    if (outs_on_device) {
      m.dst = outs[i];
      m.dst_stride = (long long)ost;
    } else {
      m.dst = e->ws + off_out[i];
      m.dst_stride = (long long)tight;
    }
    m.luts = e->ws + off_lut[i];
    m.h = h[i];
    m.w = w[i];
    m.cn = cn;
    // the library pads BOTH axes as soon as one of them is not a multiple of the grid
    int eh = h[i], ew = w[i];
    if (h[i] % gy || w[i] % gx) {
      eh = h[i] + (gy - h[i] % gy);
      ew = w[i] + (gx - w[i] % gx);
    }
    m.th = eh / gy;
    m.tw = ew / gx;
    const int area = m.th * m.tw;
    m.lut_scale = 255.0f / (float)area;
    m.clip = 0;
    if (e->cfg.clip_limit > 0) m.clip = std::max((int)((double)e->cfg.clip_limit * area / 256), 1);
    m.inv_tw = 1.0f / (float)m.tw;
    m.inv_th = 1.0f / (float)m.th;
  }
  TCD_CUDA_OK(cudaMemcpyAsync(e->desc_dev, d.data(), sizeof(EnhImg) * (size_t)n, cudaMemcpyHostToDevice, e->stream));
  if (int rc = enhance_set_constants(e->cf, e->ci)) return (tcd_status)rc;
  if (int rc = enhance_set_window(e->dy.data(), e->dx.data(), e->sw.data(), e->maxk)) return (tcd_status)rc;
  if (int rc = enhance_launch(e->desc_dev, n, cn, gx, gy, max_h, max_w, e->t, e->color_w[cn == 3], e->radius, e->maxk, e->stream))
    return (tcd_status)rc;
  if (!outs_on_device)
    for (int i = 0; i < n; ++i) {
      const size_t tight = (size_t)w[i] * cn;
      const size_t ost = out_stride && out_stride[i] > 0 ? (size_t)out_stride[i] : tight;
      TCD_CUDA_OK(cudaMemcpy2DAsync(outs[i], ost, e->ws + off_out[i], tight, tight, (size_t)h[i], cudaMemcpyDeviceToHost, e->stream));
    }
  TCD_CUDA_OK(cudaStreamSynchronize(e->stream));   // `d` (pageable descriptors) and the host crops are done with
  return TCD_OK;
}

tcd_status tcd_enhancer_stream(tcd_enhancer* e, void** stream) {
  TCD_REQUIRE(e && stream, "enhancer_stream: null argument");
  *stream = (void*)e->stream;
  return TCD_OK;
}

}  // extern "C"
