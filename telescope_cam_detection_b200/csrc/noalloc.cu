// tcd_enhance_clahe_bilateral_noalloc: the crop enhancer (enhance.cu) for ONE crop, with every byte of device scratch
// it needs -- descriptor, colour tables, bilateral colour weights, tile LUTs and the intermediate image -- in a
// file-scope static device array instead of cudaMalloc'ed workspace.  No counterpart in the reference; added
// for callers that cannot allocate (code under graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).
//
// The array costs kNoAllocScratchBytes (1 MiB) of device memory on every GPU that loads libtcd, until the process
// exits.  All calls share it and it keeps the last call's contents: nothing in it has to start at zero (the LUT and
// intermediate regions are fully rewritten for the crop at hand before they are read), but two calls that overlap on
// the GPU race on it -- the entry point synchronises the stream before it returns, so only calls from different host
// threads / streams can overlap, and the caller must serialise those.
//
// In host code the array's name is a host-side stand-in, not a device address: the address comes from
// cudaGetSymbolAddress, once per device, with that device current, in this file.
#include <cmath>
#include <cstring>
#include <mutex>

#include "kernels.cuh"
#include "tcd_common.cuh"

using namespace tcd;

namespace {

constexpr size_t kNoAllocScratchBytes = 1u << 20;
constexpr size_t kOffDesc = 0;          // EnhImg (256 B reserved)
constexpr size_t kOffGamma = 256;       // u16[256]
constexpr size_t kOffCbrt = 768;        // u16[3072]
constexpr size_t kOffYf = 6912;         // u16[512]
constexpr size_t kOffInv = 7936;        // u8[4096]
constexpr size_t kOffColorW = 12288;    // f32[768]
constexpr size_t kOffLuts = 15360;      // u8[64][256]
constexpr size_t kOffMid = 31744;       // intermediate image
constexpr int kNoAllocMaxTiles = 64;
constexpr size_t kNoAllocMaxImageBytes = 512 * 512 * 3;   // 786 432 <= kNoAllocScratchBytes - kOffMid
static_assert(kOffMid + kNoAllocMaxImageBytes <= kNoAllocScratchBytes, "scratch layout");
static_assert(sizeof(EnhImg) <= 256, "descriptor slot");

static __device__ __align__(256) uint8_t g_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
uint8_t* g_scratch_ptr[kMaxDevices] = {};
int32_t g_cf[9], g_ci[9];

// once per device, with that device current: the array's device address + the parameter-independent tables
int noalloc_setup(int device, uint8_t** out) {
  TCD_REQUIRE(device >= 0 && device < kMaxDevices, "enhance_noalloc: device %d", device);
  if (g_scratch_ptr[device]) {
    *out = g_scratch_ptr[device];
    return TCD_OK;
  }
  if (int rc = require_sm100()) return rc;
  uint8_t* ptr = nullptr;
  TCD_CUDA_OK(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
  TCD_REQUIRE(ptr != nullptr, "enhance_noalloc: cudaGetSymbolAddress returned a null address");
  static uint16_t gamma[256], cbrt[3072], yf[512];
  static uint8_t inv[4096];
  enhance_host_tables(gamma, cbrt, yf, inv, g_cf, g_ci);
  TCD_CUDA_OK(cudaMemcpy(ptr + kOffGamma, gamma, sizeof(gamma), cudaMemcpyHostToDevice));
  TCD_CUDA_OK(cudaMemcpy(ptr + kOffCbrt, cbrt, sizeof(cbrt), cudaMemcpyHostToDevice));
  TCD_CUDA_OK(cudaMemcpy(ptr + kOffYf, yf, sizeof(yf), cudaMemcpyHostToDevice));
  TCD_CUDA_OK(cudaMemcpy(ptr + kOffInv, inv, sizeof(inv), cudaMemcpyHostToDevice));
  g_scratch_ptr[device] = ptr;
  *out = ptr;
  return TCD_OK;
}

}  // namespace

extern "C" {

int64_t tcd_enhance_noalloc_max_image_bytes(void) { return (int64_t)kNoAllocMaxImageBytes; }

tcd_status tcd_enhance_clahe_bilateral_noalloc(const tcd_enhance_cfg* cfg, int32_t device, const uint8_t* img_dev, int32_t h,
                                               int32_t w, int64_t stride, int32_t channels, uint8_t* out_dev,
                                               int64_t out_stride, void* stream) {
  TCD_REQUIRE(cfg && img_dev && out_dev, "enhance_noalloc: null argument");
  TCD_REQUIRE(channels == 1 || channels == 3, "enhance_noalloc: %d channels (1 or 3)", channels);
  TCD_REQUIRE(cfg->grid_x >= 1 && cfg->grid_y >= 1 && cfg->grid_x * cfg->grid_y <= kNoAllocMaxTiles,
              "enhance_noalloc: tile grid %dx%d exceeds the %d tiles the static scratch holds", cfg->grid_x, cfg->grid_y,
              kNoAllocMaxTiles);
  TCD_REQUIRE(cfg->d >= 3 && cfg->d <= 15, "enhance_noalloc: bilateral diameter %d (3..15)", cfg->d);
  TCD_REQUIRE(cfg->sigma_color > 0 && cfg->sigma_space > 0, "enhance_noalloc: sigmas must be positive");
  TCD_REQUIRE(h >= 2 && w >= 2, "enhance_noalloc: crop is %dx%d", h, w);
  TCD_REQUIRE((size_t)h * (size_t)w * (size_t)channels <= kNoAllocMaxImageBytes,
              "enhance_noalloc: crop of %dx%dx%d = %zu bytes exceeds the static scratch maximum of %zu bytes (512x512x3); use "
              "tcd_enhance_clahe_bilateral for larger crops", h, w, channels, (size_t)h * w * channels, kNoAllocMaxImageBytes);
  const int radius = cfg->d / 2;
  TCD_REQUIRE(radius < h && radius < w, "enhance_noalloc: crop %dx%d is smaller than the bilateral radius %d", h, w, radius);
  std::lock_guard<std::mutex> lock(g_mu);
  TCD_CUDA_OK(cudaSetDevice(device));
  uint8_t* sc = nullptr;
  if (int rc = noalloc_setup(device, &sc)) return (tcd_status)rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int gx = cfg->grid_x, gy = cfg->grid_y, cn = channels;
  // bilateral tables on the stack (same construction as capi_enhance.cu: double exp, narrowed to float)
  int8_t dy[256], dx[256];
  float sw[256], cw[768];
  int maxk = 0;
  const double gcc = -0.5 / ((double)cfg->sigma_color * cfg->sigma_color), gsc = -0.5 / ((double)cfg->sigma_space * cfg->sigma_space);
  for (int i = -radius; i <= radius; ++i)
    for (int j = -radius; j <= radius; ++j) {
      const double r = std::sqrt((double)i * i + (double)j * j);
      if (r > radius || (i == 0 && j == 0)) continue;
      sw[maxk] = (float)std::exp(r * r * gsc);
      dy[maxk] = (int8_t)i;
      dx[maxk] = (int8_t)j;
      ++maxk;
    }
  for (int i = 0; i < 256 * cn; ++i) cw[i] = (float)std::exp((double)i * i * gcc);
  EnhImg m;
  memset(&m, 0, sizeof(m));
  const size_t tight = (size_t)w * cn;
  m.src = img_dev;
  m.src_stride = (long long)(stride > 0 ? (size_t)stride : tight);
  m.mid = sc + kOffMid;
  m.dst = out_dev;
  m.dst_stride = (long long)(out_stride > 0 ? (size_t)out_stride : tight);
  m.luts = sc + kOffLuts;
  m.h = h;
  m.w = w;
  m.cn = cn;
  int eh = h, ew = w;
  if (h % gy || w % gx) {   // the library pads BOTH axes as soon as one of them is not a multiple of the grid
    eh = h + (gy - h % gy);
    ew = w + (gx - w % gx);
  }
  m.th = eh / gy;
  m.tw = ew / gx;
  const int area = m.th * m.tw;
  m.lut_scale = 255.0f / (float)area;
  m.clip = 0;
  if (cfg->clip_limit > 0) {
    m.clip = (int)((double)cfg->clip_limit * area / 256);
    if (m.clip < 1) m.clip = 1;
  }
  m.inv_tw = 1.0f / (float)m.tw;
  m.inv_th = 1.0f / (float)m.th;
  TCD_CUDA_OK(cudaMemcpyAsync(sc + kOffDesc, &m, sizeof(m), cudaMemcpyHostToDevice, st));
  TCD_CUDA_OK(cudaMemcpyAsync(sc + kOffColorW, cw, sizeof(float) * 256 * cn, cudaMemcpyHostToDevice, st));
  if (int rc = enhance_set_constants(g_cf, g_ci)) return (tcd_status)rc;
  if (int rc = enhance_set_window(dy, dx, sw, maxk)) return (tcd_status)rc;
  EnhTables t;
  t.gamma = (const uint16_t*)(sc + kOffGamma);
  t.cbrt = (const uint16_t*)(sc + kOffCbrt);
  t.yf = (const uint16_t*)(sc + kOffYf);
  t.inv = sc + kOffInv;
  if (int rc = enhance_launch((const EnhImg*)(sc + kOffDesc), 1, cn, gx, gy, h, w, t, (const float*)(sc + kOffColorW), radius, maxk, st))
    return (tcd_status)rc;
  TCD_CUDA_OK(cudaStreamSynchronize(st));   // the stack tables above and the shared scratch are done with
  return TCD_OK;
}

}  // extern "C"
