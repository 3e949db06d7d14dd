// RT-DETR preprocess: Pillow's antialiased BILINEAR resize, restated for the GPU.
//
// Replaces RTDETRDetector.preprocess (src/rtdetr_detector.py:206-236): GPU tensor -> CPU copy (:217-219),
// cv2 BGR->RGB (:226), T.ToPILImage + T.Resize (PIL Image.resize, BILINEAR = triangle filter whose
// support grows with the decimation factor) + T.ToTensor (/255) (:176-180,231).  Pillow's algorithm
// (libImaging/Resample.c): per output coordinate a window [xmin, xmin+n) of normalised triangle weights
// in 22-bit fixed point; horizontal pass over all rows -> u8, then vertical pass -> u8, each rounded as
// clip8((sum + 2^21) >> 22).  The weight tables are built on the host in double precision exactly like
// Pillow does (they depend only on the two sizes); the two passes are integer kernels, so the resized
// u8 image is bit-exact against PIL.  Output: bf16 NHWC with 16 channels (R,G,B,/255 in 0..2, zeros
// above) = the layout the stem conv's TMA boxes read.
#include <math.h>
#include <map>
#include <mutex>
#include <vector>

#include "tfm_kernels.cuh"

namespace tcd {
int require_sm100();

namespace {

struct CoeffTable {
  int in_size = 0, out_size = 0, ksize = 0;
  int* dev = nullptr;  // [out][2 + ksize]: xmin, count, kk...
};

// Pillow precompute_coeffs + normalize_coeffs_8bpc, BILINEAR
std::vector<int> build_coeffs(int in_size, int out_size, int* ksize_out) {
  const double scale = (double)in_size / (double)out_size;
  const double filterscale = scale < 1.0 ? 1.0 : scale;
  const double support = 1.0 * filterscale;
  const int ksize = (int)ceil(support) * 2 + 1;
  std::vector<int> t((size_t)out_size * (2 + ksize), 0);
  std::vector<double> k(ksize);
  const double ss = 1.0 / filterscale;
  for (int xx = 0; xx < out_size; ++xx) {
    const double center = (xx + 0.5) * scale;
    int xmin = (int)(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    const int n = xmax - xmin;
    double ww = 0.0;
    for (int x = 0; x < n; ++x) {
      double a = (x + xmin - center + 0.5) * ss;
      if (a < 0) a = -a;
      const double w = a < 1.0 ? 1.0 - a : 0.0;
      k[x] = w;
      ww += w;
    }
    int* row = t.data() + (size_t)xx * (2 + ksize);
    row[0] = xmin;
    row[1] = n;
    for (int x = 0; x < n; ++x) {
      double v = ww != 0.0 ? k[x] / ww : k[x];
      v *= (double)(1 << 22);
      row[2 + x] = v < 0 ? (int)(v - 0.5) : (int)(v + 0.5);
    }
  }
  *ksize_out = ksize;
  return t;
}

std::mutex g_mu;
std::map<std::pair<int, int>, CoeffTable> g_tables;

int get_table(int in_size, int out_size, CoeffTable* out) {
  std::lock_guard<std::mutex> lock(g_mu);
  auto key = std::make_pair(in_size, out_size);
  auto it = g_tables.find(key);
  if (it == g_tables.end()) {
    CoeffTable t;
    t.in_size = in_size;
    t.out_size = out_size;
    std::vector<int> h = build_coeffs(in_size, out_size, &t.ksize);
    TCD_CUDA_OK(cudaMalloc((void**)&t.dev, h.size() * sizeof(int)));
    TCD_CUDA_OK(cudaMemcpy(t.dev, h.data(), h.size() * sizeof(int), cudaMemcpyHostToDevice));
    it = g_tables.emplace(key, t).first;
  }
  *out = it->second;
  return TCD_OK;
}

// grow-only scratch (staging of host frames + horizontal-pass image), one per calling thread
struct Scratch {
  uint8_t* p = nullptr;
  size_t cap = 0;
};
thread_local Scratch g_stage, g_tmp, g_u8;
int ensure(Scratch& s, size_t bytes) {
  if (bytes <= s.cap) return TCD_OK;
  if (s.p) cudaFree(s.p);
  s.p = nullptr;
  s.cap = 0;
  TCD_CUDA_OK(cudaMalloc((void**)&s.p, bytes));
  s.cap = bytes;
  return TCD_OK;
}

}  // namespace

// horizontal pass: src HWC u8 (3 ch) -> tmp [h][out_w][3] u8
__global__ void __launch_bounds__(256)
pil_h_kernel(const uint8_t* __restrict__ src, long long stride, int h, int out_w, const int* __restrict__ tab,
             int ksize, uint8_t* __restrict__ tmp) {
  const int xx = blockIdx.x * 64 + (threadIdx.x & 63);
  const int y = blockIdx.y * 4 + (threadIdx.x >> 6);
  if (xx >= out_w || y >= h) return;
  const int* row = tab + (size_t)xx * (2 + ksize);
  const int xmin = row[0], n = row[1];
  int a0 = 1 << 21, a1 = 1 << 21, a2 = 1 << 21;
  const uint8_t* p = src + (size_t)y * stride + (size_t)xmin * 3;
  for (int x = 0; x < n; ++x) {
    const int k = row[2 + x];
    a0 += p[3 * x] * k;
    a1 += p[3 * x + 1] * k;
    a2 += p[3 * x + 2] * k;
  }
  uint8_t* o = tmp + ((size_t)y * out_w + xx) * 3;
  o[0] = (uint8_t)min(max(a0 >> 22, 0), 255);
  o[1] = (uint8_t)min(max(a1 >> 22, 0), 255);
  o[2] = (uint8_t)min(max(a2 >> 22, 0), 255);
}

// vertical pass + BGR->RGB + /255 -> bf16 NHWC16 (and optionally the u8 RGB image)
__global__ void __launch_bounds__(256)
pil_v_kernel(const uint8_t* __restrict__ tmp, int out_h, int out_w, const int* __restrict__ tab, int ksize,
             __nv_bfloat16* __restrict__ out, uint8_t* __restrict__ out_u8) {
  const int xx = blockIdx.x * 64 + (threadIdx.x & 63);
  const int yy = blockIdx.y * 4 + (threadIdx.x >> 6);
  if (xx >= out_w || yy >= out_h) return;
  const int* row = tab + (size_t)yy * (2 + ksize);
  const int ymin = row[0], n = row[1];
  int a0 = 1 << 21, a1 = 1 << 21, a2 = 1 << 21;
  for (int y = 0; y < n; ++y) {
    const int k = row[2 + y];
    const uint8_t* p = tmp + ((size_t)(ymin + y) * out_w + xx) * 3;
    a0 += p[0] * k;
    a1 += p[1] * k;
    a2 += p[2] * k;
  }
  const int b = min(max(a0 >> 22, 0), 255), g = min(max(a1 >> 22, 0), 255), r = min(max(a2 >> 22, 0), 255);
  const size_t pix = (size_t)yy * out_w + xx;
  if (out_u8) {
    out_u8[pix * 3] = (uint8_t)r;
    out_u8[pix * 3 + 1] = (uint8_t)g;
    out_u8[pix * 3 + 2] = (uint8_t)b;
  }
  uint4 q0, q1 = make_uint4(0, 0, 0, 0);
  q0.x = pack_bf16x2((float)r / 255.0f, (float)g / 255.0f);
  q0.y = pack_bf16x2((float)b / 255.0f, 0.f);
  q0.z = 0;
  q0.w = 0;
  uint4* o = reinterpret_cast<uint4*>(out + pix * 16);
  o[0] = q0;
  o[1] = q1;
}

}  // namespace tcd

using namespace tcd;

extern "C" tcd_status tcd_pil_resize(const uint8_t* const* frames, const int32_t* h, const int32_t* w,
                                     const int64_t* row_stride, int32_t n, int32_t frames_on_device,
                                     int32_t out_h, int32_t out_w, void* out_dev, uint8_t* out_u8_host,
                                     void* stream) {
  TCD_REQUIRE(frames && h && w && out_dev && n >= 1 && out_h > 0 && out_w > 0, "pil_resize: bad argument");
  int rc = require_sm100();
  if (rc) return (tcd_status)rc;
  cudaStream_t st = (cudaStream_t)stream;
  size_t stage_bytes = 0, tmp_bytes = 0;
  for (int i = 0; i < n; ++i) {
    TCD_REQUIRE(frames[i] && h[i] > 0 && w[i] > 0, "pil_resize: frame %d is empty", i);
    if (!frames_on_device && (size_t)h[i] * w[i] * 3 > stage_bytes) stage_bytes = (size_t)h[i] * w[i] * 3;
    if ((size_t)h[i] * out_w * 3 > tmp_bytes) tmp_bytes = (size_t)h[i] * out_w * 3;
  }
  if (stage_bytes && (rc = ensure(g_stage, stage_bytes * n))) return (tcd_status)rc;
  if ((rc = ensure(g_tmp, tmp_bytes))) return (tcd_status)rc;
  const size_t u8_frame = (size_t)out_h * out_w * 3;
  if (out_u8_host && (rc = ensure(g_u8, u8_frame * n))) return (tcd_status)rc;
  for (int i = 0; i < n; ++i) {
    const uint8_t* src = frames[i];
    long long stride = row_stride ? row_stride[i] : (long long)w[i] * 3;
    if (!frames_on_device) {
      uint8_t* dst = g_stage.p + (size_t)i * stage_bytes;
      TCD_CUDA_OK(cudaMemcpy2DAsync(dst, (size_t)w[i] * 3, src, (size_t)stride, (size_t)w[i] * 3, (size_t)h[i],
                                    cudaMemcpyHostToDevice, st));
      src = dst;
      stride = (long long)w[i] * 3;
    }
    CoeffTable tx, ty;
    if ((rc = get_table(w[i], out_w, &tx)) || (rc = get_table(h[i], out_h, &ty))) return (tcd_status)rc;
    pil_h_kernel<<<dim3(cdiv(out_w, 64), cdiv(h[i], 4)), 256, 0, st>>>(src, stride, h[i], out_w, tx.dev, tx.ksize,
                                                                      g_tmp.p);
    pil_v_kernel<<<dim3(cdiv(out_w, 64), cdiv(out_h, 4)), 256, 0, st>>>(
        g_tmp.p, out_h, out_w, ty.dev, ty.ksize, (__nv_bfloat16*)out_dev + (size_t)i * out_h * out_w * 16,
        out_u8_host ? g_u8.p + (size_t)i * u8_frame : nullptr);
  }
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && out_u8_host) {
    e = cudaMemcpyAsync(out_u8_host, g_u8.p, u8_frame * n, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  }
  if (e != cudaSuccess) {
    set_last_error("pil_resize: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return TCD_OK;
}
