// CLAHE + bilateral crop enhancement.
//
// Replaces ImageEnhancer.enhance_clahe_bilateral (/root/reference/src/image_enhancement.py:146-184: cv2.cvtColor
// BGR2LAB, CLAHE(2.0, 8x8).apply on L, cv2.cvtColor LAB2BGR, cv2.bilateralFilter(9, 75, 75)) for a BATCH of crops:
// three launches whatever the number of crops (grid.y / grid.z = crop), nothing on the host in between.
//   enh_hist_kernel       one CTA per (tile, crop): L of the tile's pixels (reflect-101 padding to whole tiles) ->
//                         256-bin histogram in shared memory -> clip + redistribute -> scan -> the tile's LUT
//   enh_apply_kernel      per pixel: BGR -> Lab (integer tables), L through the float32 bilinear blend of the four
//                         neighbouring tile LUTs, Lab -> BGR (integer) -- the Lab image never exists in memory
//   enh_bilateral_kernel  32x8 output pixels per CTA, the (32+2r)x(8+2r) source patch in shared memory, 48 neighbours
//                         (d = 9) in window order, colour weights from a shared-memory exp table
// All integer work is the 8-bit fixed-point arithmetic of the library the reference calls (restated in
// oracle/enhance.py, which is pinned bit-for-bit against it); the float32 steps keep its operation order (this TU is
// compiled with -fmad=false; the one fused multiply-add the library's SIMD path has is written out as fmaf).
// HBM bytes per crop pixel (3 channels): hist 3 R, apply 3 R + 3 W, bilateral 3 R (+halo) + 3 W = 15 B/px.
#include <cmath>
#include <cstring>
#include <mutex>
#include <vector>

#include "kernels.cuh"
#include "tcd_common.cuh"

namespace tcd {

// ---------------------------------------------------------------- host tables (built once, float32 like the library)
struct EnhHostTables {
  uint16_t gamma[256];
  uint16_t cbrt[3072];
  uint16_t yf[512];     // (y, fy) per L
  uint8_t inv[4096];
  int32_t cf[9], ci[9];
};

static float trunc_to_f32(double d) {
  uint64_t b;
  memcpy(&b, &d, 8);
  b &= ~((1ull << 29) - 1);
  memcpy(&d, &b, 8);
  return (float)d;   // exact: the low 29 mantissa bits are zero
}

static float cbrt_table_f32(float x) {
  int e;
  const double m = frexp((double)x, &e);
  int shx = e % 3;
  if (shx > 0) shx -= 3;
  const int ex = (e - shx) / 3;
  double fr = ldexp(m, shx);
  fr = ((((45.2548339756803022511987494 * fr + 192.2798368355061050458134625) * fr + 119.1654824285581628956914143) * fr +
         13.43250139086239872172837314) * fr + 0.1636161226585754240958355063) /
       ((((14.80884093219134573786480845 * fr + 151.9714051044435648658557668) * fr + 168.5254414101568283957668343) * fr +
         33.9905941350215598754191872) * fr + 1.0);
  return trunc_to_f32(ldexp(fr, ex));
}

static const EnhHostTables& enh_host_tables() {
  static EnhHostTables t;
  static std::once_flag once;
  std::call_once(once, [] {
    for (int i = 0; i < 256; ++i) {
      volatile float x = (float)i / 255.0f;
      volatile float g;
      if (x <= 0.04045f) {
        volatile float k = 1.0f / 12.92f;
        g = x * k;
      } else {
        volatile float k = 1.0f / 1.055f;
        volatile float s = x + 0.055f;
        volatile float b = s * k;
        g = (float)pow((double)b, 2.4);
      }
      volatile float sc = 255.0f * 8.0f;
      t.gamma[i] = (uint16_t)lrintf(sc * g);
    }
    volatile float scale = 1.0f / (float)(255 * 8);
    for (int i = 0; i < 3072; ++i) {
      volatile float x = scale * (float)i;
      volatile float v;
      if (x < 0.008856f) {
        volatile float p = x * 7.787f;
        v = p + 0.13793103448275862f;
      } else {
        v = cbrt_table_f32(x);
      }
      t.cbrt[i] = (uint16_t)lrintf(32768.0f * v);
    }
    const double xyz[9] = {0.412453, 0.357580, 0.180423, 0.212671, 0.715160, 0.072169, 0.019334, 0.119193, 0.950227};
    const double rgb[9] = {3.240479, -1.53715, -0.498535, -0.969256, 1.875991, 0.041556, 0.055648, -0.204043, 1.057311};
    const double d65[3] = {0.950456, 1.0, 1.088754};
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) {
        t.cf[r * 3 + c] = (int32_t)lrint(4096.0 * xyz[r * 3 + c] / d65[r]);
        t.ci[r * 3 + c] = (int32_t)lrint(4096.0 * rgb[r * 3 + c] * d65[c]);
      }
    const int BASE = 1 << 14;
    for (int i = 0; i < 256; ++i) {
      volatile float a = (float)(i * 100 * BASE) / (float)(255 * 116);
      volatile float b = (float)(16 * BASE) / 116.0f;
      volatile float fy = a + b;
      int y;
      if (i <= 20) {
        y = (int)lrintf((float)(i * BASE * 20 * 9) / (float)(17 * 29 * 29 * 29));
      } else {
        volatile float f2 = fy * fy;
        volatile float f3 = f2 * fy;
        y = (int)lrintf(f3 / (float)((double)BASE * BASE));
      }
      t.yf[2 * i] = (uint16_t)y;
      t.yf[2 * i + 1] = (uint16_t)lrintf(fy);
    }
    for (int i = 0; i < 4096; ++i) {
      volatile float k = 1.0f / 4096.0f;
      volatile float x = (float)i * k;
      volatile float g;
      if (x <= 0.0031308f) {
        g = x * 12.92f;
      } else {
        volatile float p = (float)pow((double)x, 1.0 / 2.4);
        volatile float q = 1.055f * p;
        g = q - 0.055f;
      }
      t.inv[i] = (uint8_t)lrintf(255.0f * g);
    }
  });
  return t;
}

void enhance_host_tables(uint16_t* gamma, uint16_t* cbrt, uint16_t* yf, uint8_t* inv, int32_t* cf, int32_t* ci) {
  const EnhHostTables& t = enh_host_tables();
  memcpy(gamma, t.gamma, sizeof(t.gamma));
  memcpy(cbrt, t.cbrt, sizeof(t.cbrt));
  memcpy(yf, t.yf, sizeof(t.yf));
  memcpy(inv, t.inv, sizeof(t.inv));
  memcpy(cf, t.cf, sizeof(t.cf));
  memcpy(ci, t.ci, sizeof(t.ci));
}

// ---------------------------------------------------------------- device side
__constant__ int32_t c_cf[9];
__constant__ int32_t c_ci[9];
__constant__ int8_t c_bil_dy[256];
__constant__ int8_t c_bil_dx[256];
__constant__ float c_bil_w[256];

__device__ __forceinline__ int reflect101_dev(int i, int n) {
  i = i < 0 ? -i : i;
  return i >= n ? 2 * (n - 1) - i : i;
}

__device__ __forceinline__ int descale_dev(int x, int n) { return (x + (1 << (n - 1))) >> n; }

// L, a, b of one BGR pixel (8-bit fixed point)
__device__ __forceinline__ void bgr_to_lab(const EnhTables& t, int b8, int g8, int r8, int& L, int& a, int& b) {
  const int B = __ldg(t.gamma + b8), G = __ldg(t.gamma + g8), R = __ldg(t.gamma + r8);
  const int fX = __ldg(t.cbrt + descale_dev(R * c_cf[0] + G * c_cf[1] + B * c_cf[2], 12));
  const int fY = __ldg(t.cbrt + descale_dev(R * c_cf[3] + G * c_cf[4] + B * c_cf[5], 12));
  const int fZ = __ldg(t.cbrt + descale_dev(R * c_cf[6] + G * c_cf[7] + B * c_cf[8], 12));
  const int Lscale = (116 * 255 + 50) / 100;
  const int Lshift = -((16 * 255 * (1 << 15) + 50) / 100);
  L = min(255, max(0, descale_dev(Lscale * fY + Lshift, 15)));
  a = min(255, max(0, descale_dev(500 * (fX - fY) + 128 * (1 << 15), 15)));
  b = min(255, max(0, descale_dev(200 * (fY - fZ) + 128 * (1 << 15), 15)));
}

__device__ __forceinline__ int lum_only(const EnhTables& t, int b8, int g8, int r8) {
  const int B = __ldg(t.gamma + b8), G = __ldg(t.gamma + g8), R = __ldg(t.gamma + r8);
  const int fY = __ldg(t.cbrt + descale_dev(R * c_cf[3] + G * c_cf[4] + B * c_cf[5], 12));
  return min(255, max(0, descale_dev(((116 * 255 + 50) / 100) * fY - ((16 * 255 * (1 << 15) + 50) / 100), 15)));
}

__device__ __forceinline__ int ab_to_xz(int i) {
  if (i <= 3390) {
    const int n = i * 108;   // C division: toward zero
    return (n >= 0 ? n / 841 : -((-n) / 841)) - ((1 << 14) * 16 / 116) * 108 / 841;
  }
  const long long ii = (long long)i * i / (1 << 14);
  return (int)(ii * i / (1 << 14));
}

__device__ __forceinline__ void lab_to_bgr(const EnhTables& t, int L, int a, int b, int& b8, int& g8, int& r8) {
  const int BASE = 1 << 14;
  const int y = __ldg(t.yf + 2 * L), ify = __ldg(t.yf + 2 * L + 1);
  const int adiv = ((5 * a * 53687 + (1 << 7)) >> 13) - 128 * BASE / 500;
  const int bdiv = ((b * 41943 + (1 << 4)) >> 9) - 128 * BASE / 200 + 1;
  const long long x = ab_to_xz(ify + adiv), z = ab_to_xz(ify - bdiv);
  auto chan = [&](int row) -> int {
    const long long v = ((long long)c_ci[row * 3] * x + (long long)c_ci[row * 3 + 1] * y + (long long)c_ci[row * 3 + 2] * z + (1 << 13)) >> 14;
    return __ldg(t.inv + (int)min(4095ll, max(0ll, v)));
  };
  r8 = chan(0);
  g8 = chan(1);
  b8 = chan(2);
}

// one CTA of 256 threads per (tile, crop)
__global__ void __launch_bounds__(256) enh_hist_kernel(const EnhImg* __restrict__ imgs, EnhTables t, int gx, int gy) {
  const EnhImg im = imgs[blockIdx.y];
  __shared__ int hist[256];
  __shared__ int warp_tot[8];
  __shared__ int s_clipped;
  const int tid = threadIdx.x;
  hist[tid] = 0;
  if (tid == 0) s_clipped = 0;
  __syncthreads();
  const int tyi = blockIdx.x / gx, txi = blockIdx.x - tyi * gx;
  const int npx = im.th * im.tw;
  for (int p = tid; p < npx; p += 256) {
    const int ly = p / im.tw, lx = p - ly * im.tw;
    const int y = reflect101_dev(tyi * im.th + ly, im.h), x = reflect101_dev(txi * im.tw + lx, im.w);
    const uint8_t* px = im.src + (long long)y * im.src_stride + (long long)x * im.cn;
    const int v = im.cn == 3 ? lum_only(t, px[0], px[1], px[2]) : px[0];
    atomicAdd(&hist[v], 1);
  }
  __syncthreads();
  int hv = hist[tid];
  if (im.clip > 0) {
    int excess = max(hv - im.clip, 0);
    for (int o = 16; o; o >>= 1) excess += __shfl_xor_sync(0xffffffffu, excess, o);
    if ((tid & 31) == 0 && excess) atomicAdd(&s_clipped, excess);
    __syncthreads();
    const int clipped = s_clipped;
    hv = min(hv, im.clip) + clipped / 256;
    const int resid = clipped % 256;
    if (resid) {
      const int step = max(256 / resid, 1);
      if (tid % step == 0 && tid / step < resid) ++hv;
    }
  }
  // inclusive scan over the 256 bins
  int s = hv;
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, s, o);
    if ((tid & 31) >= o) s += n;
  }
  if ((tid & 31) == 31) warp_tot[tid >> 5] = s;
  __syncthreads();
  for (int wv = 0; wv < (tid >> 5); ++wv) s += warp_tot[wv];
  const int q = __float2int_rn(__fmul_rn((float)s, im.lut_scale));
  im.luts[(size_t)blockIdx.x * 256 + tid] = (uint8_t)min(255, max(0, q));
}

// CLAHE blend (+ the colour conversions around it): one pixel per thread
__global__ void __launch_bounds__(256) enh_apply_kernel(const EnhImg* __restrict__ imgs, EnhTables t, int gx, int gy) {
  const EnhImg im = imgs[blockIdx.y];
  const long long npx = (long long)im.h * im.w;
  for (long long p = (long long)blockIdx.x * 256 + threadIdx.x; p < npx; p += (long long)gridDim.x * 256) {
    const int y = (int)(p / im.w), x = (int)(p - (long long)y * im.w);
    const uint8_t* px = im.src + (long long)y * im.src_stride + (long long)x * im.cn;
    int L, a = 0, b = 0;
    if (im.cn == 3) bgr_to_lab(t, px[0], px[1], px[2], L, a, b);
    else L = px[0];
    const float txf = __fsub_rn(__fmul_rn((float)x, im.inv_tw), 0.5f);
    const float tyf = __fsub_rn(__fmul_rn((float)y, im.inv_th), 0.5f);
    int tx1 = (int)floorf(txf), ty1 = (int)floorf(tyf);
    const float xa = __fsub_rn(txf, (float)tx1), ya = __fsub_rn(tyf, (float)ty1);
    const float xa1 = __fsub_rn(1.0f, xa), ya1 = __fsub_rn(1.0f, ya);
    const int tx2 = min(tx1 + 1, gx - 1), ty2 = min(ty1 + 1, gy - 1);
    tx1 = max(tx1, 0);
    ty1 = max(ty1, 0);
    const uint8_t* lu = im.luts + L;
    const float v11 = (float)lu[(size_t)(ty1 * gx + tx1) * 256], v12 = (float)lu[(size_t)(ty1 * gx + tx2) * 256];
    const float v21 = (float)lu[(size_t)(ty2 * gx + tx1) * 256], v22 = (float)lu[(size_t)(ty2 * gx + tx2) * 256];
    const float top = __fadd_rn(__fmul_rn(v11, xa1), __fmul_rn(v12, xa));
    const float bot = __fadd_rn(__fmul_rn(v21, xa1), __fmul_rn(v22, xa));
    const float res = __fadd_rn(__fmul_rn(top, ya1), __fmul_rn(bot, ya));
    const int Lc = min(255, max(0, __float2int_rn(res)));
    uint8_t* out = im.mid + p * im.cn;
    if (im.cn == 3) {
      int b8, g8, r8;
      lab_to_bgr(t, Lc, a, b, b8, g8, r8);
      out[0] = (uint8_t)b8;
      out[1] = (uint8_t)g8;
      out[2] = (uint8_t)r8;
    } else {
      out[0] = (uint8_t)Lc;
    }
  }
}

constexpr int kBilTW = 32, kBilTH = 8, kBilMaxR = 7;

template <int CN>
__global__ void __launch_bounds__(kBilTW * kBilTH)
enh_bilateral_kernel(const EnhImg* __restrict__ imgs, const float* __restrict__ color_w, int radius, int maxk) {
  const EnhImg im = imgs[blockIdx.z];
  const int x0 = blockIdx.x * kBilTW, y0 = blockIdx.y * kBilTH;
  if (x0 >= im.w || y0 >= im.h) return;
  __shared__ float s_cw[256 * CN];
  __shared__ uint8_t s_px[(kBilTH + 2 * kBilMaxR) * (kBilTW + 2 * kBilMaxR) * CN];
  const int tid = threadIdx.y * kBilTW + threadIdx.x;
  for (int i = tid; i < 256 * CN; i += kBilTW * kBilTH) s_cw[i] = color_w[i];
  const int pw = kBilTW + 2 * radius, ph = kBilTH + 2 * radius;
  for (int i = tid; i < pw * ph; i += kBilTW * kBilTH) {
    const int py = i / pw, pxx = i - py * pw;
    const int sy = reflect101_dev(y0 + py - radius, im.h), sx = reflect101_dev(x0 + pxx - radius, im.w);
    const uint8_t* s = im.mid + ((long long)sy * im.w + sx) * CN;
#pragma unroll
    for (int c = 0; c < CN; ++c) s_px[i * CN + c] = s[c];
  }
  __syncthreads();
  const int x = x0 + threadIdx.x, y = y0 + threadIdx.y;
  if (x >= im.w || y >= im.h) return;
  const uint8_t* cpx = s_px + ((threadIdx.y + radius) * pw + threadIdx.x + radius) * CN;
  int c0[CN];
#pragma unroll
  for (int c = 0; c < CN; ++c) c0[c] = cpx[c];
  float wsum = 0.f, acc[CN];
#pragma unroll
  for (int c = 0; c < CN; ++c) acc[c] = 0.f;
  for (int k = 0; k < maxk; ++k) {
    const uint8_t* n = cpx + ((int)c_bil_dy[k] * pw + (int)c_bil_dx[k]) * CN;
    int v[CN], diff = 0;
#pragma unroll
    for (int c = 0; c < CN; ++c) {
      v[c] = n[c];
      diff += abs(v[c] - c0[c]);
    }
    const float w = __fmul_rn(c_bil_w[k], s_cw[diff]);
    wsum = __fadd_rn(wsum, w);
#pragma unroll
    for (int c = 0; c < CN; ++c) acc[c] = fmaf((float)v[c], w, acc[c]);
  }
  wsum = __fadd_rn(wsum, 1.0f);
  uint8_t* o = im.dst + (long long)y * im.dst_stride + (long long)x * CN;
#pragma unroll
  for (int c = 0; c < CN; ++c) {
    const float r = __fdiv_rn(__fadd_rn(acc[c], (float)c0[c]), wsum);
    o[c] = (uint8_t)min(255, max(0, __float2int_rn(r)));
  }
}

int enhance_set_constants(const int32_t* cf, const int32_t* ci) {
  TCD_CUDA_OK(cudaMemcpyToSymbol(c_cf, cf, 9 * sizeof(int32_t)));
  TCD_CUDA_OK(cudaMemcpyToSymbol(c_ci, ci, 9 * sizeof(int32_t)));
  return TCD_OK;
}

int enhance_set_window(const int8_t* dy, const int8_t* dx, const float* w, int maxk) {
  TCD_REQUIRE(maxk <= 256, "enhance: bilateral window of %d taps", maxk);
  TCD_CUDA_OK(cudaMemcpyToSymbol(c_bil_dy, dy, maxk));
  TCD_CUDA_OK(cudaMemcpyToSymbol(c_bil_dx, dx, maxk));
  TCD_CUDA_OK(cudaMemcpyToSymbol(c_bil_w, w, maxk * sizeof(float)));
  return TCD_OK;
}

int enhance_launch(const EnhImg* imgs_dev, int n, int cn, int gx, int gy, int max_h, int max_w, const EnhTables& t,
                   const float* color_w_dev, int radius, int maxk, cudaStream_t stream) {
  TCD_REQUIRE(radius >= 1 && radius <= kBilMaxR, "enhance: bilateral radius %d (1..%d)", radius, kBilMaxR);
  enh_hist_kernel<<<dim3(gx * gy, n), 256, 0, stream>>>(imgs_dev, t, gx, gy);
  const long long npx = (long long)max_h * max_w;
  const int blocks = (int)std::min<long long>((npx + 255) / 256, 148 * 8);
  enh_apply_kernel<<<dim3(blocks, n), 256, 0, stream>>>(imgs_dev, t, gx, gy);
  const dim3 grid(cdiv(max_w, kBilTW), cdiv(max_h, kBilTH), n);
  if (cn == 3) enh_bilateral_kernel<3><<<grid, dim3(kBilTW, kBilTH), 0, stream>>>(imgs_dev, color_w_dev, radius, maxk);
  else enh_bilateral_kernel<1><<<grid, dim3(kBilTW, kBilTH), 0, stream>>>(imgs_dev, color_w_dev, radius, maxk);
  TCD_CUDA_OK(cudaGetLastError());
  return TCD_OK;
}

}  // namespace tcd
