// Motion kernels (K8, K9, K10): MOG2 background subtraction, OpenCV-exact u8 Gaussian blur,
// thresholds, per-ROI countNonZero and the empty-frame gate.
//
// Replaces cv2.BackgroundSubtractorMOG2.apply / cv2.threshold / cv2.GaussianBlur /
// cv2.countNonZero as used by MotionFilter.has_motion_in_bbox (src/motion_filter.py:68-127)
// and cv2.cvtColor / GaussianBlur / absdiff / threshold / countNonZero in
// EmptyFrameFilter.has_motion (src/empty_frame_filter.py:51-98).
// Arithmetic follows the specification validated bit-exact against cv2 4.13 in SURVEY App. C;
// this TU is compiled with -fmad=false so every fp32 intermediate rounds like the scalar code.
#include "kernels.cuh"

#include <string.h>

#include <atomic>
#include <float.h>

namespace tcd {
extern std::atomic<uint64_t> g_launch_count;

// ------------------------------------------------------------------ MOG2
#define MOG_SWAP(i, j)                                                                     \
  do {                                                                                     \
    float _t;                                                                              \
    _t = w[i]; w[i] = w[j]; w[j] = _t;                                                     \
    _t = v[i]; v[i] = v[j]; v[j] = _t;                                                     \
    _t = m0[i]; m0[i] = m0[j]; m0[j] = _t;                                                 \
    _t = m1[i]; m1[i] = m1[j]; m1[j] = _t;                                                 \
    _t = m2[i]; m2[i] = m2[j]; m2[j] = _t;                                                 \
  } while (0)

// mask value of the current state: 0 background, 127 shadow, 255 foreground
__device__ __forceinline__ uint8_t mog2_mask_value(const Mog2Params& p, bool bg, int n, float x0, float x1, float x2,
                                                   const float (&w)[kMogModes], const float (&v)[kMogModes],
                                                   const float (&m0)[kMogModes], const float (&m1)[kMogModes],
                                                   const float (&m2)[kMogModes]) {
  uint8_t out = 0;
  if (!bg) {
    out = 255;
    if (p.detect_shadows) {
      float tW = 0.f;
      bool done = false;
#pragma unroll
      for (int mode = 0; mode < kMogModes; ++mode) {
        if (mode < n && !done) {
          const float num = ((x0 * m0[mode]) + (x1 * m1[mode])) + (x2 * m2[mode]);
          const float den = ((m0[mode] * m0[mode]) + (m1[mode] * m1[mode])) + (m2[mode] * m2[mode]);
          if (den == 0.f) {
            done = true;
          } else {
            if (num <= den && num >= (p.tau * den)) {
              const float a = num / den;
              const float e0 = (a * m0[mode]) - x0, e1 = (a * m1[mode]) - x1, e2 = (a * m2[mode]) - x2;
              const float s = ((e0 * e0) + (e1 * e1)) + (e2 * e2);
              if (s < (((p.Tb * v[mode]) * a) * a)) {
                out = 127;
                done = true;
              }
            }
            if (!done) {
              tW = tW + w[mode];
              if (tW > p.TB) done = true;
            }
          }
        }
      }
    }
  }
  return out;
}

// One thread per pixel; p.repeats model updates with the same pixel, the state staying in registers (the reference
// calls apply() once per detection with the same frame).  rois == nullptr: the mask of the LAST update goes
// to fgmask.  rois != nullptr (the batched filter): update r is only judged inside box r, so its mask is written to
// plane r of fgmask ([repeats][npix]) and only where the 21x21 clean-up of that box can see it (box +- 10 px).
__device__ __forceinline__ void mog2_body(const uint8_t* __restrict__ frame, const Mog2State& st, const Mog2Params& p,
                                          uint8_t* __restrict__ fgmask, const int idx, const int (*rois)[4]) {
  const int py = idx / p.w, px = idx - py * p.w;
  const uint8_t* pp = frame + (long long)py * p.stride + px * 3;
  const float x0 = (float)pp[0], x1 = (float)pp[1], x2 = (float)pp[2];

  float w[kMogModes], v[kMogModes], m0[kMogModes], m1[kMogModes], m2[kMogModes];
  int n = st.nmodes[idx];
#pragma unroll
  for (int k = 0; k < kMogModes; ++k) {
    if (k < n) {
      const size_t o = (size_t)k * p.npix + idx;
      w[k] = st.weight[o];
      v[k] = st.var[o];
      m0[k] = st.mean[((size_t)k * 3 + 0) * p.npix + idx];
      m1[k] = st.mean[((size_t)k * 3 + 1) * p.npix + idx];
      m2[k] = st.mean[((size_t)k * 3 + 2) * p.npix + idx];
    } else {
      w[k] = 0.f; v[k] = 0.f; m0[k] = 0.f; m1[k] = 0.f; m2[k] = 0.f;
    }
  }
  int n_in = n;
  bool bg = false;
  for (int rep = 0; rep < p.repeats; ++rep) {
    const float alpha = p.alpha[rep];
    const float a1 = 1.0f - alpha;
    const float prune = -(alpha * p.CT);
    bg = false;
    bool fits = false;
    int nmodes = n;
    float tot = 0.f;
#pragma unroll
    for (int mode = 0; mode < kMogModes; ++mode) {
      if (mode < nmodes) {
        float wt = (a1 * w[mode]) + prune;
        int swap = 0;
        if (!fits) {
          const float var = v[mode];
          const float d0 = m0[mode] - x0, d1 = m1[mode] - x1, d2 = m2[mode] - x2;
          const float dist2 = ((d0 * d0) + (d1 * d1)) + (d2 * d2);
          if (tot < p.TB && dist2 < (p.Tb * var)) bg = true;
          if (dist2 < (p.Tg * var)) {
            fits = true;
            wt = wt + alpha;
            const float k = alpha / wt;
            m0[mode] = m0[mode] - (k * d0);
            m1[mode] = m1[mode] - (k * d1);
            m2[mode] = m2[mode] - (k * d2);
            float vn = var + (k * (dist2 - var));
            vn = fmaxf(vn, p.varMin);
            vn = fminf(vn, p.varMax);
            v[mode] = vn;
            bool bubbling = true;
#pragma unroll
            for (int i = mode; i > 0; --i) {
              if (bubbling) {
                if (wt < w[i - 1]) {
                  bubbling = false;
                } else {
                  ++swap;
                  MOG_SWAP(i, i - 1);
                }
              }
            }
          }
        }
        if (wt < -prune) {
          wt = 0.f;
          --nmodes;
        }
        const int dst = mode - swap;
#pragma unroll
        for (int t = 0; t < kMogModes; ++t)
          if (t == dst) w[t] = wt;
        tot = tot + wt;
      }
    }
    const float inv = (fabsf(tot) > FLT_EPSILON) ? 1.0f / tot : 0.f;
#pragma unroll
    for (int k = 0; k < kMogModes; ++k)
      if (k < nmodes) w[k] = w[k] * inv;
    if (!fits && alpha > 0.f) {
      int mode;
      if (nmodes == kMogModes) mode = kMogModes - 1; else mode = nmodes++;
#pragma unroll
      for (int t = 0; t < kMogModes; ++t) {
        if (t == mode) {
          w[t] = (nmodes == 1) ? 1.0f : alpha;
          m0[t] = x0; m1[t] = x1; m2[t] = x2;
          v[t] = p.varInit;
        } else if (nmodes != 1 && t < nmodes - 1) {
          w[t] = w[t] * a1;
        }
      }
      bool stop = false;
#pragma unroll
      for (int i = kMogModes - 1; i > 0; --i) {
        if (i <= nmodes - 1 && !stop) {
          if (alpha < w[i - 1]) stop = true; else MOG_SWAP(i, i - 1);
        }
      }
    }
    n = nmodes;
    if (rois != nullptr) {
      const int* r = rois[rep];
      if (px >= r[0] - 10 && px < r[2] + 10 && py >= r[1] - 10 && py < r[3] + 10)
        fgmask[(size_t)rep * p.npix + idx] = mog2_mask_value(p, bg, n, x0, x1, x2, w, v, m0, m1, m2);
    }
  }
  if (rois == nullptr) fgmask[idx] = mog2_mask_value(p, bg, n, x0, x1, x2, w, v, m0, m1, m2);
  const int n_wr = n > n_in ? n : n_in;  // slots that may have changed
#pragma unroll
  for (int k = 0; k < kMogModes; ++k) {
    if (k < n_wr) {
      const size_t o = (size_t)k * p.npix + idx;
      st.weight[o] = w[k];
      st.var[o] = v[k];
      st.mean[((size_t)k * 3 + 0) * p.npix + idx] = m0[k];
      st.mean[((size_t)k * 3 + 1) * p.npix + idx] = m1[k];
      st.mean[((size_t)k * 3 + 2) * p.npix + idx] = m2[k];
    }
  }
  st.nmodes[idx] = (uint8_t)n;
}

__global__ void __launch_bounds__(256)
mog2_apply_kernel(const uint8_t* __restrict__ frame, const Mog2State st, const Mog2Params p,
                  uint8_t* __restrict__ fgmask) {
  const int idx = blockIdx.x * 256 + threadIdx.x;
  if (idx >= p.npix) return;
  mog2_body(frame, st, p, fgmask, idx, nullptr);
}

// every stream of a bank in ONE launch (grid.y = stream): per-stream frame, state, update count and boxes
__global__ void __launch_bounds__(256) bank_mog2_kernel(const BankStream* __restrict__ streams, const Mog2Params base) {
  const BankStream& s = streams[blockIdx.y];
  if (s.repeats <= 0) return;
  const int idx = blockIdx.x * 256 + threadIdx.x;
  if (idx >= base.npix) return;
  Mog2Params p = base;
  p.stride = s.stride;
  p.repeats = s.repeats;
#pragma unroll
  for (int i = 0; i < kMogMaxRepeats; ++i) p.alpha[i] = s.alpha[i];
  mog2_body(s.frame, s.st, p, s.fg, idx, s.roi);
}

// ------------------------------------------------------------------ u8 Gaussian blur (OpenCV fixed point)
constexpr int kBlurTW = 128;
constexpr int kBlurTH = 64;
constexpr int kBlurThreads = 256;

__device__ __forceinline__ int reflect101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) {
    if (i < 0) i = -i;
    if (i >= n) i = 2 * n - 2 - i;
  }
  return i;
}

__global__ void __launch_bounds__(kBlurThreads)
blur_kernel(const uint8_t* __restrict__ src, const BlurArgs a, uint8_t* __restrict__ dst,
            uint8_t* __restrict__ prev, unsigned int* __restrict__ counter) {
  extern __shared__ uint8_t sm_raw[];
  const int R = a.ksize >> 1;
  const int in_w = kBlurTW + 2 * R, in_h = kBlurTH + 2 * R;
  uint8_t* s_in = sm_raw;                                              // [in_h][in_w]
  uint16_t* s_h = reinterpret_cast<uint16_t*>(sm_raw + ((in_h * in_w + 15) & ~15));  // [in_h][TW]
  const int x0 = blockIdx.x * kBlurTW, y0 = blockIdx.y * kBlurTH;

  for (int i = threadIdx.x; i < in_h * in_w; i += kBlurThreads) {
    const int ly = i / in_w, lx = i - ly * in_w;
    const int gy = reflect101(y0 + ly - R, a.h), gx = reflect101(x0 + lx - R, a.w);
    uint8_t val;
    if (a.src_mode == 0) {
      val = src[(long long)gy * a.in_stride + gx] > a.pre_thr ? 255 : 0;
    } else if (a.src_mode == 2) {
      val = src[(long long)gy * a.in_stride + gx];
    } else {
      const uint8_t* p = src + (long long)gy * a.in_stride + gx * 3;
      // cv2.cvtColor BGR2GRAY, 15-bit coefficients (bit-exact vs cv2 4.13)
      val = (uint8_t)(((int)p[0] * 3735 + (int)p[1] * 19235 + (int)p[2] * 9798 + 16384) >> 15);
    }
    s_in[i] = val;
  }
  __syncthreads();
  // horizontal pass, exact 8.8 (no rounding)
  for (int i = threadIdx.x; i < in_h * kBlurTW; i += kBlurThreads) {
    const int ly = i / kBlurTW, lx = i - ly * kBlurTW;
    const uint8_t* r = s_in + ly * in_w + lx;
    unsigned int acc = 0;
    for (int t = 0; t < a.ksize; ++t) acc += (unsigned int)a.k[t] * r[t];
    s_h[i] = (uint16_t)acc;
  }
  __syncthreads();
  unsigned int local = 0;
  for (int i = threadIdx.x; i < kBlurTH * kBlurTW; i += kBlurThreads) {
    const int ly = i / kBlurTW, lx = i - ly * kBlurTW;
    const int gx = x0 + lx, gy = y0 + ly;
    if (gx >= a.w || gy >= a.h) continue;
    unsigned int acc = 0;
    for (int t = 0; t < a.ksize; ++t) acc += (unsigned int)a.k[t] * s_h[(ly + t) * kBlurTW + lx];
    const int v = (int)((acc + 32768u) >> 16);
    const size_t o = (size_t)gy * a.w + gx;
    if (a.post_mode == 0) {
      dst[o] = v > a.post_thr ? 255 : 0;
    } else {
      if (!a.first) {
        const int d = abs(v - (int)prev[o]);
        local += d > a.post_thr ? 1u : 0u;
      }
      prev[o] = (uint8_t)v;
      if (dst) dst[o] = (uint8_t)v;
    }
  }
  if (a.post_mode == 1 && !a.first) {
    for (int off = 16; off > 0; off >>= 1) local += __shfl_xor_sync(0xffffffffu, local, off);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(counter, local);
  }
}

// ---- 21-tap specialisation (the size both filters use: motion_blur_size = 21, blur_size = 21).
// Same arithmetic as blur_kernel, bit for bit (horizontal pass exact 8.8, vertical 16.16, +32768 >> 16),
// restructured so the integer pipes do four / two taps per instruction:
//   stage A  4 pixels per thread: aligned 32-bit loads (tile origin x0-12), gray or threshold -> bytes in smem
//   stage B  8 outputs per thread (4 columns x 2 rows): six 32-bit smem words per row, byte-shifted with
//            funnel shifts, 6 x dp4a per output; the two rows are stored as one (even,odd) u16 pair
//   stage C  8 rows x 1 column per thread over the row pairs: 11 x dp2a per output
constexpr int kB21TW = 128, kB21TH = 32, kB21R = 10;
constexpr int kB21InH = kB21TH + 2 * kB21R;          // 52 rows = 26 pairs
constexpr int kB21InWords = (kB21TW + 24) / 4;       // 38 words: pixels x0-12 .. x0+TW+11

struct Blur21Taps {
  unsigned int h4[6];     // taps 4m..4m+3 as bytes (tap 20 alone in h4[5])
  unsigned int ve[11];    // even output row: (k[2m], k[2m+1]) in bytes 0,1
  unsigned int vo[11];    // odd output row:  (k[2m-1], k[2m])
};

template <int O>
__device__ __forceinline__ unsigned int b21_window(const unsigned int (&w)[8]) {
  // four consecutive bytes starting at byte offset O of the little-endian word array
  constexpr int i = O >> 2, sh = (O & 3) * 8;
  if constexpr (sh == 0) return w[i];
  else return __funnelshift_r(w[i], w[i + 1], sh);
}
template <int J>
__device__ __forceinline__ unsigned int b21_hsum(const unsigned int (&w)[8], const Blur21Taps& t) {
  unsigned int acc = 0;
  acc = __dp4a(b21_window<2 + J + 0>(w), t.h4[0], acc);
  acc = __dp4a(b21_window<2 + J + 4>(w), t.h4[1], acc);
  acc = __dp4a(b21_window<2 + J + 8>(w), t.h4[2], acc);
  acc = __dp4a(b21_window<2 + J + 12>(w), t.h4[3], acc);
  acc = __dp4a(b21_window<2 + J + 16>(w), t.h4[4], acc);
  acc = __dp4a(b21_window<2 + J + 20>(w), t.h4[5], acc);
  return acc;
}

// post_mode 2 (batched motion filter): nothing is written; pixels with v > post_thr inside roi = (x1,y1,x2,y2) are
// counted into *counter (the cleaned mask is only ever looked at inside the detection's box).
__device__ __forceinline__ void blur21_body(const uint8_t* __restrict__ src, const BlurArgs& a, const Blur21Taps& taps,
                                            uint8_t* __restrict__ dst, uint8_t* __restrict__ prev,
                                            unsigned int* __restrict__ counter, const int bx, const int by,
                                            const int* roi) {
  __shared__ unsigned int s_in[kB21InH][kB21InWords + 2];   // +2: the last group's 7th/8th word reads
  __shared__ unsigned int s_h[kB21InH / 2][kB21TW];         // (row 2r, row 2r+1) u16 pairs
  const int x0 = bx * kB21TW, y0 = by * kB21TH;
  const bool aligned = (a.in_stride & 3) == 0 && (reinterpret_cast<uintptr_t>(src) & 3) == 0;

  // ---- stage A (three groups per thread in flight: all loads are issued before the first use)
  constexpr int kGroups = kB21InH * (kB21InWords + 2);
  for (int base = 0; base < kGroups; base += 256 * 3) {
    uint3 q[3];
    int state[3];   // 0: padding word / past the end, 1: fast path loaded, 2: border (per-pixel reflect)
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int i = base + u * 256 + (int)threadIdx.x;
      state[u] = 0;
      q[u] = make_uint3(0u, 0u, 0u);
      if (i < kGroups) {
        const int ly = i / (kB21InWords + 2), g = i - ly * (kB21InWords + 2);
        if (g < kB21InWords) {
          const int gy = reflect101(y0 + ly - kB21R, a.h);
          const int gx0 = x0 - 12 + 4 * g;
          const uint8_t* row = src + (long long)gy * a.in_stride;
          if (gx0 >= 0 && gx0 + 4 <= a.w && aligned) {
            state[u] = 1;
            if (a.src_mode == 1) q[u] = *reinterpret_cast<const uint3*>(row + (size_t)gx0 * 3);
            else q[u].x = *reinterpret_cast<const unsigned int*>(row + gx0);
          } else {
            state[u] = 2;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int i = base + u * 256 + (int)threadIdx.x;
      if (i >= kGroups) continue;
      const int ly = i / (kB21InWords + 2), g = i - ly * (kB21InWords + 2);
      unsigned int out = 0;
      if (state[u] == 1) {
        if (a.src_mode == 1) {
          const unsigned int b[3] = {q[u].x, q[u].y, q[u].z};
#pragma unroll
          for (int px = 0; px < 4; ++px) {
            const int o = px * 3;
            const unsigned int c0 = (b[o >> 2] >> ((o & 3) * 8)) & 255u;
            const unsigned int c1 = (b[(o + 1) >> 2] >> (((o + 1) & 3) * 8)) & 255u;
            const unsigned int c2 = (b[(o + 2) >> 2] >> (((o + 2) & 3) * 8)) & 255u;
            out |= ((c0 * 3735u + c1 * 19235u + c2 * 9798u + 16384u) >> 15) << (8 * px);
          }
        } else if (a.src_mode == 2) {
          out = q[u].x;
        } else {
          // per byte: (p > pre_thr) ? 255 : 0
          out = __vcmpgtu4(q[u].x, (unsigned int)a.pre_thr * 0x01010101u);
        }
      } else if (state[u] == 2) {
        const int gy = reflect101(y0 + ly - kB21R, a.h);
        const int gx0 = x0 - 12 + 4 * g;
        const uint8_t* row = src + (long long)gy * a.in_stride;
        for (int px = 0; px < 4; ++px) {
          const int gx = reflect101(gx0 + px, a.w);
          unsigned int val;
          if (a.src_mode == 0) {
            val = row[gx] > a.pre_thr ? 255u : 0u;
          } else if (a.src_mode == 2) {
            val = row[gx];
          } else {
            const uint8_t* p = row + (size_t)gx * 3;
            val = ((unsigned int)p[0] * 3735u + (unsigned int)p[1] * 19235u + (unsigned int)p[2] * 9798u + 16384u) >> 15;
          }
          out |= val << (8 * px);
        }
      }
      s_in[ly][g] = out;
    }
  }
  __syncthreads();

  // ---- stage B: task = (row pair, group of 4 columns)
  for (int i = threadIdx.x; i < (kB21InH / 2) * (kB21TW / 4); i += 256) {
    const int rp = i / (kB21TW / 4), g = i - rp * (kB21TW / 4);
    unsigned int h[2][4];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      unsigned int w[8];
#pragma unroll
      for (int m = 0; m < 8; ++m) w[m] = s_in[2 * rp + r][g + m];
      h[r][0] = b21_hsum<0>(w, taps);
      h[r][1] = b21_hsum<1>(w, taps);
      h[r][2] = b21_hsum<2>(w, taps);
      h[r][3] = b21_hsum<3>(w, taps);
    }
    uint4 o;
    o.x = h[0][0] | (h[1][0] << 16);
    o.y = h[0][1] | (h[1][1] << 16);
    o.z = h[0][2] | (h[1][2] << 16);
    o.w = h[0][3] | (h[1][3] << 16);
    *reinterpret_cast<uint4*>(&s_h[rp][4 * g]) = o;
  }
  __syncthreads();

  // ---- stage C: task = (block of 8 output rows, column)
  unsigned int local = 0;
  for (int i = threadIdx.x; i < (kB21TH / 8) * kB21TW; i += 256) {
    const int rb = i / kB21TW, lx = i - rb * kB21TW;
    const int gx = x0 + lx;
    unsigned int p[14];
#pragma unroll
    for (int m = 0; m < 14; ++m) p[m] = s_h[4 * rb + m][lx];
    if (gx >= a.w) continue;
    int pv[8];
    const bool want_prev = a.post_mode == 1 && !a.first;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int gy = y0 + 8 * rb + r;
      pv[r] = (want_prev && gy < a.h) ? (int)prev[(size_t)gy * a.w + gx] : 0;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      unsigned int ae = 0, ao = 0;
#pragma unroll
      for (int m = 0; m < 11; ++m) {
        ae = __dp2a_lo(p[q + m], taps.ve[m], ae);
        ao = __dp2a_lo(p[q + m], taps.vo[m], ao);
      }
#pragma unroll
      for (int par = 0; par < 2; ++par) {
        const int gy = y0 + 8 * rb + 2 * q + par;
        if (gy >= a.h) continue;
        const int v = (int)(((par ? ao : ae) + 32768u) >> 16);
        const size_t o = (size_t)gy * a.w + gx;
        if (a.post_mode == 0) {
          dst[o] = v > a.post_thr ? 255 : 0;
        } else if (a.post_mode == 2) {
          local += (v > a.post_thr && gx >= roi[0] && gx < roi[2] && gy >= roi[1] && gy < roi[3]) ? 1u : 0u;
        } else {
          if (!a.first) {
            const int d = abs(v - pv[2 * q + par]);
            local += d > a.post_thr ? 1u : 0u;
          }
          prev[o] = (uint8_t)v;
          if (dst) dst[o] = (uint8_t)v;
        }
      }
    }
  }
  if ((a.post_mode == 1 && !a.first) || a.post_mode == 2) {   // one atomic per CTA
    __shared__ unsigned int s_cnt;
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    for (int off = 16; off > 0; off >>= 1) local += __shfl_xor_sync(0xffffffffu, local, off);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(&s_cnt, local);
    __syncthreads();
    if (threadIdx.x == 0 && s_cnt) atomicAdd(counter, s_cnt);
  }
}

__global__ void __launch_bounds__(256)
blur21_kernel(const uint8_t* __restrict__ src, const BlurArgs a, const Blur21Taps taps,
              uint8_t* __restrict__ dst, uint8_t* __restrict__ prev, unsigned int* __restrict__ counter) {
  blur21_body(src, a, taps, dst, prev, counter, blockIdx.x, blockIdx.y, nullptr);
}

// Empty-frame gate of every stream of a bank in ONE launch (grid.z = stream): gray + 21x21 blur + absdiff against the
// stream's previous plane + count, previous plane updated in place.
__global__ void __launch_bounds__(256)
bank_gate_kernel(const BankStream* __restrict__ streams, const BlurArgs base, const Blur21Taps taps) {
  const BankStream& s = streams[blockIdx.z];
  if (!s.active) return;
  BlurArgs a = base;
  a.in_stride = s.stride;
  a.first = s.first;
  blur21_body(s.frame, a, taps, nullptr, s.prev, s.counter, blockIdx.x, blockIdx.y, nullptr);
}

// Mask clean-up + count of round r of every stream (grid.z = stream * kMogMaxRepeats + r): shadow threshold, 21x21
// blur, threshold, count inside box r -- only the tiles that touch the box do any work.
__global__ void __launch_bounds__(256)
bank_roi_kernel(const BankStream* __restrict__ streams, const BlurArgs base, const Blur21Taps taps) {
  const BankStream& s = streams[blockIdx.z / kMogMaxRepeats];
  const int r = blockIdx.z % kMogMaxRepeats;
  if (r >= s.repeats) return;
  const int* roi = s.roi[r];
  const int x0 = blockIdx.x * kB21TW, y0 = blockIdx.y * kB21TH;
  if (roi[2] <= roi[0] || roi[3] <= roi[1] || x0 >= roi[2] || x0 + kB21TW <= roi[0] || y0 >= roi[3] ||
      y0 + kB21TH <= roi[1])
    return;
  blur21_body(s.fg + (size_t)r * base.w * base.h, base, taps, nullptr, nullptr, s.roi_counts + r, blockIdx.x,
              blockIdx.y, roi);
}

// no-blur variant of the mask clean-up: (m > pre_thr) only (motion_blur_size <= 0)
__global__ void threshold_kernel(const uint8_t* __restrict__ src, int n, int thr,
                                 uint8_t* __restrict__ dst) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < n) dst[i] = src[i] > thr ? 255 : 0;
}

// countNonZero inside each ROI; grid (n_rois, 64): warps stride over the ROI rows, lanes over aligned
// 32-bit words of a row (bytes outside [x1, x2) masked off), nonzero bytes counted with SIMD compares
__global__ void __launch_bounds__(256)
roi_count_kernel(const uint8_t* __restrict__ mask, int w, int h, const int* __restrict__ rois,
                 unsigned int* __restrict__ counts) {
  const int r = blockIdx.x;
  const int x1 = rois[r * 4 + 0], y1 = rois[r * 4 + 1], x2 = rois[r * 4 + 2], y2 = rois[r * 4 + 3];
  if (x2 <= x1 || y2 <= y1) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool aligned = (w & 3) == 0 && (reinterpret_cast<uintptr_t>(mask) & 3) == 0;
  unsigned int local = 0;
  for (int y = y1 + blockIdx.y * 8 + warp; y < y2; y += gridDim.y * 8) {
    const uint8_t* row = mask + (size_t)y * w;
    if (aligned) {
      const int xa = x1 & ~3;
      for (int x = xa + 4 * lane; x < x2; x += 128) {
        unsigned int q = *reinterpret_cast<const unsigned int*>(row + x);
        unsigned int keep = 0xFFFFFFFFu;
        if (x < x1) keep &= 0xFFFFFFFFu << (8 * (x1 - x));
        if (x + 4 > x2) keep &= 0xFFFFFFFFu >> (8 * (x + 4 - x2));
        q &= keep;
        local += __popc(__vcmpne4(q, 0u) & 0x01010101u);
      }
    } else {
      for (int x = x1 + lane; x < x2; x += 32) local += row[x] != 0 ? 1u : 0u;
    }
  }
  for (int off = 16; off > 0; off >>= 1) local += __shfl_xor_sync(0xffffffffu, local, off);
  if (lane == 0 && local) atomicAdd(counts + r, local);
}

// ------------------------------------------------------------------ host launchers
static size_t blur_smem(int ksize) {
  const int R = ksize >> 1;
  const int in_w = kBlurTW + 2 * R, in_h = kBlurTH + 2 * R;
  return (size_t)((in_h * in_w + 15) & ~15) + (size_t)in_h * kBlurTW * 2;
}

int blur_launch(const uint8_t* src, const BlurArgs& a, uint8_t* dst, uint8_t* prev,
                unsigned int* counter, cudaStream_t stream) {
  if (int rc = ensure_dyn_smem((const void*)blur_kernel, blur_smem(kBlurMaxK))) return rc;
  bool small_taps = true;
  for (int t = 0; t < a.ksize; ++t) small_taps = small_taps && a.k[t] <= 255;
  if (a.ksize == 21 && small_taps) {
    Blur21Taps tp;
    memset(&tp, 0, sizeof(tp));
    auto k = [&](int t) -> unsigned int { return (t >= 0 && t < 21) ? a.k[t] : 0u; };
    for (int m = 0; m < 6; ++m)
      tp.h4[m] = k(4 * m) | (k(4 * m + 1) << 8) | (k(4 * m + 2) << 16) | (k(4 * m + 3) << 24);
    for (int m = 0; m < 11; ++m) {
      tp.ve[m] = k(2 * m) | (k(2 * m + 1) << 8);
      tp.vo[m] = k(2 * m - 1) | (k(2 * m) << 8);
    }
    dim3 g21(cdiv(a.w, kB21TW), cdiv(a.h, kB21TH));
    blur21_kernel<<<g21, 256, 0, stream>>>(src, a, tp, dst, prev, counter);
    g_launch_count.fetch_add(1, std::memory_order_relaxed);
    cudaError_t e21 = cudaGetLastError();
    if (e21 != cudaSuccess) {
      set_last_error("blur21_kernel launch: %s", cudaGetErrorString(e21));
      return TCD_ERR_CUDA;
    }
    return TCD_OK;
  }
  dim3 grid(cdiv(a.w, kBlurTW), cdiv(a.h, kBlurTH));
  blur_kernel<<<grid, kBlurThreads, blur_smem(a.ksize), stream>>>(src, a, dst, prev, counter);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("blur_kernel launch: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return TCD_OK;
}

int mog2_launch(const uint8_t* frame, const Mog2State& st, const Mog2Params& p, uint8_t* fgmask,
                cudaStream_t stream) {
  mog2_apply_kernel<<<cdiv(p.npix, 256), 256, 0, stream>>>(frame, st, p, fgmask);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("mog2_apply_kernel launch: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return TCD_OK;
}

static void blur21_taps(const BlurArgs& a, Blur21Taps* tp) {
  memset(tp, 0, sizeof(*tp));
  auto k = [&](int t) -> unsigned int { return (t >= 0 && t < 21) ? a.k[t] : 0u; };
  for (int m = 0; m < 6; ++m)
    tp->h4[m] = k(4 * m) | (k(4 * m + 1) << 8) | (k(4 * m + 2) << 16) | (k(4 * m + 3) << 24);
  for (int m = 0; m < 11; ++m) {
    tp->ve[m] = k(2 * m) | (k(2 * m + 1) << 8);
    tp->vo[m] = k(2 * m - 1) | (k(2 * m) << 8);
  }
}

int bank_gate_launch(const BankStream* streams_dev, int n_streams, const BlurArgs& a, cudaStream_t stream) {
  TCD_REQUIRE(a.ksize == 21, "stream bank: the batched gate needs blur_size 21 (got %d)", a.ksize);
  Blur21Taps tp;
  blur21_taps(a, &tp);
  bank_gate_kernel<<<dim3(cdiv(a.w, kB21TW), cdiv(a.h, kB21TH), n_streams), 256, 0, stream>>>(streams_dev, a, tp);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  return cudaGetLastError() == cudaSuccess ? TCD_OK : TCD_ERR_CUDA;
}

int bank_mog2_launch(const BankStream* streams_dev, int n_streams, const Mog2Params& p, cudaStream_t stream) {
  bank_mog2_kernel<<<dim3(cdiv(p.npix, 256), n_streams), 256, 0, stream>>>(streams_dev, p);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  return cudaGetLastError() == cudaSuccess ? TCD_OK : TCD_ERR_CUDA;
}

int bank_roi_launch(const BankStream* streams_dev, int n_streams, const BlurArgs& a, cudaStream_t stream) {
  TCD_REQUIRE(a.ksize == 21, "stream bank: the batched clean-up needs motion_blur_size 21 (got %d)", a.ksize);
  Blur21Taps tp;
  blur21_taps(a, &tp);
  bank_roi_kernel<<<dim3(cdiv(a.w, kB21TW), cdiv(a.h, kB21TH), n_streams * kMogMaxRepeats), 256, 0, stream>>>(
      streams_dev, a, tp);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  return cudaGetLastError() == cudaSuccess ? TCD_OK : TCD_ERR_CUDA;
}

int threshold_launch(const uint8_t* src, int n, int thr, uint8_t* dst, cudaStream_t stream) {
  threshold_kernel<<<cdiv(n, 256), 256, 0, stream>>>(src, n, thr, dst);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  return cudaGetLastError() == cudaSuccess ? TCD_OK : TCD_ERR_CUDA;
}

int roi_count_launch(const uint8_t* mask, int w, int h, const int* rois_dev, int n,
                     unsigned int* counts_dev, cudaStream_t stream) {
  if (n <= 0) return TCD_OK;
  roi_count_kernel<<<dim3(n, 64), 256, 0, stream>>>(mask, w, h, rois_dev, counts_dev);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("roi_count_kernel launch: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return TCD_OK;
}

}  // namespace tcd
