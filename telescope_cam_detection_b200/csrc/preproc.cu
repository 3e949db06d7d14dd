// Fused preprocess (K1+K2): batch of HWC u8 camera frames -> network input, one launch.
//
// Replaces YOLOXDetector.preprocess (src/yolox_detector.py:186-220): from_numpy/permute/
// float/.to(device) + F.interpolate(bilinear, align_corners=False) per frame + torch.cat.
// Tap rule (ATen upsample_bilinear2d): scale = in/out (float),
// s = max(scale*(d+0.5)-0.5, 0), i0 = (int)s, i1 = i0 + (i0 < in-1), l1 = s - i0, l0 = 1-l1,
// out = l0y*(l0x*p00 + l1x*p01) + l1y*(l0x*p10 + l1x*p11), no antialias.
// The frames are uploaded as u8 (the reference uploads f32, 4x the PCIe bytes).
//
// Each CTA owns a 128x16 patch of output pixels: the input rows/columns it needs are staged
// into shared memory with 16-byte coalesced loads, then every thread produces a 2x2 output
// quad = one Focus (space-to-depth) pixel of 16 bf16 channels written as two 16-byte stores.
#include "kernels.cuh"

#include <atomic>

namespace tcd {
extern std::atomic<uint64_t> g_launch_count;

constexpr int kPreTileW = 128;
constexpr int kPreTileH = 16;
constexpr int kPreThreads = 256;

struct PreArgs {
  int out_h, out_w, swap_rb, layout;
  float scale, mean[3], inv_std[3], pad_value;
  int smem_row_bytes, smem_rows;  // 0 -> direct global loads
};

// u8 -> f32 without the conversion unit: I2F issues at 16 lanes/clk/SM (the XU pipe) and the kernel converts 12 bytes
// per output pixel (39 M per step of 8 frames = 8 us of that pipe alone); 2^23 + b is exact in fp32, so one LOP3 and
// one FADD give the same value at full rate
__device__ __forceinline__ float u8f(uint8_t b) { return __uint_as_float(0x4B000000u | (unsigned)b) - 8388608.0f; }

__device__ __forceinline__ void tap(float scale, int d, int in, int* i0, int* i1, float* l1) {
  float s = scale * ((float)d + 0.5f) - 0.5f;
  s = s < 0.f ? 0.f : s;
  int a = (int)s;
  if (a > in - 1) a = in - 1;
  *i0 = a;
  *i1 = a + (a < in - 1 ? 1 : 0);
  *l1 = s - (float)a;
}

__global__ void __launch_bounds__(kPreThreads)
preproc_kernel(const PreFrame* __restrict__ frames, const PreArgs a, void* __restrict__ out) {
  extern __shared__ __align__(16) uint8_t tile[];
  const PreFrame f = frames[blockIdx.z];
  const int ox_a = blockIdx.x * kPreTileW;
  const int oy_a = blockIdx.y * kPreTileH;
  const int ox_b = min(ox_a + kPreTileW, a.out_w);
  const int oy_b = min(oy_a + kPreTileH, a.out_h);

  // input span needed by the in-content part of this tile
  const int cx_b = min(ox_b, f.content_w), cy_b = min(oy_b, f.content_h);
  const bool has_content = ox_a < cx_b && oy_a < cy_b;
  int in_x0 = 0, in_y0 = 0, in_x1 = 0, in_y1 = 0;
  long long byte0 = 0;
  int row_bytes = 0;
  const bool staged = a.smem_rows > 0;
  if (has_content) {
    int t1;
    float l;
    tap(f.scale_x, ox_a, f.w, &in_x0, &t1, &l);
    tap(f.scale_y, oy_a, f.h, &in_y0, &t1, &l);
    int t0;
    tap(f.scale_x, cx_b - 1, f.w, &t0, &in_x1, &l);
    tap(f.scale_y, cy_b - 1, f.h, &t0, &in_y1, &l);
    if (staged) {
      // aligned byte window [byte0, byte0 + row_bytes) of every needed row; with row_step > 1 the needed rows are
      // exactly one per output row (in_y0 + r * row_step), the rows between them carry weight 0 and are not read
      const int nrows = f.row_step > 1 ? (cy_b - oy_a) : in_y1 - in_y0 + 1;
      const uintptr_t rowbase = (uintptr_t)f.ptr + (uintptr_t)((long long)in_y0 * f.stride);
      const long long b_lo = (long long)in_x0 * 3;
      const long long b_hi = (long long)(in_x1 + 1) * 3;
      // stride is a multiple of 16 on the staged path (host guarantees), so one alignment fits all rows
      const int mis = (int)((rowbase + b_lo) & 15);
      byte0 = b_lo - mis;
      row_bytes = (int)(((b_hi - byte0) + 15) & ~15LL);
      const int vec_per_row = row_bytes >> 4;
      // cp.async: every 16-byte piece of the tile's source window is requested before any of them is waited for
      // (4.5 per thread at the 3:1 letterbox) and lands in shared memory without passing through registers -- the
      // load-then-store loop had ONE request in flight per thread and the kernel sat at 0.30 of the HBM rate
      const uint32_t tile_s = (uint32_t)__cvta_generic_to_shared(tile);
      // a warp per source row, lanes along the row: no divisions in the address math
      for (int r = threadIdx.x >> 5; r < nrows; r += kPreThreads / 32) {
        const uint8_t* src = f.ptr + (long long)(in_y0 + r * f.row_step) * f.stride + byte0;
        const uint32_t dst = tile_s + (uint32_t)(r * a.smem_row_bytes);
        for (int v = threadIdx.x & 31; v < vec_per_row; v += 32)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (uint32_t)(v << 4)), "l"(src + (v << 4))
                       : "memory");
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
  }
  __syncthreads();

  if (!has_content && a.layout == 1) {
    // a tile of letterbox padding only (44 % of the tiles of a 1080p frame at 640x640): constant Focus pixels
    const float pv[3] = {(a.pad_value * a.scale - a.mean[0]) * a.inv_std[0], (a.pad_value * a.scale - a.mean[1]) * a.inv_std[1],
                         (a.pad_value * a.scale - a.mean[2]) * a.inv_std[2]};
    uint4 lo, hi;
    lo.x = pack_bf16x2(pv[0], pv[1]); lo.y = pack_bf16x2(pv[2], pv[0]); lo.z = pack_bf16x2(pv[1], pv[2]);
    lo.w = pack_bf16x2(pv[0], pv[1]); hi.x = pack_bf16x2(pv[2], pv[0]); hi.y = pack_bf16x2(pv[1], pv[2]);
    hi.z = 0u; hi.w = 0u;
    for (int qi = threadIdx.x; qi < (kPreTileW / 2) * (kPreTileH / 2); qi += kPreThreads) {
      const int qy = qi / (kPreTileW / 2), qx = qi - qy * (kPreTileW / 2);
      const int ox0 = ox_a + 2 * qx, oy0 = oy_a + 2 * qy;
      if (ox0 >= a.out_w || oy0 >= a.out_h) continue;
      const size_t pix = ((size_t)blockIdx.z * (a.out_h / 2) + (oy0 >> 1)) * (a.out_w / 2) + (ox0 >> 1);
      uint4* dst = reinterpret_cast<uint4*>(out) + pix * 2;
      dst[0] = lo;
      dst[1] = hi;
    }
    return;
  }

  // 64 x 8 quads per tile, 2 per thread
  const int qw = kPreTileW / 2;
  for (int qi = threadIdx.x; qi < qw * (kPreTileH / 2); qi += kPreThreads) {
    const int qy = qi / qw, qx = qi - qy * qw;
    const int ox0 = ox_a + 2 * qx, oy0 = oy_a + 2 * qy;
    if (ox0 >= a.out_w || oy0 >= a.out_h) continue;
    float v[2][2][3];  // [dy][dx][c]
    // taps of the two columns and the two rows of the quad, once each (they were recomputed per pixel)
    int xa[2], xb[2], ya[2], yb[2];
    float lxs[2], lys[2];
#pragma unroll
    for (int d = 0; d < 2; ++d) {
      tap(f.scale_x, ox0 + d, f.w, &xa[d], &xb[d], &lxs[d]);
      tap(f.scale_y, oy0 + d, f.h, &ya[d], &yb[d], &lys[d]);
      if (f.row_step > 1) yb[d] = ya[d];  // ly == 0 exactly: the second tap is the first one again, times 0
    }
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
      const int oy = oy0 + dy;
      const int y0 = ya[dy], y1 = yb[dy];
      const float ly = lys[dy];
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int ox = ox0 + dx;
        if (ox >= f.content_w || oy >= f.content_h) {
#pragma unroll
          for (int c = 0; c < 3; ++c) v[dy][dx][c] = a.pad_value;
          continue;
        }
        const int x0 = xa[dx], x1 = xb[dx];
        const float lx = lxs[dx];
        const float hx = 1.f - lx, hy = 1.f - ly;
        const uint8_t *p00, *p01, *p10, *p11;
        if (staged) {
          const int s0 = f.row_step > 1 ? oy - oy_a : y0 - in_y0;
          const int s1 = f.row_step > 1 ? oy - oy_a : y1 - in_y0;
          const uint8_t* r0 = tile + (size_t)s0 * a.smem_row_bytes - byte0;
          const uint8_t* r1 = tile + (size_t)s1 * a.smem_row_bytes - byte0;
          p00 = r0 + x0 * 3; p01 = r0 + x1 * 3; p10 = r1 + x0 * 3; p11 = r1 + x1 * 3;
        } else {
          const uint8_t* r0 = f.ptr + (long long)y0 * f.stride;
          const uint8_t* r1 = f.ptr + (long long)y1 * f.stride;
          p00 = r0 + x0 * 3; p01 = r0 + x1 * 3; p10 = r1 + x0 * 3; p11 = r1 + x1 * 3;
        }
        // A weight that is exactly 0 (integer resize ratios: 1920 -> 640 is 3:1 on x in both modes, 1080 -> 360 on y
        // in letterbox) drops its taps: h*p + 0*q == p*h bit for bit, so the short forms below ARE the bilinear
        // formula, with a quarter / half of the loads, conversions and FMAs
        if (lx == 0.f && ly == 0.f) {
#pragma unroll
          for (int c = 0; c < 3; ++c) v[dy][dx][c] = u8f(p00[a.swap_rb ? 2 - c : c]);
        } else if (lx == 0.f) {
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const int ci = a.swap_rb ? 2 - c : c;
            v[dy][dx][c] = hy * u8f(p00[ci]) + ly * u8f(p10[ci]);
          }
        } else {
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            const int ci = a.swap_rb ? 2 - c : c;
            const float top = hx * u8f(p00[ci]) + lx * u8f(p01[ci]);
            const float bot = hx * u8f(p10[ci]) + lx * u8f(p11[ci]);
            v[dy][dx][c] = hy * top + ly * bot;
          }
        }
      }
    }
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx)
#pragma unroll
        for (int c = 0; c < 3; ++c)
          v[dy][dx][c] = (v[dy][dx][c] * a.scale - a.mean[c]) * a.inv_std[c];

    if (a.layout == 1) {
      // Focus order: TL, BL, TR, BR  (upstream Focus.forward cat order)
      uint4 lo, hi;
      lo.x = pack_bf16x2(v[0][0][0], v[0][0][1]);
      lo.y = pack_bf16x2(v[0][0][2], v[1][0][0]);
      lo.z = pack_bf16x2(v[1][0][1], v[1][0][2]);
      lo.w = pack_bf16x2(v[0][1][0], v[0][1][1]);
      hi.x = pack_bf16x2(v[0][1][2], v[1][1][0]);
      hi.y = pack_bf16x2(v[1][1][1], v[1][1][2]);
      hi.z = 0u;
      hi.w = 0u;
      const size_t pix = ((size_t)blockIdx.z * (a.out_h / 2) + (oy0 >> 1)) * (a.out_w / 2) + (ox0 >> 1);
      uint4* dst = reinterpret_cast<uint4*>(out) + pix * 2;
      dst[0] = lo;
      dst[1] = hi;
    } else {
      float* o = reinterpret_cast<float*>(out) + (size_t)blockIdx.z * 3 * a.out_h * a.out_w;
#pragma unroll
      for (int c = 0; c < 3; ++c)
#pragma unroll
        for (int dy = 0; dy < 2; ++dy)
#pragma unroll
          for (int dx = 0; dx < 2; ++dx) {
            const int oy = oy0 + dy, ox = ox0 + dx;
            if (oy < a.out_h && ox < a.out_w) o[((size_t)c * a.out_h + oy) * a.out_w + ox] = v[dy][dx][c];
          }
    }
  }
}

// frames_dev: device array of PreFrame (B entries) prepared by the caller.
int preproc_launch(const tcd_preproc_cfg* cfg, const PreFrame* frames_host, const PreFrame* frames_dev,
                   int B, int layout, void* out_dev, cudaStream_t stream) {
  PreArgs a;
  a.out_h = cfg->out_h;
  a.out_w = cfg->out_w;
  a.swap_rb = cfg->swap_rb;
  a.layout = layout;
  a.scale = cfg->scale;
  for (int c = 0; c < 3; ++c) {
    a.mean[c] = cfg->mean[c];
    a.inv_std[c] = cfg->inv_std[c];
  }
  a.pad_value = cfg->pad_value;
  // staging geometry: worst case over the batch
  int max_rows = 0, max_row_bytes = 0;
  bool can_stage = true;
  for (int b = 0; b < B; ++b) {
    const PreFrame& f = frames_host[b];
    if ((f.stride & 15) != 0 || ((uintptr_t)f.ptr & 15) != 0) can_stage = false;
    int rows = f.row_step > 1 ? kPreTileH : (int)(f.scale_y * kPreTileH) + 4;
    int rb = (((int)(f.scale_x * kPreTileW) + 4) * 3 + 47) & ~15;
    if (rows > max_rows) max_rows = rows;
    if (rb > max_row_bytes) max_row_bytes = rb;
  }
  size_t smem = (size_t)max_rows * max_row_bytes;
  if (!can_stage || smem > 160 * 1024) {
    a.smem_rows = 0;
    a.smem_row_bytes = 0;
    smem = 0;
  } else {
    a.smem_rows = max_rows;
    a.smem_row_bytes = max_row_bytes;
  }
  if (int rc = ensure_dyn_smem((const void*)preproc_kernel, smem)) return rc;
  dim3 grid(cdiv(cfg->out_w, kPreTileW), cdiv(cfg->out_h, kPreTileH), B);
  preproc_kernel<<<grid, kPreThreads, smem, stream>>>(frames_dev, a, out_dev);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("preproc_kernel launch: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return TCD_OK;
}

// fill a PreFrame from frame geometry + cfg (host helper shared with the detector)
void preproc_describe(const tcd_preproc_cfg* cfg, const uint8_t* ptr, int h, int w, long long stride,
                      PreFrame* f) {
  f->ptr = ptr;
  f->stride = stride;
  f->h = h;
  f->w = w;
  if (cfg->mode == 1) {
    // upstream YOLOX preproc: r = min(H/h, W/w); content = (int(w*r), int(h*r)), top-left
    const float r = fminf((float)cfg->out_h / (float)h, (float)cfg->out_w / (float)w);
    int cw = (int)((float)w * r), ch = (int)((float)h * r);
    if (cw < 1) cw = 1;
    if (ch < 1) ch = 1;
    if (cw > cfg->out_w) cw = cfg->out_w;
    if (ch > cfg->out_h) ch = cfg->out_h;
    f->content_w = cw;
    f->content_h = ch;
    f->scale_x = (float)w / (float)cw;
    f->scale_y = (float)h / (float)ch;
  } else {
    f->content_w = cfg->out_w;
    f->content_h = cfg->out_h;
    f->scale_x = (float)w / (float)cfg->out_w;
    f->scale_y = (float)h / (float)cfg->out_h;
  }
  // s = k*(d+0.5)-0.5 = k*d + (k-1)/2 is an integer for an odd integer ratio k: exact in fp32, FMA or not
  const int k = (int)f->scale_y;
  const bool odd_int = (float)k == f->scale_y && k >= 3 && (k & 1) && (long long)k * f->content_h <= (long long)h;
  f->row_step = odd_int ? k : 1;
  f->row_phase = odd_int ? (k - 1) / 2 : 0;
}

}  // namespace tcd
