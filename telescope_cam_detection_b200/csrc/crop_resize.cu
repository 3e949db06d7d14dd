// Stage-2 crop-and-resize gather (K13): all crops of a frame in one launch.
//
// Replaces, per detection, the slice frame[y1:y2, x1:x2] (src/two_stage_pipeline_yolox.py:289)
// followed by SpeciesClassifier.preprocess (src/species_classifier.py:311-350): BGR->RGB gather,
// bilinear resize to SxS (same ATen tap rule as preproc.cu, no antialias), /255, (x-mean)/std.
// The integer crop geometry (int() truncation, padding percent, clamps) stays on the host in
// the Python shim exactly as the reference computes it; the kernel receives final boxes.
#include "kernels.cuh"

#include <atomic>

namespace tcd {
extern std::atomic<uint64_t> g_launch_count;

struct CropArgs {
  int h, w, S, swap_rb, layout, n;
  long long stride;
  float scale, mean[3], inv_std[3];
};

__device__ __forceinline__ void ctap(float scale, int d, int in, int* i0, int* i1, float* l1) {
  float s = scale * ((float)d + 0.5f) - 0.5f;
  s = s < 0.f ? 0.f : s;
  int a = (int)s;
  if (a > in - 1) a = in - 1;
  *i0 = a;
  *i1 = a + (a < in - 1 ? 1 : 0);
  *l1 = s - (float)a;
}

// grid (ceil(S*S/256), n)
__global__ void __launch_bounds__(256)
crop_resize_kernel(const uint8_t* __restrict__ frame, const int* __restrict__ boxes, const CropArgs a,
                   void* __restrict__ out) {
  const int crop = blockIdx.y;
  const int pix = blockIdx.x * 256 + threadIdx.x;
  if (pix >= a.S * a.S) return;
  const int oy = pix / a.S, ox = pix - oy * a.S;
  const int bx1 = boxes[crop * 4 + 0], by1 = boxes[crop * 4 + 1];
  const int cw = boxes[crop * 4 + 2] - bx1, ch = boxes[crop * 4 + 3] - by1;
  float v[3] = {0.f, 0.f, 0.f};
  if (cw > 0 && ch > 0) {
    const uint8_t* base = frame + (long long)by1 * a.stride + (long long)bx1 * 3;
    if (cw == a.S && ch == a.S) {
      const uint8_t* p = base + (long long)oy * a.stride + ox * 3;
#pragma unroll
      for (int c = 0; c < 3; ++c) v[c] = (float)p[a.swap_rb ? 2 - c : c];
    } else {
      int x0, x1, y0, y1;
      float lx, ly;
      ctap((float)cw / (float)a.S, ox, cw, &x0, &x1, &lx);
      ctap((float)ch / (float)a.S, oy, ch, &y0, &y1, &ly);
      const float hx = 1.f - lx, hy = 1.f - ly;
      const uint8_t* r0 = base + (long long)y0 * a.stride;
      const uint8_t* r1 = base + (long long)y1 * a.stride;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const int ci = a.swap_rb ? 2 - c : c;
        const float top = hx * (float)r0[x0 * 3 + ci] + lx * (float)r0[x1 * 3 + ci];
        const float bot = hx * (float)r1[x0 * 3 + ci] + lx * (float)r1[x1 * 3 + ci];
        v[c] = hy * top + ly * bot;
      }
    }
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) v[c] = (v[c] * a.scale - a.mean[c]) * a.inv_std[c];
  if (a.layout == 1) {
    uint2 q;
    q.x = pack_bf16x2(v[0], v[1]);
    q.y = pack_bf16x2(v[2], 0.f);
    reinterpret_cast<uint2*>(out)[(size_t)crop * a.S * a.S + pix] = q;
  } else {
    float* o = reinterpret_cast<float*>(out) + (size_t)crop * 3 * a.S * a.S;
#pragma unroll
    for (int c = 0; c < 3; ++c) o[(size_t)c * a.S * a.S + pix] = v[c];
  }
}

int crop_resize_launch(const uint8_t* frame, int h, int w, long long stride, const int* boxes_dev,
                       int n, int S, const tcd_preproc_cfg* cfg, int layout, void* out,
                       cudaStream_t stream) {
  if (n <= 0) return TCD_OK;
  CropArgs a;
  a.h = h; a.w = w; a.S = S; a.swap_rb = cfg->swap_rb; a.layout = layout; a.n = n;
  a.stride = stride;
  a.scale = cfg->scale;
  for (int c = 0; c < 3; ++c) { a.mean[c] = cfg->mean[c]; a.inv_std[c] = cfg->inv_std[c]; }
  crop_resize_kernel<<<dim3(cdiv(S * S, 256), n), 256, 0, stream>>>(frame, boxes_dev, a, out);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("crop_resize_kernel launch: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return TCD_OK;
}

}  // namespace tcd
