// RT-DETR transformer-side kernels, part 2: row max, per-row top-k (radix select), gathers,
// reference-point updates, the postprocessor, and the two ResNet pools.
#include "tfm_kernels.cuh"

#include <atomic>

namespace tcd {
extern std::atomic<uint64_t> g_launch_count;

#define TFM_LAUNCH_OK(name)                                                  \
  do {                                                                       \
    g_launch_count.fetch_add(1, std::memory_order_relaxed);                  \
    cudaError_t _e = cudaGetLastError();                                     \
    if (_e != cudaSuccess) {                                                 \
      set_last_error(name " launch: %s", cudaGetErrorString(_e));            \
      return TCD_ERR_CUDA;                                                   \
    }                                                                        \
    return TCD_OK;                                                           \
  } while (0)

// ------------------------------------------------------------------ row max (one warp per row)
__global__ void __launch_bounds__(256) rowmax_kernel(const float* __restrict__ x, float* __restrict__ y,
                                                     int rows, int C) {
  const int row = (blockIdx.x * 256 + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  float m = -INFINITY;
  for (int c = lane; c < C; c += 32) m = fmaxf(m, x[(size_t)row * C + c]);
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
  if (lane == 0) y[row] = m;
}
int rowmax_launch(const float* x, float* y, int rows, int C, cudaStream_t st) {
  rowmax_kernel<<<cdiv(rows, 8), 256, 0, st>>>(x, y, rows, C);
  TFM_LAUNCH_OK("rowmax");
}

// ------------------------------------------------------------------ per-row top-k
// One CTA per row.  Every element gets the unique 64-bit key (orderable(value) << 32) | ~index, so
// "k largest, lowest index first on equal values" is an exact k-th-key selection: eight 8-bit radix
// passes find that key, one more pass collects the k keys >= it, a bitonic sort orders them.
__device__ __forceinline__ unsigned long long topk_key(float f, int i) {
  unsigned int u = __float_as_uint(f);
  u ^= (u >> 31) ? 0xFFFFFFFFu : 0x80000000u;
  return ((unsigned long long)u << 32) | (unsigned long long)(0xFFFFFFFFu - (unsigned int)i);
}

__global__ void __launch_bounds__(1024) topk_kernel(const float* __restrict__ keys, int n, int k,
                                                    int* __restrict__ idx_out) {
  __shared__ unsigned int hist[256];
  __shared__ unsigned long long cand[512];
  __shared__ unsigned long long s_prefix;
  __shared__ int s_krem, s_count;
  const float* row = keys + (size_t)blockIdx.x * n;
  if (threadIdx.x == 0) {
    s_prefix = 0ull;
    s_krem = k;
    s_count = 0;
  }
  for (int pass = 7; pass >= 0; --pass) {
    if (threadIdx.x < 256) hist[threadIdx.x] = 0u;
    __syncthreads();
    const unsigned long long prefix = s_prefix;
    const int shift = pass * 8;
    for (int i = threadIdx.x; i < n; i += 1024) {
      const unsigned long long key = topk_key(row[i], i);
      const bool match = (pass == 7) || ((key >> (shift + 8)) == (prefix >> (shift + 8)));
      if (match) atomicAdd(&hist[(unsigned int)(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int krem = s_krem;
      int b = 255;
      for (; b > 0; --b) {
        const int c = (int)hist[b];
        if (c >= krem) break;
        krem -= c;
      }
      s_prefix = prefix | ((unsigned long long)b << shift);
      s_krem = krem;
    }
    __syncthreads();
  }
  const unsigned long long kth = s_prefix;
  for (int i = threadIdx.x; i < 512; i += 1024) cand[i] = 0ull;
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += 1024) {
    const unsigned long long key = topk_key(row[i], i);
    if (key >= kth) {
      const int pos = atomicAdd(&s_count, 1);
      if (pos < 512) cand[pos] = key;
    }
  }
  __syncthreads();
  for (int kk = 2; kk <= 512; kk <<= 1) {
    for (int j = kk >> 1; j > 0; j >>= 1) {
      const int i = threadIdx.x;
      if (i < 512) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long x = cand[i], y = cand[ixj];
          const bool desc = ((i & kk) == 0);
          if (desc ? (x < y) : (x > y)) {
            cand[i] = y;
            cand[ixj] = x;
          }
        }
      }
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < k; i += 1024)
    idx_out[(size_t)blockIdx.x * k + i] = (int)(0xFFFFFFFFu - (unsigned int)(cand[i] & 0xFFFFFFFFull));
}
int topk_launch(const float* keys, int N, int n, int k, int* idx_out, cudaStream_t st) {
  TCD_REQUIRE(k >= 1 && k <= 512 && k <= n, "topk: k %d of %d unsupported", k, n);
  topk_kernel<<<N, 1024, 0, st>>>(keys, n, k, idx_out);
  TFM_LAUNCH_OK("topk");
}

// ------------------------------------------------------------------ gathers / reference points
__global__ void __launch_bounds__(256)
gather_rows_kernel(const uint4* __restrict__ src, const int* __restrict__ idx, uint4* __restrict__ out,
                   size_t total, int S, int K, int C8) {
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
    const int c = (int)(i % C8);
    const size_t r = i / C8;
    const int n = (int)(r / K);
    out[i] = src[((size_t)n * S + idx[r]) * C8 + c];
  }
}
int gather_rows_launch(const void* src, const int* idx, void* out, int N, int S, int K, int C,
                       cudaStream_t st) {
  TCD_REQUIRE(C % 8 == 0, "gather_rows: C %d", C);
  const size_t total = (size_t)N * K * (C / 8);
  gather_rows_kernel<<<(int)((total + 255) / 256), 256, 0, st>>>((const uint4*)src, idx, (uint4*)out, total, S,
                                                                 K, C / 8);
  TFM_LAUNCH_OK("gather_rows");
}

__device__ __forceinline__ float sigmoid_f(float x) { return 1.0f / (1.0f + expf(-x)); }
__device__ __forceinline__ void ref_store(float* ref_sig, __nv_bfloat16* ref_bf16, int row, const float (&s)[4]) {
  *reinterpret_cast<float4*>(ref_sig + (size_t)row * 4) = make_float4(s[0], s[1], s[2], s[3]);
  __nv_bfloat16* o = ref_bf16 + (size_t)row * 16;
#pragma unroll
  for (int e = 0; e < 16; ++e) o[e] = __float2bfloat16(e < 4 ? s[e] : 0.f);
}

__global__ void __launch_bounds__(256)
ref_init_kernel(const float* __restrict__ delta, int ldd, const float* __restrict__ anchors,
                const int* __restrict__ idx, float* __restrict__ ref_sig, __nv_bfloat16* __restrict__ ref_bf16,
                int rows) {
  const int row = blockIdx.x * 256 + threadIdx.x;
  if (row >= rows) return;
  const float* a = anchors + (size_t)idx[row] * 4;
  float s[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) s[e] = sigmoid_f(delta[(size_t)row * ldd + e] + a[e]);
  ref_store(ref_sig, ref_bf16, row, s);
}
int ref_init_launch(const float* delta, int ldd, const float* anchors, const int* idx, float* ref_sig,
                    void* ref_bf16, int N, int K, cudaStream_t st) {
  ref_init_kernel<<<cdiv(N * K, 256), 256, 0, st>>>(delta, ldd, anchors, idx, ref_sig, (__nv_bfloat16*)ref_bf16,
                                                    N * K);
  TFM_LAUNCH_OK("ref_init");
}

__global__ void __launch_bounds__(256)
ref_update_kernel(const float* __restrict__ delta, int ldd, float* __restrict__ ref_sig,
                  __nv_bfloat16* __restrict__ ref_bf16, int rows) {
  const int row = blockIdx.x * 256 + threadIdx.x;
  if (row >= rows) return;
  float s[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    float x = ref_sig[(size_t)row * 4 + e];
    x = fminf(fmaxf(x, 0.f), 1.f);
    const float x1 = fmaxf(x, 1e-5f), x2 = fmaxf(1.f - x, 1e-5f);
    s[e] = sigmoid_f(delta[(size_t)row * ldd + e] + logf(x1 / x2));
  }
  ref_store(ref_sig, ref_bf16, row, s);
}
int ref_update_launch(const float* delta, int ldd, float* ref_sig, void* ref_bf16, int rows,
                      cudaStream_t st) {
  ref_update_kernel<<<cdiv(rows, 256), 256, 0, st>>>(delta, ldd, ref_sig, (__nv_bfloat16*)ref_bf16, rows);
  TFM_LAUNCH_OK("ref_update");
}

__global__ void __launch_bounds__(256)
detr_post_kernel(const float* __restrict__ logits, const float* __restrict__ ref_sig,
                 const int* __restrict__ idx, const float* __restrict__ orig_wh, int Q, int C, int k,
                 int total, int* __restrict__ labels, float* __restrict__ boxes, float* __restrict__ scores) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const int n = i / k;
  const int flat = idx[i];
  const int q = flat / C;
  labels[i] = flat - q * C;
  scores[i] = sigmoid_f(logits[(size_t)n * Q * C + flat]);
  const float* r = ref_sig + ((size_t)n * Q + q) * 4;
  const float w = orig_wh[2 * n], h = orig_wh[2 * n + 1];
  boxes[4 * (size_t)i + 0] = (r[0] - 0.5f * r[2]) * w;
  boxes[4 * (size_t)i + 1] = (r[1] - 0.5f * r[3]) * h;
  boxes[4 * (size_t)i + 2] = (r[0] + 0.5f * r[2]) * w;
  boxes[4 * (size_t)i + 3] = (r[1] + 0.5f * r[3]) * h;
}
int detr_post_launch(const float* logits, const float* ref_sig, const int* idx, const float* orig_wh,
                     int N, int Q, int C, int k, int* labels, float* boxes, float* scores,
                     cudaStream_t st) {
  detr_post_kernel<<<cdiv(N * k, 256), 256, 0, st>>>(logits, ref_sig, idx, orig_wh, Q, C, k, N * k, labels,
                                                     boxes, scores);
  TFM_LAUNCH_OK("detr_post");
}

// ------------------------------------------------------------------ pools (bf16 NHWC, 8 channels/thread)
__global__ void __launch_bounds__(256)
maxpool3x3s2_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, size_t total, int H, int W, int C8) {
  const int Ho = H / 2, Wo = W / 2;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
    const int c = (int)(i % C8);
    size_t r = i / C8;
    const int ox = (int)(r % Wo);
    r /= Wo;
    const int oy = (int)(r % Ho);
    const int n = (int)(r / Ho);
    __nv_bfloat162 m[4];
    bool first = true;
    for (int dy = -1; dy <= 1; ++dy) {
      const int yy = 2 * oy + dy;
      if (yy < 0 || yy >= H) continue;
      for (int dx = -1; dx <= 1; ++dx) {
        const int xx = 2 * ox + dx;
        if (xx < 0 || xx >= W) continue;
        const uint4 v = x[((size_t)((size_t)n * H + yy) * W + xx) * C8 + c];
        const __nv_bfloat162* pv = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
        for (int e = 0; e < 4; ++e) m[e] = first ? pv[e] : __hmax2(m[e], pv[e]);
        first = false;
      }
    }
    y[i] = *reinterpret_cast<uint4*>(m);
  }
}
int maxpool3x3s2_launch(const void* x, void* y, int N, int H, int W, int C, cudaStream_t st) {
  TCD_REQUIRE(C % 8 == 0 && H % 2 == 0 && W % 2 == 0, "maxpool: shape %dx%dx%d", H, W, C);
  const size_t total = (size_t)N * (H / 2) * (W / 2) * (C / 8);
  const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  maxpool3x3s2_kernel<<<blocks, 256, 0, st>>>((const uint4*)x, (uint4*)y, total, H, W, C / 8);
  TFM_LAUNCH_OK("maxpool3x3s2");
}

__global__ void __launch_bounds__(256)
avgpool2x2_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, size_t total, int H, int W, int C8) {
  const int Ho = H / 2, Wo = W / 2;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
    const int c = (int)(i % C8);
    size_t r = i / C8;
    const int ox = (int)(r % Wo);
    r /= Wo;
    const int oy = (int)(r % Ho);
    const int n = (int)(r / Ho);
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const uint4 v = x[((size_t)((size_t)n * H + 2 * oy + (t >> 1)) * W + 2 * ox + (t & 1)) * C8 + c];
      const __nv_bfloat162* pv = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = __bfloat1622float2(pv[e]);
        acc[2 * e] += f.x;
        acc[2 * e + 1] += f.y;
      }
    }
    uint4 o;
    o.x = pack_bf16x2(acc[0] * 0.25f, acc[1] * 0.25f);
    o.y = pack_bf16x2(acc[2] * 0.25f, acc[3] * 0.25f);
    o.z = pack_bf16x2(acc[4] * 0.25f, acc[5] * 0.25f);
    o.w = pack_bf16x2(acc[6] * 0.25f, acc[7] * 0.25f);
    y[i] = o;
  }
}
int avgpool2x2_launch(const void* x, void* y, int N, int H, int W, int C, cudaStream_t st) {
  TCD_REQUIRE(C % 8 == 0 && H % 2 == 0 && W % 2 == 0, "avgpool: shape %dx%dx%d", H, W, C);
  const size_t total = (size_t)N * (H / 2) * (W / 2) * (C / 8);
  const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  avgpool2x2_kernel<<<blocks, 256, 0, st>>>((const uint4*)x, (uint4*)y, total, H, W, C / 8);
  TFM_LAUNCH_OK("avgpool2x2");
}

}  // namespace tcd
