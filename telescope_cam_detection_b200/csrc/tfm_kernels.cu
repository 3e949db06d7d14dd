// RT-DETR transformer-side kernels, part 1: LayerNorm, row adds / masks, MHA, MSDeformAttn.
// See tfm_kernels.cuh for what each one replaces.
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

__device__ __forceinline__ void bf16x8_to_f32(const uint4 q, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float2 t = __bfloat1622float2(h[e]);
    f[2 * e] = t.x;
    f[2 * e + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 f32_to_bf16x8(const float (&f)[8]) {
  uint4 q;
  q.x = pack_bf16x2(f[0], f[1]);
  q.y = pack_bf16x2(f[2], f[3]);
  q.z = pack_bf16x2(f[4], f[5]);
  q.w = pack_bf16x2(f[6], f[7]);
  return q;
}

// ------------------------------------------------------------------ LayerNorm(a + b)
// one warp per row; each lane owns C/256 chunks of 8 channels (C <= 1024)
__global__ void __launch_bounds__(256)
layernorm_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, const float* __restrict__ gamma,
                 const float* __restrict__ beta, uint4* __restrict__ y, int rows, int C8, float eps) {
  const int row = (blockIdx.x * 256 + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  float v[4][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = lane + 32 * i;
    if (c < C8) {
      bf16x8_to_f32(a[(size_t)row * C8 + c], v[i]);
      if (b != nullptr) {
        float t[8];
        bf16x8_to_f32(b[(size_t)row * C8 + c], t);
#pragma unroll
        for (int e = 0; e < 8; ++e) v[i][e] += t[e];
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) sum += v[i][e];
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
  const float mean = sum / (float)(C8 * 8);
  float var = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = lane + 32 * i;
    if (c < C8) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float d = v[i][e] - mean;
        var += d * d;
      }
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) var += __shfl_xor_sync(0xffffffffu, var, off);
  const float rstd = rsqrtf(var / (float)(C8 * 8) + eps);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = lane + 32 * i;
    if (c < C8) {
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = (v[i][e] - mean) * rstd * gamma[c * 8 + e] + beta[c * 8 + e];
      y[(size_t)row * C8 + c] = f32_to_bf16x8(o);
    }
  }
}

int layernorm_launch(const void* a, const void* b, const float* gamma, const float* beta, void* y,
                     int rows, int C, float eps, cudaStream_t st) {
  TCD_REQUIRE(C % 8 == 0 && C <= 1024, "layernorm: C %d unsupported", C);
  layernorm_kernel<<<cdiv(rows, 8), 256, 0, st>>>((const uint4*)a, (const uint4*)b, gamma, beta, (uint4*)y,
                                                 rows, C / 8, eps);
  TFM_LAUNCH_OK("layernorm");
}

// ------------------------------------------------------------------ row add / mask
__global__ void __launch_bounds__(256)
add_rows_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, uint4* __restrict__ y,
                size_t total, size_t per_img, int bcast_b) {
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
    float fa[8], fb[8];
    bf16x8_to_f32(a[i], fa);
    bf16x8_to_f32(b[bcast_b ? i % per_img : i], fb);
#pragma unroll
    for (int e = 0; e < 8; ++e) fa[e] += fb[e];
    y[i] = f32_to_bf16x8(fa);
  }
}
int add_rows_launch(const void* a, const void* b, void* y, int N, int T, int C, int bcast_b,
                    cudaStream_t st) {
  TCD_REQUIRE(C % 8 == 0, "add_rows: C %d", C);
  const size_t per_img = (size_t)T * (C / 8), total = per_img * N;
  const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  add_rows_kernel<<<blocks, 256, 0, st>>>((const uint4*)a, (const uint4*)b, (uint4*)y, total, per_img,
                                          bcast_b);
  TFM_LAUNCH_OK("add_rows");
}

__global__ void __launch_bounds__(256)
mask_rows_kernel(const uint4* __restrict__ x, const float* __restrict__ mask, uint4* __restrict__ y,
                 size_t total, int T, int C8) {
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
    const int t = (int)((i / C8) % T);
    uint4 v = x[i];
    if (mask[t] == 0.f) v = make_uint4(0, 0, 0, 0);
    y[i] = v;
  }
}
int mask_rows_launch(const void* x, const float* mask, void* y, int N, int T, int C, cudaStream_t st) {
  TCD_REQUIRE(C % 8 == 0, "mask_rows: C %d", C);
  const size_t total = (size_t)N * T * (C / 8);
  const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  mask_rows_kernel<<<blocks, 256, 0, st>>>((const uint4*)x, mask, (uint4*)y, total, T, C / 8);
  TFM_LAUNCH_OK("mask_rows");
}

// ------------------------------------------------------------------ multi-head attention
// grid (ceil(Tq/128), heads, N); one query per thread, the head's K and V (Tk x 32 bf16 each) staged
// in shared memory, streaming (online) softmax in fp32.  Tk is a few hundred here (400 AIFI tokens,
// 300 decoder queries), so the whole K/V head slice fits and every K/V row read is a broadcast.
__global__ void __launch_bounds__(128)
mha_kernel(const __nv_bfloat16* __restrict__ q, int ldq, int qoff, const __nv_bfloat16* __restrict__ k,
           int ldk, int koff, const __nv_bfloat16* __restrict__ v, int ldv, int voff,
           __nv_bfloat16* __restrict__ out, int ldo, int Tq, int Tk, float scale) {
  // K / V head slices as fp32 in shared memory: every thread reads every row (a broadcast), so converting once at
  // staging instead of once per thread per key halves the inner loop; keys are taken eight at a time so the
  // running-max rescale of the 32 accumulators is paid once per eight keys (first version: 245 us per decoder
  // self-attention at B = 32, ~240 instructions per key per thread)
  extern __shared__ float4 mha_sm[];
  float4* Ks = mha_sm;            // [Tk][8] float4 = 32 floats
  float4* Vs = mha_sm + 8 * Tk;
  const int n = blockIdx.z, h = blockIdx.y;
  const __nv_bfloat16* kb = k + (size_t)n * Tk * ldk + koff + h * 32;
  const __nv_bfloat16* vb = v + (size_t)n * Tk * ldv + voff + h * 32;
  for (int i = threadIdx.x; i < 4 * Tk; i += 128) {
    const int t = i >> 2, c = i & 3;
    float f[8];
    bf16x8_to_f32(*reinterpret_cast<const uint4*>(kb + (size_t)t * ldk + c * 8), f);
    Ks[8 * t + 2 * c] = make_float4(f[0], f[1], f[2], f[3]);
    Ks[8 * t + 2 * c + 1] = make_float4(f[4], f[5], f[6], f[7]);
    bf16x8_to_f32(*reinterpret_cast<const uint4*>(vb + (size_t)t * ldv + c * 8), f);
    Vs[8 * t + 2 * c] = make_float4(f[0], f[1], f[2], f[3]);
    Vs[8 * t + 2 * c + 1] = make_float4(f[4], f[5], f[6], f[7]);
  }
  __syncthreads();
  const int tq = blockIdx.x * 128 + threadIdx.x;
  if (tq >= Tq) return;
  float qf[32];
  const __nv_bfloat16* qp = q + ((size_t)n * Tq + tq) * ldq + qoff + h * 32;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float t[8];
    bf16x8_to_f32(*reinterpret_cast<const uint4*>(qp + c * 8), t);
#pragma unroll
    for (int e = 0; e < 8; ++e) qf[c * 8 + e] = t[e] * scale;
  }
  float acc[32];
#pragma unroll
  for (int e = 0; e < 32; ++e) acc[e] = 0.f;
  float m = -INFINITY, l = 0.f;
  for (int j0 = 0; j0 < Tk; j0 += 8) {
    float sc[8];
    float bm = -INFINITY;
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) {
      float s = -INFINITY;
      if (j0 + jj < Tk) {
        const float4* kr = Ks + 8 * (j0 + jj);
        float s0 = 0.f, s1 = 0.f;   // two chains
#pragma unroll
        for (int c = 0; c < 8; c += 2) {
          const float4 a4 = kr[c], b4 = kr[c + 1];
          s0 = fmaf(qf[4 * c + 0], a4.x, s0); s0 = fmaf(qf[4 * c + 1], a4.y, s0);
          s0 = fmaf(qf[4 * c + 2], a4.z, s0); s0 = fmaf(qf[4 * c + 3], a4.w, s0);
          s1 = fmaf(qf[4 * c + 4], b4.x, s1); s1 = fmaf(qf[4 * c + 5], b4.y, s1);
          s1 = fmaf(qf[4 * c + 6], b4.z, s1); s1 = fmaf(qf[4 * c + 7], b4.w, s1);
        }
        s = s0 + s1;
      }
      sc[jj] = s;
      bm = fmaxf(bm, s);
    }
    const float mn = fmaxf(m, bm);
    const float corr = __expf(m - mn);
    l *= corr;
#pragma unroll
    for (int e = 0; e < 32; ++e) acc[e] *= corr;
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) {
      if (j0 + jj < Tk) {
        const float p = __expf(sc[jj] - mn);
        l += p;
        const float4* vr = Vs + 8 * (j0 + jj);
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float4 t4 = vr[c];
          acc[4 * c + 0] = fmaf(p, t4.x, acc[4 * c + 0]);
          acc[4 * c + 1] = fmaf(p, t4.y, acc[4 * c + 1]);
          acc[4 * c + 2] = fmaf(p, t4.z, acc[4 * c + 2]);
          acc[4 * c + 3] = fmaf(p, t4.w, acc[4 * c + 3]);
        }
      }
    }
    m = mn;
  }
  const float inv = 1.0f / l;
  __nv_bfloat16* op = out + ((size_t)n * Tq + tq) * ldo + h * 32;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    float t[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) t[e] = acc[c * 8 + e] * inv;
    *reinterpret_cast<uint4*>(op + c * 8) = f32_to_bf16x8(t);
  }
}

int mha_launch(const void* q, int ldq, int qoff, const void* k, int ldk, int koff, const void* v, int ldv,
               int voff, void* out, int ldo, int N, int Tq, int Tk, int heads, float scale,
               cudaStream_t st) {
  const size_t smem = (size_t)16 * Tk * sizeof(float4);
  TCD_REQUIRE(smem <= 200 * 1024, "mha: %d keys do not fit shared memory", Tk);
  TCD_REQUIRE(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && ldo % 8 == 0 && qoff % 8 == 0 && koff % 8 == 0 &&
                  voff % 8 == 0, "mha: leading dims / offsets must be multiples of 8");
  static size_t attr_set = 0;
  if (smem > 48 * 1024 && smem > attr_set) {
    TCD_CUDA_OK(cudaFuncSetAttribute(mha_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    attr_set = smem;
  }
  mha_kernel<<<dim3(cdiv(Tq, 128), heads, N), 128, smem, st>>>(
      (const __nv_bfloat16*)q, ldq, qoff, (const __nv_bfloat16*)k, ldk, koff, (const __nv_bfloat16*)v, ldv,
      voff, (__nv_bfloat16*)out, ldo, Tq, Tk, scale);
  TFM_LAUNCH_OK("mha");
}

// ------------------------------------------------------------------ MSDeformAttn v2
// one warp per (image, query, head); lane = channel of the 32-wide head slice, so every bilinear tap
// is one 64-byte coalesced read of the value row.  L*P <= 32 sampling points.
__global__ void __launch_bounds__(256)
msda_kernel(const __nv_bfloat16* __restrict__ value, const float* __restrict__ offsets,
            const float* __restrict__ logits, const float* __restrict__ ref, __nv_bfloat16* __restrict__ out,
            int total_warps, int Q, int S, int heads, MsdaLevels lv, float offset_scale) {
  const int wid = (blockIdx.x * 256 + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (wid >= total_warps) return;
  const int h = wid % heads;
  const int nq = wid / heads;  // n * Q + query
  const int n = nq / Q;
  const int LP = lv.n_levels * lv.n_points;
  // softmax over the LP logits of this (query, head)
  const float lg = lane < LP ? logits[((size_t)nq * heads + h) * LP + lane] : -INFINITY;
  float mx = lg;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
  float ex = lane < LP ? expf(lg - mx) : 0.f;
  float sm = ex;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) sm += __shfl_xor_sync(0xffffffffu, sm, off);
  const float wgt = ex / sm;
  // sampling location of point `lane`
  const float* r = ref + (size_t)nq * 4;
  const float rcx = r[0], rcy = r[1], rw = r[2], rh = r[3];
  float ox = 0.f, oy = 0.f;
  if (lane < LP) {
    const float2 o = *reinterpret_cast<const float2*>(offsets + (((size_t)nq * heads + h) * LP + lane) * 2);
    ox = o.x;
    oy = o.y;
  }
  const float pscale = 1.0f / (float)lv.n_points;
  const float lx = rcx + ox * pscale * rw * offset_scale;
  const float ly = rcy + oy * pscale * rh * offset_scale;
  const __nv_bfloat16* vb = value + (size_t)n * S * heads * 32 + h * 32 + lane;
  float acc = 0.f;
  for (int p = 0; p < LP; ++p) {
    const int lvl = p / lv.n_points;
    const int H = lv.h[lvl], W = lv.w[lvl];
    const float w_p = __shfl_sync(0xffffffffu, wgt, p);
    const float px = __shfl_sync(0xffffffffu, lx, p) * (float)W - 0.5f;
    const float py = __shfl_sync(0xffffffffu, ly, p) * (float)H - 0.5f;
    const float fx = floorf(px), fy = floorf(py);
    const int x0 = (int)fx, y0 = (int)fy;
    const float ax = px - fx, ay = py - fy;
    const __nv_bfloat16* base = vb + (size_t)lv.start[lvl] * heads * 32;
    float s = 0.f;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int xx = x0 + (t & 1), yy = y0 + (t >> 1);
      if (xx >= 0 && xx < W && yy >= 0 && yy < H) {
        const float wt = ((t & 1) ? ax : 1.f - ax) * ((t >> 1) ? ay : 1.f - ay);
        s += wt * __bfloat162float(base[(size_t)(yy * W + xx) * heads * 32]);
      }
    }
    acc += w_p * s;
  }
  out[((size_t)nq * heads + h) * 32 + lane] = __float2bfloat16(acc);
}

int msda_launch(const void* value, const float* offsets, const float* logits, const float* ref, void* out,
                int N, int Q, int S, int heads, MsdaLevels lv, float offset_scale, cudaStream_t st) {
  TCD_REQUIRE(lv.n_levels >= 1 && lv.n_levels <= 4 && lv.n_levels * lv.n_points <= 32,
              "msda: %d levels x %d points unsupported", lv.n_levels, lv.n_points);
  const int warps = N * Q * heads;
  msda_kernel<<<cdiv(warps, 8), 256, 0, st>>>((const __nv_bfloat16*)value, offsets, logits, ref,
                                              (__nv_bfloat16*)out, warps, Q, S, heads, lv, offset_scale);
  TFM_LAUNCH_OK("msda");
}

}  // namespace tcd
