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
template <bool AFFINE>
__global__ void __launch_bounds__(256, AFFINE ? 4 : 5)
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
      if constexpr (AFFINE) {
        // gamma / beta as two 16-byte loads each (eight scalar loads per chunk made the kernel LSU-bound: 1.8 TB/s)
        const float4 g0 = *reinterpret_cast<const float4*>(gamma + c * 8), g1 = *reinterpret_cast<const float4*>(gamma + c * 8 + 4);
        const float4 b0 = *reinterpret_cast<const float4*>(beta + c * 8), b1 = *reinterpret_cast<const float4*>(beta + c * 8 + 4);
        const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = (v[i][e] - mean) * rstd * gg[e] + bb[e];
      } else {
        // no affine part: the caller folded gamma / beta into the Linear that reads y (W' = W diag(gamma), b' = b + W beta)
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = (v[i][e] - mean) * rstd;
      }
      y[(size_t)row * C8 + c] = f32_to_bf16x8(o);
    }
  }
}

int layernorm_launch(const void* a, const void* b, const float* gamma, const float* beta, void* y,
                     int rows, int C, float eps, cudaStream_t st) {
  TCD_REQUIRE(C % 8 == 0 && C <= 1024, "layernorm: C %d unsupported", C);
  if (gamma != nullptr)
    layernorm_kernel<true><<<cdiv(rows, 8), 256, 0, st>>>((const uint4*)a, (const uint4*)b, gamma, beta, (uint4*)y, rows, C / 8, eps);
  else   // no affine part (folded into the consumer): fewer registers, five CTAs per SM
    layernorm_kernel<false><<<cdiv(rows, 8), 256, 0, st>>>((const uint4*)a, (const uint4*)b, gamma, beta, (uint4*)y, rows, C / 8, eps);
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
// Flash-style attention on the warp-level tensor-core path (mma.sync m16n8k16, bf16 in, fp32 accumulate): head
// dim 32, a few hundred keys (400 AIFI tokens, 300 decoder queries), so the head's whole K (row-major) and V
// (transposed) slices sit in shared memory.  grid (ceil(Tq/64), heads, N); each of the 4 warps owns 16 queries
// and walks the keys 16 at a time: S = Q K^T (4 MMAs), online softmax on the accumulator fragments (row max /
// sum across the 4 lanes of a quad), O += P V (4 MMAs) with P re-used straight from the S fragments.
// (The first version was one query per thread on the CUDA cores: 0.245 ms per decoder layer at B = 32, then
// 0.152 ms with fp32 staging; tcgen05 would want 128-query tiles, which 300 queries x 8 heads do not fill.)
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

constexpr int kMhaKStride = 40;  // bf16 per staged K row (32 + 8 pad): conflict-free 32-bit fragment loads

__global__ void __launch_bounds__(128)
mha_kernel(const __nv_bfloat16* __restrict__ q, int ldq, int qoff, const __nv_bfloat16* __restrict__ k,
           int ldk, int koff, const __nv_bfloat16* __restrict__ v, int ldv, int voff,
           __nv_bfloat16* __restrict__ out, int ldo, int Tq, int Tk, float scale) {
  extern __shared__ __align__(16) uint8_t mha_sm_raw[];
  const int Tkp = (Tk + 15) & ~15;
  const int vstride = Tkp + 8;                        // bf16 per transposed V row
  __nv_bfloat16* Ks = reinterpret_cast<__nv_bfloat16*>(mha_sm_raw);   // [Tkp][40]
  __nv_bfloat16* Vt = Ks + (size_t)Tkp * kMhaKStride;                 // [32][Tkp + 8]
  const int n = blockIdx.z, h = blockIdx.y;
  const __nv_bfloat16* kb = k + (size_t)n * Tk * ldk + koff + h * 32;
  const __nv_bfloat16* vb = v + (size_t)n * Tk * ldv + voff + h * 32;
  for (int i = threadIdx.x; i < 4 * Tkp; i += 128) {
    const int t = i >> 2, c = i & 3;
    uint4 kq = make_uint4(0u, 0u, 0u, 0u), vq = kq;
    if (t < Tk) {
      kq = *reinterpret_cast<const uint4*>(kb + (size_t)t * ldk + c * 8);
      vq = *reinterpret_cast<const uint4*>(vb + (size_t)t * ldv + c * 8);
    }
    *reinterpret_cast<uint4*>(Ks + (size_t)t * kMhaKStride + c * 8) = kq;
    const __nv_bfloat16* ve = reinterpret_cast<const __nv_bfloat16*>(&vq);
#pragma unroll
    for (int e = 0; e < 8; ++e) Vt[(size_t)(c * 8 + e) * vstride + t] = ve[e];
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int q0 = blockIdx.x * 64 + warp * 16;
  if (q0 >= Tq) return;
  const int r_lo = q0 + g, r_hi = q0 + g + 8;
  // Q fragments (A operand, row-major 16 x 16 per k-step)
  uint32_t qa[2][4];
  {
    const __nv_bfloat16* qlo = q + ((size_t)n * Tq + (r_lo < Tq ? r_lo : 0)) * ldq + qoff + h * 32;
    const __nv_bfloat16* qhi = q + ((size_t)n * Tq + (r_hi < Tq ? r_hi : 0)) * ldq + qoff + h * 32;
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
      qa[ks][0] = r_lo < Tq ? *reinterpret_cast<const uint32_t*>(qlo + ks * 16 + 2 * t4) : 0u;
      qa[ks][1] = r_hi < Tq ? *reinterpret_cast<const uint32_t*>(qhi + ks * 16 + 2 * t4) : 0u;
      qa[ks][2] = r_lo < Tq ? *reinterpret_cast<const uint32_t*>(qlo + ks * 16 + 2 * t4 + 8) : 0u;
      qa[ks][3] = r_hi < Tq ? *reinterpret_cast<const uint32_t*>(qhi + ks * 16 + 2 * t4 + 8) : 0u;
    }
  }
  float o[4][4];
#pragma unroll
  for (int nb = 0; nb < 4; ++nb)
#pragma unroll
    for (int e = 0; e < 4; ++e) o[nb][e] = 0.f;
  float m_lo = -INFINITY, m_hi = -INFINITY, l_lo = 0.f, l_hi = 0.f;   // l: this lane's partial row sums
  for (int key0 = 0; key0 < Tkp; key0 += 16) {
    float sc[2][4];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
#pragma unroll
      for (int e = 0; e < 4; ++e) sc[j][e] = 0.f;
      const __nv_bfloat16* kr = Ks + (size_t)(key0 + j * 8 + g) * kMhaKStride + 2 * t4;
#pragma unroll
      for (int ks = 0; ks < 2; ++ks)
        mma_bf16_16816(sc[j], qa[ks], *reinterpret_cast<const uint32_t*>(kr + ks * 16),
                       *reinterpret_cast<const uint32_t*>(kr + ks * 16 + 8));
    }
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int key = key0 + j * 8 + 2 * t4 + (e & 1);
        sc[j][e] = key < Tk ? sc[j][e] * scale : -INFINITY;
      }
    float mx_lo = fmaxf(fmaxf(sc[0][0], sc[0][1]), fmaxf(sc[1][0], sc[1][1]));
    float mx_hi = fmaxf(fmaxf(sc[0][2], sc[0][3]), fmaxf(sc[1][2], sc[1][3]));
#pragma unroll
    for (int off = 1; off <= 2; off <<= 1) {
      mx_lo = fmaxf(mx_lo, __shfl_xor_sync(0xffffffffu, mx_lo, off));
      mx_hi = fmaxf(mx_hi, __shfl_xor_sync(0xffffffffu, mx_hi, off));
    }
    const float mn_lo = fmaxf(m_lo, mx_lo), mn_hi = fmaxf(m_hi, mx_hi);
    const float c_lo = __expf(m_lo - mn_lo), c_hi = __expf(m_hi - mn_hi);
    m_lo = mn_lo;
    m_hi = mn_hi;
    float p[2][4];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      p[j][0] = __expf(sc[j][0] - mn_lo);
      p[j][1] = __expf(sc[j][1] - mn_lo);
      p[j][2] = __expf(sc[j][2] - mn_hi);
      p[j][3] = __expf(sc[j][3] - mn_hi);
    }
    l_lo = l_lo * c_lo + ((p[0][0] + p[0][1]) + (p[1][0] + p[1][1]));
    l_hi = l_hi * c_hi + ((p[0][2] + p[0][3]) + (p[1][2] + p[1][3]));
#pragma unroll
    for (int nb = 0; nb < 4; ++nb) {
      o[nb][0] *= c_lo; o[nb][1] *= c_lo;
      o[nb][2] *= c_hi; o[nb][3] *= c_hi;
    }
    // the S accumulator fragments of two 8-key blocks ARE the A fragment of the 16-key P V step
    uint32_t pa[4];
    pa[0] = pack_bf16x2(p[0][0], p[0][1]);
    pa[1] = pack_bf16x2(p[0][2], p[0][3]);
    pa[2] = pack_bf16x2(p[1][0], p[1][1]);
    pa[3] = pack_bf16x2(p[1][2], p[1][3]);
#pragma unroll
    for (int nb = 0; nb < 4; ++nb) {
      const __nv_bfloat16* vr = Vt + (size_t)(nb * 8 + g) * vstride + key0 + 2 * t4;
      mma_bf16_16816(o[nb], pa, *reinterpret_cast<const uint32_t*>(vr), *reinterpret_cast<const uint32_t*>(vr + 8));
    }
  }
#pragma unroll
  for (int off = 1; off <= 2; off <<= 1) {
    l_lo += __shfl_xor_sync(0xffffffffu, l_lo, off);
    l_hi += __shfl_xor_sync(0xffffffffu, l_hi, off);
  }
  const float i_lo = 1.0f / l_lo, i_hi = 1.0f / l_hi;
#pragma unroll
  for (int nb = 0; nb < 4; ++nb) {
    if (r_lo < Tq)
      *reinterpret_cast<uint32_t*>(out + ((size_t)n * Tq + r_lo) * ldo + h * 32 + nb * 8 + 2 * t4) =
          pack_bf16x2(o[nb][0] * i_lo, o[nb][1] * i_lo);
    if (r_hi < Tq)
      *reinterpret_cast<uint32_t*>(out + ((size_t)n * Tq + r_hi) * ldo + h * 32 + nb * 8 + 2 * t4) =
          pack_bf16x2(o[nb][2] * i_hi, o[nb][3] * i_hi);
  }
}

int mha_launch(const void* q, int ldq, int qoff, const void* k, int ldk, int koff, const void* v, int ldv,
               int voff, void* out, int ldo, int N, int Tq, int Tk, int heads, float scale,
               cudaStream_t st) {
  // self-attention over <= 640 tokens runs on the tcgen05 kernel (fa_umma.cu; head_dim 32 zero-padded to 64 there);
  // TCD_MHA_UMMA=0 and cross-attention shapes keep the warp-level kernel below
  if (Tq == Tk && fa64_umma_eligible(Tk))
    return fa64_umma_launch(q, ldq, qoff, k, ldk, koff, v, ldv, voff, out, ldo, N, Tk, heads, 32, scale, nullptr, 0, nullptr, st);
  const int Tkp = (Tk + 15) & ~15;
  const size_t smem = ((size_t)Tkp * kMhaKStride + (size_t)32 * (Tkp + 8)) * 2;
  TCD_REQUIRE(smem <= 200 * 1024, "mha: %d keys do not fit shared memory", Tk);
  TCD_REQUIRE(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && ldo % 8 == 0 && qoff % 8 == 0 && koff % 8 == 0 &&
                  voff % 8 == 0, "mha: leading dims / offsets must be multiples of 8");
  if (int rc = ensure_dyn_smem((const void*)mha_kernel, smem)) return rc;
  mha_kernel<<<dim3(cdiv(Tq, 64), heads, N), 128, smem, st>>>(
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
