// Non-GEMM kernels of the Stage-2 classifier body: the EVA02 vision transformer that
// `timm.create_model("eva02_large_patch14_clip_336...")` builds in the reference (src/species_classifier.py:258-262,
// config/config.yaml:279-281).  The linears run on the tcgen05 conv engine (conv_umma.cuh) as 1x1 convs over token
// rows; this file holds what sits between them: the crop -> patch-row gather (the Stage-2 crop-and-resize kernel
// writing straight into the patch-embedding GEMM's A operand), token assembly (class token + position table),
// rotary multi-head attention (head_dim 64, warp-level tensor-core path), SwiGLU + LayerNorm, softmax and the
// top-k read-out.  Activations bf16 token-major, arithmetic fp32.  Declarations: tfm_kernels.cuh.
#include "tfm_kernels.cuh"

#include <atomic>

namespace tcd {
extern std::atomic<uint64_t> g_launch_count;

#define VIT_LAUNCH_OK(name)                                                  \
  do {                                                                       \
    g_launch_count.fetch_add(1, std::memory_order_relaxed);                  \
    cudaError_t _e = cudaGetLastError();                                     \
    if (_e != cudaSuccess) {                                                 \
      set_last_error(name " launch: %s", cudaGetErrorString(_e));            \
      return TCD_ERR_CUDA;                                                   \
    }                                                                        \
    return TCD_OK;                                                           \
  } while (0)

__device__ __forceinline__ void v_bf16x8_to_f32(const uint4 q, float (&f)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float2 t = __bfloat1622float2(h[e]);
    f[2 * e] = t.x;
    f[2 * e + 1] = t.y;
  }
}
__device__ __forceinline__ uint4 v_f32_to_bf16x8(const float (&f)[8]) {
  uint4 q;
  q.x = pack_bf16x2(f[0], f[1]);
  q.y = pack_bf16x2(f[2], f[3]);
  q.z = pack_bf16x2(f[4], f[5]);
  q.w = pack_bf16x2(f[6], f[7]);
  return q;
}

// ------------------------------------------------------------------ patch rows from a normalised f32 NCHW batch
// rows[(n*G + py)*G + px][c*ps*ps + ky*ps + kx] = x[n][c][py*ps + ky][px*ps + kx]  (the flattening of the
// patch-embedding Conv2d weight [D,3,ps,ps]); columns >= 3*ps*ps stay zero.  One thread per 8 output columns.
__global__ void __launch_bounds__(256)
patchify_f32_kernel(const float* __restrict__ x, uint4* __restrict__ rows, int N, int S, int ps, int G, int Kp8) {
  const size_t total = (size_t)N * G * G * Kp8;
  const int K = 3 * ps * ps;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
    const int c8 = (int)(i % Kp8);
    const size_t row = i / Kp8;
    const int px = (int)(row % G), py = (int)((row / G) % G), n = (int)(row / ((size_t)G * G));
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int k = c8 * 8 + e;
      float v = 0.f;
      if (k < K) {
        const int c = k / (ps * ps), r = k - c * ps * ps, ky = r / ps, kx = r - ky * ps;
        v = x[(((size_t)n * 3 + c) * S + (py * ps + ky)) * S + px * ps + kx];
      }
      f[e] = v;
    }
    rows[i] = v_f32_to_bf16x8(f);
  }
}

int patchify_f32_launch(const float* x, void* rows, int N, int S, int ps, int Kp, cudaStream_t st) {
  TCD_REQUIRE(ps >= 1 && S % ps == 0 && Kp % 8 == 0 && Kp >= 3 * ps * ps, "patchify: S %d ps %d Kp %d", S, ps, Kp);
  const int G = S / ps;
  const size_t total = (size_t)N * G * G * (Kp / 8);
  const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  patchify_f32_kernel<<<blocks, 256, 0, st>>>(x, (uint4*)rows, N, S, ps, G, Kp / 8);
  VIT_LAUNCH_OK("patchify_f32");
}

// ------------------------------------------------------------------ crops of a frame -> patch rows
// The Stage-2 gather of crop_resize.cu (slice + BGR->RGB + bilinear resize with the ATen tap rule + /255 +
// mean/std, src/two_stage_pipeline_yolox.py:289 + src/species_classifier.py:311-350) with the patch-row output
// layout above, so the classifier's first GEMM reads the crops without an NCHW tensor in between.
struct CropPatchArgs {
  int S, ps, G, Kp, swap_rb;
  long long stride;
  float scale, mean[3], inv_std[3];
};
__device__ __forceinline__ void vtap(float scale, int d, int in, int* i0, int* i1, float* l1) {
  float s = scale * ((float)d + 0.5f) - 0.5f;
  s = s < 0.f ? 0.f : s;
  int a = (int)s;
  if (a > in - 1) a = in - 1;
  *i0 = a;
  *i1 = a + (a < in - 1 ? 1 : 0);
  *l1 = s - (float)a;
}
__global__ void __launch_bounds__(256)
crop_patch_kernel(const uint8_t* __restrict__ frame, const int* __restrict__ boxes, const CropPatchArgs a,
                  __nv_bfloat16* __restrict__ rows) {
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
      vtap((float)cw / (float)a.S, ox, cw, &x0, &x1, &lx);
      vtap((float)ch / (float)a.S, oy, ch, &y0, &y1, &ly);
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
  const int py = oy / a.ps, ky = oy - py * a.ps, px = ox / a.ps, kx = ox - px * a.ps;
  __nv_bfloat16* r = rows + ((size_t)crop * a.G * a.G + (size_t)py * a.G + px) * a.Kp + ky * a.ps + kx;
#pragma unroll
  for (int c = 0; c < 3; ++c) r[c * a.ps * a.ps] = __float2bfloat16((v[c] * a.scale - a.mean[c]) * a.inv_std[c]);
}

int crop_patch_launch(const uint8_t* frame, long long stride, const int* boxes_dev, int n, int S, int ps, int Kp,
                      const tcd_preproc_cfg* cfg, void* rows, cudaStream_t st) {
  if (n <= 0) return TCD_OK;
  TCD_REQUIRE(ps >= 1 && S % ps == 0 && Kp >= 3 * ps * ps, "crop_patches: S %d ps %d Kp %d", S, ps, Kp);
  CropPatchArgs a;
  a.S = S; a.ps = ps; a.G = S / ps; a.Kp = Kp; a.swap_rb = cfg->swap_rb; a.stride = stride; a.scale = cfg->scale;
  for (int c = 0; c < 3; ++c) { a.mean[c] = cfg->mean[c]; a.inv_std[c] = cfg->inv_std[c]; }
  crop_patch_kernel<<<dim3(cdiv(S * S, 256), n), 256, 0, st>>>(frame, boxes_dev, a, (__nv_bfloat16*)rows);
  VIT_LAUNCH_OK("crop_patch");
}

// ------------------------------------------------------------------ token assembly
// x[n,0,:] = cls_pos0 (= cls_token + pos_embed[0], folded on the host); x[n,1+t,:] = emb[n,t,:] + pos[1+t,:]
__global__ void __launch_bounds__(256)
vit_assemble_kernel(const uint4* __restrict__ emb, const float* __restrict__ pos, uint4* __restrict__ x, int N, int T,
                    int D8) {
  const size_t total = (size_t)N * T * D8;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
    const int c8 = (int)(i % D8);
    const int t = (int)((i / D8) % T);
    const int n = (int)(i / ((size_t)D8 * T));
    float f[8];
    const float4 p0 = *reinterpret_cast<const float4*>(pos + ((size_t)t * D8 + c8) * 8);
    const float4 p1 = *reinterpret_cast<const float4*>(pos + ((size_t)t * D8 + c8) * 8 + 4);
    if (t == 0) {
#pragma unroll
      for (int e = 0; e < 8; ++e) f[e] = 0.f;
    } else {
      v_bf16x8_to_f32(emb[((size_t)n * (T - 1) + (t - 1)) * D8 + c8], f);
    }
    f[0] += p0.x; f[1] += p0.y; f[2] += p0.z; f[3] += p0.w;
    f[4] += p1.x; f[5] += p1.y; f[6] += p1.z; f[7] += p1.w;
    x[i] = v_f32_to_bf16x8(f);
  }
}
int vit_assemble_launch(const void* emb, const float* pos, void* x, int N, int T, int D, cudaStream_t st) {
  TCD_REQUIRE(D % 8 == 0 && T >= 2, "vit_assemble: T %d D %d", T, D);
  const size_t total = (size_t)N * T * (D / 8);
  const int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  vit_assemble_kernel<<<blocks, 256, 0, st>>>((const uint4*)emb, pos, (uint4*)x, N, T, D / 8);
  VIT_LAUNCH_OK("vit_assemble");
}

// ------------------------------------------------------------------ rotary multi-head attention, head_dim 64
// Same warp-level formulation as mha_kernel (tfm_kernels.cu): mma.sync m16n8k16 bf16 / fp32, online softmax on the
// accumulator fragments, P re-used from the S fragments as the A operand of P V.  Differences: head_dim 64 (four
// k-steps for S, eight 8-channel blocks for O); ONE CTA of eight warps owns a whole (image, head) -- K (row-major,
// rotated) and V (transposed) of its 577 tokens are staged ONCE (162 KB of shared memory) and the warps walk the 37
// query tiles -- and the 2-D rotary embedding of EVA02 is applied while staging K and while loading the Q fragments:
// channel pairs (2j, 2j+1) of every token after the `prefix` class tokens turn by the angle rope[token - prefix][j]
// (cos, sin; j < 16: row position, j >= 16: column position).  gridDim.x > 1 splits the query tiles of one (image,
// head) over several CTAs when the batch alone does not fill the chip.
__device__ __forceinline__ void mma_bf16_16816_v(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t rope_pair(uint32_t packed, const float2 cs) {
  const float2 x = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&packed));
  return pack_bf16x2(x.x * cs.x - x.y * cs.y, x.y * cs.x + x.x * cs.y);
}

constexpr int kMha64KStride = 72;  // bf16 per staged K row (64 + 8 pad): conflict-free 32-bit fragment loads

__global__ void __launch_bounds__(256)
mha64_rope_kernel(const __nv_bfloat16* __restrict__ q, int ldq, int qoff, const __nv_bfloat16* __restrict__ k, int ldk,
                  int koff, const __nv_bfloat16* __restrict__ v, int ldv, int voff, __nv_bfloat16* __restrict__ out,
                  int ldo, int T, float scale, const float2* __restrict__ rope, int prefix) {
  extern __shared__ __align__(16) uint8_t mha64_sm_raw[];
  const int Tkp = (T + 15) & ~15;
  const int vstride = Tkp + 8;
  __nv_bfloat16* Ks = reinterpret_cast<__nv_bfloat16*>(mha64_sm_raw);   // [Tkp][72]
  __nv_bfloat16* Vt = Ks + (size_t)Tkp * kMha64KStride;                 // [64][Tkp + 8]
  const int n = blockIdx.z, h = blockIdx.y;
  const __nv_bfloat16* kb = k + (size_t)n * T * ldk + koff + h * 64;
  const __nv_bfloat16* vb = v + (size_t)n * T * ldv + voff + h * 64;
  for (int i = threadIdx.x; i < 8 * Tkp; i += 256) {
    const int t = i >> 3, c = i & 7;
    uint4 kq = make_uint4(0u, 0u, 0u, 0u), vq = kq;
    if (t < T) {
      kq = *reinterpret_cast<const uint4*>(kb + (size_t)t * ldk + c * 8);
      vq = *reinterpret_cast<const uint4*>(vb + (size_t)t * ldv + c * 8);
      if (rope != nullptr && t >= prefix) {
        const float2* cs = rope + (size_t)(t - prefix) * 32 + c * 4;
        kq.x = rope_pair(kq.x, cs[0]);
        kq.y = rope_pair(kq.y, cs[1]);
        kq.z = rope_pair(kq.z, cs[2]);
        kq.w = rope_pair(kq.w, cs[3]);
      }
    }
    *reinterpret_cast<uint4*>(Ks + (size_t)t * kMha64KStride + c * 8) = kq;
    const __nv_bfloat16* ve = reinterpret_cast<const __nv_bfloat16*>(&vq);
#pragma unroll
    for (int e = 0; e < 8; ++e) Vt[(size_t)(c * 8 + e) * vstride + t] = ve[e];
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int n_qt = (T + 15) >> 4;
  for (int qt = blockIdx.x * 8 + warp; qt < n_qt; qt += 8 * gridDim.x) {
    const int q0 = qt * 16;
    const int r_lo = q0 + g, r_hi = q0 + g + 8;
    const bool lo_ok = r_lo < T, hi_ok = r_hi < T;
    uint32_t qa[4][4];
    {
      const __nv_bfloat16* qlo = q + ((size_t)n * T + (lo_ok ? r_lo : 0)) * ldq + qoff + h * 64;
      const __nv_bfloat16* qhi = q + ((size_t)n * T + (hi_ok ? r_hi : 0)) * ldq + qoff + h * 64;
      const bool rl = rope != nullptr && lo_ok && r_lo >= prefix, rh = rope != nullptr && hi_ok && r_hi >= prefix;
      const float2* cl = rope + (size_t)(rl ? r_lo - prefix : 0) * 32;
      const float2* chh = rope + (size_t)(rh ? r_hi - prefix : 0) * 32;
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        const int pj = ks * 8 + t4;
        uint32_t a0 = lo_ok ? *reinterpret_cast<const uint32_t*>(qlo + ks * 16 + 2 * t4) : 0u;
        uint32_t a1 = hi_ok ? *reinterpret_cast<const uint32_t*>(qhi + ks * 16 + 2 * t4) : 0u;
        uint32_t a2 = lo_ok ? *reinterpret_cast<const uint32_t*>(qlo + ks * 16 + 2 * t4 + 8) : 0u;
        uint32_t a3 = hi_ok ? *reinterpret_cast<const uint32_t*>(qhi + ks * 16 + 2 * t4 + 8) : 0u;
        if (rl) {
          a0 = rope_pair(a0, cl[pj]);
          a2 = rope_pair(a2, cl[pj + 4]);
        }
        if (rh) {
          a1 = rope_pair(a1, chh[pj]);
          a3 = rope_pair(a3, chh[pj + 4]);
        }
        qa[ks][0] = a0; qa[ks][1] = a1; qa[ks][2] = a2; qa[ks][3] = a3;
      }
    }
    float o[8][4];
#pragma unroll
    for (int nb = 0; nb < 8; ++nb)
#pragma unroll
      for (int e = 0; e < 4; ++e) o[nb][e] = 0.f;
    float m_lo = -INFINITY, m_hi = -INFINITY, l_lo = 0.f, l_hi = 0.f;
    for (int key0 = 0; key0 < Tkp; key0 += 16) {
      float sc[2][4];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
#pragma unroll
        for (int e = 0; e < 4; ++e) sc[j][e] = 0.f;
        const __nv_bfloat16* kr = Ks + (size_t)(key0 + j * 8 + g) * kMha64KStride + 2 * t4;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)
          mma_bf16_16816_v(sc[j], qa[ks], *reinterpret_cast<const uint32_t*>(kr + ks * 16),
                           *reinterpret_cast<const uint32_t*>(kr + ks * 16 + 8));
      }
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int key = key0 + j * 8 + 2 * t4 + (e & 1);
          sc[j][e] = key < T ? sc[j][e] * scale : -INFINITY;
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
      for (int nb = 0; nb < 8; ++nb) {
        o[nb][0] *= c_lo; o[nb][1] *= c_lo;
        o[nb][2] *= c_hi; o[nb][3] *= c_hi;
      }
      uint32_t pa[4];
      pa[0] = pack_bf16x2(p[0][0], p[0][1]);
      pa[1] = pack_bf16x2(p[0][2], p[0][3]);
      pa[2] = pack_bf16x2(p[1][0], p[1][1]);
      pa[3] = pack_bf16x2(p[1][2], p[1][3]);
#pragma unroll
      for (int nb = 0; nb < 8; ++nb) {
        const __nv_bfloat16* vr = Vt + (size_t)(nb * 8 + g) * vstride + key0 + 2 * t4;
        mma_bf16_16816_v(o[nb], pa, *reinterpret_cast<const uint32_t*>(vr), *reinterpret_cast<const uint32_t*>(vr + 8));
      }
    }
#pragma unroll
    for (int off = 1; off <= 2; off <<= 1) {
      l_lo += __shfl_xor_sync(0xffffffffu, l_lo, off);
      l_hi += __shfl_xor_sync(0xffffffffu, l_hi, off);
    }
    const float i_lo = 1.0f / l_lo, i_hi = 1.0f / l_hi;
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) {
      if (lo_ok)
        *reinterpret_cast<uint32_t*>(out + ((size_t)n * T + r_lo) * ldo + h * 64 + nb * 8 + 2 * t4) =
            pack_bf16x2(o[nb][0] * i_lo, o[nb][1] * i_lo);
      if (hi_ok)
        *reinterpret_cast<uint32_t*>(out + ((size_t)n * T + r_hi) * ldo + h * 64 + nb * 8 + 2 * t4) =
            pack_bf16x2(o[nb][2] * i_hi, o[nb][3] * i_hi);
    }
  }
}

int mha64_rope_launch(const void* q, int ldq, int qoff, const void* k, int ldk, int koff, const void* v, int ldv,
                      int voff, void* out, int ldo, int N, int T, int heads, float scale, const float* rope,
                      int prefix, float* stats_out, cudaStream_t st) {
  if (fa64_umma_eligible(T))
    return fa64_umma_launch(q, ldq, qoff, k, ldk, koff, v, ldv, voff, out, ldo, N, T, heads, 64, scale, rope, prefix, stats_out, st);
  TCD_REQUIRE(stats_out == nullptr, "mha64: row statistics are written by the tcgen05 kernel only (T %d)", T);
  const int Tkp = (T + 15) & ~15;
  const size_t smem = ((size_t)Tkp * kMha64KStride + (size_t)64 * (Tkp + 8)) * 2;
  TCD_REQUIRE(smem <= 220 * 1024, "mha64: %d tokens do not fit shared memory", T);
  TCD_REQUIRE(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && ldo % 8 == 0 && qoff % 8 == 0 && koff % 8 == 0 &&
                  voff % 8 == 0, "mha64: leading dims / offsets must be multiples of 8");
  TCD_REQUIRE(prefix >= 0 && prefix < T, "mha64: prefix %d of %d tokens", prefix, T);
  if (int rc = ensure_dyn_smem((const void*)mha64_rope_kernel, smem)) return rc;
  // split the query tiles of an (image, head) over CTAs only while (image, head) pairs alone leave SMs idle
  const int n_qt = (T + 15) / 16;
  int gx = 1;
  while (gx * 2 * heads * N <= 148 && gx * 8 < n_qt) gx *= 2;
  mha64_rope_kernel<<<dim3(gx, heads, N), 256, smem, st>>>(
      (const __nv_bfloat16*)q, ldq, qoff, (const __nv_bfloat16*)k, ldk, koff, (const __nv_bfloat16*)v, ldv, voff,
      (__nv_bfloat16*)out, ldo, T, scale, (const float2*)rope, prefix);
  VIT_LAUNCH_OK("mha64_rope");
}

// ------------------------------------------------------------------ SwiGLU + LayerNorm
// y[r, 0:H) = LayerNorm_H(silu(g[r,:]) * u[r,:]) * gamma + beta, columns [H, Hp) written as zeros (gamma / beta hold Hp floats).  g and u are the
// two halves of one fused fc1 output row: gu[r] = [g (Hp) | u (Hp)].  Two rows per CTA, 128 threads per row, up to
// three 8-channel chunks per thread held in registers between the statistics and the normalisation.
constexpr int kSwigluChunks = 3;   // Hp <= 3 * 128 * 8 = 3072
__global__ void __launch_bounds__(256)
swiglu_ln_kernel(const uint4* __restrict__ gu, const float* __restrict__ gamma, const float* __restrict__ beta,
                 uint4* __restrict__ y, int rows, int H, int Hp8, float eps, int gated) {
  __shared__ float red[2][2][4];
  const int half = threadIdx.x >> 7, t = threadIdx.x & 127, wq = t >> 5;
  const int row = blockIdx.x * 2 + half;
  const bool live = row < rows;
  float hv[kSwigluChunks][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < kSwigluChunks; ++i) {
    const int c = t + 128 * i;
#pragma unroll
    for (int e = 0; e < 8; ++e) hv[i][e] = 0.f;
    if (live && c < Hp8) {
      if (gated) {           // the fc1 epilogue already produced silu(g) * u (gemm_pair.cuh, ACT == 4): gu is [rows, Hp]
        float hf[8];
        v_bf16x8_to_f32(gu[(size_t)row * Hp8 + c], hf);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          hv[i][e] = (c * 8 + e < H) ? hf[e] : 0.f;
          sum += hv[i][e];
        }
      } else {
        float gf[8], uf[8];
        v_bf16x8_to_f32(gu[(size_t)row * 2 * Hp8 + c], gf);
        v_bf16x8_to_f32(gu[(size_t)row * 2 * Hp8 + Hp8 + c], uf);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float s = silu_f(gf[e]);     // h + h tanh(h), h = g / 2: one MUFU op, no division
          hv[i][e] = (c * 8 + e < H) ? s * uf[e] : 0.f;
          sum += hv[i][e];
        }
      }
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
  if ((t & 31) == 0) red[0][half][wq] = sum;
  __syncthreads();
  const float mean = (red[0][half][0] + red[0][half][1] + red[0][half][2] + red[0][half][3]) / (float)H;
  float var = 0.f;
#pragma unroll
  for (int i = 0; i < kSwigluChunks; ++i) {
    const int c = t + 128 * i;
    if (c < Hp8) {
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float d = (c * 8 + e < H) ? hv[i][e] - mean : 0.f;
        var += d * d;
      }
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) var += __shfl_xor_sync(0xffffffffu, var, off);
  if ((t & 31) == 0) red[1][half][wq] = var;
  __syncthreads();
  const float rstd = rsqrtf((red[1][half][0] + red[1][half][1] + red[1][half][2] + red[1][half][3]) / (float)H + eps);
  if (!live) return;
#pragma unroll
  for (int i = 0; i < kSwigluChunks; ++i) {
    const int c = t + 128 * i;
    if (c < Hp8) {
      float o[8];
      if (gamma != nullptr) {
        // gamma / beta are padded to Hp by the caller: two 16-byte loads each instead of eight scalar ones
        const float4 g0 = *reinterpret_cast<const float4*>(gamma + c * 8), g1 = *reinterpret_cast<const float4*>(gamma + c * 8 + 4);
        const float4 b0 = *reinterpret_cast<const float4*>(beta + c * 8), b1 = *reinterpret_cast<const float4*>(beta + c * 8 + 4);
        const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = (c * 8 + e < H) ? (hv[i][e] - mean) * rstd * gg[e] + bb[e] : 0.f;
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = (c * 8 + e < H) ? (hv[i][e] - mean) * rstd : 0.f;
      }
      y[(size_t)row * Hp8 + c] = v_f32_to_bf16x8(o);
    }
  }
}
// LayerNorm over wide rows (the gated fc1 output, 2 816 columns of which 2 730 are real): one WARP per row, up to
// twelve 8-channel chunks per lane in registers, no block barrier (the two-rows-per-CTA kernel above spent its time
// in its two __syncthreads: 2.6 TB/s whether it read 406 or 203 MB).
constexpr int kWideChunks = 12;    // Hp <= 12 * 32 * 8 = 3072
__global__ void __launch_bounds__(256, 2)
ln_wide_kernel(const uint4* __restrict__ h, const float* __restrict__ gamma, const float* __restrict__ beta,
               uint4* __restrict__ y, int rows, int H, int Hp8, float eps) {
  const int row = (blockIdx.x * 256 + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  // chunks [0, full) are entirely real columns, chunk `full` may be partial (H % 8 real ones), the rest is padding:
  // the padding and the partial chunk's tail are zeroed ONCE here, so the passes below carry no per-element masks
  const int full = H >> 3, tail = H & 7;
  uint4 raw[kWideChunks];
#pragma unroll
  for (int i = 0; i < kWideChunks; ++i) {
    const int c = lane + 32 * i;
    raw[i] = c < full + (tail ? 1 : 0) ? h[(size_t)row * Hp8 + c] : make_uint4(0u, 0u, 0u, 0u);
    if (c == full && tail) {
      uint32_t w[4] = {raw[i].x, raw[i].y, raw[i].z, raw[i].w};
#pragma unroll
      for (int e = 0; e < 8; ++e)
        if (e >= tail) w[e >> 1] &= (e & 1) ? 0x0000ffffu : 0xffff0000u;
      raw[i] = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < kWideChunks; ++i) {
    float f[8];
    v_bf16x8_to_f32(raw[i], f);
#pragma unroll
    for (int e = 0; e < 8; ++e) sum += f[e];
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
  const float mean = sum / (float)H;
  // variance about the mean (two-pass): the zeroed columns contribute (0 - mean)^2 each, taken out exactly below
  float var = 0.f;
#pragma unroll
  for (int i = 0; i < kWideChunks; ++i) {
    float f[8];
    v_bf16x8_to_f32(raw[i], f);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float d = f[e] - mean;
      var = fmaf(d, d, var);
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) var += __shfl_xor_sync(0xffffffffu, var, off);
  var -= (float)(kWideChunks * 32 * 8 - H) * mean * mean;
  const float rstd = rsqrtf(fmaxf(var, 0.f) / (float)H + eps);
#pragma unroll
  for (int i = 0; i < kWideChunks; ++i) {
    const int c = lane + 32 * i;
    if (c < Hp8) {
      float f[8], o[8];
      v_bf16x8_to_f32(raw[i], f);
      if (gamma != nullptr) {
        const float4 g0 = *reinterpret_cast<const float4*>(gamma + c * 8), g1 = *reinterpret_cast<const float4*>(gamma + c * 8 + 4);
        const float4 b0 = *reinterpret_cast<const float4*>(beta + c * 8), b1 = *reinterpret_cast<const float4*>(beta + c * 8 + 4);
        const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
        const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = (f[e] - mean) * rstd * gg[e] + bb[e];
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = (f[e] - mean) * rstd;
      }
      if (c >= full) {          // the partial chunk's tail and the padding chunks are written as zeros
#pragma unroll
        for (int e = 0; e < 8; ++e)
          if (c > full || e >= tail) o[e] = 0.f;
      }
      y[(size_t)row * Hp8 + c] = v_f32_to_bf16x8(o);
    }
  }
}

// LayerNorm statistics only: one warp per row reads the row once and writes (mean, 1 / std) -- 8 bytes -- for a
// consumer GEMM that normalises in its epilogue: y = rstd * (x W'^T - mean * colsum(W')) + b'.  The normalised
// tensor (151 MB per LayerNorm, 203 MB for the MLP's hidden one at 64 crops) is never written or re-read.
template <int CH>      // 8-channel chunks per lane: 4 covers rows of <= 1024 columns (40 registers, full occupancy), 12 up to 3072
__global__ void __launch_bounds__(256, CH <= 4 ? 6 : 2)
row_stats_kernel(const uint4* __restrict__ h, float2* __restrict__ stats, int rows, int H, int Hp8, float eps, int parts) {
  const int row = (blockIdx.x * 256 + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int full = H >> 3, tail = H & 7;
  uint4 raw[CH];
#pragma unroll
  for (int i = 0; i < CH; ++i) {
    const int c = lane + 32 * i;
    raw[i] = c < full + (tail ? 1 : 0) ? h[(size_t)row * Hp8 + c] : make_uint4(0u, 0u, 0u, 0u);
    if (c == full && tail) {
      uint32_t w[4] = {raw[i].x, raw[i].y, raw[i].z, raw[i].w};
#pragma unroll
      for (int e = 0; e < 8; ++e)
        if (e >= tail) w[e >> 1] &= (e & 1) ? 0x0000ffffu : 0xffff0000u;
      raw[i] = make_uint4(w[0], w[1], w[2], w[3]);
    }
  }
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < CH; ++i) {
    float f[8];
    v_bf16x8_to_f32(raw[i], f);
#pragma unroll
    for (int e = 0; e < 8; ++e) sum += f[e];
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
  const float mean = sum / (float)H;
  float var = 0.f;
#pragma unroll
  for (int i = 0; i < CH; ++i) {
    float f[8];
    v_bf16x8_to_f32(raw[i], f);
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const float d = f[e] - mean;
      var = fmaf(d, d, var);
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) var += __shfl_xor_sync(0xffffffffu, var, off);
  var -= (float)(CH * 32 * 8 - H) * mean * mean;     // the zeroed slots contributed (0 - mean)^2 each
  if (parts > 0) {      // partial-sum format: slot 0 = (sum, sum of squares) = (sum, var_sum + H mean^2), the other slots zero
    if (lane < parts) stats[(size_t)row * parts + lane] = lane == 0 ? make_float2(sum, fmaxf(var, 0.f) + (float)H * mean * mean) : make_float2(0.f, 0.f);
  } else if (lane == 0) {
    stats[row] = make_float2(mean, rsqrtf(fmaxf(var, 0.f) / (float)H + eps));
  }
}
int row_stats_launch(const void* x, float* stats, int rows, int H, int Hp, float eps, int parts, cudaStream_t st) {
  TCD_REQUIRE(Hp % 8 == 0 && Hp <= kWideChunks * 32 * 8 && H >= 1 && H <= Hp && parts >= 0 && parts <= 32, "row_stats: H %d Hp %d parts %d", H, Hp, parts);
  if (Hp <= 4 * 32 * 8)
    row_stats_kernel<4><<<cdiv(rows, 8), 256, 0, st>>>((const uint4*)x, (float2*)stats, rows, H, Hp / 8, eps, parts);
  else
    row_stats_kernel<kWideChunks><<<cdiv(rows, 8), 256, 0, st>>>((const uint4*)x, (float2*)stats, rows, H, Hp / 8, eps, parts);
  VIT_LAUNCH_OK("row_stats");
}

int swiglu_ln_launch(const void* gu, const float* gamma, const float* beta, void* y, int rows, int H, int Hp,
                     float eps, int gated, cudaStream_t st) {
  if (gated && Hp % 8 == 0 && Hp <= kWideChunks * 32 * 8 && H >= 1 && H <= Hp) {
    ln_wide_kernel<<<cdiv(rows, 8), 256, 0, st>>>((const uint4*)gu, gamma, beta, (uint4*)y, rows, H, Hp / 8, eps);
    VIT_LAUNCH_OK("ln_wide");
  }
  TCD_REQUIRE(Hp % 8 == 0 && H <= Hp && Hp <= kSwigluChunks * 128 * 8 && H >= 1, "swiglu_ln: H %d Hp %d", H, Hp);
  swiglu_ln_kernel<<<cdiv(rows, 2), 256, 0, st>>>((const uint4*)gu, gamma, beta, (uint4*)y, rows, H, Hp / 8, eps, gated);
  VIT_LAUNCH_OK("swiglu_ln");
}

// ------------------------------------------------------------------ classifier head: few rows x many classes
// out[b, c] = sum_k x[b, k] * w[c, k] + bias[c]   (x bf16 [B, K], w bf16 [NC, K], fp32 accumulate, f32 logits).
// With B = 64 class tokens and 10 000 classes this is a weight-streaming problem (20 MB of weights, 1.3 GFLOP): the
// conv engine's f32-row mode did it as forty 256-column launches of ONE CTA each (35 us apiece, 1.4 ms).  Here a
// CTA owns 64 classes (8 per warp) and a tile of 32 rows of x in shared memory; a lane owns a slice of k, keeps the
// weight slice of the current class in registers, and the 32 per-row partial sums of a warp are combined by a
// reduce-scatter butterfly (31 shuffles; lane r ends up with row r).
constexpr int kHeadRows = 32, kHeadMaxK = 1024;
__global__ void __launch_bounds__(256)
head_gemv_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w, const float* __restrict__ bias,
                 float* __restrict__ out, int B, int K, int NC, int ld) {
  extern __shared__ __align__(16) uint8_t head_sm[];
  uint4* xs = reinterpret_cast<uint4*>(head_sm);            // [32 rows][K / 8] uint4
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int K8 = K >> 3;
  for (int b0 = 0; b0 < B; b0 += kHeadRows) {
    __syncthreads();
    for (int i = threadIdx.x; i < kHeadRows * K8; i += 256) {
      const int rr = i / K8, c8 = i - rr * K8;
      xs[i] = (b0 + rr < B) ? reinterpret_cast<const uint4*>(x + (size_t)(b0 + rr) * K)[c8] : make_uint4(0u, 0u, 0u, 0u);
    }
    __syncthreads();
    for (int ci = 0; ci < 8; ++ci) {
      const int c = blockIdx.x * 64 + warp * 8 + ci;
      if (c >= NC) break;                                    // warp-uniform
      float wf[4][8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int k8 = lane + 32 * i;
        uint4 q = make_uint4(0u, 0u, 0u, 0u);
        if (k8 < K8) q = reinterpret_cast<const uint4*>(w + (size_t)c * K)[k8];
        v_bf16x8_to_f32(q, wf[i]);
      }
      float part[kHeadRows];
#pragma unroll
      for (int rr = 0; rr < kHeadRows; ++rr) {
        float acc = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int k8 = lane + 32 * i;
          if (k8 < K8) {
            float xf[8];
            v_bf16x8_to_f32(xs[rr * K8 + k8], xf);
#pragma unroll
            for (int e = 0; e < 8; ++e) acc = fmaf(xf[e], wf[i][e], acc);
          }
        }
        part[rr] = acc;
      }
      // reduce-scatter over the warp: after the step with offset `off`, a lane keeps the half of its rows that
      // matches its own bit `off`; lane r finishes with the full sum of row r in part[0]
#pragma unroll
      for (int off = 16; off >= 1; off >>= 1) {
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < off; ++i) {
          const float send = up ? part[i] : part[i + off];
          const float keep = up ? part[i + off] : part[i];
          part[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
      }
      if (b0 + lane < B) out[(size_t)(b0 + lane) * ld + c] = part[0] + bias[c];
    }
  }
}
int head_gemv_launch(const void* x, const void* w, const float* bias, float* out, int B, int K, int NC, int ld,
                     cudaStream_t st) {
  TCD_REQUIRE(K % 8 == 0 && K >= 8 && K <= kHeadMaxK && B >= 1 && NC >= 1 && ld >= NC, "head_gemv: B %d K %d NC %d", B, K, NC);
  const size_t smem = (size_t)kHeadRows * K * 2;
  if (int rc = ensure_dyn_smem((const void*)head_gemv_kernel, smem)) return rc;
  head_gemv_kernel<<<cdiv(NC, 64), 256, smem, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)w, bias, out, B, K, NC, ld);
  VIT_LAUNCH_OK("head_gemv");
}

// ------------------------------------------------------------------ softmax over f32 rows, top-k read-out
__global__ void __launch_bounds__(256)
softmax_rows_kernel(const float* __restrict__ x, int ldx, float* __restrict__ y, int ldy, int C) {
  __shared__ float red[8];
  __shared__ float bc;
  const float* xr = x + (size_t)blockIdx.x * ldx;
  float* yr = y + (size_t)blockIdx.x * ldy;
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  float mx = -INFINITY;
  for (int c = threadIdx.x; c < C; c += 256) mx = fmaxf(mx, xr[c]);
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
  if (l == 0) red[w] = mx;
  __syncthreads();
  if (threadIdx.x == 0) {
    float m = red[0];
    for (int i = 1; i < 8; ++i) m = fmaxf(m, red[i]);
    bc = m;
  }
  __syncthreads();
  mx = bc;
  float sum = 0.f;
  for (int c = threadIdx.x; c < C; c += 256) sum += expf(xr[c] - mx);
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
  __syncthreads();
  if (l == 0) red[w] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < 8; ++i) s += red[i];
    bc = 1.0f / s;
  }
  __syncthreads();
  const float inv = bc;
  for (int c = threadIdx.x; c < C; c += 256) yr[c] = expf(xr[c] - mx) * inv;
}
int softmax_rows_launch(const float* x, int ldx, float* y, int ldy, int rows, int C, cudaStream_t st) {
  TCD_REQUIRE(rows >= 1 && C >= 1 && ldx >= C && ldy >= C, "softmax_rows: rows %d C %d", rows, C);
  softmax_rows_kernel<<<rows, 256, 0, st>>>(x, ldx, y, ldy, C);
  VIT_LAUNCH_OK("softmax_rows");
}

__global__ void take_f32_kernel(const float* __restrict__ src, int ld, const int* __restrict__ idx, float* __restrict__ out,
                                int N, int k) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= N * k) return;
  out[i] = src[(size_t)(i / k) * ld + idx[i]];
}
int take_f32_launch(const float* src, int ld, const int* idx, float* out, int N, int k, cudaStream_t st) {
  take_f32_kernel<<<cdiv(N * k, 256), 256, 0, st>>>(src, ld, idx, out, N, k);
  VIT_LAUNCH_OK("take_f32");
}

}  // namespace tcd
