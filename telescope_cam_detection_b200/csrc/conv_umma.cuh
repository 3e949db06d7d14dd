// Implicit-GEMM convolution on tcgen05 (UMMA) with TMA-fed operands -- the K3 engine.
//
// Replaces every cudnn conv2d + BatchNorm(eval) + SiLU of upstream YOLOX-S that the
// reference runs through self.model(batch) (src/yolox_detector.py:306,378).
//
// Formulation (B200-first, not a cuDNN translation):
//   * activations are bf16 NHWC; one CTA owns a TH x TW patch of output pixels (<=128 = UMMA M)
//     and BN output channels (UMMA N, accumulators in TMEM, fp32);
//   * for every filter tap (r,s) and every chunk of ck input channels the A operand is ONE
//     4-D TMA box load (ck, TW, TH, 1) at the shifted coordinate (x0+s-1, y0+r-1): the
//     hardware zero-fills out-of-bounds pixels, which IS the conv padding.  The box lands
//     in shared memory already in the canonical K-major swizzled UMMA layout (one pixel per
//     row), so there is no im2col buffer and no gather code at all;
//   * stride-2 convs use four parity views of the input (base pointer offset by (hp,wp),
//     pixel strides doubled), so each tap is again a plain dense box;
//   * weights are pre-packed [tap][Cout][Cin] bf16 and loaded with a 3-D TMA box;
//   * warp 0 = TMA producer, warp 1 = TMEM allocator + single-thread MMA issuer,
//     warps 2-5 = epilogue (tcgen05.ld -> +bias -> SiLU -> (+residual) -> bf16 ->
//     swizzled smem -> TMA store, which clips partial tiles for free and writes straight
//     into a channel slice of a concat buffer);
//   * CTAs are PERSISTENT: each loops over output tiles (static round-robin), the smem ring
//     never drains across tile boundaries, and the accumulator is double-buffered in TMEM so
//     the epilogue of tile i overlaps the loads + MMAs of tile i+1;
//   * launched with programmatic dependent launch inside the CUDA graph: barrier init, TMEM
//     allocation, descriptor prefetch and the first weight loads overlap the previous
//     layer's tail; griddepcontrol.wait guards the first read of the previous layer's output.
#pragma once
#include "tcd_common.cuh"

namespace tcd {

constexpr int kConvMaxStages = 8;
constexpr int kConvThreads = 192;   // warp 0 TMA producer, warp 1 MMA issuer, warps 2..5 epilogue
constexpr int kConvEpiWarps = 4;
constexpr int kConvMaxAStages = 8;
constexpr int kConvMaxBias = 2048;  // every layer's bias is staged in shared memory (planner refuses wider layers)

struct ConvMaps {
  CUtensorMap a[4];
  CUtensorMap b;
  CUtensorMap o;
};

struct ConvArgs {
  int Ho, Wo, TW, TH, tiles_w, tiles_per_img;
  int taps, kchunks, ck, BN;
  int cin_off, cout_off;
  int mode;  // 0: bf16 NHWC via TMA store; 1: f32 rows (head predictions)
  int silu, out_chunk, stages, tmem_cols, n_valid;
  // ceil(2^32 / d) for d = n_tiles_n, tiles_per_img, tiles_w: t / d == __umulhi(t, magic) while t * d < 2^32
  // (checked by the planner) -- an integer divide is ~40 instructions through the MUFU pipe the SiLUs keep busy
  unsigned int magic_n, magic_img, magic_w;
  int n_tiles_n, total_tiles;   // persistent tile loop: tile t -> (m = t / n_tiles_n, n = t % n_tiles_n)
  int acc_stride;               // TMEM columns between the two accumulator buffers
  int staging_off;              // byte offset of the epilogue staging area inside dynamic smem
  int bias_off;                 // byte offset of the staged bias (bias_n floats) inside dynamic smem
  int pdl;                      // launched with programmatic stream serialization
  // halo mode (3x3 stride 1): ONE (TH+2)x(TW+2) input box per k-chunk serves all nine taps through
  // row-shifted UMMA descriptors (the swizzle is a function of the absolute shared-memory address for
  // TMA and UMMA alike, so a view may start at any 128-byte row with base-offset 0 -- measured: the
  // parity tests pass with 0 and fail with (start >> 7) & 7); the operand ring then carries weights only
  int halo, a_stages, a_stage_bytes, a_ring_off, halo_w;
  int dbg;                      // measurement only (TCD_CONV_DBG): 1 = no operand loads, 2 = no MMAs
  int tps;                      // filter taps per weight stage (halo mode: 1, 3 or 9; else 1)
  int bias_n;                   // entries of bias staged in shared memory (0: read from global)
  const float* bias;
  const __nv_bfloat16* resid;
  int r_ctot, r_coff;
  int b_resident;               // halo form: the weight stages of a tile fit the ring and every tile of the CTA uses
                                // the same ones (n_tiles_n == 1): loaded for the first tile only
  int group;                    // output tiles per accumulator round (1, 2 or 4; > 1 only with n_tiles_n == 1)
  int resid_pre;                // 1: residual added BEFORE the activation (ResNet), 0: after (CSP bottleneck)
  long long r_nstride;          // elements between images of the residual tensor
  float* out_f32;
  int out_ld, out_col, rows_per_img, row_off;
  int8_t tap_map[9], tap_dx[9], tap_dy[9];
  unsigned long long* trace;    // optional in-kernel timeline (tcd_detector_trace_op); nullptr on the product path
  int group_cin;                // grouped conv: n-tile t reads input channels [cin_off + t * group_cin, ...); else 0
  // f32-row mode, YOLOX prediction convs: only rows whose objectness can still reach the confidence threshold are
  // written, and their anchor index is appended to a per-image list the decode kernel works from
  // (score = obj * cls <= obj: a row with sigmoid(obj) < conf can never be a candidate).  nullptr: every row.
  const float* emit_thr;        // device float: confidence threshold (-inf: every row, the parity taps)
  int* maybe_count;             // [images]
  int* maybe_idx;               // [images][maybe_cap] anchor indices (row_off + y * Wo + x)
  int maybe_cap, emit_col;      // emit_col: accumulator column of the objectness logit
  // gemm_pair.cuh only: LayerNorm applied in the epilogue -- (mean, 1/std) per input row and the column sums of the weights
  const float* ln_stats;        // [rows][2] (mean, rstd), or [rows][ln_parts][2] partial (sum, sumsq); nullptr: plain linear
  const float* ln_colsum;       // [cout]
  int ln_parts, ln_count;       // ln_parts > 0: partial sums over ln_count columns
  float ln_eps;
  float* stats_out;             // [rows][stats_parts][2]: (sum, sumsq) of this launch's output tiles per row, or nullptr
  int stats_parts, stats_slot0;
};

#if defined(__CUDACC__) && defined(TCD_CONV_KERNEL_IMPL)

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void tma_store_commit() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}

__device__ __forceinline__ unsigned long long gtimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// Timeline slots (first and last CTA of the grid, 16 slots each): 0 globaltimer at entry, 1 clock at
// entry, 2 setup done, 3 first A load issued, 4 last load issued, 5 first operands landed, 6 last MMA
// committed, 7 first accumulator ready, 8 last store issued, 9 stores drained, 10 clock at exit,
// 11 globaltimer at exit.
__device__ __forceinline__ void conv_trace(const ConvArgs& p, int k, bool wall = false) {
  if (p.trace == nullptr) return;
  int sel;
  if (blockIdx.x == 0) sel = 0;
  else if (blockIdx.x == gridDim.x - 1) sel = 1;
  else return;
  p.trace[sel * 16 + k] = wall ? gtimer_ns() : (unsigned long long)clock64();
}

// per-tile stamps of CTA 0 (first 16 tiles): 32+j producer starts tile j, 48+j MMA commits tile j,
// 64+j epilogue sees accumulator j, 96+j staging free, 112+j all slabs staged, 80+j store of tile j issued
__device__ __forceinline__ void conv_trace_tile(const ConvArgs& p, int base, int j) {
  if (p.trace == nullptr || blockIdx.x != 0 || j >= 16) return;
  p.trace[base + j] = (unsigned long long)clock64();
}

// stage index + phase bit of a shared-memory ring, advanced without a divide
struct RingCursor {
  int s;
  uint32_t ph;
  int n;
  __device__ __forceinline__ void advance() {
    if (++s == n) {
      s = 0;
      ph ^= 1u;
    }
  }
};

struct ConvTile {
  int n_img, y0, x0, n_tile;
};
__device__ __forceinline__ ConvTile conv_tile(const ConvArgs& p, int t) {
  ConvTile c;
  const int m_tile = p.n_tiles_n == 1 ? t : (int)__umulhi((unsigned)t, p.magic_n);
  c.n_tile = t - m_tile * p.n_tiles_n;
  c.n_img = p.tiles_per_img == 1 ? m_tile : (int)__umulhi((unsigned)m_tile, p.magic_img);
  const int rem = m_tile - c.n_img * p.tiles_per_img;
  const int ty = p.tiles_w == 1 ? rem : (int)__umulhi((unsigned)rem, p.magic_w);
  c.y0 = ty * p.TH;
  c.x0 = (rem - ty * p.tiles_w) * p.TW;
  return c;
}

// One NC-column slab of the accumulator row owned by this thread: TMEM -> registers.
template <int NC>
__device__ __forceinline__ void conv_load_cols(uint32_t taddr, uint32_t (&r)[NC]) {
  if constexpr (NC == 32) tmem_ld32(taddr, r); else tmem_ld16(taddr, r);
}

template <int NC>
__device__ __forceinline__ void conv_add_resid(const ConvArgs& p, float (&v)[NC], int c0, const ConvTile& t,
                                               int y, int x) {
  const __nv_bfloat16* rp = p.resid + (size_t)t.n_img * p.r_nstride + ((size_t)y * p.Wo + x) * p.r_ctot +
                            p.r_coff + t.n_tile * p.BN + c0;
  const uint4* rp4 = reinterpret_cast<const uint4*>(rp);
#pragma unroll
  for (int u = 0; u < NC / 8; ++u) {
    const uint4 q = __ldg(rp4 + u);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = __bfloat1622float2(h[e]);
      v[u * 8 + e * 2] += f.x;
      v[u * 8 + e * 2 + 1] += f.y;
    }
  }
}

// SiLU + bf16 staging of one slab, the hot epilogue (every BaseConv of YOLOX): packed fp32 math.
//   x*sigmoid(x) = h + h*tanh(h), h = x/2 = 0.5*acc + 0.5*bias: s_bias holds the HALVED bias for these kernels, so a
//   slab of 32 columns costs 16 FFMA2 + 32 MUFU.TANH + 16 FFMA2 + 16 F2FP instead of 96 scalar fp32 ops + 32 MUFU
//   (measured on dark3.1.conv12, ncu warp sampling: the epilogue warps issued one instruction per 4.6 cycles and
//   spent 4 700 cycles per 128x128 tile against 512 cycles of MMA).  The residual of a CSP bottleneck is added after
//   the activation, also as FADD2.  Staging stores are explicit st.shared (the generic form compiled to ST).
template <int NC, int OFF>
__device__ __forceinline__ void conv_finish_silu(const ConvArgs& p, const uint32_t (&r)[32], int c0, uint32_t bias_saddr,
                                                 int row, bool valid, const ConvTile& t, int y, int x,
                                                 uint32_t stage_saddr) {
  const uint64_t half2 = f32x2(0.5f, 0.5f);
  uint64_t o[NC / 2];
  const uint32_t baddr = bias_saddr + (uint32_t)(t.n_tile * p.BN + c0) * 4u;
#pragma unroll
  for (int j = 0; j < NC / 4; ++j) {
    const float4 b = lds_v4f(baddr + 16u * j);
    const uint64_t h0 = fma_f32x2(f32x2_bits(r[OFF + 4 * j + 0], r[OFF + 4 * j + 1]), half2, f32x2(b.x, b.y));
    const uint64_t h1 = fma_f32x2(f32x2_bits(r[OFF + 4 * j + 2], r[OFF + 4 * j + 3]), half2, f32x2(b.z, b.w));
    float a0, a1, a2, a3;
    f32x2_split(h0, a0, a1);
    f32x2_split(h1, a2, a3);
    o[2 * j + 0] = fma_f32x2(h0, f32x2(tanh_approx(a0), tanh_approx(a1)), h0);
    o[2 * j + 1] = fma_f32x2(h1, f32x2(tanh_approx(a2), tanh_approx(a3)), h1);
  }
  if (p.resid != nullptr && valid) {   // YOLOX bottlenecks: y = act(conv(x)) + x
    const __nv_bfloat16* rp = p.resid + (size_t)t.n_img * p.r_nstride + ((size_t)y * p.Wo + x) * p.r_ctot +
                              p.r_coff + t.n_tile * p.BN + c0;
    const uint4* rp4 = reinterpret_cast<const uint4*>(rp);
#pragma unroll
    for (int u = 0; u < NC / 8; ++u) {
      const uint4 q = __ldg(rp4 + u);
      const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int e = 0; e < 4; ++e)   // bf16 pair -> two fp32: the low half shifted up, the high half masked
        o[u * 4 + e] = add_f32x2(o[u * 4 + e], f32x2_bits(w[e] << 16, w[e] & 0xffff0000u));
    }
  }
  // out_chunk is 32 or 64: shifts, not integer divides (which go through the MUFU pipe)
  const int cs = p.out_chunk == 64 ? 6 : 5;
  const int chunk = c0 >> cs;
  const int row_bytes = p.out_chunk * 2;
  const int u0 = (c0 & (p.out_chunk - 1)) >> 3;
  const int sw = (p.out_chunk == 64) ? (row & 7) : ((row >> 1) & 3);
  const uint32_t rowa = stage_saddr + (uint32_t)(chunk * 128 * row_bytes + row * row_bytes);
#pragma unroll
  for (int u = 0; u < NC / 8; ++u) {
    uint32_t w[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float lo, hi;
      f32x2_split(o[u * 4 + e], lo, hi);
      w[e] = pack_bf16x2(lo, hi);
    }
    sts_v4(rowa + (uint32_t)(((u0 + u) ^ sw) << 4), w[0], w[1], w[2], w[3]);
  }
}

template <int NC, int OFF, int ACT, int MODE>
__device__ __forceinline__ void conv_finish_cols(const ConvArgs& p, const uint32_t (&r)[32], int c0,
                                                 const float* bias_g, int row, bool valid,
                                                 const ConvTile& t, int y, int x, uint8_t* stage_base) {
  if constexpr (ACT == 1 && MODE == 0) {
    // (a residual BEFORE the activation only occurs with ReLU in the ResNet programs; SiLU layers add it after)
    conv_finish_silu<NC, OFF>(p, r, c0, smem_u32(bias_g), row, valid, t, y, x, smem_u32(stage_base));
    return;
  }
  float v[NC];
  const float4* b4 = reinterpret_cast<const float4*>(bias_g + t.n_tile * p.BN + c0);
#pragma unroll
  for (int j = 0; j < NC / 4; ++j) {
    const float4 b = b4[j];
    v[4 * j + 0] = __uint_as_float(r[OFF + 4 * j + 0]) + b.x;
    v[4 * j + 1] = __uint_as_float(r[OFF + 4 * j + 1]) + b.y;
    v[4 * j + 2] = __uint_as_float(r[OFF + 4 * j + 2]) + b.z;
    v[4 * j + 3] = __uint_as_float(r[OFF + 4 * j + 3]) + b.w;
  }
  if (MODE == 0 && p.resid != nullptr && p.resid_pre && valid) conv_add_resid<NC>(p, v, c0, t, y, x);
  if (ACT == 1) {
#pragma unroll
    for (int j = 0; j < NC; ++j) v[j] = silu_f(v[j]);
  } else if (ACT == 2) {
#pragma unroll
    for (int j = 0; j < NC; ++j) v[j] = fmaxf(v[j], 0.f);
  } else if (ACT == 3) {
#pragma unroll
    for (int j = 0; j < NC; ++j) v[j] = 0.5f * v[j] * (1.0f + erff(v[j] * 0.70710678118654752f));
  }
  if (MODE == 0) {
    if (p.resid != nullptr && !p.resid_pre && valid) conv_add_resid<NC>(p, v, c0, t, y, x);
    // out_chunk is 32 or 64: shifts, not integer divides (which go through the MUFU pipe)
    const int cs = p.out_chunk == 64 ? 6 : 5;
    const int chunk = c0 >> cs;
    const int row_bytes = p.out_chunk * 2;
    const int u0 = (c0 & (p.out_chunk - 1)) >> 3;
    const int sw = (p.out_chunk == 64) ? (row & 7) : ((row >> 1) & 3);
    uint8_t* rowp = stage_base + (size_t)chunk * 128 * row_bytes + (size_t)row * row_bytes;
#pragma unroll
    for (int u = 0; u < NC / 8; ++u) {
      uint4 q;
      q.x = pack_bf16x2(v[u * 8 + 0], v[u * 8 + 1]);
      q.y = pack_bf16x2(v[u * 8 + 2], v[u * 8 + 3]);
      q.z = pack_bf16x2(v[u * 8 + 4], v[u * 8 + 5]);
      q.w = pack_bf16x2(v[u * 8 + 6], v[u * 8 + 7]);
      *reinterpret_cast<uint4*>(rowp + (((u0 + u) ^ sw) << 4)) = q;
    }
  } else {
    // f32 prediction rows: park the slab in this warp's private staging rows (stride BN+1 words,
    // conflict-free); conv_flush_rows() writes them out coalesced
    const uint32_t srow = smem_u32(stage_base) + (uint32_t)(row * (p.BN + 1) + c0) * 4u;
#pragma unroll
    for (int j = 0; j < NC; ++j) asm volatile("st.shared.f32 [%0], %1;" ::"r"(srow + 4u * j), "f"(v[j]) : "memory");
  }
}

// mode 1: each epilogue warp streams the staged rows it keeps to out_f32, one row at a time: consecutive lanes write
// consecutive floats of the row (ceil(n_valid / 32) independent ld.shared + st.global per row).  With candidate
// emission (YOLOX prediction convs) only ~1-2 % of the rows are kept, most warps have none and skip the flush.
// (The first version walked every row with a divide and a dependent generic load per row: 440 cycles per row.)
__device__ __forceinline__ void conv_flush_rows(const ConvArgs& p, const ConvTile& t, int q, int half, int lane,
                                                const uint8_t* stage_base, bool valid, int y, int x) {
  (void)half;
  const int nv = p.n_valid;
  // element offset of the row this lane produced (lane L owns staged row q*32 + L)
  const long long my_off =
      valid ? ((long long)t.n_img * p.rows_per_img + p.row_off + (long long)y * p.Wo + x) * p.out_ld + p.out_col : 0;
  const uint32_t st = smem_u32(stage_base) + (uint32_t)(q * 32) * (uint32_t)(p.BN + 1) * 4u;
  unsigned m = __ballot_sync(0xffffffffu, valid);
  while (m) {
    const int r = __ffs(m) - 1;
    m &= m - 1;
    const long long off = __shfl_sync(0xffffffffu, my_off, r);
    const uint32_t src = st + (uint32_t)(r * (p.BN + 1)) * 4u;
    for (int c = lane; c < nv; c += 32) p.out_f32[off + c] = lds_f32(src + (uint32_t)c * 4u);
  }
}

// Specialised on (activation, halo form, f32-row output): one generic kernel was ~7 800 SASS instructions and its
// warps spent 10 % of their samples waiting for instruction fetch (ncu warp sampling, stall_no_inst).
template <int ACT, bool HALO, bool F32>
__global__ void __launch_bounds__(kConvThreads)
conv_umma_kernel(const __grid_constant__ ConvMaps maps, const __grid_constant__ ConvArgs p) {
  constexpr int kMode = F32 ? 1 : 0;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[kConvMaxStages];
  __shared__ uint64_t empty_bar[kConvMaxStages];
  __shared__ uint64_t tmem_full_bar[2];
  __shared__ uint64_t tmem_empty_bar[2];
  __shared__ uint64_t a_full_bar[kConvMaxAStages];
  __shared__ uint64_t a_empty_bar[kConvMaxAStages];
  __shared__ uint32_t tmem_base_s;

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  // the layer's bias lives behind the staging area, sized by the layer (a fixed 8 KiB array cost the prediction convs
  // their second ring stage)
  float* s_bias = reinterpret_cast<float*>(smem + p.bias_off);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int a_bytes = HALO ? 0 : 128 * p.ck * 2;
  const int b_bytes = p.BN * p.ck * 2;
  const int stage_bytes = a_bytes + b_bytes * p.tps;
  const int its_per_tile = p.taps * p.kchunks;
  if (threadIdx.x == 0) { conv_trace(p, 0, true); conv_trace(p, 1); }

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&maps.a[0]);
    prefetch_tmap(&maps.b);
    if (kMode == 0) prefetch_tmap(&maps.o);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full_bar[a], 1);
      mbar_init(&tmem_empty_bar[a], kConvEpiWarps);  // one arrive per epilogue warp
    }
    for (int a = 0; a < p.a_stages; ++a) {
      mbar_init(&a_full_bar[a], 1);
      mbar_init(&a_empty_bar[a], 1);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_s, (uint32_t)p.tmem_cols);
  // the bias never depends on the previous layer: stage it while the pipeline spins up
  if (warp >= 2)
    for (int i = threadIdx.x - 64; i < p.bias_n; i += 32 * kConvEpiWarps)
      s_bias[i] = (ACT == 1 && kMode == 0 ? 0.5f : 1.0f) * __ldg(p.bias + i);   // SiLU layers: see conv_finish_silu
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  if (threadIdx.x == 0) conv_trace(p, 2);
  // pdl == 1: the next layer's CTAs may become resident now (their prologue overlaps this layer, but they hold an SM
  // slot while they wait); pdl == 2: only once this CTA has issued its last MMA (TCD_PDL_LATE, A/B measurement)
  if (p.pdl == 1 && threadIdx.x == 0) pdl_launch_dependents();

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      bool waited = !p.pdl;
      bool first = true;
      RingCursor ring{0, 0u, p.stages};
      int jt = 0;
      if (HALO) {
        const uint32_t a_tx = (uint32_t)((p.TW + 2) * (p.TH + 2) * p.ck * 2);
        const uint32_t b_tx = (uint32_t)(b_bytes * p.tps);
        RingCursor aring{0, 0u, p.a_stages};
        for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++jt) {
          const ConvTile c = conv_tile(p, t);
          conv_trace_tile(p, 32, jt);
          const bool load_b = !p.b_resident || jt == 0;
          for (int kc = 0; kc < p.kchunks; ++kc) {
            for (int tap = 0; tap < 9; tap += p.tps) {
              if (load_b) {
                mbar_wait(&empty_bar[ring.s], ring.ph ^ 1u);
                mbar_expect_tx(&full_bar[ring.s], (p.dbg & 1) ? 0u : b_tx);
                if (!(p.dbg & 1))
                  tma_load_3d(smem + (size_t)ring.s * stage_bytes, &maps.b, &full_bar[ring.s], kc * p.ck,
                              c.n_tile * p.BN, tap);
                ring.advance();
              }
              if (tap == 0) {  // the halo box of this k-chunk, after the first weight load
                mbar_wait(&a_empty_bar[aring.s], aring.ph ^ 1u);
                mbar_expect_tx(&a_full_bar[aring.s], (p.dbg & 1) ? 0u : a_tx);
                if (!waited) {
                  pdl_wait();
                  waited = true;
                }
                if (!(p.dbg & 1))
                  tma_load_4d(smem + p.a_ring_off + (size_t)aring.s * p.a_stage_bytes, &maps.a[0],
                            &a_full_bar[aring.s], p.cin_off + c.n_tile * p.group_cin + kc * p.ck, c.x0 - 1, c.y0 - 1,
                            c.n_img);
                aring.advance();
                if (first) conv_trace(p, 3);
                first = false;
              }
            }
          }
        }
      } else {
        const uint32_t tx_bytes = (uint32_t)(p.TW * p.TH * p.ck * 2 + b_bytes);
        for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++jt) {
          const ConvTile c = conv_tile(p, t);
          conv_trace_tile(p, 32, jt);
          for (int tap = 0; tap < p.taps; ++tap) {
            const CUtensorMap* am = &maps.a[p.tap_map[tap]];
            const int cx = c.x0 + p.tap_dx[tap];
            const int cy = c.y0 + p.tap_dy[tap];
            for (int kc = 0; kc < p.kchunks; ++kc) {
              mbar_wait(&empty_bar[ring.s], ring.ph ^ 1u);
              uint8_t* sa = smem + (size_t)ring.s * stage_bytes;
              mbar_expect_tx(&full_bar[ring.s], (p.dbg & 1) ? 0u : tx_bytes);
              // weights do not depend on the previous layer: issue them before the dependency wait
              if (!(p.dbg & 1))
                tma_load_3d(sa + a_bytes, &maps.b, &full_bar[ring.s], kc * p.ck, c.n_tile * p.BN, tap);
              if (!waited) {
                pdl_wait();
                waited = true;
              }
              if (!(p.dbg & 1))
                tma_load_4d(sa, am, &full_bar[ring.s], p.cin_off + c.n_tile * p.group_cin + kc * p.ck, cx, cy, c.n_img);
              ring.advance();
              if (first) conv_trace(p, 3);
              first = false;
            }
          }
        }
      }
      conv_trace(p, 4);
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_bf16(128, p.BN);
      const uint32_t lt = (p.ck == 64) ? 2u : (p.ck == 32 ? 4u : 6u);
      const uint32_t sbo = (uint32_t)(8 * p.ck * 2);
      const int ksteps = p.ck >> 4;
      int j = 0;
      bool first = true;
      RingCursor ring{0, 0u, p.stages};
      RingCursor aring{0, 0u, p.a_stages};
      int rnd = 0, g = 0;  // accumulator round and tile within it: a round is p.group tiles, BN columns each
      for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x, ++j) {
        const int acc = rnd & 1;
        const uint32_t aph = (uint32_t)(rnd >> 1) & 1u;
        if (g == 0) {
          mbar_wait(&tmem_empty_bar[acc], aph ^ 1u);  // epilogue has drained this accumulator
          tc_fence_after();
        }
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * p.acc_stride + g * p.BN);
        if (HALO) {
          const int rowb = p.ck * 2;
          const uint32_t sbo_a = (uint32_t)(p.halo_w * rowb);
          const bool wait_b = !p.b_resident || j == 0;
          if (p.b_resident) ring.s = 0;  // resident weights: every tile walks the same stages
          for (int kc = 0; kc < p.kchunks; ++kc) {
            mbar_wait(&a_full_bar[aring.s], aring.ph);
            const uint32_t halo_addr = smem_u32(smem + p.a_ring_off + (size_t)aring.s * p.a_stage_bytes);
            for (int tap0 = 0; tap0 < 9; tap0 += p.tps) {
              if (wait_b) mbar_wait(&full_bar[ring.s], ring.ph);
              tc_fence_after();
              if (first) conv_trace(p, 5);
              first = false;
              const uint32_t b_base = smem_u32(smem + (size_t)ring.s * stage_bytes);
              for (int u = 0; u < p.tps; ++u) {
                const int tap = tap0 + u;
                const int ty = tap / 3, tx = tap - ty * 3;
                const uint32_t a_addr = halo_addr + (uint32_t)((ty * p.halo_w + tx) * rowb);
                const uint32_t b_addr = b_base + (uint32_t)(u * b_bytes);
                for (int k = 0; k < ksteps; ++k) {
                  const uint64_t ad = umma_smem_desc(a_addr + k * 32, sbo_a, lt);
                  const uint64_t bd = umma_smem_desc(b_addr + k * 32, sbo, lt);
                  if (!(p.dbg & 2)) umma_bf16(d_tmem, ad, bd, idesc, (uint32_t)((kc | tap | k) != 0));
                }
              }
              if (!p.b_resident) umma_commit(&empty_bar[ring.s]);
              ring.advance();
            }
            umma_commit(&a_empty_bar[aring.s]);
            aring.advance();
          }
        } else {
          for (int i = 0; i < its_per_tile; ++i) {
            mbar_wait(&full_bar[ring.s], ring.ph);
            tc_fence_after();
            if (first) conv_trace(p, 5);
            first = false;
            const uint32_t a_addr = smem_u32(smem + (size_t)ring.s * stage_bytes);
            const uint32_t b_addr = a_addr + (uint32_t)a_bytes;
            for (int k = 0; k < ksteps; ++k) {
              const uint64_t ad = umma_smem_desc(a_addr + k * 32, sbo, lt);
              const uint64_t bd = umma_smem_desc(b_addr + k * 32, sbo, lt);
              if (!(p.dbg & 2)) umma_bf16(d_tmem, ad, bd, idesc, (uint32_t)((i | k) != 0));
            }
            umma_commit(&empty_bar[ring.s]);
            ring.advance();
          }
        }
        if (++g == p.group || t + (int)gridDim.x >= p.total_tiles) {  // round complete (or the CTA's last tile)
          umma_commit(&tmem_full_bar[acc]);
          conv_trace_tile(p, 48, rnd);
          g = 0;
          ++rnd;
        }
      }
      conv_trace(p, 6);
      if (p.pdl == 2) pdl_launch_dependents();
    }
  } else {
    // ------------------------------------------------------------------ epilogue (kConvEpiWarps warps)
    // A warp may only read the TMEM lane quadrant (warp % 4).  With 8 warps, two share a quadrant and take
    // alternate 32-column slabs; measured, that bought nothing (tail 256 -> 231 us per step, and it costs
    // the registers that let three layers be co-resident): a tile's epilogue is a ~3000-cycle chain of
    // barrier / fence / TMA-issue latencies even with the TMEM loads, SiLU and staging stores removed
    // (TCD_CONV_DBG=28, profiles/r01_conv_timeline.md), so it is hidden by co-resident CTAs instead.
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    const int ly = row / p.TW;
    const int lx = row - ly * p.TW;
    uint8_t* staging = smem + p.staging_off;
    if (p.pdl && p.resid != nullptr) pdl_wait();
    const float* bias_src = s_bias;   // always shared: a global/shared select makes every bias read a generic load
    // candidate emission (f32 rows): the threshold is read once, long before the first accumulator is ready
    const float emit_thr = (kMode == 1 && p.emit_thr != nullptr) ? p.emit_thr[0] * 0.999f - 1e-6f : 0.f;
    const int nslab = (p.BN + 31) >> 5;
    // Accumulator rounds: p.group consecutive tiles of this CTA share one round (group * BN <= 128 columns),
    // so the fixed per-round chain (barrier waits, proxy fence, TMA-store issue, staging hand-over: ~2 000
    // cycles whatever the tile holds, profiles/r01_conv_timeline.md) is paid once per 128 columns instead of
    // once per 32/64-column tile -- the stem has 21 such tiles per CTA.
    const int n_mine = (p.total_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    const int tile_stage_bytes = 128 * p.BN * 2;
    int j = 0;  // round
    for (int i0 = 0; i0 < n_mine; i0 += p.group, ++j) {
      const int n_g = min(p.group, n_mine - i0);
      const int acc = j & 1;
      const uint32_t aph = (uint32_t)(j >> 1) & 1u;
      // the tile the slab being finished belongs to; the first one is located while the MMAs still run
      ConvTile c;
      int y = 0, x = 0;
      bool valid = false;
      uint8_t* stg = staging;
      auto set_tile = [&](int g) {
        c = conv_tile(p, (int)blockIdx.x + (i0 + g) * (int)gridDim.x);
        y = c.y0 + ly;
        x = c.x0 + lx;
        valid = (row < p.TW * p.TH) && (y < p.Ho) && (x < p.Wo);
        stg = staging + (size_t)g * tile_stage_bytes;
      };
      set_tile(0);
      mbar_wait(&tmem_full_bar[acc], aph);
      tc_fence_after();
      if (j == 0 && warp == 2 && lane == 0) conv_trace(p, 7);
      if (warp == 2 && lane == 0) conv_trace_tile(p, 64, j);
      if (j > 0) {
        // staging reuse: the previous round's TMA stores have finished READING it (mode 0) / every warp has
        // finished flushing its rows (mode 1)
        if (kMode == 0 && warp == 2 && lane == 0) tma_store_wait_read();
        asm volatile("bar.sync 1, %0;" ::"n"(32 * kConvEpiWarps) : "memory");
      }
      if (warp == 2 && lane == 0) conv_trace_tile(p, 96, j);   // staging is free
      const uint32_t taddr = tmem_base + (uint32_t)(acc * p.acc_stride) + ((uint32_t)(q * 32) << 16);
      // slabs of 32 accumulator columns (a ragged last one of 16 when group == 1), double-buffered in
      // registers: the tcgen05.ld of slab s+1 is in flight while slab s goes through bias / activation /
      // staging (167 registers: two CTAs of 192 threads per SM, which is what TCD_CONV_SHARE=2 plans for)
      const int S = n_g * nslab;      // with group > 1, BN is a multiple of 32: slab s starts at column 32*s
      uint32_t ra[32], rb[32];
      auto issue = [&](int sidx, uint32_t (&r)[32]) {
        if (p.group > 1 || p.BN - 32 * sidx >= 32) tmem_ld32(taddr + 32 * sidx, r);
        else tmem_ld16_lo(taddr + 32 * sidx, r);
      };
      auto handback = [&]() {  // every column has been read: the accumulator goes back to the MMA warp
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty_bar[acc]);
      };
      int fg = 0, fsl = 0;            // (tile, slab) of the next finish() -- finish() runs in slab order
      auto finish = [&](const uint32_t (&r)[32]) {
        if (p.BN - 32 * fsl >= 32) conv_finish_cols<32, 0, ACT, kMode>(p, r, 32 * fsl, bias_src, row, valid, c, y, x, stg);
        else conv_finish_cols<16, 0, ACT, kMode>(p, r, 32 * fsl, bias_src, row, valid, c, y, x, stg);
        if (++fsl == nslab) {  // next tile of the round (its slab is already on its way from TMEM)
          fsl = 0;
          if (++fg < n_g) set_tile(fg);
        }
      };
      issue(0, ra);
      for (int sidx = 0; sidx < S; sidx += 2) {
        tmem_ld_wait();
        if (sidx + 1 < S) issue(sidx + 1, rb); else handback();
        finish(ra);
        if (sidx + 1 < S) {
          tmem_ld_wait();
          if (sidx + 2 < S) issue(sidx + 2, ra); else handback();
          finish(rb);
        }
      }
      if (kMode == 1) {
        asm volatile("bar.sync 1, %0;" ::"n"(32 * kConvEpiWarps) : "memory");
        bool keep = valid;
        if (p.emit_thr != nullptr && valid) {
          // a superset test (the decode kernel re-evaluates the exact score): drop the row only when its
          // objectness alone is clearly below the threshold
          const float obj = reinterpret_cast<const float*>(staging)[(size_t)row * (p.BN + 1) + p.emit_col];
          keep = 1.0f / (1.0f + expf(-obj)) >= emit_thr;
        }
        conv_flush_rows(p, c, q, half, lane, staging, keep, y, x);
        if (p.emit_thr != nullptr) {
          // the kept rows' anchors go to the image's list: ONE atomic per warp and image run (a patch never spans
          // two images), issued after the rows are on their way so nothing waits for its return value
          const unsigned m = __ballot_sync(0xffffffffu, keep);
          if (m) {
            int base = 0;
            if (lane == __ffs(m) - 1) base = atomicAdd(p.maybe_count + c.n_img, __popc(m));
            base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
            const int slot = base + __popc(m & ((1u << lane) - 1u));
            if (keep && slot < p.maybe_cap)
              p.maybe_idx[(size_t)c.n_img * p.maybe_cap + slot] = p.row_off + y * p.Wo + x;
          }
        }
      } else {
        fence_proxy_async_smem();
        asm volatile("bar.sync 1, %0;" ::"n"(32 * kConvEpiWarps) : "memory");
        if (warp == 2 && lane == 0) conv_trace_tile(p, 112, j);  // every warp has staged its slabs
        if (warp == 2 && lane == 0) {
          const int nchunks = p.BN >> (p.out_chunk == 64 ? 6 : 5);
          for (int g = 0; g < n_g; ++g) {
            const ConvTile cg = conv_tile(p, (int)blockIdx.x + (i0 + g) * (int)gridDim.x);
            for (int ch = 0; ch < nchunks; ++ch)
              tma_store_4d(&maps.o, staging + (size_t)g * tile_stage_bytes + (size_t)ch * 128 * p.out_chunk * 2,
                           p.cout_off + cg.n_tile * p.BN + ch * p.out_chunk, cg.x0, cg.y0, cg.n_img);
          }
          tma_store_commit();
        }
      }
      if (warp == 2 && lane == 0) conv_trace_tile(p, 80, j);
    }
    if (warp == 2 && lane == 0) {
      conv_trace(p, 8);
      if (kMode == 0) tma_store_wait_read();
      conv_trace(p, 9);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
  if (threadIdx.x == 32) { conv_trace(p, 10); conv_trace(p, 11, true); }
}

#endif  // __CUDACC__ && TCD_CONV_KERNEL_IMPL

// ---------------------------------------------------------------- host-side plan
struct ConvDesc {
  // input view
  const void* x = nullptr;
  int N = 0, H = 0, W = 0, x_ctot = 0, x_coff = 0, cin = 0;  // cin: padded to a multiple of 16
  // weights: device, bf16 [taps][cout_pad][cin]; bias: device f32 [cout_pad]
  const void* w = nullptr;
  const float* bias = nullptr;
  int cout = 0, cout_pad = 0, k = 1, stride = 1;
  int silu = 1;  // activation: 0 none, 1 SiLU, 2 ReLU, 3 GELU (erf)
  // elements between consecutive images; 0 = dense (H*W*ctot).  Lets a conv read / write one pyramid
  // level inside a [B, sum(HW), C] token tensor (RT-DETR memory).
  long long x_nstride = 0, y_nstride = 0, r_nstride = 0;
  int resid_pre = 0;  // residual added before (1) or after (0) the activation
  int share = -1;     // plan for 1/share of an SM (co-residency with other streams' layers); -1 = TCD_CONV_SHARE
  int groups = 1;     // grouped conv: cout / groups output channels per group, each reading `cin` input channels at
                      // x_coff + g * cin (the fused cls / reg branch convs of the YOLOX head); weights [k*k][cout][cin]
  // bf16 output (mode 0)
  void* y = nullptr;
  int y_ctot = 0, y_coff = 0;
  const void* resid = nullptr;
  int r_ctot = 0, r_coff = 0;
  // f32 row output (mode 1)
  float* out_f32 = nullptr;
  int out_ld = 0, out_col = 0, rows_per_img = 0, row_off = 0;
  // f32 rows only: candidate emission (see ConvArgs)
  const float* ln_stats = nullptr;    // consumer-side LayerNorm (CTA-pair GEMM only), see ConvArgs
  const float* ln_colsum = nullptr;
  int ln_parts = 0, ln_count = 0;
  float ln_eps = 0.f;
  float* stats_out = nullptr;
  int stats_parts = 0, stats_slot0 = 0;
  const float* emit_thr = nullptr;
  int* maybe_count = nullptr;
  int* maybe_idx = nullptr;
  int maybe_cap = 0, emit_col = 0;
};

struct PairArgs;
struct ConvPlan {
  ConvMaps maps;
  ConvArgs args;
  int pair = 0;        // 1: conv_pair_kernel (cta_group::2, resident weights, conv_pair.cuh); 2: gemm_pair_kernel (gemm_pair.cuh)
  int pair_n_tiles = 0, pair_m_tiles = 0, pair_group_cin = 0;
  dim3 grid;
  int smem_bytes = 0;
  double flops = 0;
};

int conv_plan_build(const ConvDesc& d, ConvPlan* plan);
int conv_plan_launch(const ConvPlan& plan, cudaStream_t stream);
// pack f32 [cout][cin][k][k] (host) -> bf16 [k*k][cout_pad][cin_pad] (host buffer)
void conv_pack_weights(const float* w, int cout, int cin, int k, int cout_pad, int cin_pad,
                       uint16_t* out);
uint16_t f32_to_bf16_rn(float f);

}  // namespace tcd
