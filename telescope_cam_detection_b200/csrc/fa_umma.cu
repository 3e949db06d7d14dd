// Rotary multi-head attention on the tcgen05 tensor cores (head_dim 64, up to 640 tokens): the attention of the
// Stage-2 classifier body (EVA02: 577 tokens x 16 heads per crop; src/species_classifier.py:258-262,386 -> timm Eva
// attention, oracle/eva02.py::_Attn).  Same contract as mha64_rope_kernel (vit_kernels.cu, the warp-level mma.sync
// version it replaces: 0.89 ms per layer at 64 crops); this one keeps both GEMMs of attention on tcgen05.mma with
// the accumulators in tensor memory.
//
// One CTA per (image, head).  All of the head's K (rotated) and V rows are staged ONCE in shared memory in the
// 128-byte-swizzled layouts the tensor core reads: K as a K-major B operand [keys][64] for S = Q K^T, V -- the same
// row-major [keys][64] bytes -- as an MN-major B operand for O = P V (no transposition anywhere).  Then, per tile of
// 128 queries (staged + rotated by the softmax warps) and per tile of 128 keys:
//   warp 8, one thread:  S_j = Q K_j^T    (4 x tcgen05.mma M128 N128 K16)  -> TMEM, double-buffered
//                        O_j = P_j V_j    (8 x tcgen05.mma M128 N64  K16)  -> TMEM, a fresh partial per key tile
//   warps 0-7 (thread = one query row x one half of the columns; warps w and w + 4 share a TMEM lane quadrant):
//                        tcgen05.ld of the thread's 64 scores (read ONCE, kept in registers), online softmax in the
//                        exp2 domain -- only the row maximum crosses the pair, through shared memory --, P_j -> bf16
//                        -> tcgen05.st back into the first 64 columns of the score buffer (the A operand of P V is
//                        read from tensor memory: no shared-memory round trip, no proxy fence), and, one step
//                        behind, the partial O_{j-1} is read back and folded into the fp32 accumulator the thread
//                        keeps in registers: O = O * 2^(m_{j-2} - m_{j-1}) + O_{j-1}.  Nothing is rescaled in TMEM.
//   warps 9-11:          stage + rotate the NEXT tile of 128 queries into the second Q buffer.
// S_{j+1} is issued before P_j is awaited, so the tensor pipe computes the next scores while the softmax warps work.
// (First version, four softmax warps with a whole row per thread, two passes over S and Q staged by the softmax
// warps themselves: 0.555 ms per layer at 64 crops; ncu warp sampling: 26 % of the softmax warps' time in the Q
// staging, 16 % in the K / V staging, ten exposed tcgen05.ld round trips per step.  Second version, eight softmax
// warps but P through shared memory: 0.58 ms -- the chain P_j -> fence -> P V_j -> O_j -> P_{j+1} (one P buffer) was
// the critical path, 6 500 cycles per step with the issue slots 21 % busy.)
#include "tfm_kernels.cuh"

#include <atomic>
#include <stdlib.h>

namespace tcd {
extern std::atomic<uint64_t> g_launch_count;

constexpr int kFaThreads = 384;       // warps 0-7 softmax, warp 8 MMA issuer, warps 9-11 stage the query tiles
constexpr int kFaStagers = 96;
constexpr int kFaSoftmaxThreads = 256;
constexpr int kFaTile = 128;
constexpr int kFaTileBytes = kFaTile * 128;   // 128 rows x 64 bf16
constexpr int kFaMaxTiles = 5;        // <= 640 tokens; 577: 2 x 74 KiB of K / V + 2 x 16 KiB Q + 32 KiB P = 212 KiB of shared memory

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t fa_rope_pair(uint32_t packed, float c, float sn) {
  const float2 x = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&packed));
  return pack_bf16x2(x.x * c - x.y * sn, x.y * c + x.x * sn);
}
// 16-byte chunk c of row r of a [rows][128 B] tile in the 128-byte swizzle (1 KiB atoms of 8 rows)
__device__ __forceinline__ uint32_t fa_swz(int r, int c) { return (uint32_t)(r * 128 + ((c ^ (r & 7)) << 4)); }
__device__ __forceinline__ void fa_pair_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }

// 32 lanes x 16 consecutive 32-bit columns <- 16 registers per thread
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// D[tmem] (+)= A[tmem] * B[smem]: A = 128 rows (lanes) x 16 bf16, packed two per 32-bit column (8 columns per K16 step)
__device__ __forceinline__ void umma_bf16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// rotate the four channel pairs of chunk c of token t (no-op for the class tokens / without a table)
__device__ __forceinline__ void fa_rope_chunk(uint4& q, const float2* rope, int t, int c, int prefix) {
  if (rope == nullptr || t < prefix) return;
  const float4* cs = reinterpret_cast<const float4*>(rope + (size_t)(t - prefix) * 32 + c * 4);
  const float4 c01 = cs[0], c23 = cs[1];
  q.x = fa_rope_pair(q.x, c01.x, c01.y);
  q.y = fa_rope_pair(q.y, c01.z, c01.w);
  q.z = fa_rope_pair(q.z, c23.x, c23.y);
  q.w = fa_rope_pair(q.w, c23.z, c23.w);
}

__global__ void __launch_bounds__(kFaThreads, 1)
fa64_umma_kernel(const __nv_bfloat16* __restrict__ q, int ldq, int qoff, const __nv_bfloat16* __restrict__ k, int ldk,
                 int koff, const __nv_bfloat16* __restrict__ v, int ldv, int voff, __nv_bfloat16* __restrict__ out, int ldo,
                 int T, float scale_log2, const float2* __restrict__ rope, int prefix, int n_tiles, int n_last, int hd8, int trace,
                 float2* __restrict__ stats_out) {
  extern __shared__ uint8_t fa_raw[];
  __shared__ uint32_t tr[2][8][8];   // TCD_FA_TRACE=1: clock stamps of CTA (0,0,0), first query tile (printed at exit)
  const bool tracing = trace != 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
  const uint32_t t_start = (uint32_t)clock64();
#define FA_STAMP(who, step, slot) do { if (tracing && (step) < 8) tr[who][step][slot] = (uint32_t)clock64() - t_start; } while (0)
  __shared__ uint64_t bar_s_full[2], bar_o_full[2], bar_q_full[2], bar_q_free[2], bar_p_full;
  __shared__ uint32_t tmem_base_s;
  __shared__ float xch[2][kFaTile][2];          // row maxima of the two column halves, double-buffered by step
  __shared__ float xsum[kFaTile][2];            // row sums of the two halves (once per query tile)
  uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)fa_raw + 1023) & ~(uintptr_t)1023);
  // the last key tile is n_last (a multiple of 16) keys wide: 577 tokens = 4 x 128 + 80, not 5 x 128
  const int kv_rows = (n_tiles - 1) * kFaTile + n_last;
  uint8_t* Ksm = smem;
  uint8_t* Vsm = Ksm + (size_t)kv_rows * 128;            // kv_rows is a multiple of 16: 1 KiB-aligned (2 KiB steps)
  uint8_t* Qsm = Vsm + (size_t)kv_rows * 128;            // two query tiles
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.z, h = blockIdx.y;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar_s_full[i], 1);
      mbar_init(&bar_o_full[i], 1);
      mbar_init(&bar_q_full[i], kFaStagers);
      mbar_init(&bar_q_free[i], 1);
    }
    mbar_init(&bar_p_full, kFaSoftmaxThreads);
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc(&tmem_base_s, 512u);
  {
    // K (rotated) and V of this (image, head): rows >= T are zeros (a masked key must contribute exactly 0 * 0)
    const int hd = hd8 * 8;   // 64, or 32 with channels 32..63 of every staged row zero (the MMAs always run 64 wide)
    const __nv_bfloat16* kb = k + (size_t)n * T * ldk + koff + h * hd;
    const __nv_bfloat16* vb = v + (size_t)n * T * ldv + voff + h * hd;
    const int total = kv_rows * 8;
    // six chunks per thread per round, K / V data and the rotation table loads all issued before the first use
    // (twelve 16-byte loads + twelve table loads in flight: the loop is latency-bound otherwise)
    for (int i0 = threadIdx.x; i0 < total; i0 += 6 * kFaThreads) {
      uint4 kq[6], vq[6];
      float4 c01[6], c23[6];
#pragma unroll
      for (int u = 0; u < 6; ++u) {
        const int i = i0 + u * kFaThreads;
        const int t = i >> 3, c = i & 7;
        kq[u] = make_uint4(0u, 0u, 0u, 0u);
        vq[u] = kq[u];
        c01[u] = make_float4(1.f, 0.f, 1.f, 0.f);
        c23[u] = c01[u];
        if (i < total && t < T && c < hd8) {
          kq[u] = *reinterpret_cast<const uint4*>(kb + (size_t)t * ldk + c * 8);
          vq[u] = *reinterpret_cast<const uint4*>(vb + (size_t)t * ldv + c * 8);
          if (rope != nullptr && t >= prefix) {
            const float4* cs = reinterpret_cast<const float4*>(rope + (size_t)(t - prefix) * 32 + c * 4);
            c01[u] = cs[0];
            c23[u] = cs[1];
          }
        }
      }
#pragma unroll
      for (int u = 0; u < 6; ++u) {
        const int i = i0 + u * kFaThreads;
        if (i < total) {
          kq[u].x = fa_rope_pair(kq[u].x, c01[u].x, c01[u].y);
          kq[u].y = fa_rope_pair(kq[u].y, c01[u].z, c01[u].w);
          kq[u].z = fa_rope_pair(kq[u].z, c23[u].x, c23[u].y);
          kq[u].w = fa_rope_pair(kq[u].w, c23[u].z, c23[u].w);
          const uint32_t off = fa_swz(i >> 3, i & 7);
          *reinterpret_cast<uint4*>(Ksm + off) = kq[u];
          *reinterpret_cast<uint4*>(Vsm + off) = vq[u];
        }
      }
    }
  }
  {
    // the CTA's FIRST query tile goes in with the K / V rows (warp 9 alone needs ~16 000 cycles for a tile: fine one
    // tile ahead of the softmax warps, not in front of the first scores)
    const __nv_bfloat16* qb = q + (size_t)n * T * ldq + qoff + h * (hd8 * 8);
    uint4 qq[3];
    float4 c01[3], c23[3];
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int i = threadIdx.x + u * kFaThreads;
      const int tok = (int)blockIdx.x * kFaTile + (i >> 3), c = i & 7;
      qq[u] = make_uint4(0u, 0u, 0u, 0u);
      c01[u] = make_float4(1.f, 0.f, 1.f, 0.f);
      c23[u] = c01[u];
      if (i < kFaTile * 8 && tok < T && c < hd8) {
        qq[u] = *reinterpret_cast<const uint4*>(qb + (size_t)tok * ldq + c * 8);
        if (rope != nullptr && tok >= prefix) {
          const float4* cs = reinterpret_cast<const float4*>(rope + (size_t)(tok - prefix) * 32 + c * 4);
          c01[u] = cs[0];
          c23[u] = cs[1];
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const int i = threadIdx.x + u * kFaThreads;
      if (i < kFaTile * 8) {
        qq[u].x = fa_rope_pair(qq[u].x, c01[u].x, c01[u].y);
        qq[u].y = fa_rope_pair(qq[u].y, c01[u].z, c01[u].w);
        qq[u].z = fa_rope_pair(qq[u].z, c23[u].x, c23[u].y);
        qq[u].w = fa_rope_pair(qq[u].w, c23[u].z, c23[u].w);
        *reinterpret_cast<uint4*>(Qsm + fa_swz(i >> 3, i & 7)) = qq[u];
      }
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 0) FA_STAMP(0, 7, 7);    // (whole warp: a lone diverged lane would break the .sync.aligned tcgen05 ops below)
  const uint32_t tmem_base = tmem_base_s;
  const int n_qt = (T + kFaTile - 1) / kFaTile;

  if (warp >= 9) {
    // ------------------------------------------------------------------ query-tile stagers (run one tile ahead)
    // 96 threads, two rounds of up to six 16-byte chunks each with the rotation table loads issued beside the data
    // loads (one warp with dependent table loads needed 23 000 cycles per tile -- twice what the softmax warps
    // need to consume one).
    const __nv_bfloat16* qb = q + (size_t)n * T * ldq + qoff + h * (hd8 * 8);
    const int st = threadIdx.x - 9 * 32;
    int qi = 1;
    for (int qt = blockIdx.x + gridDim.x; qt < n_qt; qt += gridDim.x, ++qi) {   // tile 0 was staged with K / V
      const int buf = qi & 1;
      mbar_wait(&bar_q_free[buf], (uint32_t)(((qi >> 1) & 1) ^ 1));   // the S MMAs that read this buffer are complete
      uint8_t* dst = Qsm + buf * kFaTileBytes;
#pragma unroll 1
      for (int rnd = 0; rnd < 2; ++rnd) {
        uint4 qq[6];
        float4 c01[6], c23[6];
#pragma unroll
        for (int u = 0; u < 6; ++u) {
          const int i = st + kFaStagers * (rnd * 6 + u);      // chunk index in the tile: 8 threads per row
          const int tok = qt * kFaTile + (i >> 3), c = i & 7;
          qq[u] = make_uint4(0u, 0u, 0u, 0u);
          c01[u] = make_float4(1.f, 0.f, 1.f, 0.f);
          c23[u] = c01[u];
          if (i < kFaTile * 8 && tok < T && c < hd8) {
            qq[u] = *reinterpret_cast<const uint4*>(qb + (size_t)tok * ldq + c * 8);
            if (rope != nullptr && tok >= prefix) {
              const float4* cs = reinterpret_cast<const float4*>(rope + (size_t)(tok - prefix) * 32 + c * 4);
              c01[u] = cs[0];
              c23[u] = cs[1];
            }
          }
        }
#pragma unroll
        for (int u = 0; u < 6; ++u) {
          const int i = st + kFaStagers * (rnd * 6 + u);
          if (i < kFaTile * 8) {
            qq[u].x = fa_rope_pair(qq[u].x, c01[u].x, c01[u].y);
            qq[u].y = fa_rope_pair(qq[u].y, c01[u].z, c01[u].w);
            qq[u].z = fa_rope_pair(qq[u].z, c23[u].x, c23[u].y);
            qq[u].w = fa_rope_pair(qq[u].w, c23[u].z, c23[u].w);
            *reinterpret_cast<uint4*>(dst + fa_swz(i >> 3, i & 7)) = qq[u];
          }
        }
      }
      fence_proxy_async_smem();
      mbar_arrive(&bar_q_full[buf]);
    }
  } else if (warp == 8) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc_full = umma_idesc_bf16(128, 128), idesc_last = umma_idesc_bf16(128, n_last);
      const uint32_t idesc_o = umma_idesc_bf16(128, 64) | (1u << 16);   // B (= V) is MN-major
      const uint32_t q_addr = smem_u32(Qsm);
      const uint64_t k_desc = umma_smem_desc(smem_u32(Ksm), 1024u, 2u), v_desc = umma_smem_desc(smem_u32(Vsm), 1024u, 2u);
      uint32_t s = 0;
      int qi = 0;
      for (int qt = blockIdx.x; qt < n_qt; qt += gridDim.x, ++qi) {
        const uint32_t qa = q_addr + (uint32_t)(qi & 1) * kFaTileBytes;
        // The score buffer step & 1 last held S / P of step - 2.  Its readers are ordered before this MMA without a
        // barrier of their own: the softmax threads read S_{step-2} before they arrived at p_full(step-2), which this
        // thread has waited for, and P V_{step-2} (which read P from the buffer) was issued by this thread earlier --
        // tcgen05.mma instructions of one thread execute in issue order.
        const uint64_t q_desc = umma_smem_desc(qa, 1024u, 2u);
        auto issue_s = [&](uint32_t step, int j) {
          const uint32_t b = step & 1u;
          const uint64_t kd = k_desc + (uint64_t)((uint32_t)j * (kFaTileBytes >> 4));   // start address field, 16-byte units
          const uint32_t id = j + 1 == n_tiles ? idesc_last : idesc_full;
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) umma_bf16(tmem_base + b * 128u, q_desc + 2 * kk, kd + 2 * kk, id, (uint32_t)(kk != 0));
          umma_commit(&bar_s_full[b]);
        };
        if (qi > 0) {
          // q_full[b] completes once per tile warp 9 stages: tiles 1, 3, .. (buffer 1) and 2, 4, .. (buffer 0): use (qi-1)/2
          mbar_wait(&bar_q_full[qi & 1], (uint32_t)(((qi - 1) >> 1) & 1));
          tc_fence_after();
        }
        issue_s(s, 0);
        if (n_tiles == 1) umma_commit(&bar_q_free[qi & 1]);
        for (int j = 0; j < n_tiles; ++j, ++s) {
          if (j + 1 < n_tiles) {
            issue_s(s + 1, j + 1);
            if (j + 2 == n_tiles) umma_commit(&bar_q_free[qi & 1]);   // last scores of this query tile are issued
          }
          FA_STAMP(1, s, 0);
          mbar_wait(&bar_p_full, s & 1u);
          tc_fence_after();
          FA_STAMP(1, s, 1);
          // O_j = P_j V_j: A = P_j from tensor memory (the first 64 columns of the score buffer), B = V rows, MN-major
          const uint32_t d_o = tmem_base + 256u + (s & 1u) * 64u;
          const uint32_t p_tmem = tmem_base + (s & 1u) * 128u;
          const int ksteps = (j + 1 == n_tiles ? n_last : kFaTile) >> 4;
          const uint64_t vd = v_desc + (uint64_t)((uint32_t)j * (kFaTileBytes >> 4));   // 16 key rows = 2 KiB = 128 units
          if (ksteps == 8) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) umma_bf16_ts(d_o, p_tmem + (uint32_t)ks * 8u, vd + 128 * ks, idesc_o, (uint32_t)(ks != 0));
          } else {
            for (int ks = 0; ks < ksteps; ++ks) umma_bf16_ts(d_o, p_tmem + (uint32_t)ks * 8u, vd + 128 * ks, idesc_o, (uint32_t)(ks != 0));
          }
          umma_commit(&bar_o_full[s & 1u]);
          FA_STAMP(1, s, 2);
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax warps
    // thread = (query row r of the tile, column half): warps w and w + 4 share TMEM lane quadrant w & 3 and split
    // the 128 key columns of S (and the 64 channels of O) between them; only the row maximum crosses the pair.
    const int quad = warp & 3, half = warp >> 2;
    const int r = quad * 32 + lane;
    const uint32_t lane_base = tmem_base + ((uint32_t)(quad * 32) << 16);
    uint32_t s = 0;
    for (int qt = blockIdx.x; qt < n_qt; qt += gridDim.x) {
      const int tok = qt * kFaTile + r;
      float o[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) o[i] = 0.f;
      float m = -INFINITY, l = 0.f, c_prev = 0.f;
      for (int j = 0; j < n_tiles; ++j, ++s) {
        const uint32_t b = s & 1u;
        if (warp == 0) FA_STAMP(0, s, 0);
        mbar_wait(&bar_s_full[b], (s >> 1) & 1u);
        tc_fence_after();
        if (warp == 0) FA_STAMP(0, s, 1);
        const uint32_t s_addr = lane_base + b * 128u + (uint32_t)half * 64u;
        const int n_valid = T - j * kFaTile - half * 64;     // valid key columns of this half (may be <= 0 or >= 64)
        uint32_t sa[32], sb[32];
        tmem_ld32(s_addr, sa);
        tmem_ld32(s_addr + 32, sb);
        tmem_ld_wait();
        if (warp == 0) FA_STAMP(0, s, 2);
        if (n_valid < 64) {           // warp-uniform, last key tile only: masked keys score -inf (exp2 -> exactly 0)
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            if (i >= n_valid) sa[i] = 0xff800000u;
            if (32 + i >= n_valid) sb[i] = 0xff800000u;
          }
        }
        float mx;
        {
          float m0 = fmaxf(__uint_as_float(sa[0]), __uint_as_float(sb[0])), m1 = fmaxf(__uint_as_float(sa[1]), __uint_as_float(sb[1]));
          float m2 = fmaxf(__uint_as_float(sa[2]), __uint_as_float(sb[2])), m3 = fmaxf(__uint_as_float(sa[3]), __uint_as_float(sb[3]));
#pragma unroll
          for (int i = 4; i < 32; i += 4) {
            m0 = fmaxf(m0, fmaxf(__uint_as_float(sa[i]), __uint_as_float(sb[i])));
            m1 = fmaxf(m1, fmaxf(__uint_as_float(sa[i + 1]), __uint_as_float(sb[i + 1])));
            m2 = fmaxf(m2, fmaxf(__uint_as_float(sa[i + 2]), __uint_as_float(sb[i + 2])));
            m3 = fmaxf(m3, fmaxf(__uint_as_float(sa[i + 3]), __uint_as_float(sb[i + 3])));
          }
          mx = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
        }
        xch[b][r][half] = mx;
        fa_pair_sync(1 + quad);      // also: the partner has its scores in registers -- P may overwrite them now
        mx = fmaxf(mx, xch[b][r][half ^ 1]);
        if (warp == 0) FA_STAMP(0, s, 3);
        const float m_new = fmaxf(m, mx * scale_log2);
        const float corr = ex2_approx(m - m_new);
        // P_j: this half's 64 keys, bf16 pairs, into 32 columns of the score buffer itself (columns half*32 ..)
        const uint32_t p_addr = lane_base + b * 128u + (uint32_t)half * 32u;
        float sum;
        {
          float s0 = 0.f, s1 = 0.f;
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            uint32_t pk[16];
            if (hh * 32 < n_valid) {          // warp-uniform: 32 masked keys cost no exponentials
#pragma unroll
              for (int e = 0; e < 16; ++e) {
                const float p0 = ex2_approx(fmaf(__uint_as_float(hh == 0 ? sa[e * 2] : sb[e * 2]), scale_log2, -m_new));
                const float p1 = ex2_approx(fmaf(__uint_as_float(hh == 0 ? sa[e * 2 + 1] : sb[e * 2 + 1]), scale_log2, -m_new));
                s0 += p0;
                s1 += p1;
                pk[e] = pack_bf16x2(p0, p1);
              }
            } else {
#pragma unroll
              for (int e = 0; e < 16; ++e) pk[e] = 0u;
            }
            tmem_st16(p_addr + hh * 16, pk);
          }
          sum = s0 + s1;
        }
        l = fmaf(l, corr, sum);
        if (warp == 0) FA_STAMP(0, s, 4);
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&bar_p_full);
        if (warp == 0) FA_STAMP(0, s, 5);
        if (j > 0) {
          // the previous key tile's partial O has had this whole step to complete: fold it in
          const uint32_t sp = s - 1;
          mbar_wait(&bar_o_full[sp & 1u], (sp >> 1) & 1u);
          tc_fence_after();
          uint32_t og[32];
          tmem_ld32(lane_base + 256u + (sp & 1u) * 64u + (uint32_t)half * 32u, og);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = fmaf(o[i], c_prev, __uint_as_float(og[i]));
        }
        if (warp == 0) FA_STAMP(0, s, 6);
        c_prev = corr;
        m = m_new;
      }
      {
        const uint32_t sp = s - 1;
        mbar_wait(&bar_o_full[sp & 1u], (sp >> 1) & 1u);
        tc_fence_after();
        uint32_t og[32];
        tmem_ld32(lane_base + 256u + (sp & 1u) * 64u + (uint32_t)half * 32u, og);
        tmem_ld_wait();
        // the row sum: both halves of the row
        xsum[r][half] = l;
        fa_pair_sync(1 + quad);       // (the next write of xsum is a whole query tile, i.e. many pair barriers, away)
        const float inv = 1.0f / (l + xsum[r][half ^ 1]);
        if (tok < T && half * 4 < hd8) {
          __nv_bfloat16* orow = out + ((size_t)n * T + tok) * ldo + h * (hd8 * 8) + half * 32;
          uint32_t w[16];
          float ps = 0.f, pss = 0.f;      // LayerNorm statistics of the row for the consumer: this (head, half)'s partial
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const float a0 = fmaf(o[i], c_prev, __uint_as_float(og[i])) * inv, a1 = fmaf(o[i + 1], c_prev, __uint_as_float(og[i + 1])) * inv;
            ps += a0 + a1;
            pss = fmaf(a0, a0, fmaf(a1, a1, pss));
            w[i >> 1] = pack_bf16x2(a0, a1);
          }
          if (stats_out != nullptr) stats_out[((size_t)n * T + tok) * (gridDim.y * 2) + h * 2 + half] = make_float2(ps, pss);
#pragma unroll
          for (int c4 = 0; c4 < 4; ++c4)
            *reinterpret_cast<uint4*>(orow + c4 * 8) = make_uint4(w[c4 * 4], w[c4 * 4 + 1], w[c4 * 4 + 2], w[c4 * 4 + 3]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (tracing && threadIdx.x == 0) {
    printf("fa trace: K/V staged at %u, total %u cycles\n", tr[0][7][7], (uint32_t)clock64() - t_start);
    for (int st = 0; st < 7 && st < n_tiles * 2; ++st)
      printf(" step %d softmax: enter %u s_full %u ld %u max+xch %u exps %u st+arrive %u fold %u | mma: S issued %u p_full %u PV issued %u\n",
             st, tr[0][st][0], tr[0][st][1], tr[0][st][2], tr[0][st][3], tr[0][st][4], tr[0][st][5], tr[0][st][6], tr[1][st][0],
             tr[1][st][1], tr[1][st][2]);
  }
  if (warp == 8) tmem_dealloc(tmem_base, 512u);
}

// env TCD_MHA_UMMA=0 keeps the warp-level kernel (A/B measurements)
static const bool g_fa_enabled = [] {
  const char* e = getenv("TCD_MHA_UMMA");
  return !(e && e[0] == '0');
}();

static const bool g_fa_trace = [] {
  const char* e = getenv("TCD_FA_TRACE");
  return e && e[0] == '1';
}();

bool fa64_umma_eligible(int T) { return g_fa_enabled && T >= 1 && T <= kFaMaxTiles * kFaTile; }

int fa64_umma_launch(const void* q, int ldq, int qoff, const void* k, int ldk, int koff, const void* v, int ldv, int voff,
                     void* out, int ldo, int N, int T, int heads, int head_dim, float scale, const float* rope, int prefix,
                     float* stats_out, cudaStream_t st) {
  TCD_REQUIRE(head_dim == 64 || (head_dim == 32 && rope == nullptr), "fa64: head_dim %d unsupported", head_dim);
  TCD_REQUIRE(stats_out == nullptr || head_dim == 64, "fa64: row statistics need head_dim 64");
  const int n_tiles = (T + kFaTile - 1) / kFaTile;
  TCD_REQUIRE(n_tiles >= 1 && n_tiles <= kFaMaxTiles, "fa64: %d tokens unsupported", T);
  TCD_REQUIRE(ldq % 8 == 0 && ldk % 8 == 0 && ldv % 8 == 0 && ldo % 8 == 0 && qoff % 8 == 0 && koff % 8 == 0 &&
                  voff % 8 == 0, "fa64: leading dims / offsets must be multiples of 8");
  TCD_REQUIRE(prefix >= 0 && prefix <= T && scale > 0.f, "fa64: prefix %d of %d tokens, scale %f", prefix, T, scale);
  const int n_last = ((T - (n_tiles - 1) * kFaTile) + 15) & ~15;
  const size_t smem = (size_t)((n_tiles - 1) * kFaTile + n_last) * 256 + 2 * kFaTileBytes + 1024;
  if (int rc = ensure_dyn_smem((const void*)fa64_umma_kernel, smem)) return rc;
  // split the query tiles of an (image, head) over CTAs only while the (image, head) pairs alone leave SMs idle
  int gx = 1;
  while (gx * 2 * heads * N <= 148 && gx < n_tiles) gx *= 2;
  if (gx > n_tiles) gx = n_tiles;
  fa64_umma_kernel<<<dim3(gx, heads, N), kFaThreads, smem, st>>>(
      (const __nv_bfloat16*)q, ldq, qoff, (const __nv_bfloat16*)k, ldk, koff, (const __nv_bfloat16*)v, ldv, voff,
      (__nv_bfloat16*)out, ldo, T, scale * 1.4426950408889634f, (const float2*)rope, prefix, n_tiles, n_last, head_dim / 8, g_fa_trace ? 1 : 0, (float2*)stats_out);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("fa64_umma launch: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return TCD_OK;
}

}  // namespace tcd
