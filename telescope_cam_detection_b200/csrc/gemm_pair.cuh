// Token linears (1x1 convs over flat token rows, long K) on CTA PAIRS: tcgen05.mma.cta_group::2 with a
// 256 x 256 output tile per pair, both operands streamed through a shared-memory ring.
//
// Why: these GEMMs (the qkv / proj / fc1 / fc2 linears of the Stage-2 classifier body, 36 928 token rows x
// K >= 1024) are bound by L2 -> SM operand delivery, ~43 B/clk/SM with every SM pulling:
//   128 x 128 tile, one CTA :  32 KiB per k-chunk for 2.1 MFLOP =  64 FLOP/B  (750 TFLOP/s measured)
//   128 x 256 tile, one CTA :  48 KiB per k-chunk for 4.2 MFLOP =  87 FLOP/B  (877 TFLOP/s measured)
//   256 x 256 tile, CTA pair:  each CTA loads ITS 128 rows of A and ITS 128 of the tile's 256 weight rows
//                              = 32 KiB per k-chunk per SM for 4.2 MFLOP = 131 FLOP/B
// The leader CTA's MMA thread issues M = 256, N = 256 MMAs that read A from both CTAs (128 rows each) and B from
// both (N / 2 rows each); every CTA keeps its 128 x 256 fp32 accumulator (double-buffered: all 512 TMEM columns)
// and runs its own epilogue (bias, activation, residual, bf16, swizzled staging, four TMA stores).
// ACT == 4 is the gated form of the SwiGLU MLP's first Linear (fc1_g || fc1_x of timm's SwiGLU, oracle/eva02.py::_SwiGLU):
// the host interleaves the weight rows so that within a 256-column tile the 32-column slabs alternate gate / value;
// the epilogue then writes silu(gate + b_g) * (value + b_x) -- 128 columns per tile, in natural hidden order -- and
// the 2 x hidden intermediate never exists in memory.
// Barrier scheme as in conv_pair.cuh: the TMA loads of BOTH CTAs complete_tx on the LEADER's full barriers,
// tcgen05.commit.cta_group::2 multicasts "stage free" / "accumulator ready" to both CTAs, the peer's epilogue
// warps hand the accumulator back with a remote mbarrier.arrive on the leader.
#pragma once
#include "conv_pair.cuh"

namespace tcd {

constexpr int kGpABytes = 128 * 64 * 2;            // 128 token rows x 64 channels
constexpr int kGpStageBytes = 2 * kGpABytes;       // + this CTA's 128 weight rows x 64 channels
constexpr int kGpStages = 5;                       // 160 KiB in flight per SM: at 64 B/clk and ~2 500 cycles of L2 latency four were too few
constexpr int kGpStaging = 128 * 128 * 2;          // 128 rows x 128 output channels, two 64-channel chunks (drained in two phases)
constexpr int kGpMaxCout = kConvMaxBias;
constexpr int kGpEpiWarps = 8;                     // two per TMEM lane quadrant
constexpr int kGpThreads = 64 + 32 * kGpEpiWarps;   // warp 0 TMA producer, warp 1 MMA issuer, warps 2-9 epilogue
constexpr int kGpSmem = kGpStages * kGpStageBytes + kGpStaging + 1024;

struct GemmPairArgs {
  ConvArgs c;        // epilogue fields shared with conv_finish_cols (TW = 128, TH = 1, BN = 256, out_chunk = 64)
  int m_tiles;       // ceil(rows / 128)
  int m_pairs;       // ceil(m_tiles / 2)
  int n_tiles;       // Cout / 256
  int kchunks;       // Cin / 64
  int flags;         // TCD_GEMM_PAIR_FLAGS (default 7): 1 relaxed accumulator hand-back, 2 sixteen partial loads in flight, 4 residual by 32-byte loads, 8 per-half staging barriers and stores
};

#if defined(__CUDACC__) && defined(TCD_CONV_KERNEL_IMPL)

// 32 fp32 values of one row -> bf16 -> the 128-byte-swizzled staging tile of 64-channel chunks (what TMA stores read)
// Accumulator hand-back to the leader's MMA thread.  Nothing in generic memory travels through this barrier (the TMEM reads
// were completed by tcgen05.wait::ld and ordered by tcgen05.fence::before_thread_sync), so the arrive is relaxed: the
// default .release at cluster scope compiles to MEMBAR.ALL.GPU and waited for the staging stores before it.
__device__ __forceinline__ void gp_handback(uint32_t cluster_addr, int relaxed) {
  if (relaxed) asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
  else mbar_arrive_cluster(cluster_addr);
}
__device__ __forceinline__ void gp_stage_slab(const float (&v)[32], int c0, int row, uint8_t* stage_base) {
  const int chunk = c0 >> 6, u0 = (c0 & 63) >> 3, sw = row & 7;
  uint8_t* rowp = stage_base + (size_t)chunk * 128 * 128 + (size_t)row * 128;
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    uint4 q;
    q.x = pack_bf16x2(v[u * 8 + 0], v[u * 8 + 1]);
    q.y = pack_bf16x2(v[u * 8 + 2], v[u * 8 + 3]);
    q.z = pack_bf16x2(v[u * 8 + 4], v[u * 8 + 5]);
    q.w = pack_bf16x2(v[u * 8 + 6], v[u * 8 + 7]);
    *reinterpret_cast<uint4*>(rowp + (((u0 + u) ^ sw) << 4)) = q;
  }
}

template <int ACT>
__global__ void __launch_bounds__(kGpThreads, 1)
gemm_pair_kernel(const __grid_constant__ ConvMaps maps, const __grid_constant__ GemmPairArgs ga) {
  const ConvArgs& p = ga.c;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t full_bar[kGpStages];
  __shared__ uint64_t empty_bar[kGpStages];
  __shared__ uint64_t tmem_full_bar[2];
  __shared__ uint64_t tmem_empty_bar[2];
  __shared__ __align__(16) float s_bias[kGpMaxCout];
  __shared__ __align__(16) float s_cs[kGpMaxCout];     // column sums of the weights (consumer-side LayerNorm), else unused
  __shared__ uint32_t tmem_base_s;
  __shared__ float2 s_part[2][128][2];                // the two epilogue threads of a row exchange their statistics (per accumulator buffer)

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* staging = smem + kGpStages * kGpStageBytes;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();           // 0 = leader (issues the MMAs, owns the full barriers)
  const int n_clusters = (int)(gridDim.x >> 1);
  const int cluster_id = (int)(blockIdx.x >> 1);
  // work item w -> (row pair w / n_tiles, column tile w % n_tiles): neighbouring clusters share the A rows in L2
  // (Column tiles of one row pair back to back on a cluster -- so that the rows' statistics are read once -- was tried
  //  and lost: the 74 clusters then stream 74 different A tiles and the K = 2816 activations fall out of L2, fc2 +25 %.)
  const int n_work = ga.m_pairs * ga.n_tiles;
  const int n_mine = cluster_id < n_work ? (n_work - cluster_id + n_clusters - 1) / n_clusters : 0;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&maps.a[0]);
    prefetch_tmap(&maps.b);
    prefetch_tmap(&maps.o);
    for (int s = 0; s < kGpStages; ++s) {
      mbar_init(&full_bar[s], 1);        // leader: its own arrive.expect_tx (bytes of BOTH CTAs)
      mbar_init(&empty_bar[s], 1);       // one multicast commit
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full_bar[a], 1);                     // one multicast commit
      mbar_init(&tmem_empty_bar[a], 2 * kGpEpiWarps);      // leader: every epilogue warp of both CTAs
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc2(&tmem_base_s, 512);
  if (warp >= 2)
    for (int i = threadIdx.x - 64; i < p.bias_n; i += 32 * kGpEpiWarps)
    {
      s_bias[i] = (ACT == 1 ? 0.5f : 1.0f) * __ldg(p.bias + i);   // SiLU layers: see conv_finish_silu
      if (p.ln_stats != nullptr) s_cs[i] = __ldg(p.ln_colsum + i);
    }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                    // the peer's barriers are initialised before anything arrives on them
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  if (p.pdl == 1 && threadIdx.x == 0) pdl_launch_dependents();

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (one per CTA)
    if (lane == 0 && n_mine > 0) {
      bool waited = !p.pdl;
      RingCursor ring{0, 0u, kGpStages};
      for (int it = 0; it < n_mine; ++it) {
        const int w = cluster_id + it * n_clusters;
        const int n_tile = w % ga.n_tiles, m_pair = w / ga.n_tiles;
        const int x0 = (2 * m_pair + (int)rank) * 128;      // beyond the last row: loads zero-fill, stores are clipped
        for (int kc = 0; kc < ga.kchunks; ++kc) {
          mbar_wait(&empty_bar[ring.s], ring.ph ^ 1u);
          const uint32_t full_leader = mapa_u32(smem_u32(&full_bar[ring.s]), 0);
          if (rank == 0) mbar_expect_tx(&full_bar[ring.s], 2u * (uint32_t)kGpStageBytes);
          uint8_t* st = smem + (size_t)ring.s * kGpStageBytes;
          // weights do not depend on the previous layer: issue them before the dependency wait
          tma2_load_3d(st + kGpABytes, &maps.b, full_leader, kc * 64, n_tile * 256 + (int)rank * 128, 0);
          if (!waited) {
            pdl_wait();
            waited = true;
          }
          tma2_load_4d(st, &maps.a[0], full_leader, p.cin_off + kc * 64, x0, 0, 0);
          ring.advance();
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA only)
    if (lane == 0 && rank == 0 && n_mine > 0) {
      const uint32_t idesc = umma_idesc_bf16(256, 256);
      const uint32_t ring_base = smem_u32(smem);
      RingCursor ring{0, 0u, kGpStages};
      for (int it = 0; it < n_mine; ++it) {
        const int acc = it & 1;
        mbar_wait(&tmem_empty_bar[acc], ((uint32_t)(it >> 1) & 1u) ^ 1u);   // both epilogues have drained it
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
        for (int kc = 0; kc < ga.kchunks; ++kc) {
          mbar_wait(&full_bar[ring.s], ring.ph);
          tc_fence_after();
          const uint32_t a_addr = ring_base + (uint32_t)(ring.s * kGpStageBytes);
          const uint64_t ad = umma_smem_desc(a_addr, 1024u, 2u), bd = umma_smem_desc(a_addr + kGpABytes, 1024u, 2u);
#pragma unroll
          for (int k = 0; k < 4; ++k) umma2_bf16(d_tmem, ad + 2 * k, bd + 2 * k, idesc, (uint32_t)((kc | k) != 0));
          umma2_commit_mc(&empty_bar[ring.s], 3);     // this stage is free in both CTAs
          ring.advance();
        }
        umma2_commit_mc(&tmem_full_bar[acc], 3);      // accumulator ready in both CTAs
      }
    }
    if (lane == 0 && p.pdl == 2) pdl_launch_dependents();
  } else {
    // ------------------------------------------------------------------ epilogue (8 warps per CTA, own 128 rows)
    // Two warps per TMEM lane quadrant: thread = (row, sub); of the four 32-column slabs of a phase sub 0 takes the
    // first two and sub 1 the last two.  (With four epilogue warps a 128 x 256 tile took ~8 000 cycles to drain -- as
    // long as its K = 1024 MMA phase -- and 20 000+ with residual, normalisation and statistics work in it: proj ran
    // at 620 TFLOP/s.)
    const int q = warp & 3, sub = (warp - 2) >> 2;
    const int row = q * 32 + lane;
    if (p.pdl && (p.resid != nullptr || p.ln_stats != nullptr)) pdl_wait();   // both are the previous kernels' outputs
    const uint32_t empty_leader0 = mapa_u32(smem_u32(&tmem_empty_bar[0]), 0);
    const uint32_t empty_leader1 = mapa_u32(smem_u32(&tmem_empty_bar[1]), 0);
    constexpr int kPhases = ACT == 4 ? 1 : 2;          // the staging area holds 128 output columns (32 KiB)
    // flags & 8: the two halves of the epilogue (sub 0 / sub 1) fill different 64-column chunks of the staging area, so each
    // half syncs among its own four warps (named barriers 1 + sub) and stores its own chunk; they then drift apart and
    // one half's store wait overlaps the other's arithmetic.  Otherwise all eight warps share barrier 1 and warp 2 stores.
    const bool split = (ga.flags & 8) != 0;
    const int bar_id = split ? 1 + sub : 1, bar_n = split ? 32 * kGpEpiWarps / 2 : 32 * kGpEpiWarps;
    const bool issuer = lane == 0 && (split ? ((warp - 2) & 3) == 0 : warp == 2);
    for (int it = 0; it < n_mine; ++it) {
      const int acc = it & 1;
      const int w = cluster_id + it * n_clusters;
      ConvTile c;
      c.n_tile = w % ga.n_tiles;
      c.n_img = 0;
      c.y0 = 0;
      c.x0 = (2 * (w / ga.n_tiles) + (int)rank) * 128;
      const int x = c.x0 + row;
      const bool valid = x < p.Wo;
      // consumer-side LayerNorm: y = rstd * (acc - mean * colsum) + bias, (mean, rstd) of this thread's input row
      float ln_mean = 0.f, ln_rstd = 1.f;
      if (p.ln_stats != nullptr && valid) {
        if (p.ln_parts > 0) {
          // partial (sum, sum of squares) pairs written by the producers of x (fixed order: deterministic)
          // (ln_parts is even -- 4, 22, 32 in the classifier body: two partials per 16-byte load)
          const float4* pp = reinterpret_cast<const float4*>(reinterpret_cast<const float2*>(p.ln_stats) + (size_t)x * p.ln_parts);
          const int np = p.ln_parts >> 1;
          float su = 0.f, sq = 0.f;
          if (ga.flags & 2) {
            float4 t4[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) t4[i] = i < np ? __ldg(pp + i) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              su += t4[i].x + t4[i].z;
              sq += t4[i].y + t4[i].w;
            }
            for (int i = 16; i < np; ++i) {
              const float4 t = __ldg(pp + i);
              su += t.x + t.z;
              sq += t.y + t.w;
            }
          } else {
#pragma unroll 4
            for (int i = 0; i < np; ++i) {
              const float4 t = __ldg(pp + i);
              su += t.x + t.z;
              sq += t.y + t.w;
            }
          }
          const float inv_n = 1.0f / (float)p.ln_count;
          ln_mean = su * inv_n;
          ln_rstd = rsqrtf(fmaxf(fmaf(-ln_mean, ln_mean, sq * inv_n), 0.f) + p.ln_eps);
        } else {
          const float2 st2 = __ldg(reinterpret_cast<const float2*>(p.ln_stats) + x);
          ln_mean = st2.x;
          ln_rstd = st2.y;
        }
      }
      // Plain linears with a residual (proj, fc2): this thread's 128 residual columns (two slabs per phase) are requested
      // NOW, before the wait for the accumulator, sixteen 16-byte loads in flight (the residual stream comes from HBM).
      uint4 rres[ACT == 0 ? 16 : 1];
      const bool use_res = ACT == 0 && p.resid != nullptr && valid;
      if constexpr (ACT == 0) {
        if (use_res) {
          const __nv_bfloat16* rb0 = p.resid + (size_t)x * p.r_ctot + p.r_coff + c.n_tile * 256 + sub * 64;
          if ((ga.flags & 4) && (reinterpret_cast<uintptr_t>(rb0) & 31) == 0) {
            // whole 32-byte sectors per request (lanes read different rows: a 16-byte load uses half of its sector and
            // leaves the other half to an L1 that has 30 KiB next to this kernel's shared memory)
#pragma unroll
            for (int ph = 0; ph < 2; ++ph)
#pragma unroll
              for (int v2 = 0; v2 < 4; ++v2) {
                uint4& lo = rres[ph * 8 + 2 * v2];
                uint4& hi = rres[ph * 8 + 2 * v2 + 1];
                asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                             : "=r"(lo.x), "=r"(lo.y), "=r"(lo.z), "=r"(lo.w), "=r"(hi.x), "=r"(hi.y), "=r"(hi.z), "=r"(hi.w)
                             : "l"(rb0 + ph * 128 + v2 * 16));
              }
          } else {
#pragma unroll
            for (int ph = 0; ph < 2; ++ph)
#pragma unroll
              for (int u = 0; u < 8; ++u) rres[ph * 8 + u] = __ldg(reinterpret_cast<const uint4*>(rb0 + ph * 128) + u);
          }
        }
      }
      float ps_out = 0.f, pss_out = 0.f;     // (sum, sum of squares) of this thread's outputs in this tile (stats_out)
      mbar_wait(&tmem_full_bar[acc], (uint32_t)(it >> 1) & 1u);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t)(acc * 256) + ((uint32_t)(q * 32) << 16);
      uint32_t ra[32], rb[32];
      // this thread's slabs: plain form, phase ph: ph*4 + sub*2 and the next one; gated form: sub*4 .. sub*4 + 3
      const int first = ACT == 4 ? sub * 4 : sub * 2;
      tmem_ld32(taddr + 32 * first, ra);       // on its way while the staging area is still being read
      auto finish_plain = [&](const uint32_t (&r)[32], int sl, int rslot) {     // ACT == 0: (LN,) bias (+ residual) -> bf16 -> staging
        const float4* b4 = reinterpret_cast<const float4*>(s_bias + c.n_tile * 256 + 32 * sl);
        const float4* c4 = reinterpret_cast<const float4*>(s_cs + c.n_tile * 256 + 32 * sl);
        float v[32];
        if (p.ln_stats != nullptr) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b = b4[j], cs = c4[j];
            v[4 * j + 0] = fmaf(ln_rstd, fmaf(-ln_mean, cs.x, __uint_as_float(r[4 * j + 0])), b.x);
            v[4 * j + 1] = fmaf(ln_rstd, fmaf(-ln_mean, cs.y, __uint_as_float(r[4 * j + 1])), b.y);
            v[4 * j + 2] = fmaf(ln_rstd, fmaf(-ln_mean, cs.z, __uint_as_float(r[4 * j + 2])), b.z);
            v[4 * j + 3] = fmaf(ln_rstd, fmaf(-ln_mean, cs.w, __uint_as_float(r[4 * j + 3])), b.w);
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b = b4[j];
            v[4 * j + 0] = __uint_as_float(r[4 * j + 0]) + b.x;
            v[4 * j + 1] = __uint_as_float(r[4 * j + 1]) + b.y;
            v[4 * j + 2] = __uint_as_float(r[4 * j + 2]) + b.z;
            v[4 * j + 3] = __uint_as_float(r[4 * j + 3]) + b.w;
          }
        }
        if (use_res) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const uint4 qq = rres[ACT == 0 ? rslot * 4 + u : 0];
            const uint32_t ww[4] = {qq.x, qq.y, qq.z, qq.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              v[u * 8 + 2 * e] += __uint_as_float(ww[e] << 16);
              v[u * 8 + 2 * e + 1] += __uint_as_float(ww[e] & 0xffff0000u);
            }
          }
        }
        if (p.stats_out != nullptr) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            ps_out += v[i];
            pss_out = fmaf(v[i], v[i], pss_out);
          }
        }
        gp_stage_slab(v, 32 * (sl & 3), row, staging);
      };
#pragma unroll
      for (int ph = 0; ph < kPhases; ++ph) {
        if (it > 0 || ph > 0) {
          if (issuer) tma_store_wait_read();    // the previous stores have read the staging
          asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(bar_n) : "memory");
        }
        uint8_t* stg = staging - (size_t)ph * 2 * 128 * 128;     // conv_finish_cols places column c0 in chunk c0 / 64
        if constexpr (ACT == 4) {
          // two gate / value pairs per thread: slabs sub*4 + {0,1} and sub*4 + {2,3} -> hidden columns sub*64 + {0, 32} ..
#pragma unroll
          for (int pr = 0; pr < 2; ++pr) {
            const int sl = sub * 4 + pr * 2;
            tmem_ld_wait();
            tmem_ld32(taddr + 32 * (sl + 1), rb);
            tmem_ld_wait();
            const float* bg = s_bias + c.n_tile * 256 + 32 * sl;
            const float* cg = s_cs + c.n_tile * 256 + 32 * sl;
            float v[32];
            if (p.ln_stats != nullptr) {
#pragma unroll
              for (int i = 0; i < 32; ++i)
                v[i] = silu_f(fmaf(ln_rstd, fmaf(-ln_mean, cg[i], __uint_as_float(ra[i])), bg[i])) *
                       fmaf(ln_rstd, fmaf(-ln_mean, cg[32 + i], __uint_as_float(rb[i])), bg[32 + i]);
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i)
                v[i] = silu_f(__uint_as_float(ra[i]) + bg[i]) * (__uint_as_float(rb[i]) + bg[32 + i]);
            }
            if (pr == 0) {
              tmem_ld32(taddr + 32 * (sl + 2), ra);
            } else {
              // every column of this thread has been read: hand the accumulator back to the leader's MMA thread
              tc_fence_before();
              __syncwarp();
              if (lane == 0) gp_handback(acc ? empty_leader1 : empty_leader0, ga.flags & 1);
            }
            if (p.stats_out != nullptr) {
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                ps_out += v[i];
                pss_out = fmaf(v[i], v[i], pss_out);
              }
            }
            gp_stage_slab(v, 16 * sl, row, staging);
          }
        } else {
          const int sl = ph * 4 + sub * 2;
          tmem_ld_wait();
          tmem_ld32(taddr + 32 * (sl + 1), rb);
          if constexpr (ACT == 0) finish_plain(ra, sl, ph * 2);
          else conv_finish_cols<32, 0, ACT, 0>(p, ra, 32 * sl, s_bias, row, valid, c, 0, x, stg);
          tmem_ld_wait();
          if (ph + 1 < kPhases) {
            tmem_ld32(taddr + 32 * (sl + 4), ra);      // this thread's first slab of the next phase
          } else {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) gp_handback(acc ? empty_leader1 : empty_leader0, ga.flags & 1);
          }
          if constexpr (ACT == 0) finish_plain(rb, sl + 1, ph * 2 + 1);
          else conv_finish_cols<32, 0, ACT, 0>(p, rb, 32 * (sl + 1), s_bias, row, valid, c, 0, x, stg);
        }
        fence_proxy_async_smem();
        asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(bar_n) : "memory");
        if (issuer) {
          constexpr int kOutCols = ACT == 4 ? 128 : 256;     // the gated form writes half as many columns
          for (int ch = split ? sub : 0; ch < (split ? sub + 1 : 2); ++ch)
            tma_store_4d(&maps.o, staging + (size_t)ch * 128 * 128, p.cout_off + c.n_tile * kOutCols + (ph * 2 + ch) * 64, c.x0, 0, 0);
          tma_store_commit();
        }
      }
      if (p.stats_out != nullptr) {
        // the two threads of a row add their halves through shared memory; sub 0 writes the tile's partial
        s_part[acc][row][sub] = make_float2(ps_out, pss_out);
        asm volatile("bar.sync 3, %0;" ::"n"(32 * kGpEpiWarps) : "memory");
        if (sub == 0 && valid) {
          const float2 o2 = s_part[acc][row][1];
          reinterpret_cast<float2*>(p.stats_out)[(size_t)x * p.stats_parts + p.stats_slot0 + c.n_tile] =
              make_float2(ps_out + o2.x, pss_out + o2.y);
        }
      }
    }
    if (issuer) tma_store_wait_read();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();      // no CTA frees the pair's TMEM while its peer may still run MMAs into / read from it
  if (warp == 1) tmem_dealloc2(tmem_base, 512);
}

#endif  // __CUDACC__ && TCD_CONV_KERNEL_IMPL

}  // namespace tcd
