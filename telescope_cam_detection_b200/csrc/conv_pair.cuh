// 3x3 stride-1 convolution, 128 input channels, on CTA PAIRS: tcgen05.mma.cta_group::2 (M = 256 output pixels,
// N = 128 output channels) with the layer's weights RESIDENT in shared memory.
//
// Why (tools/conv_model.py): the one-box-per-tap form moves 128x64 A + 128x64 B operands per four
// MMAs = 64 FLOP per byte pulled from L2, and L2 -> SM delivery (~43 B/clk/SM with every SM pulling) is what the
// 128-channel 3x3 layers of the head / dark4 / PAFPN were paced by (750 cycles per stage against 304 of MMA).
// Here, per output tile of 128 pixels x 128 channels, the only traffic is the (16+2) x (8+2) x 128-channel input
// halo (46 KB for 37.7 MFLOP = 820 FLOP/B):
//   * the 3x3x128x128 filter (295 KB bf16) does not fit one SM, but a CTA pair splits it by output channel:
//     each CTA keeps the 64 rows of B it contributes to the pair's N = 128 (147 KB), loaded ONCE per launch;
//   * A is the halo form of conv_umma.cuh: one TMA box per 64-channel chunk serves all nine taps through
//     row-shifted UMMA descriptors; each CTA loads the halo of its own 8 x 16 pixel patch;
//   * the leader CTA's single MMA thread issues 2 x 9 x 4 cta_group::2 MMAs per tile pair; accumulators
//     (128 lanes x 128 columns per CTA, double buffered) stay in each CTA's own TMEM and each CTA runs its own
//     epilogue (bias, SiLU, residual, bf16, swizzled staging, TMA store into a channel slice).
// Barriers: TMA loads of BOTH CTAs complete_tx on the LEADER's full barriers (cp.async.bulk.tensor ...
// .cta_group::2, barrier address mapped into the leader with mapa); tcgen05.commit.cta_group::2 multicasts the
// "stage free" and "accumulator ready" arrivals to both CTAs; the peer's epilogue warps hand the accumulator back
// with a remote mbarrier.arrive on the leader.
#pragma once
#include "conv_umma.cuh"

namespace tcd {

constexpr int kPairAStages = 2;
constexpr int kPairBBlock = 64 * 64 * 2;                       // one tap x one 64-channel chunk of this CTA's 64 rows
constexpr int kPairBBytes = 18 * kPairBBlock;                  // 147 456
constexpr int kPairABytes = 10 * 18 * 64 * 2;                  // 23 040: (8+2) x (16+2) pixels x 64 channels
constexpr int kPairAStride = (kPairABytes + 1023) & ~1023;     // 23 552
constexpr int kPairStaging = 128 * 128 * 2;                    // 32 768
constexpr int kPairMaxCout = 512;
constexpr int kPairSmem = kPairBBytes + kPairAStages * kPairAStride + kPairStaging + 1024;

struct PairArgs {
  ConvArgs c;        // geometry / epilogue fields shared with conv_umma_kernel (TW = 8, TH = 16, BN = 128, out_chunk = 64)
  int m_tiles;       // N * tiles_per_img output patches
  int m_pairs;       // ceil(m_tiles / 2): what one cluster processes per round
  int n_tiles;       // Cout / 128
  int group_cin;     // grouped conv: n-tile t reads input channels [cin_off + t * group_cin, +128); else 0
  int n_img;         // images (an out-of-range image index makes a tile a no-op: TMA zero-fills and clips)
};

#if defined(__CUDACC__) && defined(TCD_CONV_KERNEL_IMPL)

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local` (a shared::cta address of THIS CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA loads whose completion bytes land on a barrier that may live in the peer CTA of the pair
__device__ __forceinline__ void tma2_load_3d(void* dst, const CUtensorMap* m, uint32_t bar_cluster, int c0, int c1,
                                             int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(smem_u32(dst)),
      "l"(m), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma2_load_4d(void* dst, const CUtensorMap* m, uint32_t bar_cluster, int c0, int c1,
                                             int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(dst)),
      "l"(m), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
               "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem of both CTAs] (+)= A[smem of both CTAs: 128 rows each] * B[smem of both CTAs: N/2 rows each]^T
__device__ __forceinline__ void umma2_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive (once the MMAs issued so far have completed) on the barrier at this offset in every CTA of `mask`
__device__ __forceinline__ void umma2_commit_mc(uint64_t* bar, uint16_t mask) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(mask)
      : "memory");
}

// patch `m` (0 .. m_tiles-1, or beyond: a no-op tile) -> image / origin; n-tile fixed per cluster
__device__ __forceinline__ ConvTile pair_tile(const PairArgs& pa, int m, int n_tile) {
  const ConvArgs& p = pa.c;
  ConvTile c;
  c.n_tile = n_tile;
  if (m >= pa.m_tiles) {
    c.n_img = pa.n_img;  // outside the tensor: loads zero-fill, stores are clipped
    c.y0 = 0;
    c.x0 = 0;
    return c;
  }
  c.n_img = p.tiles_per_img == 1 ? m : (int)__umulhi((unsigned)m, p.magic_img);
  const int rem = m - c.n_img * p.tiles_per_img;
  const int ty = p.tiles_w == 1 ? rem : (int)__umulhi((unsigned)rem, p.magic_w);
  c.y0 = ty * p.TH;
  c.x0 = (rem - ty * p.tiles_w) * p.TW;
  return c;
}

template <int ACT>
__global__ void __launch_bounds__(kConvThreads, 1)
conv_pair_kernel(const __grid_constant__ ConvMaps maps, const __grid_constant__ PairArgs pa) {
  const ConvArgs& p = pa.c;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t b_full_bar;
  __shared__ uint64_t a_full_bar[kPairAStages];
  __shared__ uint64_t a_empty_bar[kPairAStages];
  __shared__ uint64_t tmem_full_bar[2];
  __shared__ uint64_t tmem_empty_bar[2];
  __shared__ __align__(16) float s_bias[kPairMaxCout];
  __shared__ uint32_t tmem_base_s;

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* smem_b = smem;
  uint8_t* smem_a = smem + kPairBBytes;
  uint8_t* staging = smem_a + kPairAStages * kPairAStride;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();           // 0 = leader (issues the MMAs, owns the full barriers)
  const int n_clusters = (int)(gridDim.x >> 1);
  const int cluster_id = (int)(blockIdx.x >> 1);
  // a cluster keeps ONE n-tile (its weights stay resident) and strides over the patch pairs
  const int n_tile = cluster_id % pa.n_tiles;
  const int pair0 = cluster_id / pa.n_tiles;
  const int pair_step = n_clusters / pa.n_tiles;      // planner: n_clusters is a multiple of n_tiles
  const int n_mine = pair0 < pa.m_pairs ? (pa.m_pairs - pair0 + pair_step - 1) / pair_step : 0;
  const int cin_base = p.cin_off + n_tile * pa.group_cin;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&maps.a[0]);
    prefetch_tmap(&maps.b);
    prefetch_tmap(&maps.o);
    mbar_init(&b_full_bar, 1);
    for (int s = 0; s < kPairAStages; ++s) {
      mbar_init(&a_full_bar[s], 1);       // leader: its own arrive.expect_tx (bytes of BOTH CTAs)
      mbar_init(&a_empty_bar[s], 1);      // one multicast commit
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full_bar[a], 1);                     // one multicast commit
      mbar_init(&tmem_empty_bar[a], 2 * kConvEpiWarps);    // leader: every epilogue warp of both CTAs
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc2(&tmem_base_s, 256);
  if (warp >= 2)
    for (int i = threadIdx.x - 64; i < p.bias_n; i += 32 * kConvEpiWarps)
      s_bias[i] = (ACT == 1 ? 0.5f : 1.0f) * __ldg(p.bias + i);   // SiLU layers: see conv_finish_silu
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                    // the peer's barriers are initialised before anything arrives on them
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  if (p.pdl == 1 && threadIdx.x == 0) pdl_launch_dependents();

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (one per CTA)
    if (lane == 0 && n_mine > 0) {
      // resident weights: this CTA's 64 of the n-tile's 128 rows, two boxes of (64 ch, 64 rows, 9 taps)
      const uint32_t bfull_leader = mapa_u32(smem_u32(&b_full_bar), 0);
      if (rank == 0) mbar_expect_tx(&b_full_bar, 2u * (uint32_t)kPairBBytes);
      for (int kc = 0; kc < 2; ++kc)
        tma2_load_3d(smem_b + (size_t)kc * 9 * kPairBBlock, &maps.b, bfull_leader, kc * 64,
                     n_tile * 128 + (int)rank * 64, 0);
      if (p.pdl) pdl_wait();             // the activations are the previous layer's output
      RingCursor ar{0, 0u, kPairAStages};
      for (int j = 0; j < n_mine; ++j) {
        const ConvTile c = pair_tile(pa, 2 * (pair0 + j * pair_step) + (int)rank, n_tile);
        for (int kc = 0; kc < 2; ++kc) {
          mbar_wait(&a_empty_bar[ar.s], ar.ph ^ 1u);
          if (rank == 0) mbar_expect_tx(&a_full_bar[ar.s], 2u * (uint32_t)kPairABytes);
          tma2_load_4d(smem_a + (size_t)ar.s * kPairAStride, &maps.a[0], mapa_u32(smem_u32(&a_full_bar[ar.s]), 0),
                       cin_base + kc * 64, c.x0 - 1, c.y0 - 1, c.n_img);
          ar.advance();
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA only)
    if (lane == 0 && rank == 0 && n_mine > 0) {
      const uint32_t idesc = umma_idesc_bf16(256, 128);
      const uint32_t a_base = smem_u32(smem_a), b_base = smem_u32(smem_b);
      mbar_wait(&b_full_bar, 0u);
      tc_fence_after();
      RingCursor ar{0, 0u, kPairAStages};
      for (int j = 0; j < n_mine; ++j) {
        const int acc = j & 1;
        mbar_wait(&tmem_empty_bar[acc], ((uint32_t)(j >> 1) & 1u) ^ 1u);   // both epilogues have drained it
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 128);
        for (int kc = 0; kc < 2; ++kc) {
          mbar_wait(&a_full_bar[ar.s], ar.ph);
          tc_fence_after();
          const uint32_t halo = a_base + (uint32_t)(ar.s * kPairAStride);
#pragma unroll 1
          for (int tap = 0; tap < 9; ++tap) {
            const int ty = tap / 3, tx = tap - ty * 3;
            const uint32_t a_addr = halo + (uint32_t)((ty * 10 + tx) * 128);
            const uint32_t b_addr = b_base + (uint32_t)((kc * 9 + tap) * kPairBBlock);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint64_t ad = umma_smem_desc(a_addr + k * 32, 10 * 128, 2u);   // group stride = one halo line
              const uint64_t bd = umma_smem_desc(b_addr + k * 32, 8 * 128, 2u);
              umma2_bf16(d_tmem, ad, bd, idesc, (uint32_t)((kc | tap | k) != 0));
            }
          }
          umma2_commit_mc(&a_empty_bar[ar.s], 3);     // this halo stage is free in both CTAs
          ar.advance();
        }
        umma2_commit_mc(&tmem_full_bar[acc], 3);      // accumulator ready in both CTAs
      }
    }
    if (lane == 0 && p.pdl == 2) pdl_launch_dependents();
  } else {
    // ------------------------------------------------------------------ epilogue (4 warps per CTA, own patch)
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int ly = row >> 3, lx = row & 7;
    if (p.pdl && p.resid != nullptr) pdl_wait();
    const uint32_t empty_leader0 = mapa_u32(smem_u32(&tmem_empty_bar[0]), 0);
    const uint32_t empty_leader1 = mapa_u32(smem_u32(&tmem_empty_bar[1]), 0);
    for (int j = 0; j < n_mine; ++j) {
      const int acc = j & 1;
      const ConvTile c = pair_tile(pa, 2 * (pair0 + j * pair_step) + (int)rank, n_tile);
      const int y = c.y0 + ly, x = c.x0 + lx;
      const bool valid = c.n_img < pa.n_img && y < p.Ho && x < p.Wo;
      mbar_wait(&tmem_full_bar[acc], (uint32_t)(j >> 1) & 1u);
      tc_fence_after();
      if (j > 0) {
        if (warp == 2 && lane == 0) tma_store_wait_read();    // the previous tile's stores have read the staging
        asm volatile("bar.sync 1, %0;" ::"n"(32 * kConvEpiWarps) : "memory");
      }
      const uint32_t taddr = tmem_base + (uint32_t)(acc * 128) + ((uint32_t)(q * 32) << 16);
      uint32_t ra[32], rb[32];
      tmem_ld32(taddr, ra);
      tmem_ld_wait();
      tmem_ld32(taddr + 32, rb);
      conv_finish_cols<32, 0, ACT, 0>(p, ra, 0, s_bias, row, valid, c, y, x, staging);
      tmem_ld_wait();
      tmem_ld32(taddr + 64, ra);
      conv_finish_cols<32, 0, ACT, 0>(p, rb, 32, s_bias, row, valid, c, y, x, staging);
      tmem_ld_wait();
      tmem_ld32(taddr + 96, rb);
      conv_finish_cols<32, 0, ACT, 0>(p, ra, 64, s_bias, row, valid, c, y, x, staging);
      tmem_ld_wait();
      // every column has been read: hand the accumulator back to the leader's MMA thread
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(acc ? empty_leader1 : empty_leader0);
      conv_finish_cols<32, 0, ACT, 0>(p, rb, 96, s_bias, row, valid, c, y, x, staging);
      fence_proxy_async_smem();
      asm volatile("bar.sync 1, %0;" ::"n"(32 * kConvEpiWarps) : "memory");
      if (warp == 2 && lane == 0) {
        for (int ch = 0; ch < 2; ++ch)
          tma_store_4d(&maps.o, staging + (size_t)ch * 128 * 128, p.cout_off + n_tile * 128 + ch * 64, c.x0, c.y0,
                       c.n_img);
        tma_store_commit();
      }
    }
    if (warp == 2 && lane == 0) tma_store_wait_read();
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();      // no CTA frees the pair's TMEM while its peer may still run MMAs into / read from it
  if (warp == 1) tmem_dealloc2(tmem_base, 256);
}

#endif  // __CUDACC__ && TCD_CONV_KERNEL_IMPL

}  // namespace tcd
