// Host side of the tcgen05 implicit-GEMM conv engine: tensor-map construction, tiling
// choice, launch.  See conv_umma.cuh for the kernel and the design notes.
#define TCD_CONV_KERNEL_IMPL
#include "conv_umma.cuh"
#include "conv_pair.cuh"
#include "gemm_pair.cuh"

#include <atomic>
#include <algorithm>
#include <mutex>
#include <stdlib.h>
#include <string.h>
#include <vector>

namespace tcd {

extern std::atomic<uint64_t> g_launch_count;

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  });
  return fn;
}

// TCD_NO_PDL=1 launches every conv with plain stream ordering (debug / A-B measurements)
static const int g_use_pdl = [] {
  const char* e = getenv("TCD_NO_PDL");
  if (e && e[0] == '1') return 0;
  const char* l = getenv("TCD_PDL_LATE");
  return (l && l[0] == '1') ? 2 : 1;
}();

// Tuning knobs for A/B measurements (bench.py records the environment it ran with):
//   TCD_CONV_SMEM_KB   shared-memory budget of a one-CTA-per-SM launch (default 220; the
//                      two-per-SM budget is half of it)
//   TCD_CONV_STAGES    upper bound of the operand ring depth (default kConvMaxStages)
static int env_int(const char* name, int dflt, int min_ok = 1) {
  const char* e = getenv(name);
  if (!e || !e[0]) return dflt;
  const int v = atoi(e);
  return v >= min_ok ? v : dflt;
}
static const int g_smem_kb = env_int("TCD_CONV_SMEM_KB", 220);
static const int g_max_stages = env_int("TCD_CONV_STAGES", kConvMaxStages);
//   TCD_CONV_HALO=0    disable the halo formulation of 3x3 stride-1 convs (one box per tap instead)
static const bool g_use_halo = [] {
  const char* e = getenv("TCD_CONV_HALO");
  return !(e && e[0] == '0');
}();


using ConvKernelFn = void (*)(const ConvMaps, const ConvArgs);
// (activation 0..3, halo form, f32-row output) -> kernel instantiation
static ConvKernelFn conv_kernel_variant(int act, int halo, int f32) {
  static const ConvKernelFn tbl[4][2][2] = {
      {{conv_umma_kernel<0, false, false>, conv_umma_kernel<0, false, true>},
       {conv_umma_kernel<0, true, false>, conv_umma_kernel<0, true, true>}},
      {{conv_umma_kernel<1, false, false>, conv_umma_kernel<1, false, true>},
       {conv_umma_kernel<1, true, false>, conv_umma_kernel<1, true, true>}},
      {{conv_umma_kernel<2, false, false>, conv_umma_kernel<2, false, true>},
       {conv_umma_kernel<2, true, false>, conv_umma_kernel<2, true, true>}},
      {{conv_umma_kernel<3, false, false>, conv_umma_kernel<3, false, true>},
       {conv_umma_kernel<3, true, false>, conv_umma_kernel<3, true, true>}}};
  return tbl[act & 3][halo ? 1 : 0][f32 ? 1 : 0];
}

static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
  }
  return n;
}

static CUtensorMapSwizzle swizzle_for_inner_bytes(int bytes) {
  if (bytes == 128) return CU_TENSOR_MAP_SWIZZLE_128B;
  if (bytes == 64) return CU_TENSOR_MAP_SWIZZLE_64B;
  if (bytes == 32) return CU_TENSOR_MAP_SWIZZLE_32B;
  return CU_TENSOR_MAP_SWIZZLE_NONE;
}

static int encode_map(CUtensorMap* m, void* base, int rank, const cuuint64_t* dims,
                      const cuuint64_t* strides_bytes, const cuuint32_t* box, int inner_bytes,
                      const char* what) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) {
    set_last_error("cuTensorMapEncodeTiled entry point unavailable (no CUDA driver?)");
    return TCD_ERR_CUDA;
  }
  cuuint32_t es[5] = {1, 1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, base, dims, strides_bytes,
                   box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for_inner_bytes(inner_bytes),
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled(%s) failed: CUresult %d (dims %llu,%llu,%llu box %u,%u,%u)",
                   what, (int)r, (unsigned long long)dims[0], (unsigned long long)dims[1],
                   (unsigned long long)dims[2], box[0], box[1], box[2]);
    return TCD_ERR_CUDA;
  }
  return TCD_OK;
}

uint16_t f32_to_bf16_rn(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}

void conv_pack_weights(const float* w, int cout, int cin, int k, int cout_pad, int cin_pad,
                       uint16_t* out) {
  const int taps = k * k;
  memset(out, 0, sizeof(uint16_t) * (size_t)taps * cout_pad * cin_pad);
  for (int o = 0; o < cout; ++o)
    for (int c = 0; c < cin; ++c)
      for (int t = 0; t < taps; ++t)
        out[((size_t)t * cout_pad + o) * cin_pad + c] =
            f32_to_bf16_rn(w[((size_t)o * cin + c) * taps + t]);
}


// ---------------------------------------------------------------- CTA-pair plan (conv_pair.cuh)
//   TCD_CONV_PAIR=0 disables it (A/B measurements); TCD_CONV_PAIR_MIN = smallest map side that takes it (default 40)
using PairKernelFn = void (*)(const ConvMaps, const PairArgs);
static PairKernelFn pair_kernel_variant(int act) {
  static const PairKernelFn tbl[4] = {conv_pair_kernel<0>, conv_pair_kernel<1>, conv_pair_kernel<2>, conv_pair_kernel<3>};
  return tbl[act & 3];
}

static bool pair_eligible(const ConvDesc& d, int mode) {
  if (!env_int("TCD_CONV_PAIR", 1, 0)) return false;
  if (mode != 0 || d.k != 3 || d.stride != 1 || d.cin != 128) return false;
  if (d.cout != d.cout_pad || d.cout % 128 || d.cout > kPairMaxCout) return false;
  if (d.groups > 1 && d.cout / d.groups != 128) return false;
  const int min_side = env_int("TCD_CONV_PAIR_MIN", 40);
  if (d.H < min_side || d.W < min_side || d.W % 8) return false;
  if (d.x_nstride || d.y_nstride) return false;   // dense NHWC only (token views keep the generic kernel)
  return true;
}

static int pair_plan_build(const ConvDesc& d, ConvPlan* plan) {
  ConvArgs& a = plan->args;
  a.Ho = d.H;
  a.Wo = d.W;
  a.TW = 8;
  a.TH = 16;
  a.tiles_w = cdiv(a.Wo, 8);
  a.tiles_per_img = a.tiles_w * cdiv(a.Ho, 16);
  a.taps = 9; a.kchunks = 2; a.ck = 64; a.BN = 128;
  a.cin_off = d.x_coff; a.cout_off = d.y_coff;
  a.mode = 0; a.silu = d.silu; a.out_chunk = 64; a.n_valid = d.cout;
  a.bias = d.bias; a.bias_n = d.cout;
  a.resid = (const __nv_bfloat16*)d.resid; a.r_ctot = d.r_ctot; a.r_coff = d.r_coff; a.resid_pre = d.resid_pre;
  a.r_nstride = d.r_nstride ? d.r_nstride : (long long)a.Ho * a.Wo * d.r_ctot;
  a.pdl = g_use_pdl;
  a.halo = 1; a.halo_w = 10;
  auto magic = [](int dv) -> unsigned int { return dv <= 1 ? 0u : (unsigned int)((0x100000000ULL + dv - 1) / (unsigned)dv); };
  a.magic_img = magic(a.tiles_per_img);
  a.magic_w = magic(a.tiles_w);
  a.n_tiles_n = d.cout / 128;
  plan->pair = 1;
  plan->pair_n_tiles = d.cout / 128;
  plan->pair_m_tiles = a.tiles_per_img * d.N;
  plan->pair_group_cin = d.groups > 1 ? d.cin : 0;
  a.total_tiles = plan->pair_m_tiles * plan->pair_n_tiles;
  TCD_REQUIRE((unsigned long long)(plan->pair_m_tiles + 2) * (unsigned)std::max(a.tiles_per_img, a.tiles_w) < 0x100000000ULL,
              "conv pair: %d tiles overflow the multiply-high divide", plan->pair_m_tiles);
  const int m_pairs = cdiv(plan->pair_m_tiles, 2);
  int clusters = sm_count() / 2;
  clusters -= clusters % plan->pair_n_tiles;
  if (clusters > m_pairs * plan->pair_n_tiles) clusters = m_pairs * plan->pair_n_tiles;
  TCD_REQUIRE(clusters >= plan->pair_n_tiles, "conv pair: no cluster for %d n-tiles", plan->pair_n_tiles);
  plan->grid = dim3((unsigned)(2 * clusters), 1, 1);
  plan->smem_bytes = kPairSmem;
  plan->flops = 2.0 * d.N * a.Ho * a.Wo * (double)d.cout * d.cin * 9;
  const size_t es = 2;
  {
    cuuint64_t dims[4] = {(cuuint64_t)d.x_ctot, (cuuint64_t)d.W, (cuuint64_t)d.H, (cuuint64_t)d.N};
    cuuint64_t st[3] = {(cuuint64_t)d.x_ctot * es, (cuuint64_t)d.W * d.x_ctot * es, (cuuint64_t)d.H * d.W * d.x_ctot * es};
    cuuint32_t box[4] = {64, 10, 18, 1};
    int rc = encode_map(&plan->maps.a[0], (void*)d.x, 4, dims, st, box, 128, "A/pair");
    if (rc) return rc;
    for (int i = 1; i < 4; ++i) plan->maps.a[i] = plan->maps.a[0];
  }
  {
    cuuint64_t dims[3] = {(cuuint64_t)d.cin, (cuuint64_t)d.cout_pad, 9};
    cuuint64_t st[2] = {(cuuint64_t)d.cin * es, (cuuint64_t)d.cout_pad * d.cin * es};
    cuuint32_t box[3] = {64, 64, 9};
    int rc = encode_map(&plan->maps.b, (void*)d.w, 3, dims, st, box, 128, "B/pair");
    if (rc) return rc;
  }
  {
    cuuint64_t dims[4] = {(cuuint64_t)d.y_ctot, (cuuint64_t)a.Wo, (cuuint64_t)a.Ho, (cuuint64_t)d.N};
    cuuint64_t st[3] = {(cuuint64_t)d.y_ctot * es, (cuuint64_t)a.Wo * d.y_ctot * es, (cuuint64_t)a.Ho * a.Wo * d.y_ctot * es};
    cuuint32_t box[4] = {64, 8, 16, 1};
    int rc = encode_map(&plan->maps.o, d.y, 4, dims, st, box, 128, "O/pair");
    if (rc) return rc;
  }
  int rc = ensure_dyn_smem((const void*)pair_kernel_variant(a.silu), (size_t)kPairSmem);
  if (rc) return rc;
  return TCD_OK;
}

static int pair_plan_launch(const ConvPlan& plan, cudaStream_t stream) {
  PairArgs pa;
  memset(&pa, 0, sizeof(pa));
  pa.c = plan.args;
  pa.m_tiles = plan.pair_m_tiles;
  pa.m_pairs = cdiv(plan.pair_m_tiles, 2);
  pa.n_tiles = plan.pair_n_tiles;
  pa.group_cin = plan.pair_group_cin;
  pa.n_img = plan.pair_m_tiles / plan.args.tiles_per_img;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = plan.grid;
  cfg.blockDim = dim3(kConvThreads, 1, 1);
  cfg.dynamicSmemBytes = (size_t)plan.smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = plan.args.pdl ? 2 : 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, pair_kernel_variant(plan.args.silu), plan.maps, pa);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  if (e != cudaSuccess) {
    set_last_error("conv_pair_kernel launch: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return TCD_OK;
}

// ---------------------------------------------------------------- CTA-pair GEMM plan (gemm_pair.cuh)
//   TCD_GEMM_PAIR=0 disables it (A/B measurements)
using GemmPairKernelFn = void (*)(const ConvMaps, const GemmPairArgs);
static GemmPairKernelFn gemm_pair_kernel_variant(int act) {
  static const GemmPairKernelFn tbl[5] = {gemm_pair_kernel<0>, gemm_pair_kernel<1>, gemm_pair_kernel<2>, gemm_pair_kernel<3>,
                                          gemm_pair_kernel<4>};
  return tbl[act >= 0 && act <= 4 ? act : 0];
}

static bool gemm_pair_eligible(const ConvDesc& d, int mode) {
  if (!env_int("TCD_GEMM_PAIR", 1, 0) && d.silu != 4 && !d.ln_stats && !d.stats_out) return false;
  if (d.stats_out && d.silu != 0 && d.silu != 4) return false;
  if (d.ln_stats && d.silu != 0 && d.silu != 4) return false;   // the normalising epilogue exists for the plain and gated forms
  if (mode != 0 || d.k != 1 || d.stride != 1 || d.groups != 1) return false;
  // dense views only: a 1x1 conv over a dense NHWC tensor IS a GEMM over its N*H*W pixel rows (token rows: N = H = 1)
  if (d.x_nstride || d.y_nstride || d.r_nstride) return false;
  if (d.cin % 64 || d.cin < 512 || d.cout != d.cout_pad || d.cout % 256 || d.cout > kGpMaxCout) return false;
  const int share = d.share >= 0 ? d.share : env_int("TCD_CONV_SHARE", 2, 0);
  // whole-SM plans (the layer programs): a pair owns two SMs outright.  TCD_GEMM_PAIR_SHARED=1 also lets the layers
  // of the half-SM detector pipeline take it (A/B measurement)
  if (share != 0 && !env_int("TCD_GEMM_PAIR_SHARED", 0, 0)) return false;
  const long rows = (long)d.N * d.H * d.W;
  if (rows >= (1L << 30)) return false;
  const long work = (long)cdiv(cdiv((int)rows, 128), 2) * (d.cout / 256);
  return work >= env_int("TCD_GEMM_PAIR_MIN_WORK", sm_count() / 2);
}

static int gemm_pair_plan_build(const ConvDesc& d0, ConvPlan* plan) {
  ConvDesc d = d0;            // flattened: one "image" of N*H*W pixel rows
  d.W = d0.N * d0.H * d0.W;
  d.N = 1;
  d.H = 1;
  ConvArgs& a = plan->args;
  a.Ho = 1; a.Wo = d.W; a.TW = 128; a.TH = 1;
  a.tiles_w = cdiv(d.W, 128); a.tiles_per_img = a.tiles_w;
  a.taps = 1; a.kchunks = d.cin / 64; a.ck = 64; a.BN = 256;
  a.cin_off = d.x_coff; a.cout_off = d.y_coff;
  a.mode = 0; a.silu = d.silu; a.out_chunk = 64; a.n_valid = d.cout;
  a.bias = d.bias; a.bias_n = d.cout;
  a.resid = (const __nv_bfloat16*)d.resid; a.r_ctot = d.r_ctot; a.r_coff = d.r_coff; a.resid_pre = d.resid_pre;
  a.r_nstride = (long long)d.W * d.r_ctot;
  a.pdl = g_use_pdl;
  a.ln_stats = d.ln_stats;
  a.ln_colsum = d.ln_colsum;
  a.ln_parts = d.ln_parts; a.ln_count = d.ln_count; a.ln_eps = d.ln_eps;
  a.stats_out = d.stats_out; a.stats_parts = d.stats_parts; a.stats_slot0 = d.stats_slot0;
  a.n_tiles_n = d.cout / 256;
  plan->pair = 2;
  plan->pair_n_tiles = d.cout / 256;
  plan->pair_m_tiles = a.tiles_w;
  a.total_tiles = plan->pair_m_tiles * plan->pair_n_tiles;
  const int work = cdiv(plan->pair_m_tiles, 2) * plan->pair_n_tiles;
  const int clusters = std::min(sm_count() / 2, work);
  plan->grid = dim3((unsigned)(2 * clusters), 1, 1);
  plan->smem_bytes = kGpSmem;
  plan->flops = 2.0 * d.W * (double)d.cout * d.cin;
  const size_t es = 2;
  {
    cuuint64_t dims[4] = {(cuuint64_t)d.x_ctot, (cuuint64_t)d.W, 1, 1};
    cuuint64_t st[3] = {(cuuint64_t)d.x_ctot * es, (cuuint64_t)d.W * d.x_ctot * es, (cuuint64_t)d.W * d.x_ctot * es};
    cuuint32_t box[4] = {64, 128, 1, 1};
    int rc = encode_map(&plan->maps.a[0], (void*)d.x, 4, dims, st, box, 128, "A/gemm_pair");
    if (rc) return rc;
    for (int i = 1; i < 4; ++i) plan->maps.a[i] = plan->maps.a[0];
  }
  {
    cuuint64_t dims[3] = {(cuuint64_t)d.cin, (cuuint64_t)d.cout_pad, 1};
    cuuint64_t st[2] = {(cuuint64_t)d.cin * es, (cuuint64_t)d.cout_pad * d.cin * es};
    cuuint32_t box[3] = {64, 128, 1};
    int rc = encode_map(&plan->maps.b, (void*)d.w, 3, dims, st, box, 128, "B/gemm_pair");
    if (rc) return rc;
  }
  {
    cuuint64_t dims[4] = {(cuuint64_t)d.y_ctot, (cuuint64_t)d.W, 1, 1};
    cuuint64_t st[3] = {(cuuint64_t)d.y_ctot * es, (cuuint64_t)d.W * d.y_ctot * es, (cuuint64_t)d.W * d.y_ctot * es};
    cuuint32_t box[4] = {64, 128, 1, 1};
    int rc = encode_map(&plan->maps.o, d.y, 4, dims, st, box, 128, "O/gemm_pair");
    if (rc) return rc;
  }
  if (env_int("TCD_CONV_PLAN_DUMP", 0))
    fprintf(stderr, "convplan gemm_pair rows%d cin%d cout%d | 256x256 per CTA pair, %d stages, grid %u\n", d.W, d.cin, d.cout,
            kGpStages, plan->grid.x);
  return ensure_dyn_smem((const void*)gemm_pair_kernel_variant(a.silu), (size_t)kGpSmem);
}

static int gemm_pair_plan_launch(const ConvPlan& plan, cudaStream_t stream) {
  GemmPairArgs ga;
  memset(&ga, 0, sizeof(ga));
  ga.c = plan.args;
  ga.m_tiles = plan.pair_m_tiles;
  ga.m_pairs = cdiv(plan.pair_m_tiles, 2);
  ga.n_tiles = plan.pair_n_tiles;
  ga.kchunks = plan.args.kchunks;
  ga.flags = env_int("TCD_GEMM_PAIR_FLAGS", 7, 0);
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = plan.grid;
  cfg.blockDim = dim3(kGpThreads, 1, 1);
  cfg.dynamicSmemBytes = (size_t)plan.smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = plan.args.pdl ? 2 : 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_pair_kernel_variant(plan.args.silu), plan.maps, ga);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  if (e != cudaSuccess) {
    set_last_error("gemm_pair_kernel launch: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return TCD_OK;
}

static void choose_tile(int Ho, int Wo, int* TW, int* TH) {
  if (Wo % 16 == 0 && Ho % 8 == 0) {
    *TW = 16;
    *TH = 8;
    return;
  }
  long best = -1;
  int btw = 8, bth = 16;
  for (int tw = 4; tw <= Wo && tw <= 128; ++tw) {
    int th = 128 / tw;
    if (th > Ho) th = Ho;
    if (th < 1) continue;
    long tiles = (long)cdiv(Wo, tw) * cdiv(Ho, th);
    // fewer tiles first, then prefer wider rows (longer contiguous runs per pixel row)
    long score = tiles * 1000 - tw;
    if (best < 0 || score < best) {
      best = score;
      btw = tw;
      bth = th;
    }
  }
  *TW = btw;
  *TH = bth;
}

int conv_plan_build(const ConvDesc& d, ConvPlan* plan) {
  memset(plan, 0, sizeof(*plan));
  TCD_REQUIRE(d.k == 1 || d.k == 3, "conv: kernel size %d unsupported", d.k);
  TCD_REQUIRE(d.stride == 1 || d.stride == 2, "conv: stride %d unsupported", d.stride);
  TCD_REQUIRE(d.cin % 16 == 0 && d.cin >= 16, "conv: cin %d must be a multiple of 16", d.cin);
  TCD_REQUIRE(d.stride == 1 || (d.H % 2 == 0 && d.W % 2 == 0), "conv: stride 2 needs even H,W");
  TCD_REQUIRE(d.stride == 1 || d.k == 3, "conv: stride 2 only for 3x3");
  const int mode = d.out_f32 ? 1 : 0;
  TCD_REQUIRE(d.groups >= 1 && d.cout % d.groups == 0, "conv: %d groups do not divide cout %d", d.groups, d.cout);
  TCD_REQUIRE(!(d.silu == 1 && d.resid && d.resid_pre), "conv: SiLU layers add their residual after the activation");
  TCD_REQUIRE(!(d.silu == 4 && d.resid), "conv: the gated epilogue takes no residual");
  if (pair_eligible(d, mode)) return pair_plan_build(d, plan);
  if (gemm_pair_eligible(d, mode)) return gemm_pair_plan_build(d, plan);
  TCD_REQUIRE(d.ln_stats == nullptr && d.stats_out == nullptr,
              "conv: the normalising / statistics epilogues exist for CTA-pair GEMM shapes only");
  TCD_REQUIRE(d.silu != 4, "conv: the gated (SwiGLU) epilogue exists for CTA-pair GEMM shapes only (1x1, dense rows, Cin >= 512 in "
                           "64s, Cout in 256s, >= %d tile pairs)", sm_count() / 2);
  TCD_REQUIRE(d.silu >= 0 && d.silu <= 3, "conv: activation %d", d.silu);
  ConvArgs& a = plan->args;
  a.Ho = d.H / d.stride;
  a.Wo = d.W / d.stride;
  choose_tile(a.Ho, a.Wo, &a.TW, &a.TH);
  a.tiles_w = cdiv(a.Wo, a.TW);
  int tiles_h = cdiv(a.Ho, a.TH);
  // halo formulation (3x3, stride 1): an 8-wide x 16-tall output patch makes every 8-row core-matrix
  // group one image line, so a tap shift is a constant row offset into the (TH+2)x(TW+2) halo box and
  // the group stride (SBO) is one halo line.  Taken unless the fixed patch shape wastes > 25 % tiles.
  a.halo = 0;
  const int ck_probe = d.cin >= 64 ? 64 : d.cin;
  // Measured (profiles/r01_conv_timeline.md): the halo form wins where all nine weight taps fit one
  // <= 32 KiB stage (stem 73 -> 39 us, dark2 bottleneck 21 -> 14.5 us: one barrier round trip per
  // k-chunk instead of nine); for the 128-channel layers its 47 KiB of halo buffers cost the second
  // resident CTA per SM and it loses (head.conv_a.0 29 -> 57 us), so those keep one box per tap.
  const bool halo_pays = 9 * (d.cout_pad >= 128 ? 128 : d.cout_pad) * ck_probe * 2 <= 32768 ||
                         env_int("TCD_CONV_HALO", 1) == 2;
  if (g_use_halo && halo_pays && d.k == 3 && d.stride == 1) {
    const long t_halo = (long)cdiv(a.Wo, 8) * cdiv(a.Ho, 16);
    if (4 * t_halo <= 5 * (long)a.tiles_w * tiles_h) {
      a.halo = 1;
      a.TW = 8;
      a.TH = 16;
      a.tiles_w = cdiv(a.Wo, 8);
      tiles_h = cdiv(a.Ho, 16);
    }
  }
  a.halo_w = a.TW + 2;
  a.tiles_per_img = a.tiles_w * tiles_h;
  a.taps = d.k * d.k;
  a.ck = d.cin >= 64 ? 64 : d.cin;
  TCD_REQUIRE(d.cin % a.ck == 0 && (a.ck == 16 || a.ck == 32 || a.ck == 64),
              "conv: cin %d not chunkable", d.cin);
  a.kchunks = d.cin / a.ck;
  // N tile: the widest one that still yields enough tiles to occupy the chip
  const int n_sm = sm_count();
  const int m_tiles = a.tiles_per_img * d.N;
  // TCD_CONV_SHARE=n (default 2, 0 = off): every launch is planned for 1/n of an SM (smem, TMEM, registers) with
  // one CTA per SM, so the next layer's prologue (PDL) and the layers of the other pipeline slots are co-resident
  // with it and fill its latency bubbles (measured: 8 940 -> 11 304 frames/s at n = 2)
  int share = d.share >= 0 ? d.share : env_int("TCD_CONV_SHARE", 2, 0);
  // layers with very many tiles per CTA are paced by the per-tile epilogue chain of ONE CTA: they take whole SMs
  // (two CTAs each) even when everything else is planned for half an SM
  {
    const int thr = env_int("TCD_CONV_SHARE_MAX_TILES", 20, 0);  // measured: 20 (the stem only) +1.8 %, 8 (dark2 too) -1.4 %
    if (thr > 0 && share > 0 && (long)m_tiles >= (long)thr * n_sm) share = 0;
    // long-K layers are paced by operand delivery with the two ring stages half an SM leaves them: TCD_CONV_WHOLE_K = k
    // gives every layer with Cin * taps >= k a whole SM (deep ring) instead (0 = off)
    const int wk = env_int("TCD_CONV_WHOLE_K", 0, 0);
    if (wk > 0 && share > 0 && d.cin * d.k * d.k >= wk) share = 0;
  }
  const int bn_cap = share >= 3 ? 64 : 128;
  int BN;
  if (mode == 1) {
    BN = d.cout_pad;
    TCD_REQUIRE(BN % 16 == 0 && BN <= 256, "conv: pred cout_pad %d", BN);
  } else {
    TCD_REQUIRE(d.cout_pad == d.cout && d.cout % 32 == 0, "conv: cout %d must be a multiple of 32",
                d.cout);
    BN = 32;
    for (int cand = bn_cap; cand >= 32; cand >>= 1) {
      if (d.cout % cand) continue;
      if ((long)m_tiles * (d.cout / cand) >= (long)(n_sm * 6) / 10 || cand == 32) {
        BN = cand;
        break;
      }
    }
  }
  // GEMM-shaped layers (1x1, long K, many tiles, whole-SM plans: the token linears of the layer programs): these are
  // bound by L2 -> SM operand delivery (64 FLOP per byte for a 128 x 128 x 64 stage), and a 128 x 256 tile reads A
  // once for twice the columns (87 FLOP per byte).  Its two accumulators fill the 512 TMEM columns: one CTA per SM.
  if (mode == 0 && d.groups == 1 && share == 0 && d.k == 1 && d.cin >= 512 && d.cout % 256 == 0 &&
      (long)m_tiles * (d.cout / 256) >= n_sm && env_int("TCD_CONV_BN256", 1, 0))
    BN = 256;
  if (d.groups > 1) {   // one N tile per group: the tile index selects the group's input channels
    BN = d.cout / d.groups;
    TCD_REQUIRE(mode == 0 && BN % 32 == 0 && BN <= bn_cap && (BN & (BN - 1)) == 0,
                "conv: grouped conv needs a bf16 output and 32/64/128 channels per group (got %d)", BN);
  }
  a.BN = BN;
  a.group_cin = (d.groups > 1 && d.cout / d.groups == BN) ? d.cin : 0;
  TCD_REQUIRE(d.groups == 1 || a.group_cin > 0, "conv: grouped conv needs one N tile per group (cout/groups %d, BN %d)",
              d.cout / d.groups, BN);
  a.n_tiles_n = d.cout_pad / BN;
  a.total_tiles = m_tiles * a.n_tiles_n;
  {
    auto magic = [](int dv) -> unsigned int { return dv <= 1 ? 0u : (unsigned int)((0x100000000ULL + dv - 1) / (unsigned)dv); };
    a.magic_n = magic(a.n_tiles_n);
    a.magic_img = magic(a.tiles_per_img);
    a.magic_w = magic(a.tiles_w);
    const unsigned long long dmax = (unsigned long long)std::max(a.n_tiles_n, std::max(a.tiles_per_img, a.tiles_w));
    TCD_REQUIRE((unsigned long long)a.total_tiles * dmax < 0x100000000ULL, "conv: %d tiles overflow the multiply-high divide",
                a.total_tiles);
  }
  a.cin_off = d.x_coff;
  a.cout_off = d.y_coff;
  a.mode = mode;
  a.silu = d.silu;
  a.out_chunk = BN >= 64 ? 64 : 32;
  a.n_valid = d.cout;
  a.bias = d.bias;
  a.resid = (const __nv_bfloat16*)d.resid;
  a.r_ctot = d.r_ctot;
  a.r_coff = d.r_coff;
  a.resid_pre = d.resid_pre;
  a.r_nstride = d.r_nstride ? d.r_nstride : (long long)a.Ho * a.Wo * d.r_ctot;
  a.out_f32 = d.out_f32;
  a.out_ld = d.out_ld;
  a.out_col = d.out_col;
  a.rows_per_img = d.rows_per_img;
  a.row_off = d.row_off;
  a.emit_thr = mode == 1 ? d.emit_thr : nullptr;
  a.maybe_count = d.maybe_count;
  a.maybe_idx = d.maybe_idx;
  a.maybe_cap = d.maybe_cap;
  a.emit_col = d.emit_col;
  a.pdl = g_use_pdl;
  a.dbg = env_int("TCD_CONV_DBG", 0);
  // accumulator rounds: narrow tiles (BN 32 / 64, one N tile) are grouped so that a round holds up to 128
  // columns -- the per-round epilogue chain costs ~2 000 cycles whatever the tile holds
  a.group = 1;
  if (mode == 0 && a.n_tiles_n == 1 && BN <= 64 && BN % 32 == 0 && env_int("TCD_CONV_GROUP", 1, 0)) {
    const long ctas_est = (long)n_sm * ((share || a.total_tiles < 2 * n_sm) ? 1 : 2);
    const long tiles_per_cta_est = cdiv(a.total_tiles, ctas_est);
    while (a.group * 2 * BN <= 128 && a.group * 2 <= tiles_per_cta_est) a.group *= 2;
  }
  int tc = 32;
  while (tc < BN * a.group) tc <<= 1;
  a.acc_stride = tc;        // two accumulators, tc columns apart
  a.tmem_cols = 2 * tc;
  // taps
  for (int t = 0; t < a.taps; ++t) {
    const int r = t / d.k, s = t % d.k;
    if (d.k == 1) {
      a.tap_map[t] = 0; a.tap_dx[t] = 0; a.tap_dy[t] = 0;
    } else if (d.stride == 1) {
      a.tap_map[t] = 0; a.tap_dx[t] = (int8_t)(s - 1); a.tap_dy[t] = (int8_t)(r - 1);
    } else {
      // in = 2*out + tap - 1  ->  parity hp = (r != 1), half-res shift = (r == 0 ? -1 : 0)
      const int hp = (r != 1), wp = (s != 1);
      a.tap_map[t] = (int8_t)(hp * 2 + wp);
      a.tap_dx[t] = (int8_t)(s == 0 ? -1 : 0);
      a.tap_dy[t] = (int8_t)(r == 0 ? -1 : 0);
    }
  }
  // pipeline depth / smem: one persistent CTA per SM gets a deep ring; layers with many more
  // tiles than SMs and short K loops run two CTAs per SM (twice the epilogue warps)
  const int a_bytes = a.halo ? 0 : 128 * a.ck * 2, b_bytes = BN * a.ck * 2;
  // halo mode: a weight stage carries as many filter taps as fit 32 KiB (fewer barrier round trips
  // per tile -- the small-channel layers are otherwise paced by them, not by bytes)
  a.tps = 1;
  if (a.halo) a.tps = 9 * b_bytes <= 32768 ? 9 : (3 * b_bytes <= 32768 ? 3 : 1);
  const int stage_bytes = a_bytes + b_bytes * a.tps;
  TCD_REQUIRE(stage_bytes % 1024 == 0, "conv: stage bytes %d not 1 KiB aligned", stage_bytes);
  const int staging = mode == 0 ? a.group * 128 * BN * 2 : 128 * (BN + 1) * 4;
  const int its_per_tile = a.taps * a.kchunks / a.tps;
  const int share_div = share ? share : 2;
  bool two_per_sm = (a.total_tiles >= 2 * n_sm || share) && a.tmem_cols <= 512 / share_div;
  const int ctas_per_sm_max = share ? 1 : 2;
  const int bias_bytes = (d.cout_pad * 4 + 127) & ~127;
  int stages = 0, halo_bytes = 0;
  long its_per_cta = 0;
  for (int attempt = 0; attempt < 2; ++attempt) {
    // static smem (barriers, < 0.5 KiB), the alignment slack and the 1 KiB the SM reserves per CTA -- two CTAs must fit
    // 228 KiB -- and the layer's bias behind the staging area
    const int budget = (two_per_sm ? g_smem_kb / share_div : g_smem_kb) * 1024 - 2560 - bias_bytes;
    const long tiles_per_cta = cdiv(a.total_tiles, n_sm * (two_per_sm ? ctas_per_sm_max : 1));
    its_per_cta = (long)its_per_tile * tiles_per_cta;
    halo_bytes = 0;
    if (a.halo) {
      a.a_stage_bytes = (((a.TW + 2) * (a.TH + 2) * a.ck * 2) + 1023) & ~1023;
      const long want = (long)a.kchunks * tiles_per_cta;
      // small halo boxes (the 16-channel stem: 180 rows of 32 bytes) are paced by the TMA round trip, not by bytes:
      // per-tile stamps showed one box per ~3 400 cycles with three in flight (TCD_CONV_ASTAGES, default 6 / 3)
      const int a_cap = a.a_stage_bytes <= 8192 ? env_int("TCD_CONV_ASTAGES", 6) : 3;
      a.a_stages = (int)(want < a_cap ? want : a_cap);
      if (a.a_stages > kConvMaxAStages) a.a_stages = kConvMaxAStages;
      // a weight ring of >= 4 stages matters more than a third halo buffer
      while (a.a_stages > 2 && (budget - staging - a.a_stages * a.a_stage_bytes) / stage_bytes < 4) --a.a_stages;
      halo_bytes = a.a_stages * a.a_stage_bytes;
    }
    stages = (budget - staging - halo_bytes) / stage_bytes;
    if (stages >= (a.halo ? 4 : 2) || !two_per_sm || (long)stages >= its_per_cta) break;
    if (share && stages >= 1) break;
    two_per_sm = false;  // the half-SM budget starves the ring: one CTA per SM with a deep ring instead
  }
  if (stages > kConvMaxStages) stages = kConvMaxStages;
  if (stages > g_max_stages) stages = g_max_stages;
  if ((long)stages > its_per_cta) stages = (int)its_per_cta;
  TCD_REQUIRE(stages >= 1, "conv: no room for a pipeline stage (%d B stage, %d B staging, %d B halo)",
              stage_bytes, staging, halo_bytes);
  a.stages = stages;
  // halo form, one N tile: when a tile's weight stages all fit the ring they are loaded once per CTA -- the TMA
  // engine is paced by box ROWS (32-byte rows for the 16-channel stem), not bytes, and half of the stem's rows
  // were the same 9 KiB of weights re-fetched for each of its 43 tiles per CTA
  a.b_resident = (a.halo && a.n_tiles_n == 1 && its_per_tile <= stages && env_int("TCD_CONV_BRES", 1, 0)) ? 1 : 0;
  a.a_ring_off = stages * stage_bytes;
  a.staging_off = a.a_ring_off + halo_bytes;
  TCD_REQUIRE(d.cout_pad <= kConvMaxBias, "conv: cout_pad %d exceeds the %d staged bias entries", d.cout_pad, kConvMaxBias);
  a.bias_n = d.cout_pad;
  a.bias_off = (a.staging_off + staging + 15) & ~15;
  plan->smem_bytes = a.bias_off + bias_bytes + 1024;
  const int ctas = n_sm * (two_per_sm ? ctas_per_sm_max : 1);
  plan->grid = dim3((unsigned)(a.total_tiles < ctas ? a.total_tiles : ctas), 1, 1);
  plan->flops = 2.0 * d.N * a.Ho * a.Wo * (double)d.cout * d.cin * a.taps;

  if (env_int("TCD_CONV_PLAN_DUMP", 0))
    fprintf(stderr, "convplan N%d %dx%d cin%d cout%d k%d s%d f32:%d | halo%d BN%d group%d stages%d(%dB) a_stages%d(%dB) tps%d bres%d "
            "staging%d smem%d grid%u share%d\n", d.N, d.H, d.W, d.cin, d.cout, d.k, d.stride, mode, a.halo, BN, a.group, stages,
            stage_bytes, a.a_stages, a.a_stage_bytes, a.tps, a.b_resident, staging, plan->smem_bytes, plan->grid.x, share);
  // ---- tensor maps
  const int inner = a.ck * 2;
  const size_t es = 2;
  if (d.stride == 1) {
    cuuint64_t dims[4] = {(cuuint64_t)d.x_ctot, (cuuint64_t)d.W, (cuuint64_t)d.H, (cuuint64_t)d.N};
    cuuint64_t st[3] = {(cuuint64_t)d.x_ctot * es, (cuuint64_t)d.W * d.x_ctot * es,
                        (cuuint64_t)(d.x_nstride ? d.x_nstride : (long long)d.H * d.W * d.x_ctot) * es};
    cuuint32_t box[4] = {(cuuint32_t)a.ck, (cuuint32_t)(a.halo ? a.TW + 2 : a.TW),
                         (cuuint32_t)(a.halo ? a.TH + 2 : a.TH), 1};
    int rc = encode_map(&plan->maps.a[0], (void*)d.x, 4, dims, st, box, inner, "A");
    if (rc) return rc;
    for (int i = 1; i < 4; ++i) plan->maps.a[i] = plan->maps.a[0];
  } else {
    for (int hp = 0; hp < 2; ++hp)
      for (int wp = 0; wp < 2; ++wp) {
        uint8_t* base = (uint8_t*)d.x + ((size_t)hp * d.W + wp) * d.x_ctot * es;
        cuuint64_t dims[4] = {(cuuint64_t)d.x_ctot, (cuuint64_t)d.W / 2, (cuuint64_t)d.H / 2,
                              (cuuint64_t)d.N};
        cuuint64_t st[3] = {(cuuint64_t)2 * d.x_ctot * es, (cuuint64_t)2 * d.W * d.x_ctot * es,
                            (cuuint64_t)(d.x_nstride ? d.x_nstride : (long long)d.H * d.W * d.x_ctot) * es};
        cuuint32_t box[4] = {(cuuint32_t)a.ck, (cuuint32_t)a.TW, (cuuint32_t)a.TH, 1};
        int rc = encode_map(&plan->maps.a[hp * 2 + wp], base, 4, dims, st, box, inner, "A/s2");
        if (rc) return rc;
      }
  }
  {
    cuuint64_t dims[3] = {(cuuint64_t)d.cin, (cuuint64_t)d.cout_pad, (cuuint64_t)a.taps};
    cuuint64_t st[2] = {(cuuint64_t)d.cin * es, (cuuint64_t)d.cout_pad * d.cin * es};
    cuuint32_t box[3] = {(cuuint32_t)a.ck, (cuuint32_t)BN, (cuuint32_t)a.tps};
    int rc = encode_map(&plan->maps.b, (void*)d.w, 3, dims, st, box, inner, "B");
    if (rc) return rc;
  }
  if (mode == 0) {
    cuuint64_t dims[4] = {(cuuint64_t)d.y_ctot, (cuuint64_t)a.Wo, (cuuint64_t)a.Ho, (cuuint64_t)d.N};
    cuuint64_t st[3] = {(cuuint64_t)d.y_ctot * es, (cuuint64_t)a.Wo * d.y_ctot * es,
                        (cuuint64_t)(d.y_nstride ? d.y_nstride : (long long)a.Ho * a.Wo * d.y_ctot) * es};
    cuuint32_t box[4] = {(cuuint32_t)a.out_chunk, (cuuint32_t)a.TW, (cuuint32_t)a.TH, 1};
    int rc = encode_map(&plan->maps.o, d.y, 4, dims, st, box, a.out_chunk * 2, "O");
    if (rc) return rc;
  } else {
    plan->maps.o = plan->maps.b;
  }
  // per device, per instantiation (the plan is built on the device that will run it)
  return ensure_dyn_smem((const void*)conv_kernel_variant(a.silu, a.halo, a.mode), (size_t)plan->smem_bytes);
}

int conv_plan_launch(const ConvPlan& plan, cudaStream_t stream) {
  if (plan.pair == 2) return gemm_pair_plan_launch(plan, stream);
  if (plan.pair) return pair_plan_launch(plan, stream);
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = plan.grid;
  cfg.blockDim = dim3(kConvThreads, 1, 1);
  cfg.dynamicSmemBytes = (size_t)plan.smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = plan.args.pdl ? 1 : 0;
  cudaError_t e = cudaLaunchKernelEx(&cfg, conv_kernel_variant(plan.args.silu, plan.args.halo, plan.args.mode),
                                     plan.maps, plan.args);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  if (e != cudaSuccess) {
    set_last_error("conv_umma_kernel launch: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return TCD_OK;
}

}  // namespace tcd
