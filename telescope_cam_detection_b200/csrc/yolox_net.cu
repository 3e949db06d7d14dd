// YOLOX-S plan builder + the three non-conv ops (nearest x2 upsample into a concat slice,
// SPP max-pools).  See yolox_net.cuh.
#include "yolox_net.cuh"

#include <atomic>
#include <string.h>

namespace tcd {
extern std::atomic<uint64_t> g_launch_count;

// ------------------------------------------------------------------ weight blob
// layout (little endian), written by telescope_cam_detection_b200/weights.py:
//   char magic[8] = "TCDW0001"; u32 n_entries; u32 num_classes;
//   per entry: char name[96]; i32 cout, cin, k, reserved; f32 w[cout*cin*k*k]; f32 b[cout]
int WeightBlob::parse(const void* blob, size_t bytes) {
  const uint8_t* p = (const uint8_t*)blob;
  TCD_REQUIRE(bytes >= 16 && memcmp(p, "TCDW0001", 8) == 0, "weights: bad magic");
  uint32_t n;
  memcpy(&n, p + 8, 4);
  size_t off = 16;
  for (uint32_t i = 0; i < n; ++i) {
    TCD_REQUIRE(off + 112 <= bytes, "weights: truncated at entry %u", i);
    char name[97];
    memcpy(name, p + off, 96);
    name[96] = 0;
    int32_t hdr[4];
    memcpy(hdr, p + off + 96, 16);
    off += 112;
    WeightEntry e;
    e.cout = hdr[0];
    e.cin = hdr[1];
    e.k = hdr[2];
    TCD_REQUIRE(e.cout > 0 && e.cin > 0 && (e.k == 1 || e.k == 3), "weights: bad shape for %s", name);
    const size_t nw = (size_t)e.cout * e.cin * e.k * e.k;
    TCD_REQUIRE(off + (nw + e.cout) * 4 <= bytes, "weights: truncated data for %s", name);
    e.w = (const float*)(p + off);
    off += nw * 4;
    e.b = (const float*)(p + off);
    off += (size_t)e.cout * 4;
    entries[name] = e;
  }
  return TCD_OK;
}

static int upload_conv(const std::vector<const WeightEntry*>& parts, int cin_pad, int cout_pad,
                       DevConv* out) {
  const int k = parts[0]->k, cin = parts[0]->cin;
  int cout = 0;
  for (auto* e : parts) {
    TCD_REQUIRE(e->k == k && e->cin == cin, "weights: fused parts disagree on shape");
    cout += e->cout;
  }
  if (cout_pad < cout) cout_pad = cout;
  const int taps = k * k;
  std::vector<uint16_t> hw((size_t)taps * cout_pad * cin_pad, 0);
  std::vector<float> hb(cout_pad, 0.f);
  int o0 = 0;
  for (auto* e : parts) {
    for (int o = 0; o < e->cout; ++o) {
      hb[o0 + o] = e->b[o];
      for (int c = 0; c < cin; ++c)
        for (int t = 0; t < taps; ++t)
          hw[((size_t)t * cout_pad + o0 + o) * cin_pad + c] =
              f32_to_bf16_rn(e->w[((size_t)o * cin + c) * taps + t]);
    }
    o0 += e->cout;
  }
  TCD_CUDA_OK(cudaMalloc(&out->w, hw.size() * 2));
  TCD_CUDA_OK(cudaMalloc((void**)&out->bias, hb.size() * 4));
  TCD_CUDA_OK(cudaMemcpy(out->w, hw.data(), hw.size() * 2, cudaMemcpyHostToDevice));
  TCD_CUDA_OK(cudaMemcpy(out->bias, hb.data(), hb.size() * 4, cudaMemcpyHostToDevice));
  out->cout = cout;
  out->cout_pad = cout_pad;
  out->cin_pad = cin_pad;
  out->k = k;
  return TCD_OK;
}

int YoloxWeights::upload(const WeightBlob& blob) {
  auto get = [&](const std::string& n) -> const WeightEntry* {
    auto it = blob.entries.find(n);
    return it == blob.entries.end() ? nullptr : &it->second;
  };
  auto one = [&](const std::string& dst, const std::string& src, int cin_pad = 0) -> int {
    const WeightEntry* e = get(src);
    TCD_REQUIRE(e != nullptr, "weights: missing entry %s", src.c_str());
    return upload_conv({e}, cin_pad ? cin_pad : e->cin, 0, &convs[dst]);
  };
  auto fuse = [&](const std::string& dst, const std::string& a, const std::string& b,
                  int cout_pad = 0) -> int {
    const WeightEntry *ea = get(a), *eb = get(b);
    TCD_REQUIRE(ea && eb, "weights: missing entry %s or %s", a.c_str(), b.c_str());
    return upload_conv({ea, eb}, ea->cin, cout_pad, &convs[dst]);
  };
  int rc;
#define W_TRY(x) do { rc = (x); if (rc) return rc; } while (0)
  const std::string bb = "backbone.backbone.";
  W_TRY(one("stem", bb + "stem.conv", 16));
  auto csp = [&](const std::string& dst, const std::string& src, int n) -> int {
    int r = fuse(dst + ".conv12", src + ".conv1", src + ".conv2");
    if (r) return r;
    for (int i = 0; i < n; ++i) {
      const std::string m = ".m." + std::to_string(i);
      if ((r = one(dst + m + ".conv1", src + m + ".conv1"))) return r;
      if ((r = one(dst + m + ".conv2", src + m + ".conv2"))) return r;
    }
    return one(dst + ".conv3", src + ".conv3");
  };
  W_TRY(one("dark2.0", bb + "dark2.0"));
  W_TRY(csp("dark2.1", bb + "dark2.1", 1));
  W_TRY(one("dark3.0", bb + "dark3.0"));
  W_TRY(csp("dark3.1", bb + "dark3.1", 3));
  W_TRY(one("dark4.0", bb + "dark4.0"));
  W_TRY(csp("dark4.1", bb + "dark4.1", 3));
  W_TRY(one("dark5.0", bb + "dark5.0"));
  W_TRY(one("dark5.1.conv1", bb + "dark5.1.conv1"));
  W_TRY(one("dark5.1.conv2", bb + "dark5.1.conv2"));
  W_TRY(csp("dark5.2", bb + "dark5.2", 1));
  W_TRY(one("lateral_conv0", "backbone.lateral_conv0"));
  W_TRY(csp("C3_p4", "backbone.C3_p4", 1));
  W_TRY(one("reduce_conv1", "backbone.reduce_conv1"));
  W_TRY(csp("C3_p3", "backbone.C3_p3", 1));
  W_TRY(one("bu_conv2", "backbone.bu_conv2"));
  W_TRY(csp("C3_n3", "backbone.C3_n3", 1));
  W_TRY(one("bu_conv1", "backbone.bu_conv1"));
  W_TRY(csp("C3_n4", "backbone.C3_n4", 1));
  for (int l = 0; l < 3; ++l) {
    const std::string s = std::to_string(l);
    W_TRY(one("head.stem." + s, "head.stems." + s));
    W_TRY(fuse("head.conv_a." + s, "head.cls_convs." + s + ".0", "head.reg_convs." + s + ".0"));
    // the second 3x3 convs of the cls and reg branches as ONE grouped conv over [cls_a | reg_a] -> [cls_feat | reg_feat]
    W_TRY(fuse("head.b." + s, "head.cls_convs." + s + ".1", "head.reg_convs." + s + ".1"));
    // the three 1x1 prediction convs of a level as ONE block-diagonal GEMM over [cls_feat | reg_feat]:
    // rows 0-3 reg, 4 obj (weights on the reg_feat half), rows 5.. cls (weights on the cls_feat half), zeros
    // elsewhere.  Exact zeros add exact zeros to the fp32 accumulator, so every logit keeps its bits; the kernel
    // writes whole 85-float rows instead of a 5-float and an 80-float piece from two launches.
    {
      const WeightEntry *er = get("head.reg_preds." + s), *eo = get("head.obj_preds." + s),
                        *ec = get("head.cls_preds." + s);
      TCD_REQUIRE(er && eo && ec, "weights: missing prediction convs of level %d", l);
      TCD_REQUIRE(er->k == 1 && eo->k == 1 && ec->k == 1 && er->cin == ec->cin && eo->cin == ec->cin &&
                      er->cout == 4 && eo->cout == 1,
                  "weights: unexpected prediction conv shapes at level %d", l);
      const int hc = ec->cin, cout = 5 + ec->cout, cout_pad = (cout + 15) & ~15, cin2 = 2 * hc;
      std::vector<uint16_t> hw((size_t)cout_pad * cin2, 0);
      std::vector<float> hb(cout_pad, 0.f);
      for (int o = 0; o < 4; ++o) {
        hb[o] = er->b[o];
        for (int c = 0; c < hc; ++c) hw[(size_t)o * cin2 + hc + c] = f32_to_bf16_rn(er->w[(size_t)o * hc + c]);
      }
      hb[4] = eo->b[0];
      for (int c = 0; c < hc; ++c) hw[(size_t)4 * cin2 + hc + c] = f32_to_bf16_rn(eo->w[c]);
      for (int o = 0; o < ec->cout; ++o) {
        hb[5 + o] = ec->b[o];
        for (int c = 0; c < hc; ++c) hw[(size_t)(5 + o) * cin2 + c] = f32_to_bf16_rn(ec->w[(size_t)o * hc + c]);
      }
      DevConv* out = &convs["head.pred." + s];
      TCD_CUDA_OK(cudaMalloc(&out->w, hw.size() * 2));
      TCD_CUDA_OK(cudaMalloc((void**)&out->bias, hb.size() * 4));
      TCD_CUDA_OK(cudaMemcpy(out->w, hw.data(), hw.size() * 2, cudaMemcpyHostToDevice));
      TCD_CUDA_OK(cudaMemcpy(out->bias, hb.data(), hb.size() * 4, cudaMemcpyHostToDevice));
      out->cout = cout;
      out->cout_pad = cout_pad;
      out->cin_pad = cin2;
      out->k = 1;
    }
    const WeightEntry* cp = get("head.cls_preds." + s);
    num_classes = cp->cout;
    TCD_REQUIRE(num_classes % 16 == 0, "weights: num_classes %d must be a multiple of 16", num_classes);
  }
#undef W_TRY
  return TCD_OK;
}

void YoloxWeights::release() {
  for (auto& kv : convs) {
    if (kv.second.w) cudaFree(kv.second.w);
    if (kv.second.bias) cudaFree(kv.second.bias);
  }
  convs.clear();
}

// ------------------------------------------------------------------ aux kernels
// nearest x2 upsample of a channel slice into a channel slice (PAFPN top-down path)
__global__ void __launch_bounds__(256)
upsample2x_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, int N, int H, int W, int C8,
                  int s_ctot8, int s_coff8, int d_ctot8, int d_coff8) {
  const size_t total = (size_t)N * (2 * H) * (2 * W) * C8;
  for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
    const int c = (int)(i % C8);
    size_t r = i / C8;
    const int x = (int)(r % (2 * W));
    r /= (2 * W);
    const int y = (int)(r % (2 * H));
    const int n = (int)(r / (2 * H));
    const uint4 v = src[((size_t)((size_t)n * H + (y >> 1)) * W + (x >> 1)) * s_ctot8 + s_coff8 + c];
    dst[((size_t)((size_t)n * 2 * H + y) * 2 * W + x) * d_ctot8 + d_coff8 + c] = v;
  }
}

// SPPBottleneck pools: buf[..., 0:C] -> max 5x5 / 9x9 / 13x13 (stride 1, -inf padding) into
// buf[..., C:2C], [2C:3C], [3C:4C].  One CTA owns one image x 16 channels: the whole HxW plane
// sits in shared memory and the three pools are a cascade of separable 5-wide maxima
// (5 o 5 = 9, 5 o 5 o 5 = 13; out-of-image taps are skipped, which equals -inf padding).
__device__ __forceinline__ uint4 max_bf16x8(const uint4 a, const uint4 b) {
  uint4 r;
  const __nv_bfloat162* pa = reinterpret_cast<const __nv_bfloat162*>(&a);
  const __nv_bfloat162* pb = reinterpret_cast<const __nv_bfloat162*>(&b);
  __nv_bfloat162* pr = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) pr[i] = __hmax2(pa[i], pb[i]);
  return r;
}

__global__ void __launch_bounds__(256)
spp_pool_kernel(uint4* __restrict__ buf, int H, int W, int C8, int ctot8) {
  extern __shared__ uint4 spp_sm[];
  uint4* A = spp_sm;                 // [H*W][2]  current plane (16 channels = 2 x uint4 per pixel)
  uint4* T = spp_sm + 2 * H * W;     // [H*W][2]  horizontal pass
  const int n = blockIdx.y, cg = blockIdx.x;
  const int E = 2 * H * W;
  uint4* img = buf + (size_t)n * H * W * ctot8 + cg * 2;
  for (int i = threadIdx.x; i < E; i += 256) A[i] = img[(size_t)(i >> 1) * ctot8 + (i & 1)];
  __syncthreads();
  for (int stage = 1; stage <= 3; ++stage) {
    for (int i = threadIdx.x; i < E; i += 256) {
      const int pix = i >> 1, x = pix % W;
      uint4 m = A[i];
#pragma unroll
      for (int d = 1; d <= 2; ++d) {
        if (x - d >= 0) m = max_bf16x8(m, A[i - 2 * d]);
        if (x + d < W) m = max_bf16x8(m, A[i + 2 * d]);
      }
      T[i] = m;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < E; i += 256) {
      const int pix = i >> 1, y = pix / W;
      uint4 m = T[i];
#pragma unroll
      for (int d = 1; d <= 2; ++d) {
        if (y - d >= 0) m = max_bf16x8(m, T[i - 2 * d * W]);
        if (y + d < H) m = max_bf16x8(m, T[i + 2 * d * W]);
      }
      A[i] = m;
      img[(size_t)pix * ctot8 + stage * C8 + (i & 1)] = m;
    }
    __syncthreads();
  }
}

int upsample2x_launch(const void* src, void* dst, int N, int H, int W, int C, int s_ctot, int s_coff,
                      int d_ctot, int d_coff, cudaStream_t stream) {
  const size_t total = (size_t)N * 4 * H * W * (C / 8);
  const int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  upsample2x_kernel<<<blocks, 256, 0, stream>>>((const uint4*)src, (uint4*)dst, N, H, W, C / 8,
                                                s_ctot / 8, s_coff / 8, d_ctot / 8, d_coff / 8);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  return cudaGetLastError() == cudaSuccess ? TCD_OK : TCD_ERR_CUDA;
}

int spp_pool_launch(void* buf, int N, int H, int W, int C, int ctot, cudaStream_t stream) {
  TCD_REQUIRE(C % 16 == 0 && ctot % 8 == 0, "spp: C %d must be a multiple of 16", C);
  const size_t smem = (size_t)4 * H * W * sizeof(uint4);
  TCD_REQUIRE(smem <= 200 * 1024, "spp: %dx%d plane does not fit shared memory", H, W);
  if (int rc = ensure_dyn_smem((const void*)spp_pool_kernel, smem)) return rc;
  spp_pool_kernel<<<dim3(C / 16, N), 256, smem, stream>>>((uint4*)buf, H, W, C / 8, ctot / 8);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  return cudaGetLastError() == cudaSuccess ? TCD_OK : TCD_ERR_CUDA;
}

}  // namespace tcd
