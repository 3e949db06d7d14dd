// tcd_program: an ordered list of kernels (tcgen05 convs / linears + the transformer kernels of
// tfm_kernels.cuh + the two YOLOX aux kernels) over device buffers it owns, replayed as ONE CUDA
// graph.  The host side describes RT-DETRv2 as such a list (rtdetr_engine.py); this file owns
// weights upload, plan construction, graph capture, replay and per-op profiling.
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>

#include "conv_umma.cuh"
#include "tfm_kernels.cuh"
#include "yolox_net.cuh"

using namespace tcd;

namespace tcd {
int require_sm100();
}

namespace {

struct ProgOp {
  std::string name;
  int kind = 0;  // 0 = conv
  ConvPlan conv;
  tcd_op_desc d;
  double flops = 0;
};

int launch_small(const tcd_op_desc& d, cudaStream_t st) {
  const int32_t* i = d.i;
  void* const* p = d.p;
  switch (d.kind) {
    case TCD_OP_LAYERNORM:  // p: a, b(opt), gamma, beta, y; i: rows, C; f: eps
      return layernorm_launch(p[0], p[1], (const float*)p[2], (const float*)p[3], p[4], i[0], i[1], d.f[0], st);
    case TCD_OP_ADD_ROWS:  // p: a, b, y; i: N, T, C, bcast_b
      return add_rows_launch(p[0], p[1], p[2], i[0], i[1], i[2], i[3], st);
    case TCD_OP_MASK_ROWS:  // p: x, mask, y; i: N, T, C
      return mask_rows_launch(p[0], (const float*)p[1], p[2], i[0], i[1], i[2], st);
    case TCD_OP_MHA:  // p: q, k, v, out; i: ldq,qoff,ldk,koff,ldv,voff,ldo,N,Tq,Tk,heads; f: scale
      return mha_launch(p[0], i[0], i[1], p[1], i[2], i[3], p[2], i[4], i[5], p[3], i[6], i[7], i[8], i[9], i[10],
                        d.f[0], st);
    case TCD_OP_MSDA: {  // p: value, offsets, logits, ref, out; i: N,Q,S,heads,L,P,h0,w0,h1,w1,h2,w2; f: scale
      MsdaLevels lv;
      memset(&lv, 0, sizeof(lv));
      lv.n_levels = i[4];
      lv.n_points = i[5];
      int start = 0;
      for (int l = 0; l < lv.n_levels && l < 3; ++l) {
        lv.h[l] = i[6 + 2 * l];
        lv.w[l] = i[7 + 2 * l];
        lv.start[l] = start;
        start += lv.h[l] * lv.w[l];
      }
      return msda_launch(p[0], (const float*)p[1], (const float*)p[2], (const float*)p[3], p[4], i[0], i[1], i[2],
                         i[3], lv, d.f[0], st);
    }
    case TCD_OP_ROWMAX:  // p: x, y; i: rows, C
      return rowmax_launch((const float*)p[0], (float*)p[1], i[0], i[1], st);
    case TCD_OP_TOPK:  // p: keys, idx; i: N, n, k
      return topk_launch((const float*)p[0], i[0], i[1], i[2], (int*)p[1], st);
    case TCD_OP_GATHER:  // p: src, idx, out; i: N, S, K, C
      return gather_rows_launch(p[0], (const int*)p[1], p[2], i[0], i[1], i[2], i[3], st);
    case TCD_OP_REF_INIT:  // p: delta, anchors, idx, ref_sig, ref_bf16; i: ldd, N, K
      return ref_init_launch((const float*)p[0], i[0], (const float*)p[1], (const int*)p[2], (float*)p[3], p[4],
                             i[1], i[2], st);
    case TCD_OP_REF_UPDATE:  // p: delta, ref_sig, ref_bf16; i: ldd, rows
      return ref_update_launch((const float*)p[0], i[0], (float*)p[1], p[2], i[1], st);
    case TCD_OP_DETR_POST:  // p: logits, ref_sig, idx, orig_wh, labels, boxes, scores; i: N, Q, C, k
      return detr_post_launch((const float*)p[0], (const float*)p[1], (const int*)p[2], (const float*)p[3], i[0],
                              i[1], i[2], i[3], (int*)p[4], (float*)p[5], (float*)p[6], st);
    case TCD_OP_MAXPOOL:  // p: x, y; i: N, H, W, C
      return maxpool3x3s2_launch(p[0], p[1], i[0], i[1], i[2], i[3], st);
    case TCD_OP_AVGPOOL:
      return avgpool2x2_launch(p[0], p[1], i[0], i[1], i[2], i[3], st);
    case TCD_OP_UPSAMPLE:  // p: src, dst; i: N, H, W, C, s_ctot, s_coff, d_ctot, d_coff
      return upsample2x_launch(p[0], p[1], i[0], i[1], i[2], i[3], i[4], i[5], i[6], i[7], st);
    case TCD_OP_VIT_ASSEMBLE:  // p: emb, pos, x; i: N, T, D
      return vit_assemble_launch(p[0], (const float*)p[1], p[2], i[0], i[1], i[2], st);
    case TCD_OP_MHA_ROPE:  // p: q, k, v, out, rope, stats_out (opt); i: ldq,qoff,ldk,koff,ldv,voff,ldo,N,T,heads,prefix; f: scale
      return mha64_rope_launch(p[0], i[0], i[1], p[1], i[2], i[3], p[2], i[4], i[5], p[3], i[6], i[7], i[8], i[9],
                               d.f[0], (const float*)p[4], i[10], (float*)p[5], st);
    case TCD_OP_SWIGLU_LN:  // p: gu, gamma, beta, y; i: rows, H, Hp, gated; f: eps
      return swiglu_ln_launch(p[0], (const float*)p[1], (const float*)p[2], p[3], i[0], i[1], i[2], d.f[0], i[3], st);
    case TCD_OP_SOFTMAX:  // p: x, y; i: rows, C, ldx, ldy
      return softmax_rows_launch((const float*)p[0], i[2], (float*)p[1], i[3], i[0], i[1], st);
    case TCD_OP_HEAD_GEMV:  // p: x, w, bias, out; i: B, K, NC, ld
      return head_gemv_launch(p[0], p[1], (const float*)p[2], (float*)p[3], i[0], i[1], i[2], i[3], st);
    case TCD_OP_ROW_STATS:  // p: x, stats; i: rows, H, Hp, parts (0: (mean, rstd); > 0: (sum, sumsq) in slot 0 of `parts`); f: eps
      return row_stats_launch(p[0], (float*)p[1], i[0], i[1], i[2], d.f[0], i[3], st);
    case TCD_OP_TAKE:  // p: src, idx, out; i: N, k, ld
      return take_f32_launch((const float*)p[0], i[2], (const int*)p[1], (float*)p[2], i[0], i[1], st);
  }
  set_last_error("program: unknown op kind %d (%s)", d.kind, d.name);
  return TCD_ERR_INVALID;
}

}  // namespace

struct tcd_program {
  int device = 0;
  cudaStream_t stream = nullptr;
  std::vector<void*> buffers;
  std::vector<ProgOp> ops;
  cudaGraphExec_t graph = nullptr;
  size_t graph_ops = 0;
};

static int run_op(const ProgOp& op, cudaStream_t st) {
  return op.kind == 0 ? conv_plan_launch(op.conv, st) : launch_small(op.d, st);
}

extern "C" {

tcd_status tcd_program_create(int32_t device, tcd_program** out) {
  TCD_REQUIRE(out, "program_create: null argument");
  TCD_CUDA_OK(cudaSetDevice(device));
  int rc = require_sm100();
  if (rc) return (tcd_status)rc;
  tcd_program* p = new tcd_program();
  p->device = device;
  cudaError_t e = cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) {
    delete p;
    set_last_error("program_create: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  *out = p;
  return TCD_OK;
}

void tcd_program_destroy(tcd_program* p) {
  if (!p) return;
  cudaSetDevice(p->device);
  cudaStreamSynchronize(p->stream);
  if (p->graph) cudaGraphExecDestroy(p->graph);
  for (void* b : p->buffers) cudaFree(b);
  cudaStreamDestroy(p->stream);
  delete p;
}

tcd_status tcd_program_alloc(tcd_program* p, size_t bytes, void** out_dev) {
  TCD_REQUIRE(p && out_dev && bytes > 0, "program_alloc: bad argument");
  TCD_CUDA_OK(cudaSetDevice(p->device));
  void* d = nullptr;
  TCD_CUDA_OK(cudaMalloc(&d, (bytes + 1023) & ~(size_t)1023));
  TCD_CUDA_OK(cudaMemset(d, 0, (bytes + 1023) & ~(size_t)1023));
  p->buffers.push_back(d);
  *out_dev = d;
  return TCD_OK;
}

tcd_status tcd_program_upload(tcd_program* p, const void* host, size_t bytes, void** out_dev) {
  TCD_REQUIRE(host, "program_upload: null host pointer");
  tcd_status rc = tcd_program_alloc(p, bytes, out_dev);
  if (rc) return rc;
  TCD_CUDA_OK(cudaMemcpy(*out_dev, host, bytes, cudaMemcpyHostToDevice));
  return TCD_OK;
}

tcd_status tcd_program_add_conv(tcd_program* p, const tcd_conv_desc* c) {
  TCD_REQUIRE(p && c && c->x && c->w_host && c->bias_host && (c->y || c->out_f32), "add_conv: null argument");
  TCD_REQUIRE(p->graph == nullptr, "add_conv: program already captured");
  TCD_CUDA_OK(cudaSetDevice(p->device));
  const int cin_pad = (c->cin + 15) / 16 * 16;
  const bool f32 = c->out_f32 != nullptr;
  const int cout_pad = f32 ? (c->cout + 15) / 16 * 16 : c->cout;
  const int taps = c->k * c->k;
  std::vector<uint16_t> hw((size_t)taps * cout_pad * cin_pad);
  conv_pack_weights(c->w_host, c->cout, c->cin, c->k, cout_pad, cin_pad, hw.data());
  std::vector<float> hb(cout_pad, 0.f);
  memcpy(hb.data(), c->bias_host, sizeof(float) * c->cout);
  void *dw = nullptr, *db = nullptr;
  tcd_status rc = tcd_program_upload(p, hw.data(), hw.size() * 2, &dw);
  if (rc) return rc;
  rc = tcd_program_upload(p, hb.data(), hb.size() * 4, &db);
  if (rc) return rc;
  ConvDesc d;
  d.x = c->x; d.N = c->N; d.H = c->H; d.W = c->W; d.x_ctot = c->x_ctot; d.x_coff = c->x_coff; d.cin = cin_pad;
  d.x_nstride = c->x_nstride;
  d.w = dw; d.bias = (const float*)db; d.cout = c->cout; d.cout_pad = cout_pad; d.k = c->k; d.stride = c->stride;
  d.silu = c->act;
  d.y = c->y; d.y_ctot = c->y_ctot; d.y_coff = c->y_coff; d.y_nstride = c->y_nstride;
  d.resid = c->resid; d.r_ctot = c->r_ctot; d.r_coff = c->r_coff; d.r_nstride = c->r_nstride;
  d.resid_pre = c->resid_pre;
  // a program runs on ONE stream: nothing to be co-resident with, so each layer takes whole SMs
  {
    const char* e = getenv("TCD_PROGRAM_SHARE");
    d.share = e ? atoi(e) : 0;
  }
  if (c->ln_stats) {
    TCD_REQUIRE(c->ln_colsum_host, "add_conv: ln_stats without ln_colsum_host");
    std::vector<float> hc(cout_pad, 0.f);
    memcpy(hc.data(), c->ln_colsum_host, sizeof(float) * c->cout);
    void* dc = nullptr;
    rc = tcd_program_upload(p, hc.data(), hc.size() * 4, &dc);
    if (rc) return rc;
    d.ln_stats = c->ln_stats;
    d.ln_colsum = (const float*)dc;
    d.ln_parts = c->ln_parts; d.ln_count = c->ln_count; d.ln_eps = c->ln_eps;
    TCD_REQUIRE(c->ln_parts >= 0 && c->ln_parts <= 64 && c->ln_parts % 2 == 0 && (c->ln_parts == 0 || c->ln_count > 0),
                "add_conv: ln_parts %d (even, <= 64)", c->ln_parts);
  }
  if (c->stats_out) {
    TCD_REQUIRE(c->stats_parts > 0 && c->stats_slot0 >= 0, "add_conv: stats_out without stats_parts");
    d.stats_out = c->stats_out; d.stats_parts = c->stats_parts; d.stats_slot0 = c->stats_slot0;
  }
  d.out_f32 = c->out_f32; d.out_ld = c->out_ld; d.out_col = c->out_col; d.rows_per_img = c->rows_per_img;
  d.row_off = c->row_off;
  ProgOp op;
  op.name = std::string(c->name, strnlen(c->name, sizeof(c->name)));
  op.kind = 0;
  int r = conv_plan_build(d, &op.conv);
  if (r) {
    std::string msg = tcd_last_error();
    set_last_error("add_conv(%s): %s", op.name.c_str(), msg.c_str());
    return (tcd_status)r;
  }
  op.flops = op.conv.flops;
  p->ops.push_back(op);
  return TCD_OK;
}

tcd_status tcd_program_add_op(tcd_program* p, const tcd_op_desc* d) {
  TCD_REQUIRE(p && d, "add_op: null argument");
  TCD_REQUIRE(p->graph == nullptr, "add_op: program already captured");
  TCD_REQUIRE(d->kind >= TCD_OP_LAYERNORM && d->kind <= TCD_OP_ROW_STATS, "add_op: kind %d", d->kind);
  ProgOp op;
  op.name = std::string(d->name, strnlen(d->name, sizeof(d->name)));
  op.kind = d->kind;
  op.d = *d;
  p->ops.push_back(op);
  return TCD_OK;
}

tcd_status tcd_program_run(tcd_program* p, int32_t graph) {
  TCD_REQUIRE(p && !p->ops.empty(), "program_run: empty program");
  TCD_CUDA_OK(cudaSetDevice(p->device));
  if (!graph) {
    for (const ProgOp& op : p->ops) {
      int rc = run_op(op, p->stream);
      if (rc) return (tcd_status)rc;
    }
    return TCD_OK;
  }
  if (!p->graph || p->graph_ops != p->ops.size()) {
    if (p->graph) cudaGraphExecDestroy(p->graph);
    p->graph = nullptr;
    cudaGraph_t g = nullptr;
    TCD_CUDA_OK(cudaStreamBeginCapture(p->stream, cudaStreamCaptureModeThreadLocal));
    int rc = TCD_OK;
    for (const ProgOp& op : p->ops) {
      rc = run_op(op, p->stream);
      if (rc) break;
    }
    cudaError_t e = cudaStreamEndCapture(p->stream, &g);
    if (rc) {
      if (g) cudaGraphDestroy(g);
      return (tcd_status)rc;
    }
    if (e != cudaSuccess) {
      set_last_error("program capture: %s", cudaGetErrorString(e));
      return TCD_ERR_CUDA;
    }
    e = cudaGraphInstantiate(&p->graph, g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) {
      set_last_error("program graph instantiate: %s", cudaGetErrorString(e));
      return TCD_ERR_CUDA;
    }
    p->graph_ops = p->ops.size();
  }
  TCD_CUDA_OK(cudaGraphLaunch(p->graph, p->stream));
  return TCD_OK;
}

tcd_status tcd_program_sync(tcd_program* p) {
  TCD_REQUIRE(p, "program_sync: null program");
  TCD_CUDA_OK(cudaSetDevice(p->device));
  cudaError_t e = cudaStreamSynchronize(p->stream);
  if (e != cudaSuccess) {
    set_last_error("program_sync: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return TCD_OK;
}

void* tcd_program_stream(tcd_program* p) { return p ? (void*)p->stream : nullptr; }
int32_t tcd_program_num_ops(const tcd_program* p) { return p ? (int32_t)p->ops.size() : 0; }
const char* tcd_program_op_name(const tcd_program* p, int32_t idx) {
  if (!p || idx < 0 || idx >= (int)p->ops.size()) return "";
  return p->ops[idx].name.c_str();
}
double tcd_program_op_flops(const tcd_program* p, int32_t idx) {
  if (!p || idx < 0 || idx >= (int)p->ops.size()) return 0.0;
  return p->ops[idx].flops;
}

tcd_status tcd_program_profile(tcd_program* p, int32_t iters, float* op_ms, int32_t cap) {
  TCD_REQUIRE(p && op_ms && iters >= 1 && cap >= (int)p->ops.size(), "program_profile: bad argument");
  TCD_CUDA_OK(cudaSetDevice(p->device));
  const size_t n = p->ops.size();
  std::vector<cudaEvent_t> ev(2 * n);
  for (auto& e : ev) TCD_CUDA_OK(cudaEventCreate(&e));
  std::vector<double> acc(n, 0.0);
  int rc = TCD_OK;
  for (int it = 0; it < iters && rc == TCD_OK; ++it) {
    for (size_t i = 0; i < n && rc == TCD_OK; ++i) {
      cudaEventRecord(ev[2 * i], p->stream);
      rc = run_op(p->ops[i], p->stream);
      cudaEventRecord(ev[2 * i + 1], p->stream);
    }
    if (rc == TCD_OK && cudaStreamSynchronize(p->stream) != cudaSuccess) {
      set_last_error("program_profile: %s", cudaGetErrorString(cudaGetLastError()));
      rc = TCD_ERR_CUDA;
    }
    for (size_t i = 0; i < n && rc == TCD_OK; ++i) {
      float t = 0.f;
      cudaEventElapsedTime(&t, ev[2 * i], ev[2 * i + 1]);
      acc[i] += t;
    }
  }
  for (auto& e : ev) cudaEventDestroy(e);
  if (rc) return (tcd_status)rc;
  for (size_t i = 0; i < n; ++i) op_ms[i] = (float)(acc[i] / iters);
  return TCD_OK;
}

}  // extern "C"
