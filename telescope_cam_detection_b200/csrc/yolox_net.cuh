// YOLOX-S layer plan on the conv engine: CSPDarknet + PAFPN + decoupled head.
// The network lives here as a flat list of launches over one activation arena; concat is
// "write into a channel slice", CSP conv1/conv2 and the head's first cls/reg convs are fused
// along Cout, BN is folded by the weight packer, SiLU / residual run in the conv epilogue.
#pragma once
#include <map>
#include <string>
#include <vector>

#include "conv_umma.cuh"

namespace tcd {

struct WeightEntry {
  int cout, cin, k;
  const float* w;  // host, [cout][cin][k][k]
  const float* b;  // host, [cout]
};

struct WeightBlob {
  std::map<std::string, WeightEntry> entries;
  int parse(const void* blob, size_t bytes);
};

// device-resident packed conv weights
struct DevConv {
  void* w = nullptr;      // bf16 [taps][cout_pad][cin_pad]
  float* bias = nullptr;  // f32 [cout_pad]
  int cout = 0, cout_pad = 0, cin_pad = 0, k = 1;
};

struct Act {
  void* p = nullptr;
  int H = 0, W = 0, C = 0;
};

enum OpKind { OP_CONV = 0, OP_UPSAMPLE = 1, OP_SPP = 2 };

struct NetOp {
  OpKind kind;
  ConvPlan conv;           // OP_CONV
  // aux ops
  const void* src = nullptr;
  void* dst = nullptr;
  int N = 0, H = 0, W = 0, C = 0, s_ctot = 0, s_coff = 0, d_ctot = 0, d_coff = 0;
  std::string name;
};

struct YoloxWeights {
  std::map<std::string, DevConv> convs;  // keyed by plan-level name (fused names included)
  int num_classes = 80;
  int upload(const WeightBlob& blob);
  void release();
};

struct YoloxPlan {
  int B = 0, in_h = 0, in_w = 0, num_classes = 80, n_anchors = 0;
  std::vector<NetOp> ops;
  std::map<std::string, Act> acts;  // named activation buffers (debug taps)
  void* arena = nullptr;
  size_t arena_bytes = 0;
  bool owns_arena = true;  // false: laid out over an arena another plan of the same engine owns (smaller batch sizes)
  void* input = nullptr;   // bf16 [B,in_h/2,in_w/2,16]
  float* pred_raw = nullptr;  // f32 [B,A,5+C]
  int lvl_start[3], lvl_w[3], lvl_stride[3];
  double conv_flops = 0;

  // ext_arena: lay the plan out over an existing arena (the engine's plan for its largest batch) instead of allocating;
  // the network input is always the first buffer, so every plan of an engine shares the input address
  int build(const YoloxWeights& wts, int B, int in_h, int in_w, void* ext_arena = nullptr, size_t ext_bytes = 0);
  int run(cudaStream_t stream) const;
  int run_one(int i, cudaStream_t stream) const;
  void release();
};

int upsample2x_launch(const void* src, void* dst, int N, int H, int W, int C, int s_ctot, int s_coff,
                      int d_ctot, int d_coff, cudaStream_t stream);
int spp_pool_launch(void* buf, int N, int H, int W, int C, int ctot, cudaStream_t stream);

}  // namespace tcd
