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
  int branch = 0;        // 0: the engine's stream; 1, 2: side branches (head levels 0 / 1) that fork off the trunk
  int fork_branch = 0;   // != 0: this op's output is what branch `fork_branch` waits for
};

// side streams + events an engine lends to its plans so that independent layers overlap (YoloxPlan::run)
struct PlanBranches {
  cudaStream_t aux[2] = {nullptr, nullptr};
  cudaEvent_t fork[2] = {nullptr, nullptr};
  cudaEvent_t join[2] = {nullptr, nullptr};
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
  // candidate emission of the prediction convs (set before build; nullptr: every row is written)
  const float* emit_thr = nullptr;
  int* maybe_count = nullptr;
  int* maybe_idx = nullptr;
  int maybe_cap = 0;

  // ext_arena: lay the plan out over an existing arena (the engine's plan for its largest batch) instead of allocating;
  // the network input is always the first buffer, so every plan of an engine shares the input address
  int build(const YoloxWeights& wts, int B, int in_h, int in_w, void* ext_arena = nullptr, size_t ext_bytes = 0);
  // br != nullptr: the head of pyramid level 0 (4 launches, a third of the network's FLOPs) starts as soon as the
  // PAFPN has produced P3 and runs beside the remaining bottom-up layers (C3_n3, C3_n4: low-occupancy 40x40 / 20x20
  // maps), level 1 likewise after P4; everything joins before the decode.  Works eagerly and under stream capture
  // (the cross-stream event waits become graph edges).
  int run(cudaStream_t stream, const PlanBranches* br = nullptr) const;
  int run_one(int i, cudaStream_t stream) const;
  void release();
};

int upsample2x_launch(const void* src, void* dst, int N, int H, int W, int C, int s_ctot, int s_coff,
                      int d_ctot, int d_coff, cudaStream_t stream);
int spp_pool_launch(void* buf, int N, int H, int W, int C, int ctot, cudaStream_t stream);

}  // namespace tcd
