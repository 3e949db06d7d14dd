// C ABI glue, part 2: the YOLOX detector handle (preprocess -> network -> decode/NMS).
// The pipeline has ENGINES (own stream, activation arena, CUDA graphs, NMS scratch) and, per engine, LANES (own
// frame staging, pinned result table, device result table, events).  A pipeline slot of the C ABI is one
// (engine, lane) pair: slot v -> engine v / n_lanes, lane v % n_lanes.
//
// Dynamic batching inside the library (round 2).  A lane's submit uploads and preprocesses its frames straight into
// the next free frame positions of the engine's network input, but the network + decode/NMS launch is DEFERRED until
// the engine's group is full (no room for another max_batch submit) or somebody waits on one of its lanes.  Under
// load, n_lanes submitted batches therefore run as ONE launch per layer over n_lanes x max_batch frames -- the
// per-launch latencies (prologue, dependency wait, epilogue tail: ~10 us per layer whatever the batch) are paid once
// per group instead of once per batch: measured 12.6k -> 16.5k frames/s at 4 x 8 frames.  A lone batch is launched
// by its own wait(), at its own size: the latency of a single detect_batch does not change.  Engines still overlap
// each other (the low-occupancy layers of one group share the chip with another group's).
#include <algorithm>
#include <atomic>
#include <math.h>
#include <map>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include "kernels.cuh"
#include "yolox_net.cuh"

using namespace tcd;

namespace tcd {
extern std::atomic<uint64_t> g_launch_count;
extern std::atomic<uint64_t> g_h2d_bytes;
}

namespace {

constexpr int kEngines = 8;
constexpr int kLanes = 8;
constexpr int kProfileReps = 8;

#define TCD_USE(handle) TCD_CUDA_OK(cudaSetDevice((handle)->device))   /* the current device is per host thread */

struct Lane {                     // what ONE in-flight batch owns besides its engine
  uint8_t* staging = nullptr;     // device, max_batch frames of frame_cap bytes (grows when a larger frame arrives)
  size_t frame_cap = 0;
  FrameMeta* meta_host = nullptr; // pinned [max_batch]: filter table + box un-scaling per frame
  PreFrame* pre_dev = nullptr;    // device [max_batch]
  PreFrame* pre_host = nullptr;   // pinned [max_batch]
  float* out_pinned = nullptr;    // [max_batch,max_out,7]
  int* cnt_pinned = nullptr;      // [max_batch]
  float* tab_dets = nullptr;      // device copy of the result table (tcd_detector_result_device)
  int* tab_counts = nullptr;
  cudaEvent_t ev[6] = {};
  cudaEvent_t ev_up = nullptr;    // frames of this lane's batch are in staging
  int pending_B = 0;              // frames of the batch in flight (0: lane free)
  int frame0 = 0;                 // position of its first frame in the engine's group
  bool launched = false;          // its group's network launch has been enqueued
};

struct Slot {                     // an engine
  cudaStream_t stream = nullptr;
  cudaStream_t up = nullptr;      // frame uploads: copy-engine work off the compute stream
  FrameMeta* meta_dev = nullptr;  // [max_batch], what the captured nms_reduce reads; refilled per batch
  Lane lane[kLanes];
  int group_frames = 0;           // frames preprocessed into the network input, network not launched yet
  int group_n = 0;
  Lane* group[kLanes] = {};
  std::map<int, YoloxPlan*> plans;
  std::map<int, cudaGraphExec_t> graphs;
  std::map<int, uint64_t> graph_kernels;  // kernel nodes per captured graph (launch accounting)
  // postprocess buffers (sized for max_batch)
  PostBuffers pb;
  float* thr_dev = nullptr;
  int* maybe_count = nullptr;     // candidate emission of the prediction convs: rows written per image ...
  int* maybe_idx = nullptr;       // ... and their anchor indices [group_cap, cap]
  float* pred_in = nullptr;       // device scratch for tcd_postprocess [max_batch,A,5+C]
  const tcd_class_filter* tables = nullptr;  // the detector's table array
  const YoloxPlan* last_plan = nullptr;      // the plan whose activations the arena currently holds (debug taps)
  PlanBranches br;                           // side streams for the head levels (YoloxPlan::run)
  bool use_branches = true;                  // TCD_NO_BRANCH=1: one linear chain (A/B)
};

}  // namespace

struct tcd_detector {
  tcd_detector_cfg cfg;
  int device = 0;
  YoloxWeights wts;
  Slot slot[kEngines];
  int n_engines = 4;              // TCD_ENGINES; slots exposed = n_engines * n_lanes
  int n_lanes = 4;                // TCD_LANES (1..kLanes): batches one engine coalesces into one launch
  int group_cap = 0;              // n_lanes * max_batch frames: what an engine's arena and scratch are sized for
  float thr_host[2] = {0.25f, 0.45f};
  tcd_class_filter* tables_dev = nullptr;   // [kMaxFilterTables]
  int n_anchors = 0, cap = 0;
  float last_ms[5] = {0, 0, 0, 0, 0};
  bool use_graph = true;
};

static int slot_init(tcd_detector* d, Slot& s) {
  const tcd_detector_cfg& c = d->cfg;
  const int MB = c.max_batch;      // frames per lane (one submit)
  const int GB = d->group_cap;     // frames per engine launch
  TCD_CUDA_OK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
  TCD_CUDA_OK(cudaStreamCreateWithFlags(&s.up, cudaStreamNonBlocking));
  for (int b = 0; b < 2; ++b) {
    TCD_CUDA_OK(cudaStreamCreateWithFlags(&s.br.aux[b], cudaStreamNonBlocking));
    TCD_CUDA_OK(cudaEventCreateWithFlags(&s.br.fork[b], cudaEventDisableTiming));
    TCD_CUDA_OK(cudaEventCreateWithFlags(&s.br.join[b], cudaEventDisableTiming));
  }
  s.use_branches = getenv("TCD_NO_BRANCH") == nullptr;
  TCD_CUDA_OK(cudaMalloc((void**)&s.meta_dev, sizeof(FrameMeta) * GB));
  TCD_CUDA_OK(cudaMemset(s.meta_dev, 0, sizeof(FrameMeta) * GB));
  for (int li = 0; li < d->n_lanes; ++li) {
    Lane& l = s.lane[li];
    l.frame_cap = ((size_t)c.max_src_h * c.max_src_w * 3 + 255) & ~(size_t)255;
    TCD_CUDA_OK(cudaMalloc((void**)&l.staging, l.frame_cap * MB));
    TCD_CUDA_OK(cudaMallocHost((void**)&l.meta_host, sizeof(FrameMeta) * MB));
    TCD_CUDA_OK(cudaMalloc((void**)&l.pre_dev, sizeof(PreFrame) * MB));
    TCD_CUDA_OK(cudaMallocHost((void**)&l.pre_host, sizeof(PreFrame) * MB));
    TCD_CUDA_OK(cudaMalloc((void**)&l.tab_dets, sizeof(float) * 7 * c.max_out * MB));
    TCD_CUDA_OK(cudaMalloc((void**)&l.tab_counts, sizeof(int) * MB));
    TCD_CUDA_OK(cudaMallocHost((void**)&l.out_pinned, sizeof(float) * 7 * c.max_out * MB));
    TCD_CUDA_OK(cudaMallocHost((void**)&l.cnt_pinned, sizeof(int) * MB));
    for (auto& e : l.ev) TCD_CUDA_OK(cudaEventCreate(&e));
    TCD_CUDA_OK(cudaEventCreateWithFlags(&l.ev_up, cudaEventDisableTiming));
  }
  const int A = d->n_anchors, C = c.num_classes, cap = d->cap;
  s.tables = d->tables_dev;
  PostBuffers& pb = s.pb;
  memset(&pb, 0, sizeof(pb));
  pb.A = A; pb.C = C; pb.cap = cap; pb.max_out = c.max_out;
  TCD_CUDA_OK(cudaMalloc((void**)&s.thr_dev, 4 * sizeof(float)));   // conf, nms, emission threshold of the pred convs
  pb.thr = s.thr_dev;
  TCD_CUDA_OK(cudaMalloc((void**)&s.maybe_count, sizeof(int) * GB));
  TCD_CUDA_OK(cudaMalloc((void**)&s.maybe_idx, sizeof(int) * (size_t)cap * GB));
  TCD_CUDA_OK(cudaMalloc((void**)&pb.cand_count, sizeof(int) * GB));
  TCD_CUDA_OK(cudaMalloc((void**)&pb.cand, sizeof(float) * 8 * cap * GB));
  TCD_CUDA_OK(cudaMalloc((void**)&pb.cand_anchor, sizeof(int) * cap * GB));
  TCD_CUDA_OK(cudaMalloc((void**)&pb.srec, sizeof(float) * 8 * cap * GB));
  TCD_CUDA_OK(cudaMalloc((void**)&pb.sbox, sizeof(float4) * cap * GB));
  TCD_CUDA_OK(cudaMalloc((void**)&pb.n_sorted, sizeof(int) * GB));
  TCD_CUDA_OK(cudaMalloc((void**)&pb.mask, (size_t)GB * cap * (cap / 64) * 8));
  TCD_CUDA_OK(cudaMalloc((void**)&pb.out_dets, sizeof(float) * 7 * c.max_out * GB));
  TCD_CUDA_OK(cudaMalloc((void**)&pb.out_counts, sizeof(int) * GB));
  TCD_CUDA_OK(cudaMalloc((void**)&pb.decoded, sizeof(float) * (size_t)MB * A * (5 + C)));   // parity tap: one lane's worth
  TCD_CUDA_OK(cudaMalloc((void**)&s.pred_in, sizeof(float) * (size_t)MB * A * (5 + C)));
  TCD_CUDA_OK(cudaMemcpy(s.thr_dev, d->thr_host, 2 * sizeof(float), cudaMemcpyHostToDevice));
  TCD_CUDA_OK(cudaMemcpy(s.thr_dev + 2, d->thr_host, sizeof(float), cudaMemcpyHostToDevice));
  return TCD_OK;
}

static void slot_free(Slot& s) {
  for (auto& kv : s.graphs) cudaGraphExecDestroy(kv.second);
  for (auto& kv : s.plans) if (!kv.second->owns_arena) { kv.second->release(); delete kv.second; kv.second = nullptr; }
  for (auto& kv : s.plans) if (kv.second) { kv.second->release(); delete kv.second; }
  PostBuffers& pb = s.pb;
  void* ptrs[] = {s.maybe_count, s.maybe_idx, s.meta_dev, s.thr_dev, pb.cand_count, pb.cand, pb.cand_anchor, pb.srec,
                  pb.sbox, pb.n_sorted, pb.mask, pb.out_dets, pb.out_counts, pb.decoded, s.pred_in};
  for (void* p : ptrs) if (p) cudaFree(p);
  for (Lane& l : s.lane) {
    void* dp[] = {l.staging, l.pre_dev, l.tab_dets, l.tab_counts};
    for (void* p : dp) if (p) cudaFree(p);
    if (l.pre_host) cudaFreeHost(l.pre_host);
    if (l.meta_host) cudaFreeHost(l.meta_host);
    if (l.out_pinned) cudaFreeHost(l.out_pinned);
    if (l.cnt_pinned) cudaFreeHost(l.cnt_pinned);
    for (auto& e : l.ev) if (e) cudaEventDestroy(e);
    if (l.ev_up) cudaEventDestroy(l.ev_up);
  }
  for (int b = 0; b < 2; ++b) {
    if (s.br.aux[b]) cudaStreamDestroy(s.br.aux[b]);
    if (s.br.fork[b]) cudaEventDestroy(s.br.fork[b]);
    if (s.br.join[b]) cudaEventDestroy(s.br.join[b]);
  }
  if (s.up) cudaStreamDestroy(s.up);
  if (s.stream) cudaStreamDestroy(s.stream);
}

static int get_plan(tcd_detector* d, Slot& s, int B, YoloxPlan** out) {
  auto it = s.plans.find(B);
  if (it != s.plans.end()) { *out = it->second; return TCD_OK; }
  YoloxPlan* p = new YoloxPlan();
  if (getenv("TCD_NO_EMIT") == nullptr) {   // A/B: TCD_NO_EMIT=1 writes every prediction row and decodes densely
    p->emit_thr = s.thr_dev + 2;
    p->maybe_count = s.maybe_count;
    p->maybe_idx = s.maybe_idx;
    p->maybe_cap = d->cap;
  }
  // every plan of an engine lives in the arena of its largest one (the group capacity, built at create)
  auto big = s.plans.find(d->group_cap);
  int rc = big == s.plans.end() ? p->build(d->wts, B, d->cfg.input_h, d->cfg.input_w)
                                : p->build(d->wts, B, d->cfg.input_h, d->cfg.input_w, big->second->arena,
                                           big->second->arena_bytes);
  if (rc) { p->release(); delete p; return rc; }
  s.plans[B] = p;
  *out = p;
  return TCD_OK;
}

static PostBuffers post_for(const Slot& s, const YoloxPlan* p, int B, const float* pred, bool want_decoded) {
  PostBuffers pb = s.pb;
  pb.B = B;
  pb.pred = pred;
  if (!want_decoded) pb.decoded = nullptr;
  pb.meta = s.meta_dev;
  pb.tables = s.tables;
  pb.maybe_count = (pred == p->pred_raw && p->emit_thr != nullptr) ? s.maybe_count : nullptr;
  pb.maybe_idx = s.maybe_idx;
  pb.n_levels = 3;
  for (int l = 0; l < 3; ++l) {
    pb.lvl_start[l] = p->lvl_start[l];
    pb.lvl_w[l] = p->lvl_w[l];
    pb.lvl_stride[l] = p->lvl_stride[l];
  }
  return pb;
}

// network + postprocess for batch B on the slot's stream (graph-captured once per B)
static int run_net_post(tcd_detector* d, Slot& s, YoloxPlan* p, int B, bool want_decoded) {
  s.last_plan = p;
  if (d->use_graph && !want_decoded) {
    auto it = s.graphs.find(B);
    if (it == s.graphs.end()) {
      cudaGraph_t g = nullptr;
      const uint64_t launches_before = g_launch_count.load();
      TCD_CUDA_OK(cudaStreamBeginCapture(s.stream, cudaStreamCaptureModeThreadLocal));
      cudaMemsetAsync(s.maybe_count, 0, sizeof(int) * B, s.stream);
      int rc = p->run(s.stream, s.use_branches ? &s.br : nullptr);
      if (rc == TCD_OK) rc = postproc_launch(post_for(s, p, B, p->pred_raw, false), 1, s.stream);
      cudaError_t e = cudaStreamEndCapture(s.stream, &g);
      if (rc) { if (g) cudaGraphDestroy(g); return rc; }
      if (e != cudaSuccess) {
        set_last_error("graph capture: %s", cudaGetErrorString(e));
        return TCD_ERR_CUDA;
      }
      cudaGraphExec_t ge = nullptr;
      e = cudaGraphInstantiate(&ge, g, 0);
      cudaGraphDestroy(g);
      if (e != cudaSuccess) {
        set_last_error("graph instantiate: %s", cudaGetErrorString(e));
        return TCD_ERR_CUDA;
      }
      s.graphs[B] = ge;
      // the launches made while capturing did not execute: move them to per-replay accounting
      const uint64_t captured = g_launch_count.load() - launches_before;
      g_launch_count.fetch_sub(captured);
      s.graph_kernels[B] = captured;
      it = s.graphs.find(B);
    }
    TCD_CUDA_OK(cudaGraphLaunch(it->second, s.stream));
    g_launch_count.fetch_add(s.graph_kernels[B]);
    return TCD_OK;
  }
  // eager path.  The parity taps (want_decoded) need every row of the prediction: emission threshold -inf
  const float ninf = -INFINITY;
  if (want_decoded) TCD_CUDA_OK(cudaMemcpyAsync(s.thr_dev + 2, &ninf, sizeof(float), cudaMemcpyHostToDevice, s.stream));
  TCD_CUDA_OK(cudaMemsetAsync(s.maybe_count, 0, sizeof(int) * B, s.stream));
  int rc = p->run(s.stream, s.use_branches ? &s.br : nullptr);
  if (rc == TCD_OK) rc = postproc_launch(post_for(s, p, B, p->pred_raw, want_decoded), 1, s.stream);
  if (want_decoded) {
    TCD_CUDA_OK(cudaStreamSynchronize(s.stream));   // ninf must outlive the copy; then restore the product threshold
    TCD_CUDA_OK(cudaMemcpyAsync(s.thr_dev + 2, d->thr_host, sizeof(float), cudaMemcpyHostToDevice, s.stream));
  }
  return rc;
}

static const bool g_sparse_upload = [] {
  const char* e = getenv("TCD_SPARSE_UPLOAD");
  return !(e && e[0] == '0');
}();

// A host frame larger than the lane's staging slots (the reference accepts any frame size; its shipped config has a
// 2560x1440 camera): the lane's buffer is re-allocated.  The lane has no batch in flight (the caller checked), the
// captured graphs never see the staging address (the preprocess kernel reads it from PreFrame), and cudaFree waits
// for whatever the device still runs -- a one-off stall when a new camera size first shows up.
static int grow_staging(tcd_detector* d, Lane& l, size_t need) {
  if (need <= l.frame_cap) return TCD_OK;
  const size_t cap = (need + 255) & ~(size_t)255;
  uint8_t* fresh = nullptr;
  TCD_CUDA_OK(cudaMalloc((void**)&fresh, cap * d->cfg.max_batch));
  if (l.staging) cudaFree(l.staging);
  l.staging = fresh;
  l.frame_cap = cap;
  return TCD_OK;
}

static int enqueue_input(tcd_detector* d, Slot& s, Lane& l, YoloxPlan* p, const uint8_t* const* frames,
                         const int32_t* h, const int32_t* w, const int64_t* row_stride, int B,
                         int on_device, bool launch_pre = true, const int32_t* filter_table = nullptr,
                         int frame0 = 0) {
  // frame0: position of this batch's first frame in the engine's network input (every plan of an engine has the
  // input at the same address: bf16 [frames, in_h/2, in_w/2, 16]) and in its per-frame meta table
  uint8_t* input = (uint8_t*)p->input + (size_t)frame0 * (d->cfg.input_h / 2) * (d->cfg.input_w / 2) * 16 * 2;
  tcd_preproc_cfg pc = d->cfg.pre;
  pc.out_h = d->cfg.input_h;
  pc.out_w = d->cfg.input_w;
  // host frames travel on the engine's upload stream: the lane's staging buffer is free (the host has waited for
  // the lane's previous batch), so the copy starts now, under whatever the engine's stream is still computing
  cudaStream_t cs = on_device ? s.stream : s.up;
  if (!on_device) {
    size_t need = 0;
    for (int b = 0; b < B; ++b) need = std::max(need, (size_t)h[b] * (size_t)w[b] * 3);
    int rc = grow_staging(d, l, need);
    if (rc) return rc;
  }
  for (int b = 0; b < B; ++b) {
    const int tb = filter_table ? filter_table[b] : 0;
    TCD_REQUIRE(tb >= 0 && tb < kMaxFilterTables, "detect: filter table %d of frame %d outside [0,%d)", tb, b,
                kMaxFilterTables);
    l.meta_host[b].table = tb;
    // orig / network size evaluated like Python does (double division), then narrowed like np.float32(ratio)
    if (d->cfg.pre.mode == 1) {
      // letterbox: one isotropic ratio r = min(H/h, W/w); boxes un-scale by 1/r on both axes (the host formatter
      // passes orig = input / r, so its ratio is (input / r) / input -- evaluated the same way here)
      const double r = std::min((double)d->cfg.input_h / (double)h[b], (double)d->cfg.input_w / (double)w[b]);
      l.meta_host[b].ratio_w = (float)(((double)d->cfg.input_w / r) / (double)d->cfg.input_w);
      l.meta_host[b].ratio_h = (float)(((double)d->cfg.input_h / r) / (double)d->cfg.input_h);
    } else {
      l.meta_host[b].ratio_w = (float)((double)w[b] / (double)d->cfg.input_w);
      l.meta_host[b].ratio_h = (float)((double)h[b] / (double)d->cfg.input_h);
    }
    l.meta_host[b].pad = 0;
    const int64_t st = row_stride ? row_stride[b] : (int64_t)w[b] * 3;
    const uint8_t* ptr = frames[b];
    int64_t dst_stride = st;
    if (!on_device) {
      const size_t tight = (size_t)w[b] * 3;
      uint8_t* dst = l.staging + (size_t)b * l.frame_cap;
      // Row-sparse upload.  When the vertical resize ratio is an odd integer k (1080p -> 640 letterbox:
      // k = 3 exactly), the bilinear tap of output row d is s = k*(d+0.5)-0.5 = k*d + (k-1)/2, an integer:
      // the second tap's weight is exactly 0, so only every k-th source row contributes to the result and
      // only those rows cross PCIe (one pitched copy); preproc_kernel reads only those rows as well
      // (PreFrame::row_step), so the output is bit-identical to a full upload
      // (tests/test_gpu_yolox.py::test_row_sparse_upload_is_exact).
      PreFrame probe;
      preproc_describe(&pc, ptr, h[b], w[b], st, &probe);
      const int k = probe.row_step;
      const bool sparse = g_sparse_upload && k > 1;
      if (sparse) {
        const size_t r0 = (size_t)probe.row_phase;
        TCD_CUDA_OK(cudaMemcpy2DAsync(dst + r0 * tight, (size_t)k * tight, ptr + r0 * (size_t)st, (size_t)k * (size_t)st,
                                      tight, (size_t)probe.content_h, cudaMemcpyHostToDevice, cs));
        g_h2d_bytes.fetch_add((uint64_t)probe.content_h * tight, std::memory_order_relaxed);
      } else if ((size_t)st == tight) {
        TCD_CUDA_OK(cudaMemcpyAsync(dst, ptr, (size_t)h[b] * tight, cudaMemcpyHostToDevice, cs));
        g_h2d_bytes.fetch_add((uint64_t)h[b] * tight, std::memory_order_relaxed);
      } else {
        TCD_CUDA_OK(cudaMemcpy2DAsync(dst, tight, ptr, (size_t)st, tight, (size_t)h[b],
                                      cudaMemcpyHostToDevice, cs));
        g_h2d_bytes.fetch_add((uint64_t)h[b] * tight, std::memory_order_relaxed);
      }
      ptr = dst;
      dst_stride = (int64_t)tight;
    }
    preproc_describe(&pc, ptr, h[b], w[b], dst_stride, &l.pre_host[b]);
  }
  if (cs != s.stream) {
    TCD_CUDA_OK(cudaEventRecord(l.ev_up, cs));
    TCD_CUDA_OK(cudaStreamWaitEvent(s.stream, l.ev_up, 0));
  }
  TCD_CUDA_OK(cudaEventRecord(l.ev[1], s.stream));
  TCD_CUDA_OK(cudaMemcpyAsync(l.pre_dev, l.pre_host, sizeof(PreFrame) * B, cudaMemcpyHostToDevice, s.stream));
  TCD_CUDA_OK(cudaMemcpyAsync(s.meta_dev + frame0, l.meta_host, sizeof(FrameMeta) * B, cudaMemcpyHostToDevice,
                              s.stream));
  if (!launch_pre) return TCD_OK;
  return preproc_launch(&pc, l.pre_host, l.pre_dev, B, 1, input, s.stream);
}

static int check_batch(tcd_detector* det, const uint8_t* const* frames, const int32_t* h,
                       const int32_t* w, int B) {
  TCD_REQUIRE(det && frames && h && w, "detect: null argument");
  TCD_REQUIRE(B >= 1 && B <= det->cfg.max_batch, "detect: batch %d outside [1,%d]", B, det->cfg.max_batch);
  for (int b = 0; b < B; ++b)
    TCD_REQUIRE(frames[b] && h[b] > 0 && w[b] > 0, "detect: frame %d is empty", b);
  return TCD_OK;
}

extern "C" {

tcd_status tcd_create_detector(const tcd_detector_cfg* cfg, const void* packed_weights,
                               size_t packed_bytes, int32_t device, tcd_detector** out) {
  TCD_REQUIRE(cfg && packed_weights && out, "create_detector: null argument");
  TCD_REQUIRE(cfg->max_batch >= 1 && cfg->max_batch <= 256, "create_detector: max_batch %d", cfg->max_batch);
  TCD_REQUIRE(cfg->max_out >= 1, "create_detector: max_out %d", cfg->max_out);
  TCD_CUDA_OK(cudaSetDevice(device));
  int rc = require_sm100();
  if (rc) return (tcd_status)rc;
  tcd_detector* d = new tcd_detector();
  d->cfg = *cfg;
  d->device = device;
  WeightBlob blob;
  rc = blob.parse(packed_weights, packed_bytes);
  if (rc == TCD_OK) rc = d->wts.upload(blob);
  if (rc == TCD_OK && d->wts.num_classes != cfg->num_classes) {
    set_last_error("create_detector: weights have %d classes, cfg says %d", d->wts.num_classes,
                   cfg->num_classes);
    rc = TCD_ERR_INVALID;
  }
  if (rc == TCD_OK) {
    const int H = cfg->input_h, W = cfg->input_w;
    d->n_anchors = (H / 8) * (W / 8) + (H / 16) * (W / 16) + (H / 32) * (W / 32);
    int cap = (d->n_anchors + 63) / 64 * 64;
    if (cap > 8448) cap = 8448;
    d->cap = cap;
    if (d->n_anchors >= 131072) {
      set_last_error("create_detector: %d anchors exceed the sort key range", d->n_anchors);
      rc = TCD_ERR_UNSUPPORTED;
    }
  }
  if (const char* e = getenv("TCD_ENGINES")) {
    const int v = atoi(e);
    if (v >= 1 && v <= kEngines) d->n_engines = v;
  }
  if (const char* e = getenv("TCD_LANES")) {
    const int v = atoi(e);
    if (v >= 1 && v <= kLanes) d->n_lanes = v;
  }
  d->group_cap = d->n_lanes * cfg->max_batch;
  if (rc == TCD_OK && cfg->num_classes > TCD_MAX_CLASSES) {
    set_last_error("create_detector: %d classes exceed TCD_MAX_CLASSES", cfg->num_classes);
    rc = TCD_ERR_UNSUPPORTED;
  }
  if (rc == TCD_OK) {
    const size_t tb = sizeof(tcd_class_filter) * kMaxFilterTables;
    if (cudaMalloc((void**)&d->tables_dev, tb) != cudaSuccess || cudaMemset(d->tables_dev, 0, tb) != cudaSuccess) {
      set_last_error("create_detector: filter tables: %s", cudaGetErrorString(cudaGetLastError()));
      rc = TCD_ERR_CUDA;
    }
  }
  for (int i = 0; i < d->n_engines && rc == TCD_OK; ++i) rc = slot_init(d, d->slot[i]);
  // build the max-batch plan up front so the first frame does not pay for it
  YoloxPlan* p = nullptr;
  // the group-capacity plan owns each engine's arena; the single-batch plan is what a lone detect_batch runs
  for (int i = 0; i < d->n_engines && rc == TCD_OK; ++i) rc = get_plan(d, d->slot[i], d->group_cap, &p);
  for (int i = 0; i < d->n_engines && rc == TCD_OK; ++i) rc = get_plan(d, d->slot[i], cfg->max_batch, &p);
  if (rc) {
    tcd_destroy_detector(d);
    return (tcd_status)rc;
  }
  const char* ng = getenv("TCD_NO_GRAPH");
  d->use_graph = !(ng && ng[0] == '1');
  *out = d;
  return TCD_OK;
}

void tcd_destroy_detector(tcd_detector* det) {
  if (!det) return;
  cudaSetDevice(det->device);
  cudaDeviceSynchronize();
  for (int i = 0; i < kEngines; ++i) slot_free(det->slot[i]);
  det->wts.release();
  if (det->tables_dev) cudaFree(det->tables_dev);
  delete det;
}

tcd_status tcd_detector_set_class_filter(tcd_detector* det, int32_t table, const tcd_class_filter* f) {
  TCD_REQUIRE(det && f, "set_class_filter: null argument");
  TCD_REQUIRE(table >= 1 && table < kMaxFilterTables, "set_class_filter: table %d outside [1,%d)", table,
              kMaxFilterTables);
  TCD_USE(det);
  // ordered after the batches already queued (they keep the table they were submitted under)
  for (int i = 0; i < det->n_engines; ++i) TCD_CUDA_OK(cudaStreamSynchronize(det->slot[i].stream));
  TCD_CUDA_OK(cudaMemcpy(det->tables_dev + table, f, sizeof(*f), cudaMemcpyHostToDevice));
  return TCD_OK;
}

tcd_status tcd_detector_filter_rows(tcd_detector* det, int32_t table, const float* rows_host, int32_t n,
                                    int32_t orig_h, int32_t orig_w, int32_t* keep_host) {
  TCD_REQUIRE(det && rows_host && keep_host && n >= 0, "filter_rows: bad argument");
  TCD_REQUIRE(table >= 1 && table < kMaxFilterTables, "filter_rows: table %d outside [1,%d)", table, kMaxFilterTables);
  if (n == 0) return TCD_OK;
  TCD_USE(det);
  float* rows = nullptr;
  int* keep = nullptr;
  TCD_CUDA_OK(cudaMalloc((void**)&rows, sizeof(float) * 7 * n));
  TCD_CUDA_OK(cudaMalloc((void**)&keep, sizeof(int) * n));
  cudaStream_t st = det->slot[0].stream;
  TCD_CUDA_OK(cudaMemcpyAsync(rows, rows_host, sizeof(float) * 7 * n, cudaMemcpyHostToDevice, st));
  const float rw = (float)((double)orig_w / (double)det->cfg.input_w), rh = (float)((double)orig_h / (double)det->cfg.input_h);
  int rc = class_filter_rows_launch(det->tables_dev, table, rows, n, rw, rh, keep, st);
  cudaError_t e = cudaMemcpyAsync(keep_host, keep, sizeof(int) * n, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(rows);
  cudaFree(keep);
  if (rc == TCD_OK && e != cudaSuccess) {
    set_last_error("filter_rows: %s", cudaGetErrorString(e));
    rc = TCD_ERR_CUDA;
  }
  return (tcd_status)rc;
}

tcd_status tcd_detector_set_thresholds(tcd_detector* det, float conf_thre, float nms_thre) {
  TCD_REQUIRE(det, "set_thresholds: null detector");
  TCD_USE(det);
  det->thr_host[0] = conf_thre;
  det->thr_host[1] = nms_thre;
  for (int i = 0; i < det->n_engines; ++i) {
    // ordered after work already queued on the slot; never touches the captured graphs
    TCD_CUDA_OK(cudaStreamSynchronize(det->slot[i].stream));
    TCD_CUDA_OK(cudaMemcpy(det->slot[i].thr_dev, det->thr_host, 2 * sizeof(float), cudaMemcpyHostToDevice));
    TCD_CUDA_OK(cudaMemcpy(det->slot[i].thr_dev + 2, det->thr_host, sizeof(float), cudaMemcpyHostToDevice));
  }
  return TCD_OK;
}

int32_t tcd_detector_num_anchors(const tcd_detector* det) { return det ? det->n_anchors : 0; }

tcd_status tcd_detect_batch_async(tcd_detector* det, int32_t slot, const uint8_t* const* frames,
                                  const int32_t* h, const int32_t* w, const int64_t* row_stride,
                                  int32_t B, int32_t frames_on_device) {
  return tcd_detector_submit(det, slot, frames, h, w, row_stride, B, frames_on_device, nullptr);
}

// network + decode/NMS for everything the engine's group holds, then every lane's rows to its own tables
static int launch_group(tcd_detector* det, Slot& s) {
  if (s.group_n == 0) return TCD_OK;
  const int Bg = s.group_frames;
  YoloxPlan* p = nullptr;
  int rc = get_plan(det, s, Bg, &p);
  if (rc) return rc;
  rc = run_net_post(det, s, p, Bg, false);
  if (rc) return rc;
  const size_t row = sizeof(float) * 7 * det->cfg.max_out;
  for (int i = 0; i < s.group_n; ++i) {
    Lane& l = *s.group[i];
    const int B = l.pending_B, f0 = l.frame0;
    TCD_CUDA_OK(cudaEventRecord(l.ev[3], s.stream));
    // the engine's table belongs to the next group queued on this stream: the lane keeps its own copies
    TCD_CUDA_OK(cudaMemcpyAsync(l.out_pinned, (const uint8_t*)s.pb.out_dets + row * f0, row * B, cudaMemcpyDeviceToHost,
                                s.stream));
    TCD_CUDA_OK(cudaMemcpyAsync(l.cnt_pinned, s.pb.out_counts + f0, sizeof(int) * B, cudaMemcpyDeviceToHost, s.stream));
    TCD_CUDA_OK(cudaMemcpyAsync(l.tab_dets, (const uint8_t*)s.pb.out_dets + row * f0, row * B, cudaMemcpyDeviceToDevice,
                                s.stream));
    TCD_CUDA_OK(cudaMemcpyAsync(l.tab_counts, s.pb.out_counts + f0, sizeof(int) * B, cudaMemcpyDeviceToDevice, s.stream));
    TCD_CUDA_OK(cudaEventRecord(l.ev[4], s.stream));
    l.launched = true;
  }
  s.group_n = 0;
  s.group_frames = 0;
  return TCD_OK;
}

tcd_status tcd_detector_submit(tcd_detector* det, int32_t slot, const uint8_t* const* frames,
                               const int32_t* h, const int32_t* w, const int64_t* row_stride,
                               int32_t B, int32_t frames_on_device, const int32_t* filter_table) {
  int rc = check_batch(det, frames, h, w, B);
  if (rc) return (tcd_status)rc;
  TCD_USE(det);
  const int n_slots = det->n_engines * det->n_lanes;
  TCD_REQUIRE(slot >= 0 && slot < n_slots, "detect: slot %d outside [0,%d)", slot, n_slots);
  Slot& s = det->slot[slot / det->n_lanes];
  Lane& l = s.lane[slot % det->n_lanes];
  TCD_REQUIRE(l.pending_B == 0, "detect: slot %d still has a batch in flight (wait for it first)", slot);
  if (s.group_frames + B > det->group_cap) {   // cannot happen with <= n_lanes submits of <= max_batch; defensive
    rc = launch_group(det, s);
    if (rc) return (tcd_status)rc;
  }
  YoloxPlan* p = nullptr;
  rc = get_plan(det, s, det->group_cap, &p);   // only its input address is used here (shared by all of the engine's plans)
  if (rc) return (tcd_status)rc;
  TCD_CUDA_OK(cudaEventRecord(l.ev[0], s.stream));
  const int f0 = s.group_frames;
  rc = enqueue_input(det, s, l, p, frames, h, w, row_stride, B, frames_on_device, true, filter_table, f0);
  if (rc) return (tcd_status)rc;
  TCD_CUDA_OK(cudaEventRecord(l.ev[2], s.stream));
  l.pending_B = B;
  l.frame0 = f0;
  l.launched = false;
  s.group[s.group_n++] = &l;
  s.group_frames += B;
  // launch now when no further max_batch submit fits this group (under load: n_lanes batches per launch); otherwise
  // the launch waits for more lanes or for the first wait() on one of them
  if (s.group_n == det->n_lanes || s.group_frames + det->cfg.max_batch > det->group_cap) {
    rc = launch_group(det, s);
    if (rc) return (tcd_status)rc;
  }
  return TCD_OK;
}

tcd_status tcd_detect_batch_wait(tcd_detector* det, int32_t slot, float* out_dets, int32_t* out_counts) {
  TCD_REQUIRE(det && out_dets && out_counts && slot >= 0 && slot < det->n_engines * det->n_lanes,
              "detect_wait: bad argument");
  TCD_USE(det);
  Slot& eng = det->slot[slot / det->n_lanes];
  Lane& s = eng.lane[slot % det->n_lanes];
  TCD_REQUIRE(s.pending_B > 0, "detect_wait: nothing pending on slot %d", slot);
  if (!s.launched) {   // nobody filled the group: run what it holds
    int rc = launch_group(det, eng);
    if (rc) return (tcd_status)rc;
  }
  TCD_CUDA_OK(cudaEventSynchronize(s.ev[4]));   // not the stream: the engine's next group may be queued behind
  const int B = s.pending_B;
  s.pending_B = 0;
  const size_t row = (size_t)7 * det->cfg.max_out;
  for (int b = 0; b < B; ++b) {
    const int n = s.cnt_pinned[b];
    out_counts[b] = n;
    memcpy(out_dets + b * row, s.out_pinned + b * row, sizeof(float) * 7 * (size_t)n);
  }
  float t;
  // phases: [0]=h2d (ev0->ev1) [1]=preprocess (ev1->ev2) [2]=network+[3]post (ev2->ev3) [4]=d2h
  if (cudaEventElapsedTime(&t, s.ev[0], s.ev[1]) == cudaSuccess) det->last_ms[0] = t;
  if (cudaEventElapsedTime(&t, s.ev[1], s.ev[2]) == cudaSuccess) det->last_ms[1] = t;
  if (cudaEventElapsedTime(&t, s.ev[2], s.ev[3]) == cudaSuccess) det->last_ms[2] = t;
  det->last_ms[3] = 0.f;
  if (cudaEventElapsedTime(&t, s.ev[3], s.ev[4]) == cudaSuccess) det->last_ms[4] = t;
  return TCD_OK;
}

tcd_status tcd_detect_batch(tcd_detector* det, const uint8_t* const* frames, const int32_t* h,
                            const int32_t* w, const int64_t* row_stride, int32_t B,
                            int32_t frames_on_device, float* out_dets, int32_t* out_counts) {
  tcd_status rc = tcd_detect_batch_async(det, 0, frames, h, w, row_stride, B, frames_on_device);
  if (rc) return rc;
  return tcd_detect_batch_wait(det, 0, out_dets, out_counts);
}

tcd_status tcd_detector_last_timing(const tcd_detector* det, float out_ms[5]) {
  TCD_REQUIRE(det && out_ms, "last_timing: null argument");
  for (int i = 0; i < 5; ++i) out_ms[i] = det->last_ms[i];
  return TCD_OK;
}

tcd_status tcd_forward(tcd_detector* det, const uint8_t* const* frames, const int32_t* h,
                       const int32_t* w, const int64_t* row_stride, int32_t B,
                       int32_t frames_on_device, float* out_pred_host) {
  int rc = check_batch(det, frames, h, w, B);
  if (rc) return (tcd_status)rc;
  TCD_REQUIRE(out_pred_host, "forward: null output");
  TCD_USE(det);
  Slot& s = det->slot[0];
  TCD_REQUIRE(s.group_n == 0 && s.lane[0].pending_B == 0, "engine 0 has batches in flight: wait for them first");
  YoloxPlan* p = nullptr;
  rc = get_plan(det, s, B, &p);
  if (rc) return (tcd_status)rc;
  rc = enqueue_input(det, s, s.lane[0], p, frames, h, w, row_stride, B, frames_on_device);
  if (rc) return (tcd_status)rc;
  rc = run_net_post(det, s, p, B, true);
  if (rc) return (tcd_status)rc;
  TCD_CUDA_OK(cudaMemcpyAsync(out_pred_host, s.pb.decoded,
                              sizeof(float) * (size_t)B * det->n_anchors * (5 + det->cfg.num_classes),
                              cudaMemcpyDeviceToHost, s.stream));
  TCD_CUDA_OK(cudaStreamSynchronize(s.stream));
  return TCD_OK;
}

tcd_status tcd_postprocess(tcd_detector* det, const float* pred_host, int32_t B, float* out_dets,
                           int32_t* out_counts) {
  TCD_REQUIRE(det && pred_host && out_dets && out_counts, "postprocess: null argument");
  TCD_REQUIRE(B >= 1 && B <= det->cfg.max_batch, "postprocess: batch %d", B);
  TCD_USE(det);
  Slot& s = det->slot[0];
  TCD_REQUIRE(s.group_n == 0 && s.lane[0].pending_B == 0, "engine 0 has batches in flight: wait for them first");
  YoloxPlan* p = nullptr;
  int rc = get_plan(det, s, det->cfg.max_batch, &p);
  if (rc) return (tcd_status)rc;
  const size_t n = (size_t)B * det->n_anchors * (5 + det->cfg.num_classes);
  TCD_CUDA_OK(cudaMemcpyAsync(s.pred_in, pred_host, n * sizeof(float), cudaMemcpyHostToDevice, s.stream));
  TCD_CUDA_OK(cudaMemsetAsync(s.meta_dev, 0, sizeof(FrameMeta) * B, s.stream));   // table 0: no class filter
  rc = postproc_launch(post_for(s, p, B, s.pred_in, false), 0, s.stream);
  if (rc) return (tcd_status)rc;
  Lane& l = s.lane[0];
  TCD_REQUIRE(l.pending_B == 0, "postprocess: slot 0 has a batch in flight");
  TCD_CUDA_OK(cudaMemcpyAsync(l.out_pinned, s.pb.out_dets, sizeof(float) * 7 * det->cfg.max_out * B,
                              cudaMemcpyDeviceToHost, s.stream));
  TCD_CUDA_OK(cudaMemcpyAsync(l.cnt_pinned, s.pb.out_counts, sizeof(int) * B, cudaMemcpyDeviceToHost,
                              s.stream));
  TCD_CUDA_OK(cudaStreamSynchronize(s.stream));
  const size_t row = (size_t)7 * det->cfg.max_out;
  for (int b = 0; b < B; ++b) {
    out_counts[b] = l.cnt_pinned[b];
    memcpy(out_dets + b * row, l.out_pinned + b * row, sizeof(float) * 7 * (size_t)l.cnt_pinned[b]);
  }
  return TCD_OK;
}

void* tcd_detector_stream(tcd_detector* det, int32_t slot) {
  if (!det || slot < 0 || slot >= det->n_engines * det->n_lanes) return nullptr;
  return (void*)det->slot[slot / det->n_lanes].stream;
}

int32_t tcd_detector_num_slots(const tcd_detector* det) { return det ? det->n_engines * det->n_lanes : 0; }

tcd_status tcd_detector_result_device(tcd_detector* det, int32_t slot, float** out_dets, int32_t** out_counts) {
  TCD_REQUIRE(det && slot >= 0 && slot < det->n_engines * det->n_lanes && out_dets && out_counts,
              "result_device: bad argument");
  Lane& l = det->slot[slot / det->n_lanes].lane[slot % det->n_lanes];
  *out_dets = l.tab_dets;
  *out_counts = l.tab_counts;
  return TCD_OK;
}

// In-kernel timeline of ONE conv launch of the layer plan (op index as in tcd_detector_profile,
// i.e. 1-based over the plan): the kernel stamps clock64/globaltimer at its pipeline milestones for
// the first and the last CTA of the grid (slot meaning: conv_umma.cuh).  Runs the op `reps` times
// eagerly on slot 0 and returns the stamps of the last run.
tcd_status tcd_detector_trace_op(tcd_detector* det, int32_t B, int32_t idx, int32_t reps,
                                 uint64_t out128[128], int32_t* grid, int32_t* smem_bytes) {
  TCD_REQUIRE(det && out128, "trace_op: null argument");
  TCD_USE(det);
  Slot& s = det->slot[0];
  TCD_REQUIRE(s.group_n == 0 && s.lane[0].pending_B == 0, "engine 0 has batches in flight: wait for them first");
  YoloxPlan* p = nullptr;
  int rc = get_plan(det, s, B, &p);
  if (rc) return (tcd_status)rc;
  TCD_REQUIRE(idx >= 1 && idx <= (int)p->ops.size() && p->ops[idx - 1].kind == OP_CONV,
              "trace_op: op %d is not a conv", idx);
  ConvPlan cp = p->ops[idx - 1].conv;
  unsigned long long* buf = nullptr;
  TCD_CUDA_OK(cudaMalloc((void**)&buf, 128 * sizeof(unsigned long long)));
  TCD_CUDA_OK(cudaMemset(buf, 0, 128 * sizeof(unsigned long long)));
  cp.args.trace = buf;
  for (int r = 0; r < (reps < 1 ? 1 : reps) && rc == TCD_OK; ++r) {
    rc = conv_plan_launch(cp, s.stream);
    if (rc == TCD_OK && cudaStreamSynchronize(s.stream) != cudaSuccess) rc = TCD_ERR_CUDA;
  }
  if (rc == TCD_OK)
    cudaMemcpy(out128, buf, 128 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
  cudaFree(buf);
  if (grid) *grid = (int32_t)cp.grid.x;
  if (smem_bytes) *smem_bytes = cp.smem_bytes;
  return (tcd_status)rc;
}

// Eager (graph-less) pass with a CUDA-event pair around every kernel of one detect_batch:
// op 0 = preprocess, then the layer plan in launch order, then decode_filter / sort /
// nms_mask / nms_reduce.  op_ms[i] = mean device time of op i over `iters` passes.
tcd_status tcd_detector_profile(tcd_detector* det, const uint8_t* const* frames, const int32_t* h,
                                const int32_t* w, const int64_t* row_stride, int32_t B,
                                int32_t frames_on_device, int32_t iters, float* op_ms, int32_t cap,
                                int32_t* n_ops) {
  int rc = check_batch(det, frames, h, w, B);
  if (rc) return (tcd_status)rc;
  TCD_REQUIRE(iters >= 1 && op_ms && n_ops, "profile: bad argument");
  TCD_USE(det);
  Slot& s = det->slot[0];
  TCD_REQUIRE(s.group_n == 0 && s.lane[0].pending_B == 0, "engine 0 has batches in flight: wait for them first");
  YoloxPlan* p = nullptr;
  rc = get_plan(det, s, B, &p);
  if (rc) return (tcd_status)rc;
  const int n = 1 + (int)p->ops.size() + 4;
  *n_ops = n;
  TCD_REQUIRE(cap >= n, "profile: op_ms holds %d entries, %d needed", cap, n);
  std::vector<cudaEvent_t> ev(2 * (size_t)n);
  for (auto& e : ev) TCD_CUDA_OK(cudaEventCreate(&e));
  std::vector<double> acc(n, 0.0);
  for (int it = 0; it < iters && rc == TCD_OK; ++it) {
    // inputs are staged outside the per-op brackets; op 0 brackets the preprocess kernel only
    tcd_preproc_cfg pc = det->cfg.pre;
    pc.out_h = det->cfg.input_h;
    pc.out_w = det->cfg.input_w;
    rc = enqueue_input(det, s, s.lane[0], p, frames, h, w, row_stride, B, frames_on_device, false);
    if (rc) break;
    // like the convs: kProfileReps back-to-back launches per event pair (one launch alone measures the host's
    // launch latency as much as the 10-20 us kernel)
    cudaEventRecord(ev[0], s.stream);
    for (int r = 0; r < kProfileReps && rc == TCD_OK; ++r)
      rc = preproc_launch(&pc, s.lane[0].pre_host, s.lane[0].pre_dev, B, 1, p->input, s.stream);
    cudaEventRecord(ev[1], s.stream);
    for (size_t i = 0; i < p->ops.size() && rc == TCD_OK; ++i) {
      // conv launches are bracketed kProfileReps at a time (same launch repeated back to back, as
      // the graph issues them) so the event pair does not dominate a 5 us kernel
      const int reps = p->ops[i].kind == OP_CONV ? kProfileReps : 1;
      cudaEventRecord(ev[2 * (i + 1)], s.stream);
      for (int r = 0; r < reps && rc == TCD_OK; ++r) rc = p->run_one((int)i, s.stream);
      cudaEventRecord(ev[2 * (i + 1) + 1], s.stream);
    }
    if (rc) break;
    // layers with an in-place residual have accumulated kProfileReps times: one clean pass so that decode / NMS
    // are timed on the real prediction, not on an inflated candidate set
    cudaMemsetAsync(s.maybe_count, 0, sizeof(int) * B, s.stream);
    rc = p->run(s.stream);
    if (rc) break;
    const int base = 1 + (int)p->ops.size();
    rc = postproc_launch(post_for(s, p, B, p->pred_raw, false), 1, s.stream, &ev[2 * (size_t)base]);
    if (rc) break;
    cudaError_t e = cudaStreamSynchronize(s.stream);
    if (e != cudaSuccess) {
      set_last_error("profile: %s", cudaGetErrorString(e));
      rc = TCD_ERR_CUDA;
      break;
    }
    for (int i = 0; i < n; ++i) {
      float t = 0.f;
      cudaEventElapsedTime(&t, ev[2 * (size_t)i], ev[2 * (size_t)i + 1]);
      const bool is_conv = i >= 1 && i <= (int)p->ops.size() && p->ops[i - 1].kind == OP_CONV;
      acc[i] += (is_conv || i == 0) ? t / kProfileReps : t;
    }
  }
  for (auto& e : ev) cudaEventDestroy(e);
  if (rc) return (tcd_status)rc;
  for (int i = 0; i < n; ++i) op_ms[i] = (float)(acc[i] / iters);
  return TCD_OK;
}

static const char* kPostNames[4] = {"decode_filter", "sort", "nms_mask", "nms_reduce"};

const char* tcd_detector_op_name(tcd_detector* det, int32_t B, int32_t idx) {
  if (!det) return "";
  auto it = det->slot[0].plans.find(B);
  if (it == det->slot[0].plans.end() || idx < 0) return "";
  const YoloxPlan* p = it->second;
  if (idx == 0) return "preprocess";
  if (idx <= (int)p->ops.size()) return p->ops[idx - 1].name.c_str();
  idx -= 1 + (int)p->ops.size();
  return idx < 4 ? kPostNames[idx] : "";
}

/* kind: 0 = conv (tcgen05), 1 = other kernel */
int32_t tcd_detector_op_kind(tcd_detector* det, int32_t B, int32_t idx) {
  if (!det) return -1;
  auto it = det->slot[0].plans.find(B);
  if (it == det->slot[0].plans.end()) return -1;
  const YoloxPlan* p = it->second;
  if (idx >= 1 && idx <= (int)p->ops.size()) return p->ops[idx - 1].kind == OP_CONV ? 0 : 1;
  return 1;
}

double tcd_detector_op_flops(tcd_detector* det, int32_t B, int32_t idx) {
  if (!det) return 0.0;
  auto it = det->slot[0].plans.find(B);
  if (it == det->slot[0].plans.end()) return 0.0;
  const YoloxPlan* p = it->second;
  if (idx >= 1 && idx <= (int)p->ops.size() && p->ops[idx - 1].kind == OP_CONV) return p->ops[idx - 1].conv.flops;
  return 0.0;
}

tcd_status tcd_detector_read_activation(tcd_detector* det, const char* name, float* out_host,
                                        size_t out_elems, int32_t* shape4) {
  TCD_REQUIRE(det && name && shape4, "read_activation: null argument");
  TCD_USE(det);
  Slot& s = det->slot[0];
  const YoloxPlan* p = s.last_plan;   // the plans of an engine overlay one arena: only the last one run is readable
  TCD_REQUIRE(p != nullptr && p->acts.count(name), "read_activation: no buffer named %s in the last plan run", name);
  const Act& a = p->acts.at(name);
  shape4[0] = p->B; shape4[1] = a.H; shape4[2] = a.W; shape4[3] = a.C;
  const size_t n = (size_t)p->B * a.H * a.W * a.C;
  if (!out_host) return TCD_OK;
  TCD_REQUIRE(out_elems >= n, "read_activation: output too small (%zu < %zu)", out_elems, n);
  std::vector<uint16_t> tmp(n);
  TCD_CUDA_OK(cudaStreamSynchronize(s.stream));
  TCD_CUDA_OK(cudaMemcpy(tmp.data(), a.p, n * 2, cudaMemcpyDeviceToHost));
  for (size_t i = 0; i < n; ++i) {
    uint32_t u = (uint32_t)tmp[i] << 16;
    memcpy(out_host + i, &u, 4);
  }
  return TCD_OK;
}

}  // extern "C"
