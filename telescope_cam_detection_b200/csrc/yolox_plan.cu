// YoloxPlan::build -- lays the YOLOX network out over one activation arena for a fixed
// batch size and emits the launch list.  Topology per SURVEY App. A (upstream
// yolox/models/{darknet,yolo_pafpn,yolo_head}.py); widths/depths are read from the weights.
#include "yolox_net.cuh"

namespace tcd {

namespace {

struct Builder {
  const YoloxWeights& wts;
  YoloxPlan& plan;
  int B;
  size_t cursor = 0;
  bool dry = true;  // first pass sizes the arena
  int rc = TCD_OK;
  int cur_branch = 0;

  Builder(const YoloxWeights& w, YoloxPlan& p, int b) : wts(w), plan(p), B(b) {}

  Act alloc(const std::string& name, int H, int W, int C, size_t elt = 2) {
    Act a;
    a.H = H; a.W = W; a.C = C;
    const size_t bytes = ((size_t)B * H * W * C * elt + 1023) & ~(size_t)1023;
    a.p = dry ? nullptr : (uint8_t*)plan.arena + cursor;
    cursor += bytes;
    if (!dry) plan.acts[name] = a;
    return a;
  }

  const DevConv* dc(const std::string& n) {
    auto it = wts.convs.find(n);
    if (it == wts.convs.end()) {
      set_last_error("plan: no packed conv named %s", n.c_str());
      rc = TCD_ERR_INVALID;
      return nullptr;
    }
    return &it->second;
  }

  // y[.., ycoff:ycoff+cout] = act(conv(x[.., xcoff:xcoff+cin])) (+ resid)
  void conv(const std::string& wname, const Act& x, int xcoff, const Act& y, int ycoff, int stride,
            const Act* resid = nullptr, int rcoff = 0, int groups = 1) {
    if (rc) return;
    const DevConv* w = dc(wname);
    if (!w) return;
    if (dry) return;
    ConvDesc d;
    d.x = x.p; d.N = B; d.H = x.H; d.W = x.W; d.x_ctot = x.C; d.x_coff = xcoff; d.cin = w->cin_pad;
    d.w = w->w; d.bias = w->bias; d.cout = w->cout; d.cout_pad = w->cout_pad; d.k = w->k;
    d.stride = stride; d.silu = 1; d.groups = groups;
    d.y = y.p; d.y_ctot = y.C; d.y_coff = ycoff;
    if (resid) { d.resid = resid->p; d.r_ctot = resid->C; d.r_coff = rcoff; }
    NetOp op;
    op.kind = OP_CONV;
    op.name = wname;
    op.branch = cur_branch;
    rc = conv_plan_build(d, &op.conv);
    if (rc) return;
    plan.conv_flops += op.conv.flops;
    plan.ops.push_back(op);
  }

  void pred(const std::string& wname, const Act& x, int col, int row_off, int rows_per_img, double flops = 0) {
    if (rc) return;
    const DevConv* w = dc(wname);
    if (!w || dry) return;
    ConvDesc d;
    d.x = x.p; d.N = B; d.H = x.H; d.W = x.W; d.x_ctot = x.C; d.x_coff = 0; d.cin = w->cin_pad;
    d.w = w->w; d.bias = w->bias; d.cout = w->cout; d.cout_pad = w->cout_pad; d.k = 1; d.stride = 1;
    d.silu = 0;
    d.out_f32 = plan.pred_raw; d.out_ld = 5 + plan.num_classes; d.out_col = col;
    d.rows_per_img = rows_per_img; d.row_off = row_off;
    d.emit_thr = plan.emit_thr; d.maybe_count = plan.maybe_count; d.maybe_idx = plan.maybe_idx;
    d.maybe_cap = plan.maybe_cap; d.emit_col = 4 - col;   // the fused prediction rows are [reg 0-3 | obj 4 | cls 5..]
    NetOp op;
    op.kind = OP_CONV;
    op.name = wname;
    op.branch = cur_branch;
    rc = conv_plan_build(d, &op.conv);
    if (rc) return;
    if (flops > 0) op.conv.flops = flops;  // block-diagonal weights: count the non-zero half only
    plan.conv_flops += op.conv.flops;
    plan.ops.push_back(op);
  }

  // CSPLayer: x (cin at xcoff of xbuf) -> out slice.  h = cout/2.
  void csp(const std::string& name, const Act& x, int xcoff, int n, bool shortcut, const Act& out,
           int ocoff, int H, int W) {
    if (rc) return;
    const DevConv* c12 = dc(name + ".conv12");
    if (!c12) return;
    const int h = c12->cout / 2;
    Act T = alloc(name + ".T", H, W, 2 * h);
    Act U = alloc(name + ".U", H, W, h);
    conv(name + ".conv12", x, xcoff, T, 0, 1);
    for (int i = 0; i < n; ++i) {
      const std::string m = name + ".m." + std::to_string(i);
      conv(m + ".conv1", T, 0, U, 0, 1);
      conv(m + ".conv2", U, 0, T, 0, 1, shortcut ? &T : nullptr, 0);
    }
    conv(name + ".conv3", T, 0, out, ocoff, 1);
  }

  void upsample(const Act& s, int scoff, int C, const Act& d, int dcoff) {
    if (rc || dry) return;
    NetOp op;
    op.kind = OP_UPSAMPLE;
    op.name = "upsample";
    op.src = s.p; op.dst = d.p; op.N = B; op.H = s.H; op.W = s.W; op.C = C;
    op.s_ctot = s.C; op.s_coff = scoff; op.d_ctot = d.C; op.d_coff = dcoff;
    plan.ops.push_back(op);
  }

  int count_bottlenecks(const std::string& name) {
    int n = 0;
    while (wts.convs.count(name + ".m." + std::to_string(n) + ".conv1")) ++n;
    return n;
  }

  void run_pass() {
    cursor = 0;
    const int IH = plan.in_h, IW = plan.in_w;
    const DevConv* stem = dc("stem");
    if (!stem) return;
    const int bc = stem->cout;  // 32 for YOLOX-S
    const int H2 = IH / 2, W2 = IW / 2, H4 = IH / 4, W4 = IW / 4, H8 = IH / 8, W8 = IW / 8;
    const int H16 = IH / 16, W16 = IW / 16, H32 = IH / 32, W32 = IW / 32;
    const int A = H8 * W8 + H16 * W16 + H32 * W32;
    plan.n_anchors = A;
    plan.num_classes = wts.num_classes;

    Act in0 = alloc("input", H2, W2, 16);
    if (!dry) plan.input = in0.p;
    Act predbuf = alloc("pred_raw", 1, A, 5 + plan.num_classes, 4);
    if (!dry) plan.pred_raw = (float*)predbuf.p;

    Act a_stem = alloc("stem", H2, W2, bc);
    conv("stem", in0, 0, a_stem, 0, 1);
    Act d2a = alloc("dark2.0", H4, W4, bc * 2);
    conv("dark2.0", a_stem, 0, d2a, 0, 2);
    Act d2 = alloc("dark2", H4, W4, bc * 2);
    csp("dark2.1", d2a, 0, count_bottlenecks("dark2.1"), true, d2, 0, H4, W4);

    Act cat_p3 = alloc("cat_p3", H8, W8, bc * 8);     // [up(fpn_out1) | dark3]
    Act cat_p4 = alloc("cat_p4", H16, W16, bc * 16);  // [up(fpn_out0) | dark4]
    Act cat_n3 = alloc("cat_n3", H16, W16, bc * 8);   // [bu_conv2    | fpn_out1]
    Act cat_n4 = alloc("cat_n4", H32, W32, bc * 16);  // [bu_conv1    | fpn_out0]

    Act d3a = alloc("dark3.0", H8, W8, bc * 4);
    conv("dark3.0", d2, 0, d3a, 0, 2);
    csp("dark3.1", d3a, 0, count_bottlenecks("dark3.1"), true, cat_p3, bc * 4, H8, W8);
    Act d4a = alloc("dark4.0", H16, W16, bc * 8);
    conv("dark4.0", cat_p3, bc * 4, d4a, 0, 2);
    csp("dark4.1", d4a, 0, count_bottlenecks("dark4.1"), true, cat_p4, bc * 8, H16, W16);
    Act d5a = alloc("dark5.0", H32, W32, bc * 16);
    conv("dark5.0", cat_p4, bc * 8, d5a, 0, 2);
    Act spp = alloc("spp_cat", H32, W32, bc * 32);
    conv("dark5.1.conv1", d5a, 0, spp, 0, 1);
    if (!dry && !rc) {
      NetOp op;
      op.kind = OP_SPP;
      op.name = "spp_pool";
      op.dst = spp.p; op.N = B; op.H = H32; op.W = W32; op.C = bc * 8; op.d_ctot = bc * 32;
      plan.ops.push_back(op);
    }
    Act d5b = alloc("dark5.1", H32, W32, bc * 16);
    conv("dark5.1.conv2", spp, 0, d5b, 0, 1);
    Act d5 = alloc("dark5", H32, W32, bc * 16);
    csp("dark5.2", d5b, 0, count_bottlenecks("dark5.2"), false, d5, 0, H32, W32);

    // PAFPN
    conv("lateral_conv0", d5, 0, cat_n4, bc * 8, 1);            // fpn_out0
    upsample(cat_n4, bc * 8, bc * 8, cat_p4, 0);
    Act f_out0 = alloc("f_out0", H16, W16, bc * 8);
    csp("C3_p4", cat_p4, 0, count_bottlenecks("C3_p4"), false, f_out0, 0, H16, W16);
    conv("reduce_conv1", f_out0, 0, cat_n3, bc * 4, 1);         // fpn_out1
    upsample(cat_n3, bc * 4, bc * 4, cat_p3, 0);
    Act pan2 = alloc("pan_out2", H8, W8, bc * 4);
    csp("C3_p3", cat_p3, 0, count_bottlenecks("C3_p3"), false, pan2, 0, H8, W8);
    if (!dry && !rc && !plan.ops.empty()) plan.ops.back().fork_branch = 1;   // P3 ready: head level 0 may start
    conv("bu_conv2", pan2, 0, cat_n3, 0, 2);
    Act pan1 = alloc("pan_out1", H16, W16, bc * 8);
    csp("C3_n3", cat_n3, 0, count_bottlenecks("C3_n3"), false, pan1, 0, H16, W16);
    if (!dry && !rc && !plan.ops.empty()) plan.ops.back().fork_branch = 2;   // P4 ready: head level 1 may start
    conv("bu_conv1", pan1, 0, cat_n4, 0, 2);
    Act pan0 = alloc("pan_out0", H32, W32, bc * 16);
    csp("C3_n4", cat_n4, 0, count_bottlenecks("C3_n4"), false, pan0, 0, H32, W32);

    // decoupled head
    const Act* feats[3] = {&pan2, &pan1, &pan0};
    const int strides[3] = {8, 16, 32};
    int row_off = 0;
    for (int l = 0; l < 3; ++l) {
      const std::string s = std::to_string(l);
      const Act& f = *feats[l];
      cur_branch = l < 2 ? l + 1 : 0;
      const DevConv* hs = dc("head.stem." + s);
      if (!hs) return;
      const int hc = hs->cout;  // 128
      Act st = alloc("head.stem." + s, f.H, f.W, hc);
      conv("head.stem." + s, f, 0, st, 0, 1);
      Act ha = alloc("head.conv_a." + s, f.H, f.W, 2 * hc);
      conv("head.conv_a." + s, st, 0, ha, 0, 1);
      Act hfeat = alloc("head.feat." + s, f.H, f.W, 2 * hc);   // [cls_feat | reg_feat]
      conv("head.b." + s, ha, 0, hfeat, 0, 1, nullptr, 0, 2);
      pred("head.pred." + s, hfeat, 0, row_off, A,
           2.0 * B * f.H * f.W * (double)(5 + plan.num_classes) * hc);
      plan.lvl_start[l] = row_off;
      plan.lvl_w[l] = f.W;
      plan.lvl_stride[l] = strides[l];
      row_off += f.H * f.W;
    }
  }
};

}  // namespace

int YoloxPlan::build(const YoloxWeights& wts, int B_, int in_h_, int in_w_, void* ext_arena, size_t ext_bytes) {
  TCD_REQUIRE(in_h_ % 32 == 0 && in_w_ % 32 == 0 && in_h_ >= 64 && in_w_ >= 64,
              "plan: input size %dx%d must be a multiple of 32", in_h_, in_w_);
  B = B_;
  in_h = in_h_;
  in_w = in_w_;
  ops.clear();
  acts.clear();
  conv_flops = 0;
  Builder b(wts, *this, B_);
  b.dry = true;
  b.run_pass();
  if (b.rc) return b.rc;
  arena_bytes = b.cursor;
  if (ext_arena) {
    TCD_REQUIRE(arena_bytes <= ext_bytes, "plan: batch %d needs %zu B of arena, the engine's holds %zu", B_, arena_bytes,
                ext_bytes);
    arena = ext_arena;
    owns_arena = false;
  } else {
    TCD_CUDA_OK(cudaMalloc(&arena, arena_bytes));
    TCD_CUDA_OK(cudaMemset(arena, 0, arena_bytes));
    owns_arena = true;
  }
  b.dry = false;
  b.run_pass();
  return b.rc;
}

int YoloxPlan::run(cudaStream_t stream, const PlanBranches* br) const {
  bool used[3] = {false, false, false};
  for (size_t i = 0; i < ops.size(); ++i) {
    const NetOp& op = ops[i];
    cudaStream_t st = stream;
    if (br && op.branch > 0) {
      st = br->aux[op.branch - 1];
      if (!used[op.branch]) {   // first op of the branch: wait for the trunk op that produced its input
        TCD_CUDA_OK(cudaStreamWaitEvent(st, br->fork[op.branch - 1], 0));
        used[op.branch] = true;
      }
    }
    int rc = run_one((int)i, st);
    if (rc) return rc;
    if (br && op.fork_branch > 0) TCD_CUDA_OK(cudaEventRecord(br->fork[op.fork_branch - 1], stream));
  }
  if (br)
    for (int b = 1; b <= 2; ++b)
      if (used[b]) {
        TCD_CUDA_OK(cudaEventRecord(br->join[b - 1], br->aux[b - 1]));
        TCD_CUDA_OK(cudaStreamWaitEvent(stream, br->join[b - 1], 0));
      }
  return TCD_OK;
}

int YoloxPlan::run_one(int i, cudaStream_t stream) const {
  {
    const NetOp& op = ops[(size_t)i];
    int rc = TCD_OK;
    switch (op.kind) {
      case OP_CONV:
        rc = conv_plan_launch(op.conv, stream);
        break;
      case OP_UPSAMPLE:
        rc = upsample2x_launch(op.src, op.dst, op.N, op.H, op.W, op.C, op.s_ctot, op.s_coff,
                               op.d_ctot, op.d_coff, stream);
        break;
      case OP_SPP:
        rc = spp_pool_launch(op.dst, op.N, op.H, op.W, op.C, op.d_ctot, stream);
        break;
    }
    return rc;
  }
}

void YoloxPlan::release() {
  if (arena && owns_arena) cudaFree(arena);
  arena = nullptr;
  ops.clear();
  acts.clear();
}

}  // namespace tcd
