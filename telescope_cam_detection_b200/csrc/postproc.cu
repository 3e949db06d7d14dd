// Decode + confidence filter + class-aware NMS (K4-K6), batched, warp-level.
//
// Replaces the YOLOX head decode [upstream yolo_head.py], yolox.utils.postprocess
// [upstream boxes.py, SURVEY App. B] as called at src/yolox_detector.py:309-314,381-386, and
// torchvision.ops.batched_nms in its coordinate-offset form (torchvision/ops/boxes.py:80-83):
//   boxes_for_nms = boxes + class_id * (max_coordinate + 1)   (fp32)
//   suppress j (lower score) when IoU(i, j) > nms_thre (strict), scores stable-descending.
// This TU is compiled with -fmad=false so the IoU / decode arithmetic keeps the same
// rounding sequence as the scalar reference code.
//
// Pipeline (fixed grids, thresholds read from device memory -> CUDA-graph friendly):
//   1. decode_filter_kernel  one warp per anchor: sigmoid, arg-max over classes (first index
//                            on ties), obj*cls >= conf -> append candidate
//   2. sort_kernel           one CTA per frame: bitonic sort of 64-bit keys
//                            (score desc, anchor asc), gather rows, offset boxes
//   3. nms_mask_kernel       64x64 IoU tiles -> suppression bit matrix (upper triangle)
//   4. nms_reduce_kernel     one CTA per frame: serial scan over the bit matrix staged in
//                            shared memory, then parallel write of the kept rows
#include "kernels.cuh"

#include <atomic>

namespace tcd {
extern std::atomic<uint64_t> g_launch_count;

__device__ __forceinline__ float sigmoid_acc(float x) { return 1.0f / (1.0f + expf(-x)); }

// one warp decodes anchor `a` of image `b`: sigmoid / exp / grid, arg-max over classes, candidate append
__device__ __forceinline__ void decode_one(const PostBuffers& pb, const int raw, const int b, const int a,
                                           const int lane) {
  const int warp_global = b * pb.A + a;
  const int E = 5 + pb.C;
  const float* row = pb.pred + (size_t)warp_global * E;
  float best = -1.0f;
  int best_j = 0x7fffffff;
  float head[3] = {0.f, 0.f, 0.f};
  float dec[3];
  int lvl = 0;
  if (raw) {
    while (lvl + 1 < pb.n_levels && a >= pb.lvl_start[lvl + 1]) ++lvl;
  }
#pragma unroll
  for (int s = 0; s < 3; ++s) {
    const int e = lane + 32 * s;
    float v = 0.f;
    if (e < E) {
      v = row[e];
      if (raw) {
        if (e < 2) {
          const int cell = a - pb.lvl_start[lvl];
          const int gy = cell / pb.lvl_w[lvl], gx = cell - gy * pb.lvl_w[lvl];
          v = (v + (float)(e == 0 ? gx : gy)) * (float)pb.lvl_stride[lvl];
        } else if (e < 4) {
          v = expf(v) * (float)pb.lvl_stride[lvl];
        } else {
          v = sigmoid_acc(v);
        }
      }
      if (e >= 5) {
        const int j = e - 5;
        if (v > best) { best = v; best_j = j; }
      }
    }
    dec[s] = v;
    if (s == 0) head[0] = v;
  }
  if (pb.decoded != nullptr) {
    float* drow = pb.decoded + (size_t)warp_global * E;
#pragma unroll
    for (int s = 0; s < 3; ++s) {
      const int e = lane + 32 * s;
      if (e < E) drow[e] = dec[s];
    }
  }
  // arg-max over the warp, first index wins ties (torch.max semantics)
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, off);
    const int oj = __shfl_xor_sync(0xffffffffu, best_j, off);
    if (ov > best || (ov == best && oj < best_j)) { best = ov; best_j = oj; }
  }
  const float cx = __shfl_sync(0xffffffffu, head[0], 0);
  const float cy = __shfl_sync(0xffffffffu, head[0], 1);
  const float w = __shfl_sync(0xffffffffu, head[0], 2);
  const float h = __shfl_sync(0xffffffffu, head[0], 3);
  const float obj = __shfl_sync(0xffffffffu, head[0], 4);
  if (lane == 0) {
    const float score = obj * best;
    if (score >= pb.thr[0]) {
      const int slot = atomicAdd(pb.cand_count + b, 1);
      if (slot < pb.cap) {
        float* c = pb.cand + ((size_t)b * pb.cap + slot) * 8;
        c[0] = cx - w / 2.0f;
        c[1] = cy - h / 2.0f;
        c[2] = cx + w / 2.0f;
        c[3] = cy + h / 2.0f;
        c[4] = obj;
        c[5] = best;
        c[6] = (float)best_j;
        c[7] = score;
        pb.cand_anchor[(size_t)b * pb.cap + slot] = a;
      }
    }
  }
}

// SPARSE: the per-image lists the prediction convs emitted (grid.y = image, warps stride over the list: a fixed small
// grid -- one CTA per 8 possible entries was 34 000 mostly empty CTAs at 32 frames); otherwise one warp per anchor of
// the dense prediction.
template <bool SPARSE>
__global__ void __launch_bounds__(256)
decode_filter_kernel(const PostBuffers pb, const int raw) {
  const int lane = threadIdx.x & 31;
  if (SPARSE) {
    const int b = blockIdx.y;
    const int n = min(pb.maybe_count[b], pb.cap);
    for (int e = blockIdx.x * 8 + (threadIdx.x >> 5); e < n; e += gridDim.x * 8)
      decode_one(pb, raw, b, pb.maybe_idx[(size_t)b * pb.cap + e], lane);
  } else {
    const int wg = (blockIdx.x * 256 + threadIdx.x) >> 5;
    if (wg >= pb.B * pb.A) return;
    decode_one(pb, raw, wg / pb.A, wg % pb.A, lane);
  }
}

__global__ void __launch_bounds__(1024) sort_kernel(const PostBuffers pb) {
  extern __shared__ unsigned long long keys[];
  __shared__ float s_max[32];
  const int b = blockIdx.x;
  int n = pb.cand_count[b];
  if (n > pb.cap) n = pb.cap;
  int P = 1;
  while (P < n) P <<= 1;
  const float* cand = pb.cand + (size_t)b * pb.cap * 8;
  float lmax = -3.0e38f;
  for (int i = threadIdx.x; i < P; i += 1024) {
    unsigned long long k = 0ull;
    if (i < n) {
      const float* c = cand + (size_t)i * 8;
      const unsigned int sb = __float_as_uint(c[7]);
      const unsigned int an = (unsigned int)pb.cand_anchor[(size_t)b * pb.cap + i];
      k = ((unsigned long long)sb << 32) | (unsigned long long)(((131071u - an) << 15) | (unsigned)i);
      lmax = fmaxf(fmaxf(fmaxf(lmax, c[0]), fmaxf(c[1], c[2])), c[3]);
    }
    keys[i] = k;
  }
  // block max of coordinates
  for (int off = 16; off > 0; off >>= 1) lmax = fmaxf(lmax, __shfl_xor_sync(0xffffffffu, lmax, off));
  if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = lmax;
  __syncthreads();
  float maxc = s_max[0];
  for (int i = 1; i < 32; ++i) maxc = fmaxf(maxc, s_max[i]);
  // bitonic sort, descending
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < P; i += 1024) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long x = keys[i], y = keys[ixj];
          const bool desc = ((i & k) == 0);
          if (desc ? (x < y) : (x > y)) { keys[i] = y; keys[ixj] = x; }
        }
      }
      __syncthreads();
    }
  }
  const float off1 = maxc + 1.0f;
  for (int i = threadIdx.x; i < n; i += 1024) {
    const int slot = (int)(keys[i] & 0x7fffu);
    const float* c = cand + (size_t)slot * 8;
    float* o = pb.srec + ((size_t)b * pb.cap + i) * 8;
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] = c[e];
    const float offs = c[6] * off1;
    pb.sbox[(size_t)b * pb.cap + i] = make_float4(c[0] + offs, c[1] + offs, c[2] + offs, c[3] + offs);
  }
  if (threadIdx.x == 0) pb.n_sorted[b] = n;
}

__device__ __forceinline__ bool iou_gt(const float4 a, const float4 b, const float thr) {
  const float left = fmaxf(a.x, b.x), right = fminf(a.z, b.z);
  const float top = fmaxf(a.y, b.y), bottom = fminf(a.w, b.w);
  const float width = fmaxf(right - left, 0.f), height = fmaxf(bottom - top, 0.f);
  const float inter = width * height;
  const float sa = (a.z - a.x) * (a.w - a.y);
  const float sb = (b.z - b.x) * (b.w - b.y);
  return (inter / (sa + sb - inter)) > thr;
}

__global__ void __launch_bounds__(64) nms_mask_kernel(const PostBuffers pb) {
  __shared__ float4 cols[64];
  const int b = blockIdx.y;
  const int n = pb.n_sorted[b];
  const int nb = (n + 63) >> 6;
  const int words = pb.cap >> 6;
  const int ntiles = nb * (nb + 1) / 2;
  const float thr = pb.thr[1];
  const float4* box = pb.sbox + (size_t)b * pb.cap;
  for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
    // decode (rb, cb), cb >= rb, row-major over the upper triangle
    int rb = 0, rem = t;
    while (rem >= nb - rb) { rem -= nb - rb; ++rb; }
    const int cb = rb + rem;
    const int j0 = cb * 64;
    __syncthreads();
    if (j0 + (int)threadIdx.x < n) cols[threadIdx.x] = box[j0 + threadIdx.x];
    __syncthreads();
    const int i = rb * 64 + threadIdx.x;
    if (i < n) {
      const float4 me = box[i];
      const int ncol = min(64, n - j0);
      unsigned long long bits = 0ull;
      const int start = (rb == cb) ? (int)threadIdx.x + 1 : 0;
      for (int j = start; j < ncol; ++j)
        if (iou_gt(me, cols[j], thr)) bits |= 1ull << j;
      pb.mask[((size_t)b * pb.cap + i) * words + cb] = bits;
    }
  }
}

// Per-class filter of InferenceEngine._post_process_detections (src/inference_engine_yolox.py:505-534) and the
// wildlife mask of YOLOXDetector._format_model_output_to_detections (src/yolox_detector.py:258), evaluated on a
// kept row exactly as the Python code does: coordinates scaled in fp32 (np.float32 * weak Python float), then
// widened; ensure_valid_bbox (swap, minimum size 1), area and every comparison in double.
__device__ __forceinline__ bool class_filter_pass(const tcd_class_filter& f, const float* row, const float rw,
                                                  const float rh) {
  const int cid = (int)row[6];
  if (cid < 0 || cid >= TCD_MAX_CLASSES) return false;
  if (!((f.class_mask[cid >> 6] >> (cid & 63)) & 1ull)) return false;
  double x1 = (double)(row[0] * rw), y1 = (double)(row[1] * rh);
  double x2 = (double)(row[2] * rw), y2 = (double)(row[3] * rh);
  const double conf = (double)(row[4] * row[5]);
  if (x1 > x2) { const double t = x1; x1 = x2; x2 = t; }
  if (x2 - x1 < 1.0) x2 = x1 + 1.0;
  if (y1 > y2) { const double t = y1; y1 = y2; y2 = t; }
  if (y2 - y1 < 1.0) y2 = y1 + 1.0;
  const double area = (x2 - x1) * (y2 - y1);
  if (conf < f.cls_conf[cid]) return false;
  if (f.min_area > 0.0 && area < f.min_area) return false;
  if (area < f.cls_min_area[cid] || area > f.cls_max_area[cid]) return false;
  return true;
}

// One CTA per frame.  The greedy scan of torchvision's nms (row i is kept unless an earlier kept row
// suppresses it) is serial only INSIDE a block of 64 rows, where it is 64 register steps of one warp on the
// block's diagonal mask words; the effect of a block's kept rows on all later blocks is then ORed into the
// removal words by the whole CTA in parallel.  (Round 1 scanned all n rows with one warp: 67 us for ~1000
// candidates per frame.)  Kept rows then go through the per-camera class filter (table 0: none) and the
// first max_det survivors are written in score order.
__global__ void __launch_bounds__(256) nms_reduce_kernel(const PostBuffers pb) {
  __shared__ unsigned long long remv[kPostMaxWords];
  __shared__ unsigned long long keepw[kPostMaxWords];
  __shared__ int base[kPostMaxWords + 1];
  __shared__ int s_cnt[8];
  const int b = blockIdx.x;
  const int n = pb.n_sorted[b];
  const int nb = (n + 63) >> 6;
  const int words = pb.cap >> 6;
  const unsigned long long* gmask = pb.mask + (size_t)b * pb.cap * words;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < nb; i += 256) { remv[i] = 0ull; keepw[i] = 0ull; }
  __syncthreads();
  for (int kb = 0; kb < nb; ++kb) {
    if (warp == 0) {
      const int r0 = kb * 64 + lane, r1 = r0 + 32;
      const unsigned long long d0 = r0 < n ? gmask[(size_t)r0 * words + kb] : 0ull;
      const unsigned long long d1 = r1 < n ? gmask[(size_t)r1 * words + kb] : 0ull;
      unsigned long long cur = remv[kb], kept = 0ull;
      const int lim = min(64, n - kb * 64);
      for (int i = 0; i < lim; ++i) {
        if (!((cur >> i) & 1ull)) {          // warp-uniform
          kept |= 1ull << i;
          cur |= __shfl_sync(0xffffffffu, i < 32 ? d0 : d1, i & 31);
        }
      }
      if (lane == 0) keepw[kb] = kept;
    }
    __syncthreads();
    const unsigned long long kept = keepw[kb];
    const int nw = nb - kb - 1;              // later words this block's kept rows may mark
    if (nw > 0) {
      for (int e = tid; e < 64 * nw; e += 256) {
        const int r = e / nw, w = kb + 1 + (e - r * nw);
        if ((kept >> r) & 1ull) {
          const unsigned long long m = gmask[(size_t)(kb * 64 + r) * words + w];
          if (m) atomicOr(&remv[w], m);
        }
      }
    }
    __syncthreads();
  }
  // class filter on the kept rows (in place: keepw loses the rows that fail)
  const float* srec = pb.srec + (size_t)b * pb.cap * 8;
  int limit = pb.max_out;
  if (pb.meta != nullptr && pb.meta[b].table > 0) {
    const tcd_class_filter& f = pb.tables[pb.meta[b].table];
    const float rw = pb.meta[b].ratio_w, rh = pb.meta[b].ratio_h;
    for (int i = tid; i < n; i += 256)
      if (((keepw[i >> 6] >> (i & 63)) & 1ull) && !class_filter_pass(f, srec + (size_t)i * 8, rw, rh))
        atomicAnd(&keepw[i >> 6], ~(1ull << (i & 63)));
    const int md = f.max_det < 1 ? 1 : f.max_det;   // `if len(kept) >= max_det: break` runs after the first append
    if (md < limit) limit = md;
    __syncthreads();
  }
  if (warp == 0) {
    int run = 0;
    for (int w0 = 0; w0 < nb; w0 += 32) {
      const int w = w0 + lane;
      const int c = w < nb ? __popcll(keepw[w]) : 0;
      int inc = c;
#pragma unroll
      for (int off = 1; off < 32; off <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, off);
        if (lane >= off) inc += v;
      }
      if (w < nb) base[w] = run + inc - c;
      run += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) s_cnt[0] = run;
  }
  __syncthreads();
  int nk = s_cnt[0];
  if (nk > limit) nk = limit;
  float* out = pb.out_dets + (size_t)b * pb.max_out * 7;
  for (int i = tid; i < n; i += 256) {
    const unsigned long long kw = keepw[i >> 6];
    if ((kw >> (i & 63)) & 1ull) {
      const int pos = base[i >> 6] + __popcll(kw & ((1ull << (i & 63)) - 1ull));
      if (pos < nk) {
        const float* r = srec + (size_t)i * 8;
        float* o = out + (size_t)pos * 7;
#pragma unroll
        for (int e = 0; e < 7; ++e) o[e] = r[e];
      }
    }
  }
  if (tid == 0) pb.out_counts[b] = nk;
}

// Parity tap: the class filter alone on caller-supplied rows [n,7] (one CTA; same predicate, same order-preserving
// cut at max_det as the tail of nms_reduce_kernel).  keep[i] = 1 when row i survives.
__global__ void __launch_bounds__(256)
class_filter_rows_kernel(const tcd_class_filter* tables, int table, const float* rows, int n, float rw, float rh,
                         int* keep) {
  __shared__ int s_run;
  const tcd_class_filter& f = tables[table];
  if (threadIdx.x == 0) s_run = 0;
  __syncthreads();
  const int md = f.max_det < 1 ? 1 : f.max_det;
  for (int i0 = 0; i0 < n; i0 += 256) {
    const int i = i0 + threadIdx.x;
    const bool pass = i < n && class_filter_pass(f, rows + (size_t)i * 7, rw, rh);
    // order-preserving rank of this row among the survivors: ballot per warp + serial add over the 8 warps
    __shared__ int wcnt[8];
    const unsigned bal = __ballot_sync(0xffffffffu, pass);
    if ((threadIdx.x & 31) == 0) wcnt[threadIdx.x >> 5] = __popc(bal);
    __syncthreads();
    int before = s_run;
    for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) before += wcnt[w];
    before += __popc(bal & ((1u << (threadIdx.x & 31)) - 1u));
    if (i < n) keep[i] = (pass && before < md) ? 1 : 0;
    __syncthreads();
    if (threadIdx.x == 0) { int t = 0; for (int w = 0; w < 8; ++w) t += wcnt[w]; s_run += t; }
    __syncthreads();
  }
}

int class_filter_rows_launch(const tcd_class_filter* tables, int table, const float* rows_dev, int n, float rw, float rh,
                             int* keep_dev, cudaStream_t stream) {
  class_filter_rows_kernel<<<1, 256, 0, stream>>>(tables, table, rows_dev, n, rw, rh, keep_dev);
  g_launch_count.fetch_add(1, std::memory_order_relaxed);
  return cudaGetLastError() == cudaSuccess ? TCD_OK : TCD_ERR_CUDA;
}

int postproc_launch(const PostBuffers& pb, int raw, cudaStream_t stream, cudaEvent_t* evs) {
  TCD_REQUIRE(pb.cap % 64 == 0 && (pb.cap >> 6) <= kPostMaxWords, "postproc: cap %d outside the mask word budget", pb.cap);
  int P = 1;
  while (P < pb.cap) P <<= 1;
  const size_t sort_smem = (size_t)P * 8;
  if (int rc = ensure_dyn_smem((const void*)sort_kernel, sort_smem)) return rc;
  cudaMemsetAsync(pb.cand_count, 0, sizeof(int) * pb.B, stream);
  const int warps = pb.B * pb.A;
  if (evs) cudaEventRecord(evs[0], stream);
  if (pb.maybe_count != nullptr)
    decode_filter_kernel<true><<<dim3(64, pb.B), 256, 0, stream>>>(pb, raw);
  else
    decode_filter_kernel<false><<<cdiv(warps, 8), 256, 0, stream>>>(pb, raw);
  if (evs) { cudaEventRecord(evs[1], stream); cudaEventRecord(evs[2], stream); }
  sort_kernel<<<pb.B, 1024, sort_smem, stream>>>(pb);
  if (evs) { cudaEventRecord(evs[3], stream); cudaEventRecord(evs[4], stream); }
  nms_mask_kernel<<<dim3(296, pb.B), 64, 0, stream>>>(pb);
  if (evs) { cudaEventRecord(evs[5], stream); cudaEventRecord(evs[6], stream); }
  nms_reduce_kernel<<<pb.B, 256, 0, stream>>>(pb);
  if (evs) cudaEventRecord(evs[7], stream);
  g_launch_count.fetch_add(4, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_last_error("postproc launch: %s", cudaGetErrorString(e));
    return TCD_ERR_CUDA;
  }
  return TCD_OK;
}

}  // namespace tcd
