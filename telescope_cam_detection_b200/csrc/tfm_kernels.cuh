// Non-conv kernels of the RT-DETR path (src/rtdetr_detector.py:238-403 -> upstream RT-DETRv2,
// SURVEY App. D): LayerNorm(+residual), multi-head attention over a few hundred tokens,
// multi-scale deformable attention (gather-and-weight), per-row top-k, pooling, and the small
// reference-point / gather glue of the decoder.  All activations bf16 token-major / NHWC,
// arithmetic fp32; reference points, sampling offsets, logits stay fp32.
#pragma once
#include "tcd_common.cuh"

namespace tcd {

// y[r,:] = LayerNorm(a[r,:] + (b ? b[r,:] : 0)) * gamma + beta      (C <= 1024, C % 8 == 0); gamma == NULL: no affine part
int layernorm_launch(const void* a, const void* b, const float* gamma, const float* beta, void* y,
                     int rows, int C, float eps, cudaStream_t st);
// y[n,t,:] = a[n,t,:] + b[(bcast ? 0 : n),t,:]        (bf16, row length C % 8 == 0)
int add_rows_launch(const void* a, const void* b, void* y, int N, int T, int C, int bcast_b,
                    cudaStream_t st);
// y[n,t,:] = x[n,t,:] * mask[t]   (mask f32 0/1 per token)
int mask_rows_launch(const void* x, const float* mask, void* y, int N, int T, int C, cudaStream_t st);
// softmax(q k^T * scale) v per head; q/k/v: bf16 rows with leading dims ldq/ldk/ldv (elements) and
// column offsets; out bf16 [N,Tq,heads*32].  head_dim fixed at 32 (256 / 8).
int mha_launch(const void* q, int ldq, int qoff, const void* k, int ldk, int koff, const void* v, int ldv,
               int voff, void* out, int ldo, int N, int Tq, int Tk, int heads, float scale,
               cudaStream_t st);
// Multi-scale deformable attention v2 (HF modeling_rt_detr_v2.py:44-115,166-225): value bf16
// [N,S,heads*32]; offsets f32 [N,Q,heads,L*P,2]; logits f32 [N,Q,heads,L*P]; ref f32 [N,Q,4] (sigmoid
// space cx,cy,w,h); out bf16 [N,Q,heads*32].  grid_sample bilinear, zeros padding, align_corners=False.
struct MsdaLevels {
  int n_levels, n_points;
  int h[4], w[4], start[4];
};
int msda_launch(const void* value, const float* offsets, const float* logits, const float* ref, void* out,
                int N, int Q, int S, int heads, MsdaLevels lv, float offset_scale, cudaStream_t st);
// per-row max of f32 [rows, C] -> f32 [rows]
int rowmax_launch(const float* x, float* y, int rows, int C, cudaStream_t st);
// per-row top-k (descending by value, ascending index on ties) of f32 [N, n] -> int32 idx [N, k]; k <= 512
int topk_launch(const float* keys, int N, int n, int k, int* idx_out, cudaStream_t st);
// out[n,i,:] = src[n, idx[n,i], :]   bf16 rows of C
int gather_rows_launch(const void* src, const int* idx, void* out, int N, int S, int K, int C,
                       cudaStream_t st);
// ref_unact[n,i,:] = delta[n,i,0:4] (f32, leading dim ldd) + anchors[idx[n,i],:]; ref_sig = sigmoid(.)
// also writes the bf16 [N,K,16] copy (4 real + 12 zero channels) the query-pos MLP reads
int ref_init_launch(const float* delta, int ldd, const float* anchors, const int* idx, float* ref_sig,
                    void* ref_bf16, int N, int K, cudaStream_t st);
// ref_sig = sigmoid(delta + inverse_sigmoid(ref_sig)), eps 1e-5 (HF :551-556, :637-640)
int ref_update_launch(const float* delta, int ldd, float* ref_sig, void* ref_bf16, int rows,
                      cudaStream_t st);
// RT-DETR postprocessor (use_focal_loss, src/rtdetr_detector.py:257 deploy tuple): idx [N,k] over the
// flattened [Q*C] logits -> labels, boxes xyxy scaled by orig (w,h), scores = sigmoid(logit)
int detr_post_launch(const float* logits, const float* ref_sig, const int* idx, const float* orig_wh,
                     int N, int Q, int C, int k, int* labels, float* boxes, float* scores,
                     cudaStream_t st);
// bf16 NHWC pools: 3x3 stride 2 pad 1 max (ResNet stem), 2x2 stride 2 average (ResNet-vd shortcut)
int maxpool3x3s2_launch(const void* x, void* y, int N, int H, int W, int C, cudaStream_t st);
int avgpool2x2_launch(const void* x, void* y, int N, int H, int W, int C, cudaStream_t st);

// ---- Stage-2 classifier body (EVA02 vision transformer, vit_kernels.cu; src/species_classifier.py:258-262)
// rows[(n*G+py)*G+px][c*ps*ps+ky*ps+kx] = x[n][c][py*ps+ky][px*ps+kx], bf16, Kp >= 3*ps*ps columns (pad stays zero)
int patchify_f32_launch(const float* x, void* rows, int N, int S, int ps, int Kp, cudaStream_t st);
// the Stage-2 crop gather (crop_resize.cu) writing the same patch rows straight from a u8 BGR frame
int crop_patch_launch(const uint8_t* frame, long long stride, const int* boxes_dev, int n, int S, int ps, int Kp,
                      const tcd_preproc_cfg* cfg, void* rows, cudaStream_t st);
// x[n,0,:] = pos[0,:] (class token already folded in); x[n,1+t,:] = emb[n,t,:] + pos[1+t,:];  pos f32 [T,D]
int vit_assemble_launch(const void* emb, const float* pos, void* x, int N, int T, int D, cudaStream_t st);
// softmax(rot(q) rot(k)^T * scale) v, head_dim 64; rope f32 [(T-prefix), 32, 2] (cos, sin) or NULL (no rotation)
// stats_out (optional, tcgen05 kernel only): [N*T][heads*2][2] partial (sum, sum of squares) of every output row
int mha64_rope_launch(const void* q, int ldq, int qoff, const void* k, int ldk, int koff, const void* v, int ldv,
                      int voff, void* out, int ldo, int N, int T, int heads, float scale, const float* rope,
                      int prefix, float* stats_out, cudaStream_t st);
// the same attention on tcgen05 / TMEM (fa_umma.cu) for T <= 640 tokens; mha64_rope_launch takes it when eligible
bool fa64_umma_eligible(int T);
// head_dim 64, or 32 (rows zero-padded to 64 channels in shared memory; rope must be NULL)
int fa64_umma_launch(const void* q, int ldq, int qoff, const void* k, int ldk, int koff, const void* v, int ldv,
                     int voff, void* out, int ldo, int N, int T, int heads, int head_dim, float scale,
                     const float* rope, int prefix, float* stats_out, cudaStream_t st);
// y[r,0:H) = LayerNorm(silu(gu[r,0:H)) * gu[r,Hp:Hp+H)); gu bf16 [rows, 2*Hp], y bf16 [rows, Hp] (pad zero);
// gamma / beta: Hp floats each (16-byte aligned, padding ignored), or both NULL (no affine part)
// gated != 0: gu is already silu(g) * u, bf16 [rows, Hp] (written by the gated GEMM epilogue): LayerNorm only
int swiglu_ln_launch(const void* gu, const float* gamma, const float* beta, void* y, int rows, int H, int Hp,
                     float eps, int gated, cudaStream_t st);
// out[b,c] = x[b,:] . w[c,:] + bias[c]: x bf16 [B,K], w bf16 [NC,K] (K <= 1024, K % 8 == 0), out f32 [B, ld]
int head_gemv_launch(const void* x, const void* w, const float* bias, float* out, int B, int K, int NC, int ld,
                     cudaStream_t st);
// stats[r] = (mean, 1/sqrt(var + eps)) over the first H columns of bf16 row r (row stride Hp, Hp <= 3072): the
// LayerNorm statistics for a consumer that normalises in its own epilogue (gemm_pair.cuh)
// parts > 0: stats[r][0] = (sum, sum of squares), stats[r][1 .. parts) = 0 -- the partial-sum format the producers write
int row_stats_launch(const void* x, float* stats, int rows, int H, int Hp, float eps, int parts, cudaStream_t st);
int softmax_rows_launch(const float* x, int ldx, float* y, int ldy, int rows, int C, cudaStream_t st);
// out[n,i] = src[n*ld + idx[n*k+i]]
int take_f32_launch(const float* src, int ld, const int* idx, float* out, int N, int k, cudaStream_t st);
}  // namespace tcd
