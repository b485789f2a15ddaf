// Internal model state of libcvt_b200: parameter arenas, per-stage activation buffers, scratch.
#pragma once
#include <string>
#include <vector>

#include "../../include/cvt_b200.h"
#include "common.cuh"
#include "gemm.cuh"
#include "layers.cuh"

namespace cvt {

struct ParamInfo {
  std::string name;
  int rank = 0;
  int64_t shape[4] = {1, 1, 1, 1};
  bool trainable = true;
  int64_t numel = 0;
  int64_t offset = 0;          // floats; into the trainable arena or the state arena
};

// offsets (floats) of one stage's weights inside the trainable arena P (and G / Adam m / v)
struct StageW {
  int64_t conv_w, conv_b, cls, ln_g, ln_b;
  int64_t dw, bn_g, bn_b;                 // [3][9][D], [3][D], [3][D]       (dw_bn only)
  int64_t lw, lb;                         // Projection Dense of the 'linear_dense' method [3][D][D], [3][D]  (tools/grad_cam_CvT.py:103)
  int64_t pw, pb, mw, mb;                 // proj_{q,k,v} [3][D][D],[3][D] (-1 when skip_proj_qkv); mha {query,key,value} [3][D][D],[3][D]
  int64_t ow, ob, prw, prb, f1w, f1b, f2w, f2b;
  int64_t bn_mm, bn_mv;                   // state arena: moving mean / variance [3][D]
};

// folded-weight scratch (TF32 mode): offsets (floats) into the fold arena Fd; Zq / Zo are offsets into the gradient arena past n_train
struct StageFold { int64_t Wc, bc, Wft, bf, Wo, Wot, bo, Zq, Zo; int64_t convt; int Kc, Kp; };   // convt: K-major conv kernel [D][Kp] in Pt

struct StageGeo {
  int hin, win, cin, H, W, D, T, Tc, heads, has_cls, k, s, pt, pl, method;
};

// activations kept from the forward pass (training keeps everything the backward reads)
struct StageBuf {
  bool conv_implicit = false;            // the last forward ran this stage's ConvEmbed as an implicit GEMM: `col` holds no patches
  float* col;                            // TF32 mode: im2col patch matrix of this stage's ConvEmbed [B*T, Kp] (kept for the weight gradient)
  float *x0, *xn, *conv, *qkv1, *qkv2, *attn_o, *a1, *x1, *y, *hpre, *hact, *x2;
  float *lna_mean, *lna_rstd, *lnb_mean, *lnb_rstd;
  float *bn_stats, *bn_scale, *bn_shift, *bn_mean, *bn_rstd, *bn_coef;
  float *scores;                         // fp32 attention path: P [B*heads, Tc, Tc]
  float *lse;                            // tensor-core attention path: log-sum-exp [B*heads, Tc]
  uint8_t *mask_proj, *mask1, *mask2, *mask_attn;      // mask_attn: attention-output dropout of the Grad-CAM variant (null otherwise)
};

struct Model {
  cvt_config cfg;
  int n_stages = 0;
  StageGeo geo[CVT_MAX_STAGES];
  StageW w[CVT_MAX_STAGES];
  StageBuf buf[CVT_MAX_STAGES];
  int64_t hn_g = 0, hn_b = 0, pd1w = 0, pd1b = 0, pd2w = 0, pd2b = 0, fw = 0, fb = 0;   // head weights
  int head_kind = CVT_HEAD_CLS_LN, feat_dim = 0, d_last = 0;

  std::vector<ParamInfo> params;
  int64_t n_train = 0, n_state = 0;       // arena lengths (floats, multiples of 4)
  float *P = nullptr, *G = nullptr, *Mo = nullptr, *Vo = nullptr, *St = nullptr;
  // K-major (transposed) copies of the dense kernels the forward GEMMs read, same offsets as P; refreshed at the start of a forward pass
  float* Fd = nullptr; int64_t n_fold = 0, n_z = 0, n_pt_extra = 0; StageFold fold[CVT_MAX_STAGES]; bool folded = false;
  bool x3 = false;                        // CVT_PREC_TF32X3: dense products error-compensated, nothing is pre-rounded to TF32 except the attention operands
  MgDesc* mg_fwd_dev = nullptr; int mg_fwd_n = 0, mg_fwd_tiles = 0; MgDesc* mg_bwd_dev = nullptr; int mg_bwd_n = 0, mg_bwd_tiles = 0;
  float* Pr = nullptr;                    // TF32-rounded straight copies of the dense kernels the dgrad GEMMs read (same offsets as P)
  float* Pt = nullptr; std::vector<TrDesc> tr_descs; TrDesc* tr_descs_dev = nullptr; int tr_tiles = 0; bool pt_valid = false;

  // head buffers
  float *feat = nullptr, *hln = nullptr, *h_mean = nullptr, *h_rstd = nullptr, *p1 = nullptr, *pred = nullptr;
  float *dpred = nullptr, *dfeat = nullptr, *dp1 = nullptr, *dhln = nullptr;
  float *metrics = nullptr;               // device [2]: loss, mae
  float *dz2 = nullptr, *head_partial = nullptr; unsigned* head_ticket = nullptr;   // fused head (head.cu)
  const float* cur_y = nullptr;           // targets of the train step being issued: the fused head also produces loss + d pred
  // gradient scratch shared by the stages
  float *gX[2] = {nullptr, nullptr};      // gradient w.r.t. a block output [B,Tc,D] (ping-pong across stages)
  float* cls_scratch = nullptr;           // Z_cls [3][D][D] + s_cls [3][D] of the folded q/k/v backward (fold.cu)
  float* dw_part = nullptr;               // per-CTA partials of the depthwise-weight gradient (dwproj.cu)
  float* attn_dd = nullptr;               // tensor-core attention backward: rowsum(dO * O) [B*heads, Tc]
  float *gH = nullptr, *gC = nullptr, *gD = nullptr, *gE = nullptr, *gM = nullptr, *gQ = nullptr, *gQ2 = nullptr, *gS = nullptr;
  // staging for host-buffer entry points
  float *in_img = nullptr, *in_proc = nullptr, *in_y = nullptr, *out_pred = nullptr;
  float *pin_metrics = nullptr;           // pinned host [2]
  // double-buffered host -> device staging of the epoch / predict loops: copies of batch i+1 overlap the step on batch i
  float *in_img2 = nullptr, *in_proc2 = nullptr, *in_y2 = nullptr;
  cudaStream_t copy_stream = nullptr; cudaEvent_t ev_copy[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr};
  float* pin_series = nullptr; int64_t pin_series_cap = 0;     // pinned host: per-batch {loss, mae} of an epoch

  // resident datasets (cvt_dataset_*): the training / validation sets live in HBM, epochs gather their batches by index
  struct Dataset { float* img = nullptr; float* proc = nullptr; float* y = nullptr; int64_t n = 0; };
  Dataset ds[CVT_DATASET_SLOTS];
  int32_t* order_dev = nullptr; int64_t order_cap = 0;      // the epoch permutation
  // pinned ring that pageable host rows are gathered into before their asynchronous copy (shuffle and staging are one pass)
  static constexpr int kRing = 6;
  uint8_t* ring[kRing] = {}; size_t ring_cap = 0; cudaEvent_t ring_ev[kRing] = {};
  bool ring_busy[kRing] = {}; int ring_next = 0;
  double* raw64[2] = {nullptr, nullptr}; size_t raw64_cap = 0;      // landing buffers of float64 host rows (converted on the device)
  uint8_t* pp_pix = nullptr; size_t pp_pix_cap = 0; int32_t* pp_tab = nullptr; float* pp_lut = nullptr; int pp_key[4] = {0, 0, 0, 0};

  void* workspace = nullptr;
  int64_t workspace_bytes = 0;

  cudaStream_t stream = nullptr;
  // independent kernels of one step may run side by side (attention backward: the dQ kernel fills the SMs the dK/dV kernel's last, partial wave
  // leaves idle); forked and joined with events, also under graph capture
  cudaStream_t side_stream = nullptr; cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  int last_batch = 0, last_training = 0, have_forward = 0;
  const float *last_img = nullptr, *last_proc = nullptr;   // device inputs of the last forward (read again by the backward pass)
  int dropout_mode = CVT_DROPOUT_OFF;
  uint64_t dropout_seed = 0;
  uint32_t dropout_step = 0;
  int64_t adam_t = 0;
  long long launches = 0;
  std::vector<std::pair<const char*, int>> prof_groups;   // (group, kernels launched) of the last cvt_profile_step

  // data parallel
  void* nccl_comm = nullptr;
  int rank = 0, world = 1;
  // peer-memory step tail (dp_peer.cu): every rank's workspace mapped through cudaIpc; flags live in the workspace too
  void* peer_base[CVT_MAX_PEERS] = {}; bool peer_ready = false; uint32_t* dp_flags = nullptr; uint32_t dp_epoch = 0; int dp_sync_grads = 0;
  // gradient buckets (one per stage, the last one with the head): reduced on comm_stream while the earlier stages' backward still runs
  cudaStream_t comm_stream = nullptr; cudaEvent_t ev_seg[CVT_MAX_STAGES] = {}, ev_red = nullptr;
  int mg_bwd_first[CVT_MAX_STAGES] = {}, mg_bwd_cnt[CVT_MAX_STAGES] = {}, mg_bwd_tile0[CVT_MAX_STAGES] = {}, mg_bwd_ntiles[CVT_MAX_STAGES] = {};

  // CUDA graphs
  // A step is first run eagerly for a given (kind, buffers, batch); the second time it is captured, afterwards replayed.
  // Per-step scalars (Adam step size, gradient scale, dropout counter) live in lr_dev so that replays see fresh values.
  struct GraphEntry { int kind; const float* img; const float* proc; const float* y; int batch; cudaGraphExec_t exec; long long launches;
                      cudaGraphExec_t seg[CVT_MAX_STAGES]; };      // data parallel: exec = forward + head + last stage's backward, seg[k] = the k-th earlier stage
  int use_graphs = 1;
  bool premasked = false;                 // backward: gM already holds dX2 * mask2 / keep of the stage about to run (written by col2im)
  bool mlp_fused_last = false;            // the last forward ran the fused inference Mlp: y / hpre / hact were not materialised
  bool infer_prep_valid = false;          // the weight-only products of an inference forward (K-major / TF32 copies, folded kernels, BN scale / shift,
                                          // bn_fold) are current: no weight, moving statistic or optimizer step has touched them since they were made
  bool arena_exposed = false;             // cvt_param_arena handed the weight arena to the caller: weights may change behind the library's back
  bool skip_prep = false;                 // set while a predict-loop chunk after the first is issued: weight-only kernels are skipped
  std::vector<GraphEntry> graphs;
  float* lr_dev = nullptr;
  const float* dyn = nullptr;             // == lr_dev while a train step is being issued (eagerly or under capture)
};

// attention dispatch (attention.cu)
cudaError_t attention_forward(Model* m, int si, int B, cudaStream_t s);
cudaError_t attention_backward(Model* m, int si, int B, const float* dO, float* dQKV, cudaStream_t s);

// fused inference Mlp (mlp_fused.cu): x2 = x1 + fc2(gelu(fc1(norm1(x1)))) in one kernel, the hidden activation stays in TMEM
bool mlp_fused_supports(const Model* m, int si);
cudaError_t launch_mlp_fused(const Model* m, int si, int R, int training, cudaStream_t s);

// gradient reduce-scatter + Adam + weight all-gather over peer memory, one kernel (dp_peer.cu)
cudaError_t launch_dp_reduce_adam(Model* m, cudaStream_t s);

// contraction dispatch by precision (fp32 FFMA vs tcgen05)
cudaError_t run_gemm(const Model* m, const GemmP& p, cudaStream_t s);

}  // namespace cvt
