// C-ABI of libcvt_b200 (include/cvt_b200.h): model construction, weights, forward / backward / Adam orchestration.
// The network built here is the one create_cvt_model builds in the reference (models/CvT(Par).py:292-354,
// models/CvT(Img).py:291-341); layer semantics are the Keras-2 ones restated in SURVEY.md section 9.
#include <dlfcn.h>
#include <math.h>
#include <string.h>
#include <stdlib.h>

#include <algorithm>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <random>
#include <thread>

#include <nvtx3/nvToolsExt.h>

#include "model.cuh"

namespace cvt {

static thread_local std::string g_err;
thread_local long long g_launch_count = 0;
void set_error(const std::string& msg) { g_err = msg; }
bool pdl_enabled() { static const bool on = getenv("CVT_NO_PDL") == nullptr; return on; }

#define API_CUDA(expr)                                   \
  do {                                                   \
    cudaError_t _e = (expr);                             \
    if (_e != cudaSuccess) {                             \
      if (g_err.empty() || g_err.find(cudaGetErrorString(_e)) == std::string::npos) \
        set_error(std::string(#expr) + ": " + cudaGetErrorString(_e)); \
      return CVT_ERR_CUDA;                               \
    }                                                    \
  } while (0)

// ---- profiling marks (cvt_profile_step) -----------------------------------------------------------------
struct ProfMark { const char* name; cudaEvent_t ev; long long launches; };
static thread_local std::vector<ProfMark>* g_prof = nullptr;
static thread_local bool g_nvtx_open = false;
// group boundary: a CUDA event when cvt_profile_step is timing, and (CVT_NVTX=1) an NVTX range named after the kernel group, so that
// nsys / ncu --nvtx attribute kernels to "s1.attention", "s2.bwd_mlp", ... (tracing)
static inline void prof_mark(cudaStream_t s, const char* name) {
  static const bool nvtx = getenv("CVT_NVTX") != nullptr;
  if (nvtx) {
    if (g_nvtx_open) { nvtxRangePop(); g_nvtx_open = false; }
    if (strncmp(name, "end", 3) != 0) { nvtxRangePushA(name); g_nvtx_open = true; }
  }
  if (!g_prof) return;
  cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, s);
  g_prof->push_back({name, e, g_launch_count});
}
static const char* kStageNames[CVT_MAX_STAGES][12] = {
  {"s1.conv_embed", "s1.ln_proj_bn", "s1.qkv_gemm", "s1.attention", "s1.out_gemm", "s1.mlp", "s1.bwd_mlp", "s1.bwd_out", "s1.bwd_attention", "s1.bwd_qkv", "s1.bwd_proj_ln", "s1.bwd_conv"},
  {"s2.conv_embed", "s2.ln_proj_bn", "s2.qkv_gemm", "s2.attention", "s2.out_gemm", "s2.mlp", "s2.bwd_mlp", "s2.bwd_out", "s2.bwd_attention", "s2.bwd_qkv", "s2.bwd_proj_ln", "s2.bwd_conv"},
  {"s3.conv_embed", "s3.ln_proj_bn", "s3.qkv_gemm", "s3.attention", "s3.out_gemm", "s3.mlp", "s3.bwd_mlp", "s3.bwd_out", "s3.bwd_attention", "s3.bwd_qkv", "s3.bwd_proj_ln", "s3.bwd_conv"},
  {"s4.conv_embed", "s4.ln_proj_bn", "s4.qkv_gemm", "s4.attention", "s4.out_gemm", "s4.mlp", "s4.bwd_mlp", "s4.bwd_out", "s4.bwd_attention", "s4.bwd_qkv", "s4.bwd_proj_ln", "s4.bwd_conv"}};

// ---- contraction dispatch ----------------------------------------------------------------------------------
// which contraction kernel every run_gemm call went to (cvt_gemm_dispatch_counts): a TF32-mode product that quietly lands on the FFMA
// kernel is a 5x slowdown nobody asked for, so tests pin the counts for the reference and the widened configuration
thread_local long long g_gemm_dispatch[3] = {0, 0, 0};      // gemm_tma, gemm_tc, gemm_simt
cudaError_t run_gemm(const Model* m, const GemmP& p_in, cudaStream_t s) {
  struct Count { int k; ~Count() { g_gemm_dispatch[k]++; } } count{2};
  GemmP p = p_in;
  if (m->x3) {
    p.x3 = 1;                                  // the TMA-fed kernel runs three MMAs per k-step on split operand tiles
    if (!(p.epi & EPI_KEEP_ROUND)) p.epi &= ~EPI_ROUND;      // nothing is pre-rounded in this mode except what only the attention kernels read
  }
  p.epi &= ~EPI_KEEP_ROUND;
  if (m->cfg.precision != CVT_PREC_FP32) {
    static const bool no_tma = getenv("CVT_NO_TMA") != nullptr;
    if (!no_tma) {
      if (p.b_mode == B_KN && p.a_mode == A_PLAIN && m->pt_valid && p.B >= m->P && p.B < m->P + m->n_train && p.ldb == p.N) {
        // a dense kernel [K,N] of the parameter arena: read its K-major copy (same offset in Pt)
        GemmP q = p; q.B = m->Pt + (p.B - m->P); q.b_mode = B_NK; q.ldb = p.K;
        if (gemm_tma_supports(q)) { count.k = 0; return launch_gemm_tma(q, s); }
      }
      if (p.b_mode == B_NK && p.a_mode == A_PLAIN && m->pt_valid && p.B >= m->P && p.B < m->P + m->n_train) {
        // dgrad through a dense kernel of the parameter arena: read its TF32-rounded copy when one is kept
        const long long off = p.B - m->P;
        for (const TrDesc& d : m->tr_descs)
          if (d.copy && d.off == off) { GemmP q = p; q.B = m->Pr + off; if (gemm_tma_supports(q)) { count.k = 0; return launch_gemm_tma(q, s); } }
      }
      if (gemm_tma_supports(p)) { count.k = 0; return launch_gemm_tma(p, s); }
    }
    if (p.epi & EPI_GELU_DSAVE) { set_error("run_gemm: the saved-GELU-derivative epilogue exists only in the TMA kernel"); return cudaErrorNotSupported; }
    if (!m->x3 && gemm_tc_supports(p)) { count.k = 1; return launch_gemm_tc(p, s); }      // (a single-pass TF32 kernel would break the TF32X3 contract)
  }
  if (p.epi & EPI_GELU_DSAVE) { set_error("run_gemm: the saved-GELU-derivative epilogue exists only in the TMA kernel"); return cudaErrorNotSupported; }
  if (p.epi & EPI_PHILOX) { set_error("run_gemm: in-epilogue dropout needs a tensor-core shape"); return cudaErrorNotSupported; }
  return launch_gemm_simt(p, s);
}

// producers round what only feeds single-pass TF32 products; CVT_PREC_TF32X3 keeps everything in full fp32 (its GEMM kernels split operands)
static inline bool round_ops(const Model* m) { return m->folded && !m->x3; }

// TF32 folded mode on the TMA kernel: the Mlp forward epilogue stores the GELU derivative instead of the pre-activation (gemm.cuh)
static bool gelu_dsave(const Model* m, int D) {
  static const bool off = getenv("CVT_NO_TMA") != nullptr || getenv("CVT_NO_DSAVE") != nullptr;
  return !off && m->folded && m->cfg.precision != CVT_PREC_FP32 && D % 16 == 0;
}

// tc: the model uses the tcgen05 contraction kernel (128-row tiles, 2 CTAs per SM)
static int auto_splitk(bool tc, int M, int N, int K, int batch) {
  if (tc) {
    const int tiles = cdiv(M, 128) * cdiv(N, 64) * batch;
    const int sk = cdiv(kNumSMs * 2, tiles);
    const int maxsk = std::max(1, K / 256);              // at least 8 k-tiles of 32 per CTA
    return std::max(1, std::min(sk, maxsk));
  }
  const int tiles = cdiv(M, 64) * cdiv(N, 64) * batch;
  int sk = cdiv(kNumSMs * 4, tiles);
  const int maxsk = std::max(1, K / 128);
  return std::max(1, std::min(sk, maxsk));
}

static GemmP dense_fwd(const float* A, long long lda, int M, int K, const float* W, int N, const float* bias, float* C, long long ldc) {
  GemmP p; p.M = M; p.N = N; p.K = K; p.A = A; p.lda = lda; p.B = W; p.ldb = N; p.b_mode = B_KN; p.C = C; p.ldc = ldc;
  if (bias) { p.bias = bias; p.epi |= EPI_BIAS; }
  return p;
}
// dX[M,Kin] = dY[M,Nout] * W[Kin,Nout]^T
static GemmP dense_dgrad(const float* dY, long long lddy, int M, int Nout, const float* W, int Kin, float* dX, long long lddx) {
  GemmP p; p.M = M; p.N = Kin; p.K = Nout; p.A = dY; p.lda = lddy; p.B = W; p.ldb = Nout; p.b_mode = B_NK; p.C = dX; p.ldc = lddx;
  return p;
}
// dW[Kin,Nout] += X[tokens,Kin]^T * dY[tokens,Nout]
static GemmP dense_wgrad(bool tc, const float* X, long long ldx, int Kin, const float* dY, long long lddy, int Nout, int tokens, float* dW) {
  GemmP p; p.M = Kin; p.N = Nout; p.K = tokens; p.A = X; p.lda = ldx; p.a_mode = A_TRANS; p.B = dY; p.ldb = lddy; p.b_mode = B_KN;
  p.C = dW; p.ldc = Nout; p.epi = EPI_ATOMIC; p.splitk = auto_splitk(tc, Kin, Nout, tokens, 1);
  return p;
}

// Dense backward w.r.t. its kernel and bias: dW[z] += X[z]^T dY[z], db[z] += colsum(dY[z]).  TF32 mode: one TMA-fed kernel does both.
static cudaError_t wgrad_bias(const Model* m, const float* X, long long ldx, int Kin, const float* dY, long long lddy, int Nout, int tokens,
                              float* dW, float* db, int batch, long long x_bs, long long dy_bs, cudaStream_t s) {
  static const bool no_tma = getenv("CVT_NO_TMA") != nullptr;
  if (m->cfg.precision != CVT_PREC_FP32 && !no_tma) {
    WgradP w; w.X = X; w.ldx = ldx; w.x_bs = x_bs; w.Kin = Kin; w.dY = dY; w.lddy = lddy; w.dy_bs = dy_bs; w.N = Nout; w.tokens = tokens;
    w.batch = batch; w.dW = dW; w.dw_bs = (long long)Kin * Nout; w.db = db; w.db_bs = Nout; w.x3 = m->x3 ? 1 : 0;
    if (wgrad_tma_supports(w)) return launch_wgrad_tma(w, s);
  }
  if (db) CVT_CUDA(launch_colsum(dY, lddy, rowmap_none(), tokens, Nout, batch, dy_bs, db, Nout, s));
  const bool tc = m->cfg.precision != CVT_PREC_FP32;
  GemmP p = dense_wgrad(tc, X, ldx, Kin, dY, lddy, Nout, tokens, dW);
  if (batch > 1) { p.batch = batch; p.a_bs0 = x_bs; p.b_bs0 = dy_bs; p.c_bs0 = (long long)Kin * Nout; p.splitk = auto_splitk(tc, Kin, Nout, tokens, batch); }
  return run_gemm(m, p, s);
}

// ---- parameters ------------------------------------------------------------------------------------------------
static void clear_graphs(Model* m) {
  for (auto& g : m->graphs) { if (g.exec) cudaGraphExecDestroy(g.exec); for (auto e : g.seg) if (e) cudaGraphExecDestroy(e); }
  m->graphs.clear();
}

static int64_t add_param(Model* m, const std::string& name, std::initializer_list<int64_t> shape, bool trainable) {
  ParamInfo pi; pi.name = name; pi.rank = (int)shape.size(); pi.trainable = trainable;
  int64_t n = 1; int i = 0;
  for (int64_t d : shape) { pi.shape[i++] = d; n *= d; }
  pi.numel = n;
  int64_t& cursor = trainable ? m->n_train : m->n_state;
  pi.offset = cursor;
  cursor += (n + 3) / 4 * 4;
  m->params.push_back(pi);
  return pi.offset;
}

static void build_params(Model* m) {
  const cvt_config& c = m->cfg;
  int cin = c.in_chans;
  for (int i = 0; i < m->n_stages; ++i) {
    const StageGeo& g = m->geo[i];
    StageW& w = m->w[i];
    const int64_t D = g.D, N = g.heads, hd = D / N;
    const std::string ps = "stage" + std::to_string(i + 1);
    const std::string t = ps + "_transformer";
    w.conv_w = add_param(m, ps + "_ConvEmbed/kernel", {g.k, g.k, cin, D}, true);
    w.conv_b = add_param(m, ps + "_ConvEmbed/bias", {D}, true);
    w.cls = g.has_cls ? add_param(m, t + "/cls_token", {1, 1, 1, D}, true) : -1;
    w.ln_g = add_param(m, t + "/norm1/gamma", {D}, true);
    w.ln_b = add_param(m, t + "/norm1/beta", {D}, true);
    w.dw = w.bn_g = w.bn_b = w.bn_mm = w.bn_mv = -1;
    const char* qkv[3] = {"q", "k", "v"};
    if (g.method == CVT_PROJ_DW_BN) {
      for (int z = 0; z < 3; ++z) { int64_t o = add_param(m, t + "/attn/" + qkv[z] + "_proj/depthwise_kernel", {3, 3, D, 1}, true); if (z == 0) w.dw = o; }
      for (int z = 0; z < 3; ++z) { int64_t o = add_param(m, t + "/attn/" + qkv[z] + "_proj/bn/gamma", {D}, true); if (z == 0) w.bn_g = o; }
      for (int z = 0; z < 3; ++z) { int64_t o = add_param(m, t + "/attn/" + qkv[z] + "_proj/bn/beta", {D}, true); if (z == 0) w.bn_b = o; }
      for (int z = 0; z < 3; ++z) { int64_t o = add_param(m, t + "/attn/" + qkv[z] + "_proj/bn/moving_mean", {D}, false); if (z == 0) w.bn_mm = o; }
      for (int z = 0; z < 3; ++z) { int64_t o = add_param(m, t + "/attn/" + qkv[z] + "_proj/bn/moving_variance", {D}, false); if (z == 0) w.bn_mv = o; }
    }
    w.lw = w.lb = -1;
    if (g.method == CVT_PROJ_LINEAR_DENSE) {      // Projection(method='linear') of tools/grad_cam_CvT.py:103-104: Dense(dim)
      for (int z = 0; z < 3; ++z) { int64_t o = add_param(m, t + "/attn/" + qkv[z] + "_proj/dense/kernel", {D, D}, true); if (z == 0) w.lw = o; }
      for (int z = 0; z < 3; ++z) { int64_t o = add_param(m, t + "/attn/" + qkv[z] + "_proj/dense/bias", {D}, true); if (z == 0) w.lb = o; }
    }
    w.pw = w.pb = -1;
    if (!c.skip_proj_qkv) {                       // tools/grad_cam_CvT.py builds proj_q/k/v but never calls them: no weights
      for (int z = 0; z < 3; ++z) { int64_t o = add_param(m, t + "/attn/proj_" + qkv[z] + "/kernel", {D, D}, true); if (z == 0) w.pw = o; }
      for (int z = 0; z < 3; ++z) { int64_t o = add_param(m, t + "/attn/proj_" + qkv[z] + "/bias", {D}, true); if (z == 0) w.pb = o; }
    }
    const char* mh[3] = {"query", "key", "value"};
    for (int z = 0; z < 3; ++z) { int64_t o = add_param(m, t + "/attn/mha/" + mh[z] + "/kernel", {D, N, hd}, true); if (z == 0) w.mw = o; }
    for (int z = 0; z < 3; ++z) { int64_t o = add_param(m, t + "/attn/mha/" + mh[z] + "/bias", {N, hd}, true); if (z == 0) w.mb = o; }
    w.ow = add_param(m, t + "/attn/mha/attention_output/kernel", {N, hd, D}, true);
    w.ob = add_param(m, t + "/attn/mha/attention_output/bias", {D}, true);
    w.prw = add_param(m, t + "/attn/proj/kernel", {D, D}, true);
    w.prb = add_param(m, t + "/attn/proj/bias", {D}, true);
    w.f1w = add_param(m, t + "/mlp/fc1/kernel", {D, c.mlp_ratio * D}, true);
    w.f1b = add_param(m, t + "/mlp/fc1/bias", {c.mlp_ratio * D}, true);
    w.f2w = add_param(m, t + "/mlp/fc2/kernel", {c.mlp_ratio * D, D}, true);
    w.f2b = add_param(m, t + "/mlp/fc2/bias", {D}, true);
    cin = (int)D;
  }
  const int64_t D = m->d_last;
  m->hn_g = m->hn_b = -1;
  if (m->head_kind == CVT_HEAD_CLS_LN || m->head_kind == CVT_HEAD_LN_MEAN) {
    m->hn_g = add_param(m, "head_norm/gamma", {D}, true);
    m->hn_b = add_param(m, "head_norm/beta", {D}, true);
  }
  if (c.proc_dim > 0) {
    m->pd1w = add_param(m, "Proc_Dense_1/kernel", {c.proc_dim, c.proc_hidden}, true);
    m->pd1b = add_param(m, "Proc_Dense_1/bias", {c.proc_hidden}, true);
    m->pd2w = add_param(m, "Proc_Dense_2/kernel", {c.proc_hidden, c.proc_hidden}, true);
    m->pd2b = add_param(m, "Proc_Dense_2/bias", {c.proc_hidden}, true);
  }
  m->fw = add_param(m, "Final_Dense/kernel", {m->feat_dim, c.num_classes}, true);
  m->fb = add_param(m, "Final_Dense/bias", {c.num_classes}, true);
}

// ---- workspace ---------------------------------------------------------------------------------------------------
constexpr int kStatParts = 2 * kNumSMs;      // per-CTA partial slices of the BN statistics (dwproj.cu)
struct Bump {
  char* base = nullptr; int64_t off = 0;
  template <typename T> T* take(int64_t count) {
    off = (off + 255) / 256 * 256;
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += count * (int64_t)sizeof(T);
    return p;
  }
};

static void carve(Model* m, Bump& b) {
  const cvt_config& c = m->cfg;
  const int64_t MB = c.max_batch;
  int64_t maxRD = 0, maxRH = 0, maxS = 0, maxIn = 0;
  const bool need_scores = c.precision == CVT_PREC_FP32;   // only the fp32 attention path materialises P
  int64_t maxL = 0;
  for (int i = 0; i < m->n_stages; ++i) {
    const StageGeo& g = m->geo[i];
    StageBuf& s = m->buf[i];
    const int64_t R = MB * g.Tc, RD = R * g.D, RH = RD * c.mlp_ratio;
    const int64_t S = MB * g.heads * (int64_t)g.Tc * g.Tc;
    maxRD = std::max(maxRD, RD); maxRH = std::max(maxRH, RH); maxS = std::max(maxS, S);
    maxIn = std::max(maxIn, MB * (int64_t)g.hin * g.win * g.cin);
    maxL = std::max(maxL, MB * g.heads * (int64_t)g.Tc);
    s.col = m->folded ? b.take<float>(MB * g.T * (int64_t)m->fold[i].Kp) : nullptr;
    s.x0 = b.take<float>(RD); s.xn = b.take<float>(RD); s.conv = b.take<float>(3 * RD);
    s.qkv1 = b.take<float>(3 * RD); s.qkv2 = b.take<float>(3 * RD);
    s.attn_o = b.take<float>(RD); s.a1 = b.take<float>(RD); s.x1 = b.take<float>(RD); s.y = b.take<float>(RD);
    s.hpre = b.take<float>(RH); s.hact = b.take<float>(RH); s.x2 = b.take<float>(RD);
    s.lna_mean = b.take<float>(R); s.lna_rstd = b.take<float>(R); s.lnb_mean = b.take<float>(R); s.lnb_rstd = b.take<float>(R);
    s.bn_stats = b.take<float>((int64_t)kStatParts * 6 * g.D); s.bn_scale = b.take<float>(3 * g.D); s.bn_shift = b.take<float>(3 * g.D);
    s.bn_mean = b.take<float>(3 * g.D); s.bn_rstd = b.take<float>(3 * g.D); s.bn_coef = b.take<float>(9 * g.D);
    s.scores = need_scores ? b.take<float>(S) : nullptr;
    s.lse = b.take<float>(MB * g.heads * (int64_t)g.Tc);
    s.mask_proj = b.take<uint8_t>(RD); s.mask1 = b.take<uint8_t>(RH); s.mask2 = b.take<uint8_t>(RD);
    s.mask_attn = c.attn_drop_rate > 0.f ? b.take<uint8_t>(RD) : nullptr;
  }
  const StageGeo& gl = m->geo[m->n_stages - 1];
  m->feat = b.take<float>(MB * m->feat_dim);
  m->hln = b.take<float>(MB * (int64_t)gl.Tc * gl.D);
  m->h_mean = b.take<float>(MB * (int64_t)gl.Tc); m->h_rstd = b.take<float>(MB * (int64_t)gl.Tc);
  m->p1 = b.take<float>(MB * std::max(1, c.proc_hidden));
  m->pred = b.take<float>(MB * c.num_classes);
  m->dpred = b.take<float>(MB * c.num_classes);
  m->dfeat = b.take<float>(MB * m->feat_dim);
  m->dp1 = b.take<float>(MB * std::max(1, c.proc_hidden));
  m->dhln = b.take<float>(MB * (int64_t)gl.Tc * gl.D);
  m->metrics = b.take<float>(4);
  m->dz2 = b.take<float>(MB * std::max(1, c.proc_hidden));
  m->head_partial = b.take<float>(2 * MB + 4);
  m->head_ticket = reinterpret_cast<unsigned*>(b.take<float>(4));
  m->gX[0] = b.take<float>(maxRD); m->gX[1] = b.take<float>(maxRD);
  m->gH = b.take<float>(maxRH); m->gC = b.take<float>(maxRD); m->gD = b.take<float>(maxRD); m->gE = b.take<float>(maxRD);
  m->gM = b.take<float>(maxRD); m->gQ = b.take<float>(3 * maxRD); m->gQ2 = b.take<float>(3 * maxRD);
  m->gS = need_scores ? b.take<float>(maxS) : nullptr;
  m->attn_dd = b.take<float>(maxL);
  { int64_t maxD = 0; for (int i = 0; i < m->n_stages; ++i) maxD = std::max<int64_t>(maxD, m->geo[i].D); m->dw_part = b.take<float>((int64_t)2 * kNumSMs * 27 * maxD); m->cls_scratch = b.take<float>(3 * maxD * maxD + 3 * maxD); }
  m->in_img = b.take<float>(maxIn > 0 ? MB * (int64_t)c.image_h * c.image_w * c.in_chans : 0);
  m->in_proc = b.take<float>(MB * std::max(1, c.proc_dim));
  m->in_y = b.take<float>(MB * c.num_classes);
  m->out_pred = b.take<float>(MB * c.num_classes);
  m->in_img2 = b.take<float>(maxIn > 0 ? MB * (int64_t)c.image_h * c.image_w * c.in_chans : 0);
  m->in_proc2 = b.take<float>(MB * std::max(1, c.proc_dim));
  m->in_y2 = b.take<float>(MB * c.num_classes);
  m->lr_dev = b.take<float>(4);
  m->dp_flags = b.take<uint32_t>(64);
  m->Pt = b.take<float>(m->n_train + m->n_pt_extra);
  m->Pr = b.take<float>(m->n_train);
  m->Fd = b.take<float>(m->n_fold);
  m->mg_fwd_dev = b.take<MgDesc>(64); m->mg_bwd_dev = b.take<MgDesc>(96);
  m->tr_descs_dev = b.take<TrDesc>(64);
}

static int same_pad(int n, int k, int s, int* before) {
  const int out = (n + s - 1) / s;
  const int total = std::max((out - 1) * s + k - n, 0);
  *before = total / 2;
  return out;
}

}  // namespace cvt

using namespace cvt;

struct cvt_model { Model m; };

extern "C" {

int cvt_abi_version(void) { return CVT_ABI_VERSION; }
const char* cvt_last_error(void) { return g_err.c_str(); }

void cvt_default_config(cvt_config* c) {
  memset(c, 0, sizeof(*c));
  c->image_h = 128; c->image_w = 128; c->in_chans = 1; c->n_stages = 3;
  const int dims[3] = {64, 128, 256}, patch[3] = {7, 3, 3}, stride[3] = {4, 2, 2}, heads[3] = {1, 2, 4};
  for (int i = 0; i < 3; ++i) {
    c->stage[i].embed_dim = dims[i]; c->stage[i].patch_size = patch[i]; c->stage[i].stride = stride[i];
    c->stage[i].num_heads = heads[i]; c->stage[i].qkv_method = CVT_PROJ_DW_BN; c->stage[i].with_cls_token = (i == 2);
  }
  c->proc_dim = 5; c->proc_hidden = 256; c->num_classes = 1; c->mlp_ratio = 4;
  c->drop_rate = 0.1f; c->ln_eps = 1e-6f; c->bn_eps = 1e-3f; c->bn_momentum = 0.99f;
  c->precision = CVT_PREC_FP32; c->max_batch = 64; c->device = 0;
}

int cvt_create(const cvt_config* cfg, cvt_model** out) {
  if (!cfg || !out) { set_error("cvt_create: null argument"); return CVT_ERR_INVALID; }
  *out = nullptr;
  const cvt_config& c = *cfg;
  if (c.n_stages < 1 || c.n_stages > CVT_MAX_STAGES || c.image_h < 1 || c.image_w < 1 || c.in_chans < 1 || c.num_classes < 1 ||
      c.max_batch < 1 || c.mlp_ratio < 1 || c.proc_dim < 0 || (c.proc_dim > 0 && c.proc_hidden < 1)) {
    set_error("cvt_create: invalid configuration"); return CVT_ERR_INVALID;
  }
  if (c.precision == CVT_PREC_BF16) {
    set_error("cvt_create: CVT_PREC_BF16 is not implemented (fp32 and tf32 are); refusing rather than running another precision under that name");
    return CVT_ERR_INVALID;
  }
  if (c.precision != CVT_PREC_FP32 && c.precision != CVT_PREC_TF32 && c.precision != CVT_PREC_TF32X3) { set_error("cvt_create: unknown precision"); return CVT_ERR_INVALID; }
  if (c.drop_rate < 0.f || c.drop_rate >= 1.f) { set_error("cvt_create: drop_rate must be in [0,1)"); return CVT_ERR_INVALID; }
  for (int i = 0; i < c.n_stages; ++i) {
    const cvt_stage_config& s = c.stage[i];
    if (s.embed_dim < 64 || s.embed_dim % 64 != 0 || s.embed_dim > 512 || s.num_heads * 64 != s.embed_dim) {
      set_error("cvt_create: embed_dim must be a multiple of 64 (<= 512) with num_heads = embed_dim / 64 (key_dim 64)"); return CVT_ERR_INVALID;
    }
    if (s.patch_size < 1 || s.stride < 1 || s.qkv_method < 0 || s.qkv_method > CVT_PROJ_LINEAR_DENSE) { set_error("cvt_create: bad stage spec"); return CVT_ERR_INVALID; }
  }
  if (c.head_kind < CVT_HEAD_AUTO || c.head_kind > CVT_HEAD_MEAN) { set_error("cvt_create: unknown head_kind"); return CVT_ERR_INVALID; }
  if ((c.head_kind == CVT_HEAD_CLS_LN || c.head_kind == CVT_HEAD_CLS) && !c.stage[c.n_stages - 1].with_cls_token) {
    set_error("cvt_create: a cls head needs with_cls_token on the last stage"); return CVT_ERR_INVALID;
  }
  if (c.attn_drop_rate < 0.f || c.attn_drop_rate >= 1.f) { set_error("cvt_create: attn_drop_rate must be in [0,1)"); return CVT_ERR_INVALID; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= c.device) {
    (void)cudaGetLastError();
    set_error("cvt_create: no CUDA device " + std::to_string(c.device) + " (libcvt_b200 has no CPU fallback)");
    return CVT_ERR_NO_DEVICE;
  }
  cudaDeviceProp prop;
  API_CUDA(cudaGetDeviceProperties(&prop, c.device));
  if (prop.major != 10) {
    set_error(std::string("cvt_create: device is sm_") + std::to_string(prop.major) + std::to_string(prop.minor) + ", this library is built for sm_100a only");
    return CVT_ERR_NO_DEVICE;
  }
  API_CUDA(cudaSetDevice(c.device));

#define CREATE_CUDA(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { set_error(std::string(#expr) + ": " + cudaGetErrorString(_e)); cvt_destroy(h); return CVT_ERR_CUDA; } } while (0)
  cvt_model* h = new cvt_model();
  Model* m = &h->m;
  m->cfg = c; m->n_stages = c.n_stages;
  int hh = c.image_h, ww = c.image_w, cin = c.in_chans;
  for (int i = 0; i < c.n_stages; ++i) {
    StageGeo& g = m->geo[i];
    const cvt_stage_config& s = c.stage[i];
    g.hin = hh; g.win = ww; g.cin = cin; g.k = s.patch_size; g.s = s.stride;
    g.H = same_pad(hh, g.k, g.s, &g.pt); g.W = same_pad(ww, g.k, g.s, &g.pl);
    g.D = s.embed_dim; g.T = g.H * g.W; g.has_cls = s.with_cls_token ? 1 : 0; g.Tc = g.T + g.has_cls;
    g.heads = s.num_heads; g.method = s.qkv_method;
    hh = g.H; ww = g.W; cin = g.D;
  }
  m->d_last = m->geo[c.n_stages - 1].D;
  m->head_kind = c.head_kind != CVT_HEAD_AUTO ? c.head_kind : (m->geo[c.n_stages - 1].has_cls ? CVT_HEAD_CLS_LN : CVT_HEAD_LN_MEAN);
  m->feat_dim = m->d_last + (c.proc_dim > 0 ? c.proc_hidden : 0);
  build_params(m);

  {  // folded-weight scratch layout (TF32 / BF16 modes only)
    // the Grad-CAM script's variants (no proj_q/k/v, dropout between attention_output and proj, Dense projections) break the
    // "two stacked affine maps" algebra the folding relies on: those models run the layer-by-layer path
    bool variant = c.skip_proj_qkv != 0 || c.attn_drop_rate > 0.f;
    for (int i = 0; i < c.n_stages; ++i) variant = variant || c.stage[i].qkv_method == CVT_PROJ_LINEAR_DENSE;
    m->folded = c.precision != CVT_PREC_FP32 && getenv("CVT_NO_FOLD") == nullptr && !variant;
    m->x3 = c.precision == CVT_PREC_TF32X3;
    if (m->x3 && !m->folded) { set_error("cvt_create: CVT_PREC_TF32X3 is implemented for the folded path only (reference network, any head / cls layout)"); delete h; return CVT_ERR_INVALID; }
    if (getenv("CVT_NO_GRAPH")) m->use_graphs = 0;
    int64_t f = 0, zc = 0, pt_extra = 0;
    auto al = [](int64_t v) { return (v + 63) / 64 * 64; };
    for (int i = 0; i < c.n_stages; ++i) {
      const int64_t D = m->geo[i].D; StageFold& q = m->fold[i];
      q.Wc = f; f += al(3 * D * D); q.bc = f; f += al(3 * D); q.Wft = f; f += al(3 * D * D); q.bf = f; f += al(3 * D);
      q.Wo = f; f += al(D * D); q.Wot = f; f += al(D * D); q.bo = f; f += al(D);
      q.Zq = m->n_train + zc; zc += al(3 * D * D); q.Zo = m->n_train + zc; zc += al(D * D);
      q.Kc = m->geo[i].k * m->geo[i].k * m->geo[i].cin; q.Kp = (q.Kc + 3) / 4 * 4;
      q.convt = m->n_train + pt_extra; pt_extra += al(D * q.Kp);
    }
    m->n_fold = m->folded ? f : 0; m->n_z = m->folded ? zc : 0; m->n_pt_extra = m->folded ? pt_extra : 0;
  }
  Bump measure; carve(m, measure);
  const int64_t arena_bytes = (4 * m->n_train + m->n_z + std::max<int64_t>(m->n_state, 4)) * (int64_t)sizeof(float);
  m->workspace_bytes = measure.off + 256 + arena_bytes + 8192;
  cudaError_t e = cudaMalloc(&m->workspace, (size_t)m->workspace_bytes);
  if (e != cudaSuccess) { set_error(std::string("cvt_create: cudaMalloc of ") + std::to_string(m->workspace_bytes) + " bytes: " + cudaGetErrorString(e)); m->workspace = nullptr; cvt_destroy(h); return CVT_ERR_CUDA; }
  Bump b; b.base = reinterpret_cast<char*>(m->workspace);
  m->P = b.take<float>(m->n_train); m->G = b.take<float>(m->n_train + m->n_z); m->Mo = b.take<float>(m->n_train); m->Vo = b.take<float>(m->n_train);
  m->St = b.take<float>(std::max<int64_t>(m->n_state, 4));
  carve(m, b);
  CREATE_CUDA(cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking));
  CREATE_CUDA(cudaMemsetAsync(m->workspace, 0, (size_t)m->workspace_bytes, m->stream));
  {  // dense kernels the forward GEMMs read K-major (transposed copies in Pt)
    m->tr_descs.clear(); int tiles = 0;
    auto add = [&](int64_t off, int K, int N, int copy = 0, int64_t toff = -1, int ldt = 0) {
      m->tr_descs.push_back(TrDesc{off, K, N, tiles, copy, toff < 0 ? off : toff, ldt ? ldt : K}); tiles += cdiv(K, 32) * cdiv(N, 32); };
    for (int i = 0; i < m->n_stages; ++i) {
      const StageW& w = m->w[i]; const int D = m->geo[i].D, H4 = c.mlp_ratio * D;
      if (!m->folded) {
        for (int z = 0; z < 3; ++z) { if (w.pw >= 0) add(w.pw + (int64_t)z * D * D, D, D); add(w.mw + (int64_t)z * D * D, D, D); if (w.lw >= 0) add(w.lw + (int64_t)z * D * D, D, D); }
        add(w.ow, D, D); add(w.prw, D, D);
      }
      add(w.f1w, D, H4, 1); add(w.f2w, H4, D, 1);
      if (m->folded) add(w.conv_w, m->fold[i].Kc, D, i > 0 ? 1 : 0, m->fold[i].convt, m->fold[i].Kp);      // ConvEmbed kernel [K][D] -> [D][Kp]
    }
    m->tr_tiles = tiles;
    CREATE_CUDA(cudaMemcpyAsync(m->tr_descs_dev, m->tr_descs.data(), m->tr_descs.size() * sizeof(TrDesc), cudaMemcpyHostToDevice, m->stream));
  }
  if (m->folded) {  // multi-GEMM lists that fold the stacked linear maps (forward) and unfold their gradients (backward), see fold.cu
    std::vector<MgDesc> fw, bw; int tf = 0, tb = 0;
    const bool round_fold = !m->x3;          // TF32X3 keeps the folded kernels in full fp32: the GEMM kernels split them into high / low parts
    auto push = [round_fold](std::vector<MgDesc>& v, int& tiles, const float* A, int ta, const float* B, int tbb, float* C, int M, int N, int K,
                   int lda, int ldb, int ldc, int acc, const float* u, const float* vv, int u1) {
      MgDesc d{A, B, C, M, N, K, lda, ldb, ldc, ta, tbb, 1.f, acc, u, vv, u1, tiles, (M > 1 && !acc && round_fold) ? 1 : 0};   // folded kernels are MMA operands
      tiles += mg_tiles(d); v.push_back(d);
    };
    float* P = m->P; float* G = m->G; float* F = m->Fd;
    for (int i = 0; i < m->n_stages; ++i) {
      const StageW& w = m->w[i]; const StageFold& q = m->fold[i]; const int D = m->geo[i].D; const int64_t DD = (int64_t)D * D;
      m->mg_bwd_first[i] = (int)bw.size(); m->mg_bwd_tile0[i] = tb;
      for (int z = 0; z < 3; ++z) {
        const float* pw = P + w.pw + z * DD; const float* mw = P + w.mw + z * DD;
        push(fw, tf, pw, 0, mw, 0, F + q.Wc + z * DD, D, D, D, D, D, D, 0, nullptr, nullptr, 0);                                // Wc = Wp Wm
        push(fw, tf, P + w.pb + z * D, 0, mw, 0, F + q.bc + z * D, 1, D, D, D, D, D, 0, nullptr, P + w.mb + z * D, 1);            // bc = bp Wm + bm
        float* N = G + q.Zq + z * DD; float* sv = G + w.mb + z * D;                                                               // s = colsum dq = dbm
        push(bw, tb, pw, 1, N, 0, G + w.mw + z * DD, D, D, D, D, D, D, 1, P + w.pb + z * D, sv, 0);                               // dWm = Wp^T N + bp (x) s
        push(bw, tb, N, 0, mw, 1, G + w.pw + z * DD, D, D, D, D, D, D, 1, nullptr, nullptr, 0);                                   // dWp = N Wm^T
        push(bw, tb, sv, 0, mw, 1, G + w.pb + z * D, 1, D, D, D, D, D, 1, nullptr, nullptr, 0);                                   // dbp = s Wm^T
      }
      const float* ow = P + w.ow; const float* prw = P + w.prw;
      push(fw, tf, ow, 0, prw, 0, F + q.Wo, D, D, D, D, D, D, 0, nullptr, nullptr, 0);                                            // Wo = Wout Wproj
      push(fw, tf, prw, 1, ow, 1, F + q.Wot, D, D, D, D, D, D, 0, nullptr, nullptr, 0);                                           // Wo^T (K-major copy)
      push(fw, tf, P + w.ob, 0, prw, 0, F + q.bo, 1, D, D, D, D, D, 0, nullptr, P + w.prb, 1);                                    // bo = bout Wproj + bproj
      float* Zo = G + q.Zo; float* so = G + w.prb;
      push(bw, tb, ow, 1, Zo, 0, G + w.prw, D, D, D, D, D, D, 1, P + w.ob, so, 0);                                                // dWproj = Wout^T Zo + bout (x) so
      push(bw, tb, Zo, 0, prw, 1, G + w.ow, D, D, D, D, D, D, 1, nullptr, nullptr, 0);                                            // dWout = Zo Wproj^T
      push(bw, tb, so, 0, prw, 1, G + w.ob, 1, D, D, D, D, D, 1, nullptr, nullptr, 0);                                            // dbout = so Wproj^T
      m->mg_bwd_cnt[i] = (int)bw.size() - m->mg_bwd_first[i]; m->mg_bwd_ntiles[i] = tb - m->mg_bwd_tile0[i];
    }
    if (fw.size() > 64 || bw.size() > 96) { set_error("cvt_create: too many stages for the fold descriptor tables"); cvt_destroy(h); return CVT_ERR_INVALID; }
    m->mg_fwd_n = (int)fw.size(); m->mg_fwd_tiles = tf; m->mg_bwd_n = (int)bw.size(); m->mg_bwd_tiles = tb;
    CREATE_CUDA(cudaMemcpyAsync(m->mg_fwd_dev, fw.data(), fw.size() * sizeof(MgDesc), cudaMemcpyHostToDevice, m->stream));
    CREATE_CUDA(cudaMemcpyAsync(m->mg_bwd_dev, bw.data(), bw.size() * sizeof(MgDesc), cudaMemcpyHostToDevice, m->stream));
    CREATE_CUDA(cudaStreamSynchronize(m->stream));      // fw / bw are locals
  }
  CREATE_CUDA(cudaMemsetAsync(m->head_ticket, 0, 16, m->stream));
  CREATE_CUDA(cudaMallocHost(&m->pin_metrics, 4 * sizeof(float)));
  CREATE_CUDA(cudaStreamCreateWithFlags(&m->copy_stream, cudaStreamNonBlocking));
  CREATE_CUDA(cudaStreamCreateWithFlags(&m->side_stream, cudaStreamNonBlocking));
  CREATE_CUDA(cudaEventCreateWithFlags(&m->ev_fork, cudaEventDisableTiming)); CREATE_CUDA(cudaEventCreateWithFlags(&m->ev_join, cudaEventDisableTiming));
  for (int i = 0; i < 2; ++i) { CREATE_CUDA(cudaEventCreateWithFlags(&m->ev_copy[i], cudaEventDisableTiming)); CREATE_CUDA(cudaEventCreateWithFlags(&m->ev_done[i], cudaEventDisableTiming)); }
  CREATE_CUDA(cudaStreamSynchronize(m->stream));
  *out = h;
  return CVT_OK;
}
#undef CREATE_CUDA

void cvt_destroy(cvt_model* h) {
  if (!h) return;
  Model* m = &h->m;
  cvt_comm_destroy(h);
  clear_graphs(m);
  if (m->stream) cudaStreamSynchronize(m->stream);
  if (m->copy_stream) cudaStreamSynchronize(m->copy_stream);
  for (int i = 0; i < CVT_DATASET_SLOTS; ++i) cvt_dataset_free(h, i);
  for (int i = 0; i < Model::kRing; ++i) { if (m->ring[i]) cudaFreeHost(m->ring[i]); if (m->ring_ev[i]) cudaEventDestroy(m->ring_ev[i]); }
  for (int i = 0; i < 2; ++i) if (m->raw64[i]) cudaFree(m->raw64[i]);
  if (m->order_dev) cudaFree(m->order_dev);
  if (m->pp_pix) cudaFree(m->pp_pix);
  if (m->pp_tab) cudaFree(m->pp_tab);
  if (m->pp_lut) cudaFree(m->pp_lut);
  if (m->stream) cudaStreamDestroy(m->stream);
  if (m->copy_stream) cudaStreamDestroy(m->copy_stream);
  if (m->side_stream) { cudaStreamSynchronize(m->side_stream); cudaStreamDestroy(m->side_stream); }
  if (m->ev_fork) cudaEventDestroy(m->ev_fork);
  if (m->ev_join) cudaEventDestroy(m->ev_join);
  if (m->workspace) cudaFree(m->workspace);
  if (m->pin_metrics) cudaFreeHost(m->pin_metrics);
  if (m->pin_series) cudaFreeHost(m->pin_series);
  for (int i = 0; i < 2; ++i) { if (m->ev_copy[i]) cudaEventDestroy(m->ev_copy[i]); if (m->ev_done[i]) cudaEventDestroy(m->ev_done[i]); }
  (void)cudaGetLastError();
  delete h;
}

int cvt_param_count(const cvt_model* h) { return h ? (int)h->m.params.size() : CVT_ERR_INVALID; }

int cvt_param_info(const cvt_model* h, int index, const char** name, int32_t* rank, int64_t shape[4], int32_t* trainable, int64_t* numel) {
  if (!h || index < 0 || index >= (int)h->m.params.size()) { set_error("cvt_param_info: bad index"); return CVT_ERR_INVALID; }
  const ParamInfo& p = h->m.params[index];
  if (name) *name = p.name.c_str();
  if (rank) *rank = p.rank;
  if (shape) for (int i = 0; i < 4; ++i) shape[i] = p.shape[i];
  if (trainable) *trainable = p.trainable ? 1 : 0;
  if (numel) *numel = p.numel;
  return CVT_OK;
}

int cvt_param_index(const cvt_model* h, const char* name) {
  if (!h || !name) return CVT_ERR_INVALID;
  for (size_t i = 0; i < h->m.params.size(); ++i) if (h->m.params[i].name == name) return (int)i;
  return CVT_ERR_INVALID;
}

static float* param_ptr(Model* m, const ParamInfo& p, float* train_arena) { return (p.trainable ? train_arena : m->St) + p.offset; }

int cvt_set_param(cvt_model* h, int index, const float* src) {
  if (!h || !src || index < 0 || index >= (int)h->m.params.size()) { set_error("cvt_set_param: bad argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const ParamInfo& p = m->params[index];
  m->infer_prep_valid = false;
  API_CUDA(cudaMemcpyAsync(param_ptr(m, p, m->P), src, p.numel * sizeof(float), cudaMemcpyHostToDevice, m->stream));
  API_CUDA(cudaStreamSynchronize(m->stream));
  return CVT_OK;
}
int cvt_get_param(cvt_model* h, int index, float* dst) {
  if (!h || !dst || index < 0 || index >= (int)h->m.params.size()) { set_error("cvt_get_param: bad argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const ParamInfo& p = m->params[index];
  API_CUDA(cudaMemcpyAsync(dst, param_ptr(m, p, m->P), p.numel * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
  API_CUDA(cudaStreamSynchronize(m->stream));
  return CVT_OK;
}
int cvt_get_grad(cvt_model* h, int index, float* dst) {
  if (!h || !dst || index < 0 || index >= (int)h->m.params.size()) { set_error("cvt_get_grad: bad argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const ParamInfo& p = m->params[index];
  if (!p.trainable) { set_error("cvt_get_grad: parameter is not trainable"); return CVT_ERR_INVALID; }
  API_CUDA(cudaMemcpyAsync(dst, m->G + p.offset, p.numel * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
  API_CUDA(cudaStreamSynchronize(m->stream));
  return CVT_OK;
}

int cvt_init_params(cvt_model* h, uint64_t seed) {
  if (!h) return CVT_ERR_INVALID;
  Model* m = &h->m;
  std::mt19937_64 rng(seed);
  std::vector<float> host;
  for (size_t i = 0; i < m->params.size(); ++i) {
    const ParamInfo& p = m->params[i];
    host.assign(p.numel, 0.f);
    const std::string leaf = p.name.substr(p.name.rfind('/') + 1);
    if (leaf == "kernel" || leaf == "depthwise_kernel") {
      // keras glorot_uniform: receptive = prod(shape[:-2]); fan_in = shape[-2]*rec; fan_out = shape[-1]*rec
      double fi, fo;
      if (p.rank == 2) { fi = (double)p.shape[0]; fo = (double)p.shape[1]; }
      else { double rec = 1; for (int d = 0; d < p.rank - 2; ++d) rec *= (double)p.shape[d]; fi = p.shape[p.rank - 2] * rec; fo = p.shape[p.rank - 1] * rec; }
      const double lim = sqrt(6.0 / (fi + fo));
      std::uniform_real_distribution<double> U(-lim, lim);
      for (auto& v : host) v = (float)U(rng);
    } else if (leaf == "gamma" || leaf == "moving_variance") {
      for (auto& v : host) v = 1.f;
    }
    int r = cvt_set_param(h, (int)i, host.data());
    if (r != CVT_OK) return r;
  }
  return CVT_OK;
}

// =====================================================================================================================
// forward
// =====================================================================================================================
// TF32 mode draws the dropout masks inside the GEMM epilogues that apply them (every such product has a tensor-core shape:
// N = D or 4D with D a multiple of 64); the fp32 FFMA path keeps the separate philox_mask_kernel launches.
static bool inline_philox(const Model* m) {
  static const bool no_tma = getenv("CVT_NO_TMA") != nullptr;
  if (no_tma || m->cfg.precision == CVT_PREC_FP32 || m->dropout_mode != CVT_DROPOUT_PHILOX) return false;
  for (int i = 0; i < m->n_stages; ++i) if (m->geo[i].D % 64 != 0) return false;
  return true;
}
static void epi_dropout(const Model* m, GemmP& q, uint8_t* mask, int site, float keep_scale, float rate = -1.f) {
  q.epi |= EPI_MASK; q.mask = mask; q.mask_scale = keep_scale;
  if (inline_philox(m)) {
    const double t = (double)(rate >= 0.f ? rate : m->cfg.drop_rate) * 4294967296.0;
    q.epi |= EPI_PHILOX; q.mask_gen = mask; q.ph_thr = t >= 4294967295.0 ? 0xFFFFFFFFu : (uint32_t)t;
    q.ph_k0 = (uint32_t)(m->dropout_seed & 0xFFFFFFFFu); q.ph_k1 = (uint32_t)(m->dropout_seed >> 32);
    q.ph_site = (uint32_t)site; q.ph_step = m->dropout_step; q.ph_dyn = m->dyn;
  }
}

// ConvEmbed without a patch matrix (CVT_NO_IMPLICIT_CONV=1 restores im2col + GEMM); forward and weight gradient must agree on it
static bool conv_implicit(const Model* m, int si, int B, long long xin_bstride, int xin_boff, const float* xin) {
  static const bool off = getenv("CVT_NO_IMPLICIT_CONV") != nullptr || getenv("CVT_NO_TMA") != nullptr;
  if (off || !m->folded) return false;
  const StageGeo& g = m->geo[si]; const StageFold& f = m->fold[si];
  if (f.Kc != f.Kp) return false;
  const ConvGeom cv{g.hin, g.win, g.cin, g.H, g.W, g.k, g.s, g.pt, g.pl, xin_bstride, xin_boff};
  GemmP p; p.M = B * g.T; p.N = g.D; p.K = f.Kc; p.A = xin; p.a_mode = A_CONV; p.cv = cv; p.B = m->Pt + f.convt; p.ldb = f.Kp; p.b_mode = B_NK;
  WgradP w; w.X = xin; w.conv = 1; w.cv = cv; w.Kin = f.Kc; w.dY = m->gD; w.lddy = g.D; w.N = g.D; w.tokens = B * g.T; w.dW = m->G;
  return gemm_tma_conv_supports(p) && wgrad_tma_supports(w);
}

static cudaError_t stage_forward(Model* m, int si, const float* xin, long long xin_bstride, int xin_boff, int B, int training) {
  const StageGeo& g = m->geo[si]; const StageW& w = m->w[si]; StageBuf& a = m->buf[si];
  const cvt_config& c = m->cfg;
  cudaStream_t s = m->stream;
  const int D = g.D, R = B * g.Tc, H4 = c.mlp_ratio * D;
  const long long RD = (long long)R * D;
  const bool drop = training && m->dropout_mode != CVT_DROPOUT_OFF && c.drop_rate > 0.f;
  const float keep_scale = 1.f / (1.f - c.drop_rate);
  float* P = m->P;

  prof_mark(s, kStageNames[si][0]);
  if (m->folded) {
    // ConvEmbed as explicit im2col (kept for the weight gradient) + TMA-fed GEMM against the K-major kernel copy — im2col.cu
    const StageFold& f = m->fold[si];
    const ConvGeom cv{g.hin, g.win, g.cin, g.H, g.W, g.k, g.s, g.pt, g.pl, xin_bstride, xin_boff, m->x3 ? 1 : 0};
    GemmP p; p.M = B * g.T; p.N = D; p.K = f.Kp; p.A = a.col; p.lda = f.Kp; p.B = m->Pt + f.convt; p.ldb = f.Kp; p.b_mode = B_NK;
    p.C = a.x0; p.ldc = D; if (g.has_cls) p.c_map = RowMap{g.T, g.Tc, 1};
    p.bias = P + w.conv_b; p.epi = EPI_BIAS;
    GemmP ic = p; ic.A = xin; ic.a_mode = A_CONV; ic.cv = cv; ic.K = f.Kc; ic.lda = 0; ic.x3 = m->x3 ? 1 : 0;
    // measured at batch 64: the forward gains (s2 34 -> 25 us, s3 34 -> 28 us) but the weight gradient over 32-token strided boxes is slower than
    // over the patch matrix (s2.bwd_conv 58 -> 71 us), so training keeps im2col (CVT_IMPLICIT_CONV_TRAIN=1 switches both)
    static const bool train_too = getenv("CVT_IMPLICIT_CONV_TRAIN") != nullptr;
    a.conv_implicit = (!training || train_too) && conv_implicit(m, si, B, xin_bstride, xin_boff, xin);
    if (a.conv_implicit) {
      // implicit GEMM: the A operand is gathered tap by tap straight from the NHWC activation by strided TMA boxes (zero fill = SAME
      // padding); no patch matrix is written or read
      CVT_CUDA(launch_gemm_tma(ic, s));
    } else {
      CVT_CUDA(launch_im2col(xin, cv, B * g.T, f.Kc, f.Kp, a.col, s));
      CVT_CUDA(run_gemm(m, p, s));
    }
    if (g.has_cls) CVT_CUDA(launch_fill_rows(a.x0, RowMap{1, g.Tc, 0}, D, P + w.cls, B, D, s));   // :264-268
  } else
  {  // ConvEmbed: Conv2D(D, patch, stride, 'same') + bias — models/CvT(Par).py:203-212 (its LayerNorm is dead code)
    GemmP p; p.M = B * g.T; p.N = D; p.K = g.k * g.k * g.cin;
    p.A = xin; p.a_mode = A_CONV; p.cv = ConvGeom{g.hin, g.win, g.cin, g.H, g.W, g.k, g.s, g.pt, g.pl, xin_bstride, xin_boff};
    p.B = P + w.conv_w; p.ldb = D; p.b_mode = B_KN; p.C = a.x0; p.ldc = D;
    if (g.has_cls) p.c_map = RowMap{g.T, g.Tc, 1};
    p.bias = P + w.conv_b; p.epi = EPI_BIAS;
    CVT_CUDA(run_gemm(m, p, s));
    if (g.has_cls) CVT_CUDA(launch_fill_rows(a.x0, RowMap{1, g.Tc, 0}, D, P + w.cls, B, D, s));   // :264-268
  }
  prof_mark(s, kStageNames[si][1]);
  // x = norm1(inputs)  :272
  CVT_CUDA(launch_ln_fwd(a.x0, rowmap_none(), D, P + w.ln_g, P + w.ln_b, c.ln_eps, a.xn, D, a.lna_mean, a.lna_rstd, R, D, s));
  {  // q/k/v Projection  :157-170
    const bool bn = g.method == CVT_PROJ_DW_BN;
    DwProjP dp; dp.xn = a.xn; dp.out = a.conv; dp.out_bs = RD; dp.w = bn ? P + w.dw : nullptr;
    const int mz = g.method == CVT_PROJ_DW_BN ? 0 : (g.method == CVT_PROJ_AVG ? 1 : 2);
    dp.method[0] = (g.method == CVT_PROJ_AVG) ? 2 : mz;      // q is 'linear' when the stage method is 'avg'  (:127)
    dp.method[1] = mz; dp.method[2] = mz;
    dp.B = B; dp.H = g.H; dp.W = g.W; dp.D = D; dp.Tc = g.Tc; dp.has_cls = g.has_cls;
    dp.stats = (bn && training) ? a.bn_stats : nullptr;
    dp.cls_zero = m->folded && bn;
    dp.round_out = round_ops(m);
    const bool fast = g.D <= 256 && (g.D & 3) == 0;
    int stat_parts = fast ? kStatParts : 1;
    if (g.method == CVT_PROJ_LINEAR_DENSE) {
      // Projection(method='linear') of tools/grad_cam_CvT.py:103-104,117: q/k/v = Dense_z(x) on the spatial tokens; the cls row is
      // concatenated back unchanged (:168-170)
      GemmP p = dense_fwd(a.xn, D, R, D, P + w.lw, D, P + w.lb, a.conv, D);
      p.batch = 3; p.a_bs0 = 0; p.b_bs0 = (long long)D * D; p.c_bs0 = RD; p.bias_bs = D;
      CVT_CUDA(run_gemm(m, p, s));
      if (g.has_cls)
        for (int z = 0; z < 3; ++z)
          CVT_CUDA(launch_copy_rows(a.conv + z * RD, RowMap{1, g.Tc, 0}, D, a.xn, RowMap{1, g.Tc, 0}, D, B, D, s));
    }
    else if (fast) {
      dp.nparts = kStatParts;
      // dw_bn in the folded path: the column walkers (dwcol.cu); avg / linear projections and the layer-by-layer path: the register-window kernels
      cudaError_t e = (m->folded && bn) ? launch_dwcol_fwd(dp, s, &stat_parts) : cudaErrorNotSupported;
      if (e == cudaErrorNotSupported) e = launch_dwproj_fwd2(dp, s);
      CVT_CUDA(e);
    }
    else {
      if (dp.stats) CVT_CUDA(cudaMemsetAsync(a.bn_stats, 0, 6 * D * sizeof(float), s));
      CVT_CUDA(launch_dwproj_fwd(dp, s));
    }
    if (bn) {
      BnFinalizeP f; f.stats = a.bn_stats; f.nparts = stat_parts; f.count = (long long)B * g.T; f.gamma = P + w.bn_g; f.beta = P + w.bn_b;
      f.moving_mean = m->St + w.bn_mm; f.moving_var = m->St + w.bn_mv; f.training = training; f.update_moving = training;
      f.eps = c.bn_eps; f.momentum = c.bn_momentum; f.scale = a.bn_scale; f.shift = a.bn_shift; f.save_mean = a.bn_mean; f.save_rstd = a.bn_rstd; f.D = D;
      if (training || !m->skip_prep) CVT_CUDA(launch_bn_finalize(f, s));      // inference statistics are the moving ones: weight-only
    }
  }
  prof_mark(s, kStageNames[si][2]);
  if (m->folded) {
    // proj_q/k/v Dense (:180-182) + MHA query/key/value EinsumDense + BatchNorm affine + 1/sqrt(key_dim), folded into one GEMM (fold.cu)
    const StageFold& f = m->fold[si]; float* F = m->Fd;
    const bool bn = g.method == CVT_PROJ_DW_BN;
    if (training || !m->skip_prep)
      CVT_CUDA(launch_bn_fold(F + f.Wc, F + f.bc, bn ? a.bn_scale : nullptr, bn ? a.bn_shift : nullptr, 0.125f, D, F + f.Wft, F + f.bf, s, m->x3 ? 0 : 1));
    GemmP p; p.M = R; p.N = D; p.K = D; p.A = a.conv; p.lda = D; p.B = F + f.Wft; p.ldb = D; p.b_mode = B_NK; p.C = a.qkv2; p.ldc = D;
    p.bias = F + f.bf; p.epi = EPI_BIAS | EPI_ROUND | EPI_KEEP_ROUND; p.batch = 3;      // Q, K, V: operands of the (single-pass TF32) attention kernels
    p.batch = 3; p.a_bs0 = RD; p.b_bs0 = (long long)D * D; p.c_bs0 = RD; p.bias_bs = D;
    CVT_CUDA(run_gemm(m, p, s));
    if (bn && g.has_cls) CVT_CUDA(launch_cls_fix(a.xn, RD, F + f.Wc, F + f.bc, 0.125f, B, g.Tc, D, a.qkv2, s));   // cls rows skip BatchNorm
  } else
  {  // proj_q/k/v = Dense(dim)  :180-182  (BatchNorm applied while the A operand is loaded)
    const bool skip = c.skip_proj_qkv != 0;     // tools/grad_cam_CvT.py:199: q, k, v enter the attention without proj_q/k/v
    if (!skip) {
      GemmP p = dense_fwd(a.conv, D, R, D, P + w.pw, D, P + w.pb, a.qkv1, D);
      p.batch = 3; p.a_bs0 = RD; p.b_bs0 = (long long)D * D; p.c_bs0 = RD; p.bias_bs = D;
      if (g.method == CVT_PROJ_DW_BN) { p.a_scale = a.bn_scale; p.a_shift = a.bn_shift; p.a_aff_bs = D; p.a_skip_period = g.has_cls ? g.Tc : 0; }
      CVT_CUDA(run_gemm(m, p, s));
    }
    // keras MultiHeadAttention query/key/value EinsumDense; query scaled by 1/sqrt(key_dim) after its bias
    GemmP q = dense_fwd(skip ? a.conv : a.qkv1, D, R, D, P + w.mw, D, P + w.mb, a.qkv2, D);
    q.batch = 3; q.a_bs0 = RD; q.b_bs0 = (long long)D * D; q.c_bs0 = RD; q.bias_bs = D;
    if (skip && g.method == CVT_PROJ_DW_BN) { q.a_scale = a.bn_scale; q.a_shift = a.bn_shift; q.a_aff_bs = D; q.a_skip_period = g.has_cls ? g.Tc : 0; }
    q.epi |= EPI_ALPHA0; q.alpha0 = 0.125f;
    if (c.precision != CVT_PREC_FP32) q.epi |= EPI_ROUND;
    CVT_CUDA(run_gemm(m, q, s));
  }
  prof_mark(s, kStageNames[si][3]);
  CVT_CUDA(attention_forward(m, si, B, s));                                 // self.attention(q, v, k)  :185
  prof_mark(s, kStageNames[si][4]);
  if (m->folded) {
    // MHA attention_output EinsumDense + self.proj Dense (:188) as one GEMM, dropout (:189) and residual (:274) in its epilogue
    const StageFold& f = m->fold[si]; float* F = m->Fd;
    GemmP q; q.M = R; q.N = D; q.K = D; q.A = a.attn_o; q.lda = D; q.B = F + f.Wot; q.ldb = D; q.b_mode = B_NK; q.C = a.x1; q.ldc = D;
    q.bias = F + f.bo; q.epi = EPI_BIAS | EPI_RESID; q.resid = a.x0; q.ldr = D;
    if (drop) epi_dropout(m, q, a.mask_proj, si * 4 + 0, keep_scale);
    CVT_CUDA(run_gemm(m, q, s));
  } else
  {
    GemmP p = dense_fwd(a.attn_o, D, R, D, P + w.ow, D, P + w.ob, a.a1, D);      // MHA attention_output
    if (training && m->dropout_mode != CVT_DROPOUT_OFF && c.attn_drop_rate > 0.f)   // self.attn_dropout, tools/grad_cam_CvT.py:200
      epi_dropout(m, p, a.mask_attn, si * 4 + 3, 1.f / (1.f - c.attn_drop_rate), c.attn_drop_rate);
    CVT_CUDA(run_gemm(m, p, s));
    GemmP q = dense_fwd(a.a1, D, R, D, P + w.prw, D, P + w.prb, a.x1, D);        // self.proj :188, dropout :189, residual :274
    q.epi |= EPI_RESID; q.resid = a.x0; q.ldr = D;
    if (drop) epi_dropout(m, q, a.mask_proj, si * 4 + 0, keep_scale);
    CVT_CUDA(run_gemm(m, q, s));
  }
  prof_mark(s, kStageNames[si][5]);
  if (!training && mlp_fused_supports(m, si)) {
    // inference: norm1 (:278) + Mlp (:253-258) + residual (:281) as one kernel, the 4D-wide hidden activation never leaves the SM
    m->mlp_fused_last = true;
    return launch_mlp_fused(m, si, R, 0, s);
  }
  {
    // training (TF32 folded mode, saved-derivative backward): the same kernel also leaves y, the LayerNorm statistics, the dropped-out
    // hidden activation, d hact / d hpre and the fc2 dropout mask for stage_backward
    static const bool train_off = getenv("CVT_NO_FUSED_MLP_TRAIN") != nullptr;
    const bool philox_or_masks = !drop || m->dropout_mode == CVT_DROPOUT_MASKS || inline_philox(m);
    if (training && !train_off && gelu_dsave(m, D) && philox_or_masks && mlp_fused_supports(m, si))
      return launch_mlp_fused(m, si, R, 1, s);
  }
  // y = norm1(x)  :278  (same gamma/beta)
  CVT_CUDA(launch_ln_fwd(a.x1, rowmap_none(), D, P + w.ln_g, P + w.ln_b, c.ln_eps, a.y, D, a.lnb_mean, a.lnb_rstd, R, D, s, round_ops(m)));
  {  // Mlp :253-258
    GemmP p = dense_fwd(a.y, D, R, D, P + w.f1w, H4, P + w.f1b, a.hact, H4);
    p.epi |= EPI_GELU; p.aux = a.hpre;
    if (m->folded) p.epi |= EPI_ROUND;
    if (training && gelu_dsave(m, D)) p.epi |= EPI_GELU_DSAVE;      // hpre then holds d hact / d hpre (dropout included) for the backward epilogue
    if (drop) epi_dropout(m, p, a.mask1, si * 4 + 1, keep_scale);
    CVT_CUDA(run_gemm(m, p, s));
    GemmP q = dense_fwd(a.hact, H4, R, H4, P + w.f2w, D, P + w.f2b, a.x2, D);
    q.epi |= EPI_RESID; q.resid = a.x1; q.ldr = D;                                // output = x + Mlp_output :281
    if (drop) epi_dropout(m, q, a.mask2, si * 4 + 2, keep_scale);
    CVT_CUDA(run_gemm(m, q, s));
  }
  return cudaSuccess;
}

static HeadP head_params(Model* m, const float* proc, int B) {
  const cvt_config& c = m->cfg;
  const StageGeo& g = m->geo[m->n_stages - 1]; StageBuf& a = m->buf[m->n_stages - 1];
  float* P = m->P; float* G = m->G;
  HeadP p{};
  p.B = B; p.D = g.D; p.F = m->feat_dim; p.PH = c.proc_dim > 0 ? c.proc_hidden : 0; p.proc_dim = c.proc_dim; p.nc = c.num_classes;
  p.cls_ln = m->head_kind == CVT_HEAD_CLS_LN; p.eps = c.ln_eps;
  p.x2 = a.x2; p.x2_bs = (long long)g.Tc * g.D;
  p.hn_g = m->hn_g >= 0 ? P + m->hn_g : nullptr; p.hn_b = m->hn_b >= 0 ? P + m->hn_b : nullptr; p.h_mean = m->h_mean; p.h_rstd = m->h_rstd;
  p.feat = m->feat; p.proc = proc; p.p1 = m->p1;
  if (c.proc_dim > 0) { p.w1 = P + m->pd1w; p.b1 = P + m->pd1b; p.w2 = P + m->pd2w; p.b2 = P + m->pd2b; }
  p.fw = P + m->fw; p.fb = P + m->fb; p.pred = m->pred;
  p.dpred = m->dpred; p.metrics = m->metrics; p.partial = m->head_partial; p.ticket = m->head_ticket;
  p.dfeat = m->dfeat; p.dz2 = m->dz2; p.dz1 = m->dp1;
  if (c.proc_dim > 0) { p.g_w1 = G + m->pd1w; p.g_b1 = G + m->pd1b; p.g_w2 = G + m->pd2w; p.g_b2 = G + m->pd2b; }
  p.g_fw = G + m->fw; p.g_fb = G + m->fb; p.g_hn_g = m->hn_g >= 0 ? G + m->hn_g : nullptr; p.g_hn_b = m->hn_b >= 0 ? G + m->hn_b : nullptr;
  return p;
}

// head of create_cvt_model :325-350 as one launch (head.cu); with m->cur_y set it also produces the loss, the MAE and d pred
static cudaError_t head_forward(Model* m, const float* proc, int B) {
  const cvt_config& c = m->cfg; cudaStream_t s = m->stream;
  const StageGeo& g = m->geo[m->n_stages - 1]; StageBuf& a = m->buf[m->n_stages - 1];
  const int D = g.D, F = m->feat_dim; float* P = m->P;
  prof_mark(s, "head");
  const RowMap spatial = g.has_cls ? RowMap{g.T, g.Tc, 1} : rowmap_none();      // the block output without its cls rows (:283-286)
  if (m->head_kind == CVT_HEAD_LN_MEAN) {         // reshape -> LN -> reduce_mean over tokens  :334-337
    CVT_CUDA(launch_ln_fwd(a.x2, spatial, D, P + m->hn_g, P + m->hn_b, c.ln_eps, m->hln, D, m->h_mean, m->h_rstd, B * g.T, D, s));
    CVT_CUDA(launch_token_mean(m->hln, B, g.T, D, m->feat, F, s));
  } else if (m->head_kind == CVT_HEAD_MEAN) {     // tf.reduce_mean(x, [1, 2]) (tools/model_plot.py:292) / GlobalAveragePooling2D (tools/grad_cam_CvT.py:336)
    const float* src = a.x2;
    if (g.has_cls) { CVT_CUDA(launch_copy_rows(m->hln, rowmap_none(), D, a.x2, spatial, D, B * g.T, D, s)); src = m->hln; }
    CVT_CUDA(launch_token_mean(src, B, g.T, D, m->feat, F, s));
  } else if (m->head_kind == CVT_HEAD_CLS) {      // tf.squeeze(cls_tokens, axis=1) without LayerNorm (tools/model_plot.py:290)
    CVT_CUDA(launch_copy_rows(m->feat, rowmap_none(), F, a.x2, RowMap{1, g.Tc, 0}, D, B, D, s));
  }
  HeadP p = head_params(m, proc, B);
  p.y = m->cur_y;
  return launch_head_fwd(p, s);
}

static cudaError_t forward_impl(Model* m, const float* img, const float* proc, int B, int training) {
  cudaStream_t s = m->stream;
  const cvt_config& c = m->cfg;
  const bool drop = training && m->dropout_mode == CVT_DROPOUT_PHILOX && c.drop_rate > 0.f;
  if (drop && !inline_philox(m)) {
    for (int i = 0; i < m->n_stages; ++i) {
      const StageGeo& g = m->geo[i]; StageBuf& a = m->buf[i];
      const long long RD = (long long)B * g.Tc * g.D;
      CVT_CUDA(launch_philox_mask(a.mask_proj, RD, c.drop_rate, m->dropout_seed, m->dropout_step, i * 4 + 0, s, m->dyn));
      CVT_CUDA(launch_philox_mask(a.mask1, RD * c.mlp_ratio, c.drop_rate, m->dropout_seed, m->dropout_step, i * 4 + 1, s, m->dyn));
      CVT_CUDA(launch_philox_mask(a.mask2, RD, c.drop_rate, m->dropout_seed, m->dropout_step, i * 4 + 2, s, m->dyn));
    }
  }
  if (training && m->dropout_mode == CVT_DROPOUT_PHILOX && c.attn_drop_rate > 0.f && !inline_philox(m))
    for (int i = 0; i < m->n_stages; ++i)
      CVT_CUDA(launch_philox_mask(m->buf[i].mask_attn, (long long)B * m->geo[i].Tc * m->geo[i].D, c.attn_drop_rate, m->dropout_seed, m->dropout_step,
                                  i * 4 + 3, s, m->dyn));
  m->mlp_fused_last = false;
  if (training) m->infer_prep_valid = false;            // the BatchNorm moving statistics move, and a weight update usually follows
  const bool prep = training || !m->skip_prep;           // predict loop: weight-only work was done by the first chunk of the call
  if (c.precision != CVT_PREC_FP32 && prep) {
    CVT_CUDA(launch_transpose_weights(m->P, m->Pt, m->Pr, m->tr_descs_dev, (int)m->tr_descs.size(), m->tr_tiles, s, m->x3 ? 0 : 1));
    m->pt_valid = true;
    if (m->folded) CVT_CUDA(launch_multigemm(m->mg_fwd_dev, m->mg_fwd_n, m->mg_fwd_tiles, s));
  }
  const float* x = img; long long bstride = (long long)c.image_h * c.image_w * c.in_chans; int boff = 0;
  for (int i = 0; i < m->n_stages; ++i) {
    CVT_CUDA(stage_forward(m, i, x, bstride, boff, B, training));
    const StageGeo& g = m->geo[i];
    x = m->buf[i].x2; bstride = (long long)g.Tc * g.D; boff = g.has_cls * g.D;
  }
  CVT_CUDA(head_forward(m, proc, B));
  prof_mark(s, "end_forward");
  m->last_batch = B; m->last_training = training; m->have_forward = 1;
  if (training && m->dropout_mode == CVT_DROPOUT_PHILOX) m->dropout_step++;
  return cudaSuccess;
}

// =====================================================================================================================
// backward
// =====================================================================================================================
static cudaError_t head_backward(Model* m, const float* proc, int B, float* dX2) {
  const cvt_config& c = m->cfg; cudaStream_t s = m->stream;
  const StageGeo& g = m->geo[m->n_stages - 1]; StageBuf& a = m->buf[m->n_stages - 1];
  const int D = g.D, F = m->feat_dim; float* P = m->P; float* G = m->G;
  prof_mark(s, "bwd_head");
  // gradient w.r.t. the last block's output: the cls row (cls head) or every row (LN-mean head); zero elsewhere
  CVT_CUDA(cudaMemsetAsync(dX2, 0, (size_t)B * g.Tc * D * sizeof(float), s));
  HeadP p = head_params(m, proc, B);
  p.dx2 = dX2;
  CVT_CUDA(launch_head_bwd(p, s));                 // Final_Dense, Proc_Dense_1/2, cls LayerNormalization: input and parameter gradients
  const RowMap spatial = g.has_cls ? RowMap{g.T, g.Tc, 1} : rowmap_none();
  if (m->head_kind == CVT_HEAD_LN_MEAN) {
    CVT_CUDA(launch_token_mean_bwd(m->dfeat, F, B, g.T, D, m->dhln, s));
    CVT_CUDA(launch_ln_bwd(m->dhln, D, a.x2, spatial, D, m->h_mean, m->h_rstd, P + m->hn_g, nullptr, dX2, spatial, D,
                           G + m->hn_g, G + m->hn_b, B * g.T, D, s));
  } else if (m->head_kind == CVT_HEAD_MEAN) {
    if (g.has_cls) {
      CVT_CUDA(launch_token_mean_bwd(m->dfeat, F, B, g.T, D, m->dhln, s));
      CVT_CUDA(launch_copy_rows(dX2, spatial, D, m->dhln, rowmap_none(), D, B * g.T, D, s));
    } else CVT_CUDA(launch_token_mean_bwd(m->dfeat, F, B, g.T, D, dX2, s));
  } else if (m->head_kind == CVT_HEAD_CLS) {
    CVT_CUDA(launch_copy_rows(dX2, RowMap{1, g.Tc, 0}, D, m->dfeat, rowmap_none(), F, B, D, s));
  }
  return cudaSuccess;
}

// dX2: gradient w.r.t. this block's output [B,Tc,D]; dXprev: gradient buffer of the previous block's output (nullptr for stage 0)
static cudaError_t stage_backward(Model* m, int si, const float* xin, long long xin_bstride, int xin_boff, int B, int training,
                                  const float* dX2, float* dXprev, bool param_grads_only_to_input) {
  (void)param_grads_only_to_input;
  const StageGeo& g = m->geo[si]; const StageW& w = m->w[si]; StageBuf& a = m->buf[si];
  const cvt_config& c = m->cfg; cudaStream_t s = m->stream;
  const int D = g.D, R = B * g.Tc, H4 = c.mlp_ratio * D;
  const long long RD = (long long)R * D;
  const bool drop = training && m->dropout_mode != CVT_DROPOUT_OFF && c.drop_rate > 0.f;
  const float keep_scale = 1.f / (1.f - c.drop_rate);
  float* P = m->P; float* G = m->G;

  prof_mark(s, kStageNames[si][6]);
  // ---- Mlp ----
  const float* dM = dX2;
  if (drop) { if (!m->premasked) CVT_CUDA(launch_mask_scale(m->gM, dX2, a.mask2, keep_scale, RD, s, round_ops(m))); dM = m->gM; }
  m->premasked = false;
  CVT_CUDA(wgrad_bias(m, a.hact, H4, H4, dM, D, D, R, G + w.f2w, G + w.f2b, 1, 0, 0, s));
  {
    GemmP p = dense_dgrad(dM, D, R, D, P + w.f2w, H4, m->gH, H4);      // d hpre = (dM W2^T) * gelu'(hpre) * mask1/keep
    p.epi |= EPI_GELU_BWD; p.aux = a.hpre;
    if (m->folded) p.epi |= EPI_ROUND;
    if (training && gelu_dsave(m, D)) p.epi |= EPI_GELU_DSAVE;      // the forward epilogue left gelu'(hpre) * mask1 / keep in hpre
    else if (drop) { p.epi |= EPI_MASK; p.mask = a.mask1; p.mask_scale = keep_scale; }
    CVT_CUDA(run_gemm(m, p, s));
  }
  CVT_CUDA(wgrad_bias(m, a.y, D, D, m->gH, H4, H4, R, G + w.f1w, G + w.f1b, 1, 0, 0, s));
  CVT_CUDA(run_gemm(m, dense_dgrad(m->gH, H4, R, H4, P + w.f1w, D, m->gC, D), s));          // dY
  // second use of norm1 (:278): dX1 = dX2 + LN'(dY)
  // ... and, in the same pass, the gradient through the dropout of the attention output projection (:189): gM = gD * mask_proj / keep
  CVT_CUDA(launch_ln_bwd(m->gC, D, a.x1, rowmap_none(), D, a.lnb_mean, a.lnb_rstd, P + w.ln_g, dX2, m->gD, rowmap_none(), D,
                         G + w.ln_g, G + w.ln_b, R, D, s, drop ? a.mask_proj : nullptr, keep_scale, m->gM, round_ops(m) ? 1 : 0));
  prof_mark(s, kStageNames[si][7]);
  if (m->folded) {
    // ---- folded path (fold.cu): one wgrad + one dgrad per stacked pair; individual gradients are unfolded at the end of backward ----
    const StageFold& f = m->fold[si]; float* F = m->Fd;
    const float* dA2 = m->gD;
    if (drop) dA2 = m->gM;
    CVT_CUDA(wgrad_bias(m, a.attn_o, D, D, dA2, D, D, R, G + f.Zo, G + w.prb, 1, 0, 0, s));        // Zo = o^T dA2, so = colsum dA2 = dbproj
    { GemmP q = dense_dgrad(dA2, D, R, D, F + f.Wo, D, m->gE, D); q.epi |= EPI_ROUND | EPI_KEEP_ROUND; CVT_CUDA(run_gemm(m, q, s)); }   // dO = dA2 Wo^T (TMA operand of attention backward)
    prof_mark(s, kStageNames[si][8]);
    CVT_CUDA(attention_backward(m, si, B, m->gE, m->gQ, s));                                        // gQ = [dQ(pre-scale), dK, dV]
    prof_mark(s, kStageNames[si][9]);
    CVT_CUDA(wgrad_bias(m, a.conv, D, D, m->gQ, D, D, R, G + f.Zq, G + w.mb, 3, RD, RD, s));         // Z_z = c_z^T dq_z, s_z = dbm_z
    const bool bn = g.method == CVT_PROJ_DW_BN;
    if (bn)
      CVT_CUDA(launch_qkv_unfold(G + f.Zq, G + w.mb, a.xn, m->gQ, RD, B, g.T, g.Tc, g.has_cls, D, F + f.Wc, a.bn_scale, a.bn_shift, P + w.bn_g,
                                 a.bn_mean, a.bn_rstd, training, G + w.bn_g, G + w.bn_b, a.bn_coef, m->cls_scratch, s));
    {
      GemmP q = dense_dgrad(m->gQ, D, R, D, F + f.Wc, D, m->gQ2, D);                                 // dn = dq Wc^T, BatchNorm backward in the epilogue
      q.batch = 3; q.a_bs0 = RD; q.b_bs0 = (long long)D * D; q.c_bs0 = RD;
      if (bn) { q.epi |= EPI_BN_BWD; q.aux = a.conv; q.epi_coef = a.bn_coef; q.epi_coef_bs = 3 * D; q.epi_skip_period = g.has_cls ? g.Tc : 0; }
      CVT_CUDA(run_gemm(m, q, s));
    }
    prof_mark(s, kStageNames[si][10]);
    DwProjBwdP dp; dp.dconv = m->gQ2; dp.bs = RD; dp.xn = a.xn; dp.w = bn ? P + w.dw : nullptr;
    const int mz = bn ? 0 : (g.method == CVT_PROJ_AVG ? 1 : 2);
    dp.method[0] = (g.method == CVT_PROJ_AVG) ? 2 : mz; dp.method[1] = mz; dp.method[2] = mz;
    dp.B = B; dp.H = g.H; dp.W = g.W; dp.D = D; dp.Tc = g.Tc; dp.has_cls = g.has_cls; dp.dxn = m->gE; dp.ddw = bn ? G + w.dw : nullptr;
    if (g.D <= 256 && (g.D & 3) == 0) {
      dp.ddw_part = m->dw_part;
      cudaError_t e = bn ? launch_dwcol_bwd(dp, s) : cudaErrorNotSupported;
      if (e == cudaErrorNotSupported) e = launch_dwproj_bwd2(dp, s);
      CVT_CUDA(e);
    }
    else CVT_CUDA(launch_dwproj_bwd(dp, s));
  } else {
  // ---- attention output projections ----
  const float* dA2 = m->gD;
  if (drop) dA2 = m->gM;
  CVT_CUDA(wgrad_bias(m, a.a1, D, D, dA2, D, D, R, G + w.prw, G + w.prb, 1, 0, 0, s));
  CVT_CUDA(run_gemm(m, dense_dgrad(dA2, D, R, D, P + w.prw, D, m->gC, D), s));              // dA1
  if (drop && c.attn_drop_rate > 0.f)                                                       // through self.attn_dropout (tools/grad_cam_CvT.py:200)
    CVT_CUDA(launch_mask_scale(m->gC, m->gC, a.mask_attn, 1.f / (1.f - c.attn_drop_rate), RD, s, 0));
  CVT_CUDA(wgrad_bias(m, a.attn_o, D, D, m->gC, D, D, R, G + w.ow, G + w.ob, 1, 0, 0, s));
  { GemmP q = dense_dgrad(m->gC, D, R, D, P + w.ow, D, m->gE, D); if (c.precision != CVT_PREC_FP32) q.epi |= EPI_ROUND; CVT_CUDA(run_gemm(m, q, s)); }   // dO
  prof_mark(s, kStageNames[si][8]);
  CVT_CUDA(attention_backward(m, si, B, m->gE, m->gQ, s));                                  // gQ = [dQ(pre-scale), dK, dV]
  prof_mark(s, kStageNames[si][9]);
  // ---- MHA q/k/v EinsumDense ----
  const bool skip = c.skip_proj_qkv != 0;
  float* dq0 = m->gQ;                      // d(q0, k0, v0): gradient w.r.t. the three Projection outputs
  if (!skip) {
    CVT_CUDA(wgrad_bias(m, a.qkv1, D, D, m->gQ, D, D, R, G + w.mw, G + w.mb, 3, RD, RD, s));
  } else {
    // proj_q/k/v never ran (tools/grad_cam_CvT.py:199): the EinsumDense input is the Projection output itself (BatchNorm recomputed on load)
    CVT_CUDA(launch_colsum(m->gQ, D, rowmap_none(), R, D, 3, RD, G + w.mb, D, s));
    GemmP p = dense_wgrad(c.precision != CVT_PREC_FP32, a.conv, D, D, m->gQ, D, D, R, G + w.mw);
    p.batch = 3; p.a_bs0 = RD; p.b_bs0 = RD; p.c_bs0 = (long long)D * D; p.splitk = auto_splitk(c.precision != CVT_PREC_FP32, D, D, R, 3);
    if (g.method == CVT_PROJ_DW_BN) { p.a_scale = a.bn_scale; p.a_shift = a.bn_shift; p.a_aff_bs = D; p.a_skip_period = g.has_cls ? g.Tc : 0; }
    CVT_CUDA(run_gemm(m, p, s));
  }
  {
    GemmP q = dense_dgrad(m->gQ, D, R, D, P + w.mw, D, m->gQ2, D);
    q.batch = 3; q.a_bs0 = RD; q.b_bs0 = (long long)D * D; q.c_bs0 = RD;
    CVT_CUDA(run_gemm(m, q, s));
  }
  if (skip) dq0 = m->gQ2;
  else {
  // ---- proj_q/k/v Dense (input = BN(conv) recomputed on load) ----
  CVT_CUDA(launch_colsum(m->gQ2, D, rowmap_none(), R, D, 3, RD, G + w.pb, D, s));
  {
    GemmP p = dense_wgrad(c.precision != CVT_PREC_FP32, a.conv, D, D, m->gQ2, D, D, R, G + w.pw);
    p.batch = 3; p.a_bs0 = RD; p.b_bs0 = RD; p.c_bs0 = (long long)D * D; p.splitk = auto_splitk(c.precision != CVT_PREC_FP32, D, D, R, 3);
    if (g.method == CVT_PROJ_DW_BN) { p.a_scale = a.bn_scale; p.a_shift = a.bn_shift; p.a_aff_bs = D; p.a_skip_period = g.has_cls ? g.Tc : 0; }
    CVT_CUDA(run_gemm(m, p, s));
    GemmP q = dense_dgrad(m->gQ2, D, R, D, P + w.pw, D, m->gQ, D);
    q.batch = 3; q.a_bs0 = RD; q.b_bs0 = (long long)D * D; q.c_bs0 = RD;
    CVT_CUDA(run_gemm(m, q, s));                                                            // gQ = d(q0,k0,v0)
  }
  }
  float* dq_other = dq0 == m->gQ ? m->gQ2 : m->gQ;      // the free one of the two [3][R][D] scratch buffers
  prof_mark(s, kStageNames[si][10]);
  // ---- Projection backward (BatchNorm + depthwise conv / avg pool / identity) ----
  {
    const bool bn = g.method == CVT_PROJ_DW_BN;
    const bool fast = g.D <= 256 && (g.D & 3) == 0;
    if (bn) {
      if (training)
        CVT_CUDA(launch_bn_bwd_reduce(dq0, a.conv, RD, a.bn_mean, a.bn_rstd, G + w.bn_g, G + w.bn_b, B, g.T, g.Tc, g.has_cls, D, s));
      if (fast)
        CVT_CUDA(launch_bn_bwd_coef(P + w.bn_g, a.bn_mean, a.bn_rstd, a.bn_scale, G + w.bn_g, G + w.bn_b, (long long)B * g.T, training, D, a.bn_coef, s));
      else
        CVT_CUDA(launch_bn_bwd_apply(dq0, a.conv, RD, a.bn_mean, a.bn_rstd, P + w.bn_g, a.bn_scale, G + w.bn_g, G + w.bn_b, training,
                                     B, g.T, g.Tc, g.has_cls, D, s));
    }
    if (g.method == CVT_PROJ_LINEAR_DENSE) {
      // Projection Dense backward (tools/grad_cam_CvT.py:103,117): weights see the spatial tokens only; the cls rows pass through
      const RowMap sp = g.has_cls ? RowMap{g.T, g.Tc, 1} : rowmap_none();
      CVT_CUDA(launch_colsum(dq0, D, sp, B * g.T, D, 3, RD, G + w.lb, D, s));
      GemmP p = dense_wgrad(false, a.xn, D, D, dq0, D, D, B * g.T, G + w.lw);
      p.a_map = sp; p.b_map = sp; p.batch = 3; p.a_bs0 = 0; p.b_bs0 = RD; p.c_bs0 = (long long)D * D; p.splitk = auto_splitk(false, D, D, B * g.T, 3);
      CVT_CUDA(launch_gemm_simt(p, s));       // row-mapped operands: the fp32 kernel is the one that takes them (a tool-side variant, not the benched path)
      GemmP q = dense_dgrad(dq0, D, R, D, P + w.lw, D, dq_other, D);
      q.batch = 3; q.a_bs0 = RD; q.b_bs0 = (long long)D * D; q.c_bs0 = RD;
      CVT_CUDA(run_gemm(m, q, s));
      if (g.has_cls)
        for (int z = 0; z < 3; ++z)
          CVT_CUDA(launch_copy_rows(dq_other + z * RD, RowMap{1, g.Tc, 0}, D, dq0 + z * RD, RowMap{1, g.Tc, 0}, D, B, D, s));
      dq0 = dq_other;                          // ... and the three branches are summed by the identity Projection backward below
    }
    DwProjBwdP dp; dp.dconv = dq0; dp.bs = RD; dp.xn = a.xn; dp.w = bn ? P + w.dw : nullptr;
    const int mz = bn ? 0 : (g.method == CVT_PROJ_AVG ? 1 : 2);
    dp.method[0] = (g.method == CVT_PROJ_AVG) ? 2 : mz; dp.method[1] = mz; dp.method[2] = mz;
    dp.B = B; dp.H = g.H; dp.W = g.W; dp.D = D; dp.Tc = g.Tc; dp.has_cls = g.has_cls; dp.dxn = m->gE; dp.ddw = bn ? G + w.dw : nullptr;
    if (fast) {
      if (bn) { dp.conv = a.conv; dp.coef = a.bn_coef; }
      dp.ddw_part = m->dw_part;
      CVT_CUDA(launch_dwproj_bwd2(dp, s));
    } else {
      CVT_CUDA(launch_dwproj_bwd(dp, s));
    }
  }
  }
  // first use of norm1 (:272): dX0 = dX1 + LN'(dXn)   (in place on gD)
  CVT_CUDA(launch_ln_bwd(m->gE, D, a.x0, rowmap_none(), D, a.lna_mean, a.lna_rstd, P + w.ln_g, m->gD, m->gD, rowmap_none(), D,
                         G + w.ln_g, G + w.ln_b, R, D, s));
  prof_mark(s, kStageNames[si][11]);
  // ---- cls token + ConvEmbed ----
  if (g.has_cls) CVT_CUDA(launch_colsum(m->gD, D, RowMap{1, g.Tc, 0}, B, D, 1, 0, G + w.cls, 0, s));
  if (m->folded) {
    const StageFold& f = m->fold[si];
    const float* dY = m->gD;                       // gradient w.r.t. the embed output, spatial rows only
    if (g.has_cls) { CVT_CUDA(launch_copy_rows(m->gC, rowmap_none(), D, m->gD, RowMap{g.T, g.Tc, 1}, D, B * g.T, D, s)); dY = m->gC; }
    WgradP wg; wg.X = a.col; wg.ldx = f.Kp; wg.Kin = f.Kc; wg.dY = dY; wg.lddy = D; wg.N = D; wg.tokens = B * g.T;
    wg.dW = G + w.conv_w; wg.dw_bs = 0; wg.db = G + w.conv_b; wg.db_bs = 0; wg.x3 = m->x3 ? 1 : 0;
    {
      WgradP ic = wg; ic.X = xin; ic.conv = 1; ic.cv = ConvGeom{g.hin, g.win, g.cin, g.H, g.W, g.k, g.s, g.pt, g.pl, xin_bstride, xin_boff};
      if (a.conv_implicit) wg = ic;      // the patches are read from the activation itself (the forward wrote no patch matrix)
    }
    CVT_CUDA(launch_wgrad_tma(wg, s));             // dW = col^T dY, db = colsum dY
    if (dXprev) {
      const StageGeo& gp = m->geo[si - 1];
      const ConvGeom cv{g.hin, g.win, g.cin, g.H, g.W, g.k, g.s, g.pt, g.pl, xin_bstride, xin_boff};
      if (gp.has_cls) CVT_CUDA(cudaMemsetAsync(dXprev, 0, (size_t)B * gp.Tc * gp.D * sizeof(float), s));
      GemmP q; q.M = B * g.T; q.N = f.Kc; q.K = D; q.A = dY; q.lda = D; q.B = m->Pr + w.conv_w; q.ldb = D; q.b_mode = B_NK;
      q.C = a.col; q.ldc = f.Kp;                   // dcol = dY W^T overwrites the patch matrix (its last reader was the wgrad above)
      CVT_CUDA(run_gemm(m, q, s));
      // the previous block ends in the Mlp dropout (:257): its masked gradient is written in the same pass (no cls rows there: dense offsets)
      const bool pm = drop && !gp.has_cls && xin_boff == 0 && xin_bstride == (long long)gp.Tc * gp.D;
      CVT_CUDA(launch_col2im(a.col, cv, B, f.Kp, dXprev, s, pm ? m->buf[si - 1].mask2 : nullptr, keep_scale, m->gM, round_ops(m) ? 1 : 0));
      m->premasked = pm;
    }
    return cudaSuccess;
  }
  const RowMap emap = g.has_cls ? RowMap{g.T, g.Tc, 1} : rowmap_none();
  CVT_CUDA(launch_colsum(m->gD, D, emap, B * g.T, D, 1, 0, G + w.conv_b, 0, s));
  const ConvGeom cv{g.hin, g.win, g.cin, g.H, g.W, g.k, g.s, g.pt, g.pl, xin_bstride, xin_boff};
  {
    GemmP p; p.M = g.k * g.k * g.cin; p.N = D; p.K = B * g.T; p.A = xin; p.a_mode = A_CONVT; p.cv = cv;
    p.B = m->gD; p.ldb = D; p.b_mode = B_KN; p.b_map = emap; p.C = G + w.conv_w; p.ldc = D; p.epi = EPI_ATOMIC;
    p.splitk = auto_splitk(c.precision != CVT_PREC_FP32, p.M, p.N, p.K, 1);
    CVT_CUDA(run_gemm(m, p, s));
  }
  if (dXprev) {
    const StageGeo& gp = m->geo[si - 1];
    CVT_CUDA(cudaMemsetAsync(dXprev, 0, (size_t)B * gp.Tc * gp.D * sizeof(float), s));
    GemmP p; p.M = B * g.T; p.N = g.k * g.k * g.cin; p.K = D; p.A = m->gD; p.lda = D; p.a_map = emap;
    p.B = P + w.conv_w; p.ldb = D; p.b_mode = B_NK; p.C = dXprev; p.cv = cv; p.epi = EPI_SCATTER;
    CVT_CUDA(run_gemm(m, p, s));
  }
  return cudaSuccess;
}


// backward of stages hi, hi-1, ..., lo (hi == n_stages-1: preceded by the gradient-arena reset and the head).  The gradient w.r.t. the
// output of block i lives in gX[i & 1].  stage_unfold: the folded attention weights get their individual gradients right after their
// stage (data parallel: the stage's bucket can then be reduced while the earlier stages still run) instead of once at the end.
// tap_grad != nullptr: stop before stage `stop_after_stage` and return the buffer holding the gradient w.r.t. its output (grad-cam tap).
static cudaError_t backward_range(Model* m, const float* img, const float* proc, int B, int training, int hi, int lo, bool stage_unfold,
                                  int stop_after_stage, float** tap_grad) {
  cudaStream_t s = m->stream;
  if (hi == m->n_stages - 1) {
    m->premasked = false;
    CVT_CUDA(cudaMemsetAsync(m->G, 0, (size_t)(m->n_train + m->n_z) * sizeof(float), s));
    CVT_CUDA(head_backward(m, proc, B, m->gX[hi & 1]));
  }
  for (int i = hi; i >= lo; --i) {
    if (tap_grad && i == stop_after_stage) { *tap_grad = m->gX[i & 1]; return cudaSuccess; }
    const float* xin; long long bstride; int boff;
    if (i == 0) { xin = img; bstride = (long long)m->cfg.image_h * m->cfg.image_w * m->cfg.in_chans; boff = 0; }
    else { const StageGeo& gp = m->geo[i - 1]; xin = m->buf[i - 1].x2; bstride = (long long)gp.Tc * gp.D; boff = gp.has_cls * gp.D; }
    float* dprev = i > 0 ? m->gX[(i & 1) ^ 1] : nullptr;
    CVT_CUDA(stage_backward(m, i, xin, bstride, boff, B, training, m->gX[i & 1], dprev, false));
    if (m->folded && stage_unfold && !tap_grad)
      CVT_CUDA(launch_multigemm(m->mg_bwd_dev + m->mg_bwd_first[i], m->mg_bwd_cnt[i], m->mg_bwd_ntiles[i], s, m->mg_bwd_tile0[i]));
  }
  if (lo == 0) {
    if (m->folded && !stage_unfold && !tap_grad) CVT_CUDA(launch_multigemm(m->mg_bwd_dev, m->mg_bwd_n, m->mg_bwd_tiles, s));   // unfold dWp, dbp, dWm, dWout, dbout, dWproj
    prof_mark(s, "end_backward");
  }
  return cudaSuccess;
}
static cudaError_t backward_impl(Model* m, const float* img, const float* proc, int B, int training, int stop_after_stage, float** tap_grad) {
  return backward_range(m, img, proc, B, training, m->n_stages - 1, 0, false, stop_after_stage, tap_grad);
}

static int check_batch(Model* m, int batch, const char* who) {
  if (batch < 1 || batch > m->cfg.max_batch) { set_error(std::string(who) + ": batch " + std::to_string(batch) + " outside [1, max_batch=" + std::to_string(m->cfg.max_batch) + "]"); return CVT_ERR_INVALID; }
  return CVT_OK;
}

static int infer_forward(Model* m, const float* img, const float* proc, int B, bool cached = false);

int cvt_forward(cvt_model* h, const float* img_dev, const float* proc_dev, int batch, int training, float* out_dev) {
  if (!h || !img_dev) { set_error("cvt_forward: null argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m;
  if (m->cfg.proc_dim > 0 && !proc_dev) { set_error("cvt_forward: this model needs process parameters"); return CVT_ERR_INVALID; }
  int r = check_batch(m, batch, "cvt_forward"); if (r) return r;
  API_CUDA(cudaSetDevice(m->cfg.device));
  const long long l0 = g_launch_count;
  if (training) API_CUDA(forward_impl(m, img_dev, proc_dev, batch, training));
  else { const int rf = infer_forward(m, img_dev, proc_dev, batch); if (rf) return rf; }
  m->last_img = img_dev; m->last_proc = proc_dev;
  if (out_dev) API_CUDA(cudaMemcpyAsync(out_dev, m->pred, (size_t)batch * m->cfg.num_classes * sizeof(float), cudaMemcpyDeviceToDevice, m->stream));
  m->launches += g_launch_count - l0;
  return CVT_OK;
}

int cvt_loss_backward(cvt_model* h, const float* y_dev, float* loss_host, float* mae_host) {
  if (!h || !y_dev) { set_error("cvt_loss_backward: null argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m;
  if (!m->have_forward || !m->last_img) { set_error("cvt_loss_backward: call cvt_forward first"); return CVT_ERR_STATE; }
  const int B = m->last_batch;
  const long long l0 = g_launch_count;
  const float gs = 1.f;
  API_CUDA(launch_mse_loss(m->pred, y_dev, B * m->cfg.num_classes, m->metrics, m->dpred, gs, 0, m->stream));
  API_CUDA(backward_impl(m, m->last_img, m->last_proc, B, m->last_training, -1, nullptr));
  m->launches += g_launch_count - l0;
  if (loss_host || mae_host) {
    float l, a; int r = cvt_read_metrics(h, &l, &a); if (r) return r;
    if (loss_host) *loss_host = l;
    if (mae_host) *mae_host = a;
  }
  return CVT_OK;
}

int cvt_read_metrics(cvt_model* h, float* loss, float* mae) {
  if (!h) return CVT_ERR_INVALID;
  Model* m = &h->m;
  API_CUDA(cudaMemcpyAsync(m->pin_metrics, m->metrics, 2 * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
  API_CUDA(cudaStreamSynchronize(m->stream));
  if (loss) *loss = m->pin_metrics[0];
  if (mae) *mae = m->pin_metrics[1];
  return CVT_OK;
}

// ---- NCCL (dlopen'ed so that the library loads on machines without it) -------------------------------------------
typedef int (*nccl_get_uid_t)(void*);
struct NcclUid { char internal[128]; };   // ncclUniqueId is passed by value (128 bytes)
typedef int (*nccl_init_rank_fn)(void**, int, NcclUid, int);
typedef int (*nccl_allreduce_fn)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*nccl_destroy_fn)(void*);
typedef const char* (*nccl_errstr_fn)(int);
static struct { void* lib; nccl_get_uid_t get_uid; nccl_init_rank_fn init_rank; nccl_allreduce_fn allreduce; nccl_destroy_fn destroy; nccl_errstr_fn errstr; } g_nccl = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};

static int nccl_load() {
  if (g_nccl.lib) return CVT_OK;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) { g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (g_nccl.lib) break; }
  if (!g_nccl.lib) { set_error("NCCL: libnccl.so.2 not found"); return CVT_ERR_NCCL; }
  g_nccl.get_uid = (nccl_get_uid_t)dlsym(g_nccl.lib, "ncclGetUniqueId");
  g_nccl.init_rank = (nccl_init_rank_fn)dlsym(g_nccl.lib, "ncclCommInitRank");
  g_nccl.allreduce = (nccl_allreduce_fn)dlsym(g_nccl.lib, "ncclAllReduce");
  g_nccl.destroy = (nccl_destroy_fn)dlsym(g_nccl.lib, "ncclCommDestroy");
  g_nccl.errstr = (nccl_errstr_fn)dlsym(g_nccl.lib, "ncclGetErrorString");
  if (!g_nccl.get_uid || !g_nccl.init_rank || !g_nccl.allreduce || !g_nccl.destroy) { set_error("NCCL: missing symbols"); return CVT_ERR_NCCL; }
  return CVT_OK;
}
#define API_NCCL(expr) do { int _r = (expr); if (_r != 0) { set_error(std::string(#expr) + ": " + (g_nccl.errstr ? g_nccl.errstr(_r) : "nccl error")); return CVT_ERR_NCCL; } } while (0)

int cvt_comm_unique_id(uint8_t id_out[128]) {
  int r = nccl_load(); if (r) return r;
  API_NCCL(g_nccl.get_uid(id_out));
  return CVT_OK;
}
int cvt_comm_init(cvt_model* h, const uint8_t unique_id[128], int rank, int world) {
  if (!h || !unique_id || world < 1 || rank < 0 || rank >= world) { set_error("cvt_comm_init: bad argument"); return CVT_ERR_INVALID; }
  int r = nccl_load(); if (r) return r;
  Model* m = &h->m;
  API_CUDA(cudaSetDevice(m->cfg.device));
  NcclUid uid; memcpy(uid.internal, unique_id, 128);
  API_NCCL(g_nccl.init_rank(&m->nccl_comm, world, uid, rank));
  m->rank = rank; m->world = world;
  if (!m->comm_stream) {
    int lo = 0, hi = 0;
    API_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));      // the collective's CTAs should be placed as soon as an SM frees up
    API_CUDA(cudaStreamCreateWithPriority(&m->comm_stream, cudaStreamNonBlocking, hi));
    for (int i = 0; i < CVT_MAX_STAGES; ++i) API_CUDA(cudaEventCreateWithFlags(&m->ev_seg[i], cudaEventDisableTiming));
    API_CUDA(cudaEventCreateWithFlags(&m->ev_red, cudaEventDisableTiming));
  }
  clear_graphs(m);
  return CVT_OK;
}
int cvt_comm_destroy(cvt_model* h) {
  if (!h) return CVT_ERR_INVALID;
  Model* m = &h->m;
  if (m->comm_stream) cudaStreamSynchronize(m->comm_stream);
  if (m->peer_ready) {
    if (m->stream) cudaStreamSynchronize(m->stream);
    for (int p = 0; p < CVT_MAX_PEERS; ++p) if (m->peer_base[p] && p != m->rank) cudaIpcCloseMemHandle(m->peer_base[p]);
    for (int p = 0; p < CVT_MAX_PEERS; ++p) m->peer_base[p] = nullptr;
    m->peer_ready = false;
  }
  if (m->nccl_comm && g_nccl.destroy) { g_nccl.destroy(m->nccl_comm); m->nccl_comm = nullptr; }
  if (m->comm_stream) { cudaStreamDestroy(m->comm_stream); m->comm_stream = nullptr; }
  for (int i = 0; i < CVT_MAX_STAGES; ++i) if (m->ev_seg[i]) { cudaEventDestroy(m->ev_seg[i]); m->ev_seg[i] = nullptr; }
  if (m->ev_red) { cudaEventDestroy(m->ev_red); m->ev_red = nullptr; }
  m->world = 1; m->rank = 0;
  clear_graphs(m);
  return CVT_OK;
}
int cvt_comm_peer_export(cvt_model* h, uint8_t handle_out[64]) {
  if (!h || !handle_out) { set_error("cvt_comm_peer_export: null argument"); return CVT_ERR_INVALID; }
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  Model* m = &h->m;
  API_CUDA(cudaSetDevice(m->cfg.device));
  cudaIpcMemHandle_t hd;
  API_CUDA(cudaIpcGetMemHandle(&hd, m->workspace));
  memcpy(handle_out, &hd, 64);
  return CVT_OK;
}
int cvt_comm_peer_open(cvt_model* h, const uint8_t* handles, int sync_grads) {
  if (!h || !handles) { set_error("cvt_comm_peer_open: null argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m;
  if (m->world < 2 || m->world > CVT_MAX_PEERS) { set_error("cvt_comm_peer_open: needs cvt_comm_init with 2.." + std::to_string(CVT_MAX_PEERS) + " ranks first"); return CVT_ERR_STATE; }
  if (m->peer_ready) return CVT_OK;
  API_CUDA(cudaSetDevice(m->cfg.device));
  for (int p = 0; p < m->world; ++p) {
    if (p == m->rank) { m->peer_base[p] = m->workspace; continue; }
    cudaIpcMemHandle_t hd; memcpy(&hd, handles + 64 * p, 64);
    const cudaError_t e = cudaIpcOpenMemHandle(&m->peer_base[p], hd, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      set_error(std::string("cvt_comm_peer_open: cudaIpcOpenMemHandle of rank ") + std::to_string(p) + ": " + cudaGetErrorString(e));
      for (int q = 0; q < p; ++q) if (q != m->rank && m->peer_base[q]) cudaIpcCloseMemHandle(m->peer_base[q]);
      for (int q = 0; q < CVT_MAX_PEERS; ++q) m->peer_base[q] = nullptr;
      (void)cudaGetLastError();
      return CVT_ERR_CUDA;
    }
  }
  m->dp_sync_grads = sync_grads ? 1 : 0; m->dp_epoch = 0;
  API_CUDA(cudaMemsetAsync(m->dp_flags, 0, 64 * sizeof(uint32_t), m->stream));
  API_CUDA(cudaStreamSynchronize(m->stream));
  m->peer_ready = true;
  clear_graphs(m);
  return CVT_OK;
}
int cvt_comm_peer_close(cvt_model* h) {
  if (!h) return CVT_ERR_INVALID;
  Model* m = &h->m;
  if (!m->peer_ready) return CVT_OK;
  API_CUDA(cudaSetDevice(m->cfg.device));
  if (m->stream) API_CUDA(cudaStreamSynchronize(m->stream));
  for (int p = 0; p < CVT_MAX_PEERS; ++p) if (m->peer_base[p] && p != m->rank) cudaIpcCloseMemHandle(m->peer_base[p]);
  for (int p = 0; p < CVT_MAX_PEERS; ++p) m->peer_base[p] = nullptr;
  m->peer_ready = false;
  clear_graphs(m);
  (void)cudaGetLastError();
  return CVT_OK;
}
int cvt_gemm_dispatch_counts(int64_t out[3], int reset) {
  if (out) for (int i = 0; i < 3; ++i) out[i] = g_gemm_dispatch[i];
  if (reset) for (int i = 0; i < 3; ++i) g_gemm_dispatch[i] = 0;
  return CVT_OK;
}
int cvt_grad_arena(cvt_model* h, float** g, int64_t* count) {
  if (!h) return CVT_ERR_INVALID;
  if (g) *g = h->m.G; if (count) *count = h->m.n_train;
  return CVT_OK;
}
int cvt_param_arena(cvt_model* h, float** p, int64_t* count) {
  if (!h) return CVT_ERR_INVALID;
  h->m.arena_exposed = true;      // the caller may now write weights directly: inference never trusts cached weight-only products again
  if (p) *p = h->m.P; if (count) *count = h->m.n_train;
  return CVT_OK;
}

static float adam_alpha(Model* m, float lr) {
  const double b1 = 0.9, b2 = 0.999;
  const double t = (double)m->adam_t;
  return (float)((double)lr * sqrt(1.0 - pow(b2, t)) / (1.0 - pow(b1, t)));
}
static cudaError_t adam_impl(Model* m, float lr, float grad_scale) {
  m->adam_t += 1; m->infer_prep_valid = false;
  return launch_adam(m->P, m->G, m->Mo, m->Vo, m->n_train, adam_alpha(m, lr), 0.9f, 0.999f, 1e-7f, grad_scale, m->stream);
}

int cvt_adam_step(cvt_model* h, float lr) {
  if (!h) return CVT_ERR_INVALID;
  Model* m = &h->m;
  const long long l0 = g_launch_count;
  API_CUDA(adam_impl(m, lr, 1.f));
  m->launches += g_launch_count - l0;
  return CVT_OK;
}
int cvt_optimizer_reset(cvt_model* h) {
  if (!h) return CVT_ERR_INVALID;
  Model* m = &h->m;
  API_CUDA(cudaMemsetAsync(m->Mo, 0, (size_t)m->n_train * sizeof(float), m->stream));
  API_CUDA(cudaMemsetAsync(m->Vo, 0, (size_t)m->n_train * sizeof(float), m->stream));
  m->adam_t = 0;
  return CVT_OK;
}
int64_t cvt_optimizer_step_count(const cvt_model* h) { return h ? h->m.adam_t : -1; }

// ---- CUDA graphs: one captured graph per (kind, input buffers, batch) ----------------------------------------------
enum { GRAPH_TRAIN = 0, GRAPH_INFER = 1, GRAPH_INFER_CACHED = 2 };
static Model::GraphEntry* graph_slot(Model* m, int kind, const float* img, const float* proc, const float* y, int B) {
  for (auto& g : m->graphs) if (g.kind == kind && g.img == img && g.proc == proc && g.y == y && g.batch == B) return &g;
  if (m->graphs.size() >= 16) {
    if (m->graphs.front().exec) cudaGraphExecDestroy(m->graphs.front().exec);
    for (auto e : m->graphs.front().seg) if (e) cudaGraphExecDestroy(e);
    m->graphs.erase(m->graphs.begin());
  }
  m->graphs.push_back(Model::GraphEntry{kind, img, proc, y, B, nullptr, 0, {}});
  return nullptr;                                        // first sighting: the caller runs eagerly (also warms every lazy initialiser)
}

// the stream work of one train step; per-step scalars come from m->lr_dev (set by the caller), so the sequence can be replayed.
static int train_step_body(Model* m, const float* img, const float* proc, const float* y, int B) {
  m->dyn = m->lr_dev;
  int rc = CVT_OK;
  do {
    m->cur_y = y;                                      // the head kernel also emits loss, MAE and d pred (mse_loss_kernel otherwise)
    const cudaError_t ef = forward_impl(m, img, proc, B, 1);
    m->cur_y = nullptr;
    if (ef != cudaSuccess) { rc = CVT_ERR_CUDA; break; }
    if (backward_impl(m, img, proc, B, 1, -1, nullptr) != cudaSuccess) { rc = CVT_ERR_CUDA; break; }
    if (launch_adam(m->P, m->G, m->Mo, m->Vo, m->n_train, 0.f, 0.9f, 0.999f, 1e-7f, 1.f, m->stream, m->dyn) != cudaSuccess) { rc = CVT_ERR_CUDA; break; }
  } while (0);
  m->dyn = nullptr;
  return rc;
}

// ---- data parallel: the step in n_stages segments, one gradient bucket each ------------------------------------------------------------
// segment 0 = forward + loss + head backward + backward of the last stage, segment k = backward of stage n-1-k.  When segment k has been
// issued its stage's gradients are final (the folded attention weights are unfolded per stage), so bucket k - the stage's slice of the
// flat gradient arena, the head riding with the last stage: 77 % / 19 % / 4 % of the bytes at the reference sizes - is all-reduced on
// comm_stream while the main stream runs the earlier stages.  Adam waits for the last bucket.
// number of segments / buckets: one per stage unless CVT_DP_SEGMENTS asks for fewer (1 = a single reduction after the backward pass);
// segment k covers stages hi..lo
static int dp_segments(const Model* m) {
  static const int env = getenv("CVT_DP_SEGMENTS") ? atoi(getenv("CVT_DP_SEGMENTS")) : 0;
  if (m->peer_ready) return 1;                         // peer-memory tail: nothing to overlap, one forward + backward graph
  return env >= 1 && env < m->n_stages ? env : m->n_stages;
}
static void dp_segment_range(const Model* m, int k, int* hi, int* lo) {
  const int nseg = dp_segments(m);
  *hi = m->n_stages - 1 - k;
  *lo = k == nseg - 1 ? 0 : *hi;
}
static int train_segment(Model* m, const float* img, const float* proc, const float* y, int B, int k) {
  m->dyn = m->lr_dev;
  int rc = CVT_OK;
  int hi_s, lo_s; dp_segment_range(m, k, &hi_s, &lo_s);
  if (k == 0) {
    m->cur_y = y;
    const cudaError_t ef = forward_impl(m, img, proc, B, 1);
    m->cur_y = nullptr;
    if (ef != cudaSuccess) rc = CVT_ERR_CUDA;
  }
  if (rc == CVT_OK && backward_range(m, img, proc, B, 1, hi_s, lo_s, dp_segments(m) > 1, -1, nullptr) != cudaSuccess) rc = CVT_ERR_CUDA;
  m->dyn = nullptr;
  return rc;
}
static int reduce_bucket(Model* m, int k) {
  int hi_s, lo_s; dp_segment_range(m, k, &hi_s, &lo_s);
  const int64_t lo = m->w[lo_s].conv_w, hi = k == 0 ? m->n_train : m->w[hi_s + 1].conv_w;      // a stage's parameters are contiguous, ConvEmbed kernel first
  API_CUDA(cudaEventRecord(m->ev_seg[k], m->stream));
  API_CUDA(cudaStreamWaitEvent(m->comm_stream, m->ev_seg[k], 0));
  // sum of per-rank mean-loss gradients; the 1 / world factor is folded into the Adam update (gradient scale in lr_dev)
  const int r = g_nccl.allreduce(m->G + lo, m->G + lo, (size_t)(hi - lo), /*ncclFloat32*/ 7, /*ncclSum*/ 0, m->nccl_comm, m->comm_stream);
  if (r != 0) { set_error(std::string("ncclAllReduce: ") + (g_nccl.errstr ? g_nccl.errstr(r) : "nccl error")); return CVT_ERR_NCCL; }
  // ... and the bucket's Adam update follows on the same stream: the remaining backward reads neither these weights nor these gradients,
  // so only the first stage's small bucket (4 % of the bytes) is reduced and stepped after the backward pass has ended
  API_CUDA(launch_adam(m->P + lo, m->G + lo, m->Mo + lo, m->Vo + lo, hi - lo, 0.f, 0.9f, 0.999f, 1e-7f, 1.f, m->comm_stream, m->lr_dev));
  return CVT_OK;
}
static int train_step_dp(Model* m, const float* img, const float* proc, const float* y, int B, Model::GraphEntry* g, bool philox, uint32_t step) {
  const int nseg = dp_segments(m);
  if (g && !g->exec) {                                  // second sighting: capture every segment into its own graph
    const long long l0 = g_launch_count;
    cudaGraphExec_t ex[CVT_MAX_STAGES] = {};
    int rc = CVT_OK;
    for (int k = 0; k < nseg && rc == CVT_OK; ++k) {
      cudaGraph_t graph = nullptr;
      API_CUDA(cudaStreamBeginCapture(m->stream, cudaStreamCaptureModeThreadLocal));
      rc = train_segment(m, img, proc, y, B, k);
      const cudaError_t e = cudaStreamEndCapture(m->stream, &graph);
      if (rc == CVT_OK && (e != cudaSuccess || !graph)) { set_error(std::string("train segment capture: ") + cudaGetErrorString(e)); rc = CVT_ERR_CUDA; }
      if (rc == CVT_OK && cudaGraphInstantiate(&ex[k], graph, 0) != cudaSuccess) { set_error("cudaGraphInstantiate (train segment) failed"); rc = CVT_ERR_CUDA; }
      if (graph) cudaGraphDestroy(graph);
    }
    m->dropout_step = step;                             // forward_impl advanced it while capturing; the replay below counts
    if (rc != CVT_OK) { for (int k = 0; k < nseg; ++k) if (ex[k]) cudaGraphExecDestroy(ex[k]); cudaGetLastError(); return rc; }
    g->exec = ex[0];
    for (int k = 1; k < nseg; ++k) g->seg[k] = ex[k];
    g->launches = g_launch_count - l0;
    g_launch_count = l0;
  }
  const bool replay = g && g->exec;
  for (int k = 0; k < nseg; ++k) {
    if (replay) API_CUDA(cudaGraphLaunch(k == 0 ? g->exec : g->seg[k], m->stream));
    else { const int rc = train_segment(m, img, proc, y, B, k); if (rc != CVT_OK) return rc; }
    if (!m->peer_ready) { const int rr = reduce_bucket(m, k); if (rr != CVT_OK) return rr; }
  }
  if (replay) {
    g_launch_count += g->launches;
    if (philox) m->dropout_step = step + 1;
    m->last_batch = B; m->last_training = 1; m->have_forward = 1;
  }
  if (m->peer_ready) { API_CUDA(launch_dp_reduce_adam(m, m->stream)); return CVT_OK; }
  API_CUDA(cudaEventRecord(m->ev_red, m->comm_stream));
  API_CUDA(cudaStreamWaitEvent(m->stream, m->ev_red, 0));      // the next step (and any reader of the weights) follows the last bucket's update
  return CVT_OK;
}

static int train_step_dev(Model* m, const float* img, const float* proc, const float* y, int B, float lr) {
  const bool philox = m->dropout_mode == CVT_DROPOUT_PHILOX && m->cfg.drop_rate > 0.f;
  const uint32_t step = m->dropout_step;
  m->adam_t += 1; m->infer_prep_valid = false;
  const bool dp = m->nccl_comm && m->world > 1;
  const float gs = dp ? 1.f / (float)m->world : 1.f;
  API_CUDA(launch_set_step_params(m->lr_dev, adam_alpha(m, lr), gs, step, m->stream));
  Model::GraphEntry* g = (m->use_graphs && !g_prof) ? graph_slot(m, GRAPH_TRAIN, img, proc, y, B) : nullptr;
  if (dp) return train_step_dp(m, img, proc, y, B, g, philox, step);
  if (g && !g->exec) {
    const long long l0 = g_launch_count;
    cudaGraph_t graph = nullptr;
    API_CUDA(cudaStreamBeginCapture(m->stream, cudaStreamCaptureModeThreadLocal));
    const int rc = train_step_body(m, img, proc, y, B);
    const cudaError_t e = cudaStreamEndCapture(m->stream, &graph);
    m->dropout_step = step;                              // forward_impl advanced it while capturing; the replay below counts
    if (rc != CVT_OK || e != cudaSuccess || !graph) {
      if (graph) cudaGraphDestroy(graph);
      if (rc == CVT_OK) set_error(std::string("train step capture: ") + cudaGetErrorString(e));
      cudaGetLastError();
      return rc != CVT_OK ? rc : CVT_ERR_CUDA;
    }
    const cudaError_t ei = cudaGraphInstantiate(&g->exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ei != cudaSuccess) { g->exec = nullptr; set_error(std::string("cudaGraphInstantiate: ") + cudaGetErrorString(ei)); return CVT_ERR_CUDA; }
    g->launches = g_launch_count - l0;
    g_launch_count = l0;
  }
  if (g && g->exec) {
    API_CUDA(cudaGraphLaunch(g->exec, m->stream));
    g_launch_count += g->launches;
    if (philox) m->dropout_step = step + 1;
    m->last_batch = B; m->last_training = 1; m->have_forward = 1;
    return CVT_OK;
  }
  return train_step_body(m, img, proc, y, B);
}

// inference forward (training = 0) through the graph cache; falls back to eager issue on first sighting.
// cached: the weights, BatchNorm moving statistics and everything derived from them alone (K-major / TF32 copies, folded q/k/v
// kernels, BatchNorm scale / shift) are as the previous forward of this call left them: the predict loop's chunks after the first.
static int infer_forward(Model* m, const float* img, const float* proc, int B, bool cached) {
  struct Scope { Model* m; ~Scope() { m->skip_prep = false; } } scope{m};
  // also across calls: repeated inference on unchanged weights (model.predict after model.predict) finds the weight-only work done
  if (m->infer_prep_valid && !m->arena_exposed && m->pt_valid) cached = true;
  m->skip_prep = cached;
  Model::GraphEntry* g = (m->use_graphs && !g_prof) ? graph_slot(m, cached ? GRAPH_INFER_CACHED : GRAPH_INFER, img, proc, nullptr, B) : nullptr;
  if (g && !g->exec) {
    const long long l0 = g_launch_count;
    cudaGraph_t graph = nullptr;
    API_CUDA(cudaStreamBeginCapture(m->stream, cudaStreamCaptureModeThreadLocal));
    const cudaError_t ef = forward_impl(m, img, proc, B, 0);
    const cudaError_t e = cudaStreamEndCapture(m->stream, &graph);
    if (ef != cudaSuccess || e != cudaSuccess || !graph) {
      if (graph) cudaGraphDestroy(graph);
      if (ef == cudaSuccess) set_error(std::string("forward capture: ") + cudaGetErrorString(e));
      cudaGetLastError();
      return CVT_ERR_CUDA;
    }
    const cudaError_t ei = cudaGraphInstantiate(&g->exec, graph, 0);
    cudaGraphDestroy(graph);
    if (ei != cudaSuccess) { g->exec = nullptr; set_error(std::string("cudaGraphInstantiate: ") + cudaGetErrorString(ei)); return CVT_ERR_CUDA; }
    g->launches = g_launch_count - l0;
    g_launch_count = l0;
  }
  if (g && g->exec) {
    API_CUDA(cudaGraphLaunch(g->exec, m->stream));
    g_launch_count += g->launches;
    m->last_batch = B; m->last_training = 0; m->have_forward = 1;
    m->infer_prep_valid = true;
    return CVT_OK;
  }
  API_CUDA(forward_impl(m, img, proc, B, 0));
  m->infer_prep_valid = true;
  return CVT_OK;
}

int cvt_train_step(cvt_model* h, const float* img_dev, const float* proc_dev, const float* y_dev, int batch, float lr) {
  if (!h || !img_dev || !y_dev) { set_error("cvt_train_step: null argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m;
  if (m->cfg.proc_dim > 0 && !proc_dev) { set_error("cvt_train_step: this model needs process parameters"); return CVT_ERR_INVALID; }
  int r = check_batch(m, batch, "cvt_train_step"); if (r) return r;
  API_CUDA(cudaSetDevice(m->cfg.device));
  const long long l0 = g_launch_count;
  r = train_step_dev(m, img_dev, proc_dev, y_dev, batch, lr);
  m->last_img = img_dev; m->last_proc = proc_dev;
  m->launches += g_launch_count - l0;
  return r;
}

int cvt_train_step_host(cvt_model* h, const float* img, const float* proc, const float* y, int batch, float lr, float* loss, float* mae) {
  if (!h || !img || !y) { set_error("cvt_train_step_host: null argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const cvt_config& c = m->cfg;
  if (c.proc_dim > 0 && !proc) { set_error("cvt_train_step_host: this model needs process parameters"); return CVT_ERR_INVALID; }
  int r = check_batch(m, batch, "cvt_train_step_host"); if (r) return r;
  API_CUDA(cudaSetDevice(c.device));
  const size_t per = (size_t)c.image_h * c.image_w * c.in_chans;
  API_CUDA(cudaMemcpyAsync(m->in_img, img, batch * per * sizeof(float), cudaMemcpyHostToDevice, m->stream));
  if (c.proc_dim > 0) API_CUDA(cudaMemcpyAsync(m->in_proc, proc, (size_t)batch * c.proc_dim * sizeof(float), cudaMemcpyHostToDevice, m->stream));
  API_CUDA(cudaMemcpyAsync(m->in_y, y, (size_t)batch * c.num_classes * sizeof(float), cudaMemcpyHostToDevice, m->stream));
  r = cvt_train_step(h, m->in_img, c.proc_dim > 0 ? m->in_proc : nullptr, m->in_y, batch, lr);
  if (r) return r;
  if (loss || mae) return cvt_read_metrics(h, loss, mae);
  return CVT_OK;
}

// =====================================================================================================================
// input pipeline: pinned ring, resident datasets, epochs by permutation (models/CvT(Par).py:418-423,475)
// =====================================================================================================================
static bool host_is_pinned(const void* p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { (void)cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}
// A small persistent pool for the host-side gathers (pageable numpy rows -> pinned ring).  One core copies pageable memory at 5-10 GB/s,
// less than a B200 consumes (64 float64 images = 8.4 MB every 0.5 ms at inference), and spawning threads per batch costs as much as the
// copy itself: the workers are created once and woken per job.  CVT_COPY_THREADS overrides the count (default: half the cores, at most 8).
class CopyPool {
 public:
  static CopyPool& get() { static CopyPool p; return p; }
  int threads() const { return nthreads_; }
  // runs fn(part, nparts) for part = 0 .. nparts-1 (part 0 on the caller's thread); returns when all parts are done
  void run(int nparts, const std::function<void(int, int)>& fn) {
    if (nparts <= 1) { fn(0, 1); return; }
    std::lock_guard<std::mutex> job(job_mu_);            // one job at a time (models on different host threads share the pool)
    {
      std::lock_guard<std::mutex> lk(mu_);
      fn_ = &fn; nparts_ = nparts; next_ = 1; pending_ = nparts - 1; ++epoch_;
    }
    cv_.notify_all();
    fn(0, nparts);
    std::unique_lock<std::mutex> lk(mu_);
    done_.wait(lk, [&] { return pending_ == 0; });
    fn_ = nullptr;
  }
 private:
  CopyPool() {
    const char* e = getenv("CVT_COPY_THREADS");
    int t = e ? atoi(e) : (int)std::min<unsigned>(8, std::max<unsigned>(1, std::thread::hardware_concurrency() / 2));
    nthreads_ = std::max(1, std::min(t, 16));
    for (int i = 1; i < nthreads_; ++i) workers_.emplace_back([this] { loop(); });
  }
  ~CopyPool() {
    { std::lock_guard<std::mutex> lk(mu_); stop_ = true; }
    cv_.notify_all();
    for (auto& w : workers_) w.join();
  }
  void loop() {
    uint64_t seen = 0;
    for (;;) {
      std::unique_lock<std::mutex> lk(mu_);
      cv_.wait(lk, [&] { return stop_ || (epoch_ != seen && fn_ && next_ < nparts_); });
      if (stop_) return;
      while (fn_ && next_ < nparts_) {
        const int part = next_++; const int np = nparts_; const auto* fn = fn_;
        lk.unlock();
        (*fn)(part, np);
        lk.lock();
        if (--pending_ == 0) done_.notify_all();
      }
      seen = epoch_;
    }
  }
  std::mutex mu_, job_mu_; std::condition_variable cv_, done_;
  std::vector<std::thread> workers_;
  const std::function<void(int, int)>* fn_ = nullptr;
  int nthreads_ = 1, nparts_ = 0, next_ = 0, pending_ = 0; uint64_t epoch_ = 0; bool stop_ = false;
};

// dst[i] <- src[idx ? idx[first + i] : first + i], rows of row_bytes; copies of a megabyte or more are split over the pool
static void host_gather(uint8_t* dst, const uint8_t* src, const int32_t* idx, int64_t first, int64_t cnt, size_t row_bytes) {
  auto work = [=](int64_t lo, int64_t hi) {
    if (!idx) { memcpy(dst + lo * row_bytes, src + (first + lo) * row_bytes, (size_t)(hi - lo) * row_bytes); return; }
    for (int64_t i = lo; i < hi; ++i) memcpy(dst + i * row_bytes, src + (int64_t)idx[first + i] * row_bytes, row_bytes);
  };
  const size_t bytes = (size_t)cnt * row_bytes;
  CopyPool& pool = CopyPool::get();
  const int nt = (int)std::min<int64_t>(std::min<int64_t>(pool.threads(), cnt), (int64_t)(bytes >> 19));      // >= 512 KB per thread
  if (nt <= 1) { work(0, cnt); return; }
  pool.run(nt, [&](int part, int np) { work(cnt * part / np, cnt * (part + 1) / np); });
}
static int ring_init(Model* m) {
  if (m->ring_cap) return CVT_OK;
  const size_t cap = (size_t)8 << 20;
  for (int i = 0; i < Model::kRing; ++i) {
    API_CUDA(cudaMallocHost(reinterpret_cast<void**>(&m->ring[i]), cap));
    API_CUDA(cudaEventCreateWithFlags(&m->ring_ev[i], cudaEventDisableTiming));
    m->ring_busy[i] = false;
  }
  m->ring_cap = cap; m->ring_next = 0;
  return CVT_OK;
}
static int ring_acquire(Model* m, int* slot) {
  int r = ring_init(m); if (r) return r;
  *slot = m->ring_next; m->ring_next = (m->ring_next + 1) % Model::kRing;
  if (m->ring_busy[*slot]) { API_CUDA(cudaEventSynchronize(m->ring_ev[*slot])); m->ring_busy[*slot] = false; }
  return CVT_OK;
}
// dst_dev[r] (dense rows of row_bytes) <- src[(idx ? idx[first + r] : first + r)] for r < nrows, asynchronously on stream s.  Pinned
// contiguous sources are copied directly; everything else is gathered row by row into the pinned ring (the only host pass over the data).
static int h2d_rows(Model* m, void* dst_dev, const void* src_host, const int32_t* idx, int64_t first, int64_t nrows, size_t row_bytes, cudaStream_t s) {
  if (nrows <= 0 || row_bytes == 0) return CVT_OK;
  const uint8_t* src = static_cast<const uint8_t*>(src_host);
  if (!idx && host_is_pinned(src + first * row_bytes)) {
    API_CUDA(cudaMemcpyAsync(dst_dev, src + first * row_bytes, (size_t)nrows * row_bytes, cudaMemcpyHostToDevice, s));
    return CVT_OK;
  }
  int r = ring_init(m); if (r) return r;
  if (row_bytes > m->ring_cap) { set_error("h2d_rows: one row exceeds the pinned ring slot (8 MB)"); return CVT_ERR_INVALID; }
  const int64_t per_slot = (int64_t)(m->ring_cap / row_bytes);
  for (int64_t done = 0; done < nrows; done += per_slot) {
    const int64_t cnt = std::min(per_slot, nrows - done);
    int slot; r = ring_acquire(m, &slot); if (r) return r;
    host_gather(m->ring[slot], src, idx, first + done, cnt, row_bytes);
    API_CUDA(cudaMemcpyAsync(static_cast<uint8_t*>(dst_dev) + done * row_bytes, m->ring[slot], (size_t)cnt * row_bytes, cudaMemcpyHostToDevice, s));
    API_CUDA(cudaEventRecord(m->ring_ev[slot], s)); m->ring_busy[slot] = true;
  }
  return CVT_OK;
}
// small float rows (process parameters, targets): gathered and, when the host holds float64, converted on the host into a ring slot
static int h2d_small(Model* m, float* dst_dev, const void* src_host, int f64, const int32_t* idx, int64_t first, int64_t nrows, int width, cudaStream_t s) {
  if (nrows <= 0 || width <= 0) return CVT_OK;
  if (!f64) return h2d_rows(m, dst_dev, src_host, idx, first, nrows, (size_t)width * sizeof(float), s);
  int r = ring_init(m); if (r) return r;
  const int64_t per_slot = (int64_t)(m->ring_cap / ((size_t)width * sizeof(float)));
  const double* src = static_cast<const double*>(src_host);
  for (int64_t done = 0; done < nrows; done += per_slot) {
    const int64_t cnt = std::min(per_slot, nrows - done);
    int slot; r = ring_acquire(m, &slot); if (r) return r;
    float* d = reinterpret_cast<float*>(m->ring[slot]);
    for (int64_t i = 0; i < cnt; ++i) {
      const double* row = src + (idx ? (int64_t)idx[first + done + i] : first + done + i) * width;
      for (int c = 0; c < width; ++c) d[i * width + c] = (float)row[c];
    }
    API_CUDA(cudaMemcpyAsync(dst_dev + done * width, d, (size_t)cnt * width * sizeof(float), cudaMemcpyHostToDevice, s));
    API_CUDA(cudaEventRecord(m->ring_ev[slot], s)); m->ring_busy[slot] = true;
  }
  return CVT_OK;
}
static int ensure_raw64(Model* m) {
  const size_t need = (size_t)m->cfg.max_batch * m->cfg.image_h * m->cfg.image_w * m->cfg.in_chans * sizeof(double);
  if (m->raw64_cap >= need) return CVT_OK;
  for (int i = 0; i < 2; ++i) { if (m->raw64[i]) cudaFree(m->raw64[i]); m->raw64[i] = nullptr; }
  m->raw64_cap = 0;
  for (int i = 0; i < 2; ++i) API_CUDA(cudaMalloc(reinterpret_cast<void**>(&m->raw64[i]), need));
  m->raw64_cap = need;
  return CVT_OK;
}
// image rows -> fp32 device rows; float64 host rows land in raw64[lane] and are converted by a kernel on the same stream (nrows <= max_batch)
static int h2d_images(Model* m, float* dst_dev, const void* src_host, int f64, const int32_t* idx, int64_t first, int nrows, int lane, cudaStream_t s) {
  const cvt_config& c = m->cfg;
  const size_t per = (size_t)c.image_h * c.image_w * c.in_chans;
  if (!f64) return h2d_rows(m, dst_dev, src_host, idx, first, nrows, per * sizeof(float), s);
  int r = ensure_raw64(m); if (r) return r;
  r = h2d_rows(m, m->raw64[lane], src_host, idx, first, nrows, per * sizeof(double), s); if (r) return r;
  API_CUDA(launch_f64_to_f32(dst_dev, m->raw64[lane], (long long)nrows * (long long)per, s));
  return CVT_OK;
}
// ---- double-buffered host -> device staging ----------------------------------------------------------------------------
struct Staging { float* img; float* proc; float* y; };
static Staging staging_of(Model* m, int slot) {
  return slot ? Staging{m->in_img2, m->in_proc2, m->in_y2} : Staging{m->in_img, m->in_proc, m->in_y};
}
// enqueue the copies of rows [o, o + B) of `a` on the copy stream (after the step that last read this staging slot) and make the model
// stream wait for them.  Pinned float32 sources are copied in place; pageable and float64 ones (what numpy hands over) pass through the
// library's pinned ring, so that the copy of chunk i+1 is asynchronous and really overlaps the step on chunk i.
static int stage_chunk(Model* m, int64_t i, const cvt_host_arrays& a, int64_t o, int B, bool with_y) {
  const cvt_config& c = m->cfg;
  const int slot = (int)(i & 1); const Staging st = staging_of(m, slot);
  if (i >= 2) API_CUDA(cudaStreamWaitEvent(m->copy_stream, m->ev_done[slot], 0));
  { const int r = h2d_images(m, st.img, a.img, a.img_f64, nullptr, o, B, slot, m->copy_stream); if (r) return r; }
  if (c.proc_dim > 0) { const int r = h2d_small(m, st.proc, a.proc, a.proc_f64, nullptr, o, B, c.proc_dim, m->copy_stream); if (r) return r; }
  if (with_y) { const int r = h2d_small(m, st.y, a.y, a.y_f64, nullptr, o, B, c.num_classes, m->copy_stream); if (r) return r; }
  API_CUDA(cudaEventRecord(m->ev_copy[slot], m->copy_stream));
  API_CUDA(cudaStreamWaitEvent(m->stream, m->ev_copy[slot], 0));
  return CVT_OK;
}

static int ensure_pin_series(Model* m, int64_t floats) {
  if (m->pin_series_cap >= floats) return CVT_OK;
  API_CUDA(cudaStreamSynchronize(m->stream));
  if (m->pin_series) cudaFreeHost(m->pin_series);
  m->pin_series = nullptr; m->pin_series_cap = 0;
  API_CUDA(cudaMallocHost(&m->pin_series, (size_t)(floats + 64) * sizeof(float)));
  m->pin_series_cap = floats + 64;
  return CVT_OK;
}

// The batch loop of one `model.fit` epoch (models/CvT(Par).py:475): n samples in the caller's (already shuffled) order, batches of
// `batch` with the last partial batch kept, per-batch loss / MAE returned.  Batch i+1 is copied while batch i trains.
int cvt_train_epoch_host(cvt_model* h, const float* img, const float* proc, const float* y, int64_t n, int batch, float lr,
                         float* loss_out, float* mae_out) {
  if (!h || !img || !y || n < 1) { set_error("cvt_train_epoch_host: bad argument"); return CVT_ERR_INVALID; }
  const cvt_host_arrays a{img, proc, y, 0, 0, 0, n};
  return cvt_train_epoch_stream(h, &a, nullptr, 0, batch, lr, -1, loss_out, mae_out);
}

int cvt_predict_arrays(cvt_model* h, const cvt_host_arrays* a, float* out) {
  if (!h || !a || !a->img || !out || a->n < 0) { set_error("cvt_predict: bad argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const cvt_config& c = m->cfg;
  if (c.proc_dim > 0 && !a->proc) { set_error("cvt_predict: this model needs process parameters"); return CVT_ERR_INVALID; }
  API_CUDA(cudaSetDevice(c.device));
  const int64_t n = a->n;
  const long long l0 = g_launch_count;
  { const int rp = ensure_pin_series(m, n * c.num_classes); if (rp) return rp; }       // device -> pinned -> out: the loop never blocks
  Staging last = staging_of(m, 0);
  int64_t i = 0;
  for (int64_t o = 0; o < n; o += c.max_batch, ++i) {        // chunk i+1 is copied while chunk i runs
    const int B = (int)std::min<int64_t>(c.max_batch, n - o);
    { const int rs = stage_chunk(m, i, *a, o, B, false); if (rs) return rs; }
    last = staging_of(m, (int)(i & 1));
    { const int rf = infer_forward(m, last.img, c.proc_dim > 0 ? last.proc : nullptr, B, i > 0); if (rf) return rf; }
    API_CUDA(cudaMemcpyAsync(m->pin_series + o * c.num_classes, m->pred, (size_t)B * c.num_classes * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
    API_CUDA(cudaEventRecord(m->ev_done[i & 1], m->stream));
  }
  API_CUDA(cudaStreamSynchronize(m->stream));
  memcpy(out, m->pin_series, (size_t)n * c.num_classes * sizeof(float));
  m->last_img = last.img; m->last_proc = c.proc_dim > 0 ? last.proc : nullptr;
  m->launches += g_launch_count - l0;
  return CVT_OK;
}
int cvt_predict_host(cvt_model* h, const float* img, const float* proc, int64_t n, float* out) {
  if (!h || !img || !out || n < 0) { set_error("cvt_predict_host: bad argument"); return CVT_ERR_INVALID; }
  const cvt_host_arrays a{img, proc, nullptr, 0, 0, 0, n};
  return cvt_predict_arrays(h, &a, out);
}

int cvt_evaluate_arrays(cvt_model* h, const cvt_host_arrays* a, float* loss, float* mae) {
  if (!h || !a || !a->img || !a->y || a->n < 1) { set_error("cvt_evaluate: bad argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const cvt_config& c = m->cfg;
  if (c.proc_dim > 0 && !a->proc) { set_error("cvt_evaluate: this model needs process parameters"); return CVT_ERR_INVALID; }
  API_CUDA(cudaSetDevice(c.device));
  const int64_t n = a->n;
  const long long l0 = g_launch_count;
  API_CUDA(cudaMemsetAsync(m->metrics + 2, 0, 2 * sizeof(float), m->stream));
  Staging last = staging_of(m, 0);
  int64_t i = 0;
  for (int64_t o = 0; o < n; o += c.max_batch, ++i) {        // as cvt_predict_arrays: chunk i+1 is copied while chunk i runs
    const int B = (int)std::min<int64_t>(c.max_batch, n - o);
    { const int rs = stage_chunk(m, i, *a, o, B, true); if (rs) return rs; }
    last = staging_of(m, (int)(i & 1));
    { const int rf = infer_forward(m, last.img, c.proc_dim > 0 ? last.proc : nullptr, B, i > 0); if (rf) return rf; }
    API_CUDA(launch_mse_loss(m->pred, last.y, B * c.num_classes, m->metrics + 2, nullptr, 1.f, 1, m->stream));
    API_CUDA(cudaEventRecord(m->ev_done[i & 1], m->stream));
  }
  m->last_img = last.img; m->last_proc = c.proc_dim > 0 ? last.proc : nullptr;
  API_CUDA(cudaMemcpyAsync(m->pin_metrics + 2, m->metrics + 2, 2 * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
  API_CUDA(cudaStreamSynchronize(m->stream));
  const double cnt = (double)n * c.num_classes;
  if (loss) *loss = (float)(m->pin_metrics[2] / cnt);
  if (mae) *mae = (float)(m->pin_metrics[3] / cnt);
  m->launches += g_launch_count - l0;
  return CVT_OK;
}
int cvt_evaluate_host(cvt_model* h, const float* img, const float* proc, const float* y, int64_t n, float* loss, float* mae) {
  if (!h || !img || !y || n < 1) { set_error("cvt_evaluate_host: bad argument"); return CVT_ERR_INVALID; }
  const cvt_host_arrays a{img, proc, y, 0, 0, 0, n};
  return cvt_evaluate_arrays(h, &a, loss, mae);
}

// ---- resident datasets ----------------------------------------------------------------------------------------------------------
static int check_slot(Model* m, int slot, const char* who, bool need_data) {
  if (slot < 0 || slot >= CVT_DATASET_SLOTS) { set_error(std::string(who) + ": slot outside [0, CVT_DATASET_SLOTS)"); return CVT_ERR_INVALID; }
  if (need_data && m->ds[slot].n < 1) { set_error(std::string(who) + ": dataset slot " + std::to_string(slot) + " is empty (cvt_dataset_alloc first)"); return CVT_ERR_STATE; }
  return CVT_OK;
}
static int upload_order(Model* m, const int32_t* order, int64_t n, int64_t limit, const char* who) {
  if (!order) return CVT_OK;
  for (int64_t i = 0; i < n; ++i) if (order[i] < 0 || order[i] >= limit) { set_error(std::string(who) + ": order[" + std::to_string(i) + "] outside the data set"); return CVT_ERR_INVALID; }
  if (m->order_cap < n) {
    API_CUDA(cudaStreamSynchronize(m->stream)); API_CUDA(cudaStreamSynchronize(m->copy_stream));
    if (m->order_dev) cudaFree(m->order_dev);
    m->order_dev = nullptr; m->order_cap = 0;
    API_CUDA(cudaMalloc(reinterpret_cast<void**>(&m->order_dev), (size_t)n * sizeof(int32_t)));
    m->order_cap = n;
  }
  // (the previous epoch's gathers have been issued on m->stream / copy_stream: order the overwrite after both)
  API_CUDA(cudaStreamSynchronize(m->copy_stream));
  API_CUDA(cudaMemcpyAsync(m->order_dev, order, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, m->stream));
  API_CUDA(cudaStreamSynchronize(m->stream));
  return CVT_OK;
}

int cvt_dataset_free(cvt_model* h, int slot) {
  if (!h) return CVT_ERR_INVALID;
  Model* m = &h->m;
  int r = check_slot(m, slot, "cvt_dataset_free", false); if (r) return r;
  if (m->stream) cudaStreamSynchronize(m->stream);
  if (m->copy_stream) cudaStreamSynchronize(m->copy_stream);
  Model::Dataset& d = m->ds[slot];
  if (d.img) cudaFree(d.img); if (d.proc) cudaFree(d.proc); if (d.y) cudaFree(d.y);
  d = Model::Dataset{};
  return CVT_OK;
}
int cvt_dataset_alloc(cvt_model* h, int slot, int64_t n) {
  if (!h || n < 1) { set_error("cvt_dataset_alloc: bad argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const cvt_config& c = m->cfg;
  int r = check_slot(m, slot, "cvt_dataset_alloc", false); if (r) return r;
  API_CUDA(cudaSetDevice(c.device));
  cvt_dataset_free(h, slot);
  Model::Dataset& d = m->ds[slot];
  const size_t per = (size_t)c.image_h * c.image_w * c.in_chans;
  cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&d.img), (size_t)n * per * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&d.proc), (size_t)n * std::max(1, c.proc_dim) * sizeof(float));
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&d.y), (size_t)n * c.num_classes * sizeof(float));
  if (e != cudaSuccess) {
    set_error(std::string("cvt_dataset_alloc: ") + std::to_string(n) + " samples do not fit in device memory: " + cudaGetErrorString(e));
    (void)cudaGetLastError(); cvt_dataset_free(h, slot); return CVT_ERR_CUDA;
  }
  d.n = n;
  API_CUDA(cudaMemsetAsync(d.proc, 0, (size_t)n * std::max(1, c.proc_dim) * sizeof(float), m->stream));
  API_CUDA(cudaMemsetAsync(d.y, 0, (size_t)n * c.num_classes * sizeof(float), m->stream));
  return CVT_OK;
}
int64_t cvt_dataset_count(const cvt_model* h, int slot) {
  if (!h || slot < 0 || slot >= CVT_DATASET_SLOTS) return CVT_ERR_INVALID;
  return h->m.ds[slot].n;
}
int cvt_dataset_write(cvt_model* h, int slot, int64_t offset, const cvt_host_arrays* a) {
  if (!h || !a || a->n < 0 || offset < 0) { set_error("cvt_dataset_write: bad argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const cvt_config& c = m->cfg;
  int r = check_slot(m, slot, "cvt_dataset_write", true); if (r) return r;
  Model::Dataset& d = m->ds[slot];
  if (offset + a->n > d.n) { set_error("cvt_dataset_write: rows past the end of the dataset"); return CVT_ERR_INVALID; }
  API_CUDA(cudaSetDevice(c.device));
  const size_t per = (size_t)c.image_h * c.image_w * c.in_chans;
  const long long l0 = g_launch_count;
  if (a->img) {
    for (int64_t o = 0; o < a->n; o += c.max_batch) {
      const int cnt = (int)std::min<int64_t>(c.max_batch, a->n - o);
      r = h2d_images(m, d.img + (offset + o) * per, a->img, a->img_f64, nullptr, o, cnt, 0, m->stream); if (r) return r;
    }
  }
  if (a->proc && c.proc_dim > 0) { r = h2d_small(m, d.proc + offset * c.proc_dim, a->proc, a->proc_f64, nullptr, 0, a->n, c.proc_dim, m->stream); if (r) return r; }
  if (a->y) { r = h2d_small(m, d.y + offset * c.num_classes, a->y, a->y_f64, nullptr, 0, a->n, c.num_classes, m->stream); if (r) return r; }
  API_CUDA(cudaStreamSynchronize(m->stream));
  m->launches += g_launch_count - l0;
  return CVT_OK;
}
int cvt_dataset_write_u8(cvt_model* h, int slot, int64_t offset, int64_t count, const uint8_t* pixels, int src_h, int src_w, int src_c) {
  if (!h || !pixels || count < 0 || offset < 0 || src_h < 1 || src_w < 1 || (src_c != 1 && src_c != 3)) { set_error("cvt_dataset_write_u8: bad argument (src_c must be 1 or 3)"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const cvt_config& c = m->cfg;
  int r = check_slot(m, slot, "cvt_dataset_write_u8", true); if (r) return r;
  if (c.in_chans != 1) { set_error("cvt_dataset_write_u8: the grey conversion needs a model with in_chans = 1"); return CVT_ERR_INVALID; }
  Model::Dataset& d = m->ds[slot];
  if (offset + count > d.n) { set_error("cvt_dataset_write_u8: rows past the end of the dataset"); return CVT_ERR_INVALID; }
  API_CUDA(cudaSetDevice(c.device));
  const int dh = c.image_h, dw = c.image_w;
  if (!m->pp_tab) {
    API_CUDA(cudaMalloc(reinterpret_cast<void**>(&m->pp_tab), (size_t)3 * (8192 + 8192) * sizeof(int32_t)));
    API_CUDA(cudaMalloc(reinterpret_cast<void**>(&m->pp_lut), 256 * sizeof(float)));
    float lut[256];
    for (int g = 0; g < 256; ++g) lut[g] = (float)((double)g / 255.0);      // image / 255.0 in float64, then Keras' cast to float32
    API_CUDA(cudaMemcpy(m->pp_lut, lut, sizeof(lut), cudaMemcpyHostToDevice));
  }
  if (dh > 8192 || dw > 8192) { set_error("cvt_dataset_write_u8: image larger than 8192"); return CVT_ERR_INVALID; }
  if (m->pp_key[0] != src_h || m->pp_key[1] != src_w || m->pp_key[2] != dh || m->pp_key[3] != dw) {
    std::vector<int32_t> tab((size_t)3 * (dw + dh));
    resize_axis_table(src_w, dw, 1, tab.data(), tab.data() + dw, tab.data() + 2 * dw);
    resize_axis_table(src_h, dh, 0, tab.data() + 3 * dw, tab.data() + 3 * dw + dh, tab.data() + 3 * dw + 2 * dh);
    API_CUDA(cudaStreamSynchronize(m->stream));
    API_CUDA(cudaMemcpy(m->pp_tab, tab.data(), tab.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    m->pp_key[0] = src_h; m->pp_key[1] = src_w; m->pp_key[2] = dh; m->pp_key[3] = dw;
  }
  const size_t src_bytes = (size_t)src_h * src_w * src_c;
  const int64_t chunk = std::max<int64_t>(1, ((int64_t)64 << 20) / (int64_t)src_bytes);
  if (m->pp_pix_cap < (size_t)std::min<int64_t>(chunk, count) * src_bytes) {
    API_CUDA(cudaStreamSynchronize(m->stream));
    if (m->pp_pix) cudaFree(m->pp_pix);
    m->pp_pix = nullptr; m->pp_pix_cap = 0;
    API_CUDA(cudaMalloc(reinterpret_cast<void**>(&m->pp_pix), (size_t)chunk * src_bytes));
    m->pp_pix_cap = (size_t)chunk * src_bytes;
  }
  const long long l0 = g_launch_count;
  const size_t per = (size_t)dh * dw;
  for (int64_t o = 0; o < count; o += chunk) {
    const int64_t cnt = std::min(chunk, count - o);
    r = h2d_rows(m, m->pp_pix, pixels, nullptr, o, cnt, src_bytes, m->stream); if (r) return r;
    API_CUDA(launch_preprocess_u8(m->pp_pix, src_h, src_w, src_c, m->pp_tab, m->pp_lut, dh, dw, cnt, d.img + (offset + o) * per, m->stream));
  }
  API_CUDA(cudaStreamSynchronize(m->stream));
  m->launches += g_launch_count - l0;
  return CVT_OK;
}
int cvt_dataset_read(cvt_model* h, int slot, int64_t offset, int64_t count, float* img, float* proc, float* y) {
  if (!h || offset < 0 || count < 0) { set_error("cvt_dataset_read: bad argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const cvt_config& c = m->cfg;
  int r = check_slot(m, slot, "cvt_dataset_read", true); if (r) return r;
  Model::Dataset& d = m->ds[slot];
  if (offset + count > d.n) { set_error("cvt_dataset_read: rows past the end of the dataset"); return CVT_ERR_INVALID; }
  const size_t per = (size_t)c.image_h * c.image_w * c.in_chans;
  if (img) API_CUDA(cudaMemcpyAsync(img, d.img + offset * per, (size_t)count * per * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
  if (proc && c.proc_dim > 0) API_CUDA(cudaMemcpyAsync(proc, d.proc + offset * c.proc_dim, (size_t)count * c.proc_dim * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
  if (y) API_CUDA(cudaMemcpyAsync(y, d.y + offset * c.num_classes, (size_t)count * c.num_classes * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
  API_CUDA(cudaStreamSynchronize(m->stream));
  return CVT_OK;
}

int cvt_train_epoch_stream(cvt_model* h, const cvt_host_arrays* a, const int32_t* order, int64_t n_order, int batch, float lr, int fill_slot,
                           float* loss_out, float* mae_out) {
  if (!h || !a || !a->img || !a->y || a->n < 1) { set_error("cvt_train_epoch_stream: bad argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const cvt_config& c = m->cfg;
  if (c.proc_dim > 0 && !a->proc) { set_error("cvt_train_epoch_stream: this model needs process parameters"); return CVT_ERR_INVALID; }
  int r = check_batch(m, batch, "cvt_train_epoch_stream"); if (r) return r;
  const int64_t n = order ? n_order : a->n;
  if (n < 1) { set_error("cvt_train_epoch_stream: empty epoch"); return CVT_ERR_INVALID; }
  if (fill_slot >= 0) {
    r = check_slot(m, fill_slot, "cvt_train_epoch_stream", true); if (r) return r;
    if (m->ds[fill_slot].n != a->n) { set_error("cvt_train_epoch_stream: the dataset to fill must have as many rows as the host arrays"); return CVT_ERR_INVALID; }
  }
  API_CUDA(cudaSetDevice(c.device));
  r = upload_order(m, order, n, a->n, "cvt_train_epoch_stream"); if (r) return r;
  const int64_t nb = (n + batch - 1) / batch;
  r = ensure_pin_series(m, 2 * nb); if (r) return r;
  const size_t per = (size_t)c.image_h * c.image_w * c.in_chans;
  const long long l0 = g_launch_count;
  for (int64_t i = 0; i < nb; ++i) {
    const int64_t o = i * batch; const int B = (int)std::min<int64_t>(batch, n - o);
    const int slot = (int)(i & 1); const Staging st = staging_of(m, slot);
    if (i >= 2) API_CUDA(cudaStreamWaitEvent(m->copy_stream, m->ev_done[slot], 0));
    r = h2d_images(m, st.img, a->img, a->img_f64, order, o, B, slot, m->copy_stream); if (r) return r;
    if (c.proc_dim > 0) { r = h2d_small(m, st.proc, a->proc, a->proc_f64, order, o, B, c.proc_dim, m->copy_stream); if (r) return r; }
    r = h2d_small(m, st.y, a->y, a->y_f64, order, o, B, c.num_classes, m->copy_stream); if (r) return r;
    if (fill_slot >= 0) {      // the batch also lands at its own rows of the resident dataset
      Model::Dataset& d = m->ds[fill_slot];
      const int32_t* di = order ? m->order_dev + o : nullptr;
      const int64_t base = order ? 0 : o;
      API_CUDA(launch_move_rows(d.img + base * per, di, st.img, nullptr, B, (long long)per, d.proc + base * c.proc_dim, c.proc_dim > 0 ? st.proc : nullptr,
                                c.proc_dim, d.y + base * c.num_classes, st.y, c.num_classes, m->copy_stream));
    }
    API_CUDA(cudaEventRecord(m->ev_copy[slot], m->copy_stream));
    API_CUDA(cudaStreamWaitEvent(m->stream, m->ev_copy[slot], 0));
    r = train_step_dev(m, st.img, c.proc_dim > 0 ? st.proc : nullptr, st.y, B, lr); if (r) return r;
    API_CUDA(cudaMemcpyAsync(m->pin_series + 2 * i, m->metrics, 2 * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
    API_CUDA(cudaEventRecord(m->ev_done[slot], m->stream));
    m->last_img = st.img; m->last_proc = c.proc_dim > 0 ? st.proc : nullptr;
  }
  API_CUDA(cudaStreamSynchronize(m->copy_stream));
  API_CUDA(cudaStreamSynchronize(m->stream));
  for (int64_t i = 0; i < nb; ++i) { if (loss_out) loss_out[i] = m->pin_series[2 * i]; if (mae_out) mae_out[i] = m->pin_series[2 * i + 1]; }
  m->launches += g_launch_count - l0;
  return CVT_OK;
}

int cvt_train_epoch_resident(cvt_model* h, int slot, const int32_t* order, int64_t n_order, int batch, float lr, float* loss_out, float* mae_out) {
  if (!h) { set_error("cvt_train_epoch_resident: null model"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const cvt_config& c = m->cfg;
  int r = check_slot(m, slot, "cvt_train_epoch_resident", true); if (r) return r;
  r = check_batch(m, batch, "cvt_train_epoch_resident"); if (r) return r;
  Model::Dataset& d = m->ds[slot];
  const int64_t n = order ? n_order : d.n;
  if (n < 1) { set_error("cvt_train_epoch_resident: empty epoch"); return CVT_ERR_INVALID; }
  API_CUDA(cudaSetDevice(c.device));
  r = upload_order(m, order, n, d.n, "cvt_train_epoch_resident"); if (r) return r;
  const int64_t nb = (n + batch - 1) / batch;
  r = ensure_pin_series(m, 2 * nb); if (r) return r;
  const size_t per = (size_t)c.image_h * c.image_w * c.in_chans;
  const Staging st = staging_of(m, 0);
  const long long l0 = g_launch_count;
  for (int64_t i = 0; i < nb; ++i) {
    const int64_t o = i * batch; const int B = (int)std::min<int64_t>(batch, n - o);
    const int32_t* si = order ? m->order_dev + o : nullptr;
    const int64_t base = order ? 0 : o;
    API_CUDA(launch_move_rows(st.img, nullptr, d.img + base * per, si, B, (long long)per, st.proc, c.proc_dim > 0 ? d.proc + base * c.proc_dim : nullptr,
                              c.proc_dim, st.y, d.y + base * c.num_classes, c.num_classes, m->stream));
    r = train_step_dev(m, st.img, c.proc_dim > 0 ? st.proc : nullptr, st.y, B, lr); if (r) return r;
    API_CUDA(cudaMemcpyAsync(m->pin_series + 2 * i, m->metrics, 2 * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
  }
  m->last_img = st.img; m->last_proc = c.proc_dim > 0 ? st.proc : nullptr;
  API_CUDA(cudaStreamSynchronize(m->stream));
  for (int64_t i = 0; i < nb; ++i) { if (loss_out) loss_out[i] = m->pin_series[2 * i]; if (mae_out) mae_out[i] = m->pin_series[2 * i + 1]; }
  m->launches += g_launch_count - l0;
  return CVT_OK;
}

// forward(training = 0) over a resident dataset in chunks of max_batch; with_loss: MSE / MAE sums, else predictions into pin_series
static int resident_forward(Model* m, int slot, bool with_loss) {
  const cvt_config& c = m->cfg; Model::Dataset& d = m->ds[slot];
  const size_t per = (size_t)c.image_h * c.image_w * c.in_chans;
  const Staging st = staging_of(m, 0);
  int64_t i = 0;
  for (int64_t o = 0; o < d.n; o += c.max_batch, ++i) {
    const int B = (int)std::min<int64_t>(c.max_batch, d.n - o);
    // the chunk is copied into the staging buffers so that every chunk replays the same captured graph
    API_CUDA(launch_move_rows(st.img, nullptr, d.img + o * per, nullptr, B, (long long)per, st.proc, c.proc_dim > 0 ? d.proc + o * c.proc_dim : nullptr,
                              c.proc_dim, st.y, d.y + o * c.num_classes, c.num_classes, m->stream));
    { const int rf = infer_forward(m, st.img, c.proc_dim > 0 ? st.proc : nullptr, B, i > 0); if (rf) return rf; }
    if (with_loss) API_CUDA(launch_mse_loss(m->pred, st.y, B * c.num_classes, m->metrics + 2, nullptr, 1.f, 1, m->stream));
    else API_CUDA(cudaMemcpyAsync(m->pin_series + o * c.num_classes, m->pred, (size_t)B * c.num_classes * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
  }
  m->last_img = st.img; m->last_proc = c.proc_dim > 0 ? st.proc : nullptr;
  return CVT_OK;
}
int cvt_evaluate_resident(cvt_model* h, int slot, float* loss, float* mae) {
  if (!h) { set_error("cvt_evaluate_resident: null model"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const cvt_config& c = m->cfg;
  int r = check_slot(m, slot, "cvt_evaluate_resident", true); if (r) return r;
  API_CUDA(cudaSetDevice(c.device));
  const long long l0 = g_launch_count;
  API_CUDA(cudaMemsetAsync(m->metrics + 2, 0, 2 * sizeof(float), m->stream));
  r = resident_forward(m, slot, true); if (r) return r;
  API_CUDA(cudaMemcpyAsync(m->pin_metrics + 2, m->metrics + 2, 2 * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
  API_CUDA(cudaStreamSynchronize(m->stream));
  const double cnt = (double)m->ds[slot].n * c.num_classes;
  if (loss) *loss = (float)(m->pin_metrics[2] / cnt);
  if (mae) *mae = (float)(m->pin_metrics[3] / cnt);
  m->launches += g_launch_count - l0;
  return CVT_OK;
}
int cvt_predict_resident(cvt_model* h, int slot, float* out) {
  if (!h || !out) { set_error("cvt_predict_resident: null argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const cvt_config& c = m->cfg;
  int r = check_slot(m, slot, "cvt_predict_resident", true); if (r) return r;
  API_CUDA(cudaSetDevice(c.device));
  const long long l0 = g_launch_count;
  r = ensure_pin_series(m, m->ds[slot].n * c.num_classes); if (r) return r;
  r = resident_forward(m, slot, false); if (r) return r;
  API_CUDA(cudaStreamSynchronize(m->stream));
  memcpy(out, m->pin_series, (size_t)m->ds[slot].n * c.num_classes * sizeof(float));
  m->launches += g_launch_count - l0;
  return CVT_OK;
}

int cvt_set_dropout(cvt_model* h, int mode, uint64_t seed) {
  if (!h || mode < 0 || mode > 2) { set_error("cvt_set_dropout: bad mode"); return CVT_ERR_INVALID; }
  if (h->m.dropout_mode == mode && h->m.dropout_seed == seed) return CVT_OK;      // same stream: the step counter keeps running
  h->m.dropout_mode = mode; h->m.dropout_seed = seed; h->m.dropout_step = 0;
  clear_graphs(&h->m);                                   // mode and seed are baked into captured launches
  return CVT_OK;
}
int cvt_set_dropout_mask(cvt_model* h, int site, const uint8_t* host_mask, int64_t count) {
  if (!h || !host_mask || site < 0 || site / 4 >= h->m.n_stages || (site % 4 == 3 && !(h->m.cfg.attn_drop_rate > 0.f))) { set_error("cvt_set_dropout_mask: bad site"); return CVT_ERR_INVALID; }
  Model* m = &h->m; const StageGeo& g = m->geo[site / 4]; StageBuf& a = m->buf[site / 4];
  const int64_t width = (site % 4 == 1) ? (int64_t)m->cfg.mlp_ratio * g.D : g.D;
  const int64_t cap = (int64_t)m->cfg.max_batch * g.Tc * width;
  if (count < 0 || count > cap) { set_error("cvt_set_dropout_mask: count exceeds max_batch*Tc*width"); return CVT_ERR_INVALID; }
  uint8_t* dst = site % 4 == 0 ? a.mask_proj : (site % 4 == 1 ? a.mask1 : (site % 4 == 2 ? a.mask2 : a.mask_attn));
  API_CUDA(cudaMemcpyAsync(dst, host_mask, (size_t)count, cudaMemcpyHostToDevice, m->stream));
  API_CUDA(cudaStreamSynchronize(m->stream));
  return CVT_OK;
}

// ---- taps -------------------------------------------------------------------------------------------------------------
static int parse_stage(const char* name, int* stage, const char** field) {
  if (name[0] != 's' || name[1] < '1' || name[1] > '9' || name[2] != '.') return -1;
  *stage = name[1] - '1'; *field = name + 3;
  return 0;
}

int64_t cvt_get_tap(cvt_model* h, const char* name, float* dst, int64_t capacity) {
  if (!h || !name) { set_error("cvt_get_tap: null argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m;
  if (!m->have_forward) { set_error("cvt_get_tap: no forward pass has run"); return CVT_ERR_STATE; }
  const int B = m->last_batch; cudaStream_t s = m->stream;
  const float* src = nullptr; int64_t n = 0;
  int si; const char* f;
  if (!strcmp(name, "pred")) { src = m->pred; n = (int64_t)B * m->cfg.num_classes; }
  else if (!strcmp(name, "head.feat")) { src = m->feat; n = (int64_t)B * m->feat_dim; }
  else if (!strcmp(name, "head.proc1")) { src = m->p1; n = (int64_t)B * m->cfg.proc_hidden; }
  else if (parse_stage(name, &si, &f) == 0 && si < m->n_stages) {
    const StageGeo& g = m->geo[si]; StageBuf& a = m->buf[si];
    const int64_t RD = (int64_t)B * g.Tc * g.D, TD = (int64_t)B * g.T * g.D;
    const RowMap strip = g.has_cls ? RowMap{g.T, g.Tc, 1} : rowmap_none();
    if (!strcmp(f, "x0")) { src = a.x0; n = RD; }
    else if (!strcmp(f, "xn")) { src = a.xn; n = RD; }
    else if (!strcmp(f, "qkv1") || !strcmp(f, "a1")) {
      if (m->folded) { set_error(std::string("cvt_get_tap: '") + name + "' does not exist in TF32 mode (the stacked linear maps are folded into one GEMM)"); return CVT_ERR_STATE; }
      if (!strcmp(f, "qkv1") && m->cfg.skip_proj_qkv) { set_error("cvt_get_tap: this model does not apply proj_q/k/v"); return CVT_ERR_STATE; }
      if (!strcmp(f, "qkv1")) { src = a.qkv1; n = 3 * RD; } else { src = a.a1; n = RD; }
    }
    else if (!strcmp(f, "qkv2")) { src = a.qkv2; n = 3 * RD; }
    else if (!strcmp(f, "attn_o")) { src = a.attn_o; n = RD; }
    else if (!strcmp(f, "x1")) { src = a.x1; n = RD; }
    else if (m->mlp_fused_last && mlp_fused_supports(m, si) && (!strcmp(f, "y") || !strcmp(f, "hpre") || !strcmp(f, "hact"))) {
      set_error(std::string("cvt_get_tap: '") + name + "' is not materialised by the fused inference Mlp kernel (set CVT_NO_FUSED_MLP=1 to get it)");
      return CVT_ERR_STATE;
    }
    else if (!strcmp(f, "y")) { src = a.y; n = RD; }
    else if (!strcmp(f, "hpre")) { src = a.hpre; n = RD * m->cfg.mlp_ratio; }
    else if (!strcmp(f, "hact")) { src = a.hact; n = RD * m->cfg.mlp_ratio; }
    else if (!strcmp(f, "x2")) { src = a.x2; n = RD; }
    else if (!strcmp(f, "embed") || !strcmp(f, "out")) {
      const float* from = !strcmp(f, "embed") ? a.x0 : a.x2;
      API_CUDA(launch_copy_rows(m->gQ2, rowmap_none(), g.D, from, strip, g.D, B * g.T, g.D, s));
      src = m->gQ2; n = TD;
    } else if (!strcmp(f, "conv_qkv")) {
      for (int z = 0; z < 3; ++z)
        API_CUDA(launch_copy_rows(m->gQ2 + z * TD, rowmap_none(), g.D, a.conv + z * RD, strip, g.D, B * g.T, g.D, s));
      src = m->gQ2; n = 3 * TD;
    }
  }
  if (!src) { set_error(std::string("cvt_get_tap: unknown tap '") + name + "'"); return CVT_ERR_INVALID; }
  if (dst) {
    if (capacity < n) { set_error("cvt_get_tap: destination too small"); return CVT_ERR_INVALID; }
    API_CUDA(cudaMemcpyAsync(dst, src, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, s));
    API_CUDA(cudaStreamSynchronize(s));
  }
  return n;
}

int cvt_backward_to_tap(cvt_model* h, const float* dout_host, const char* name, float* dst, int64_t capacity) {
  if (!h || !dout_host || !name || !dst) { set_error("cvt_backward_to_tap: null argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m;
  if (!m->have_forward || !m->last_img) { set_error("cvt_backward_to_tap: call cvt_forward first"); return CVT_ERR_STATE; }
  int si; const char* f;
  if (parse_stage(name, &si, &f) != 0 || si >= m->n_stages || strcmp(f, "x2") != 0) { set_error("cvt_backward_to_tap: tap must be s{i}.x2"); return CVT_ERR_INVALID; }
  const int B = m->last_batch; const StageGeo& g = m->geo[si];
  const int64_t n = (int64_t)B * g.Tc * g.D;
  if (capacity < n) { set_error("cvt_backward_to_tap: destination too small"); return CVT_ERR_INVALID; }
  API_CUDA(cudaMemcpyAsync(m->dpred, dout_host, (size_t)B * m->cfg.num_classes * sizeof(float), cudaMemcpyHostToDevice, m->stream));
  float* gptr = nullptr;
  API_CUDA(backward_impl(m, m->last_img, m->last_proc, B, m->last_training, si, &gptr));
  API_CUDA(cudaMemcpyAsync(dst, gptr, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
  API_CUDA(cudaStreamSynchronize(m->stream));
  return CVT_OK;
}

int cvt_synchronize(cvt_model* h) {
  if (!h) return CVT_ERR_INVALID;
  API_CUDA(cudaStreamSynchronize(h->m.stream));
  return CVT_OK;
}
void* cvt_stream(cvt_model* h) { return h ? (void*)h->m.stream : nullptr; }
int cvt_set_graphs(cvt_model* h, int enable) { if (!h) return CVT_ERR_INVALID; h->m.use_graphs = enable; if (!enable) clear_graphs(&h->m); return CVT_OK; }
int64_t cvt_kernel_launches(const cvt_model* h) { return h ? h->m.launches : -1; }
int64_t cvt_workspace_bytes(const cvt_model* h) { return h ? h->m.workspace_bytes : -1; }

int cvt_profile_step(cvt_model* h, const float* img_dev, const float* proc_dev, const float* y_dev, int batch, int training,
                     int max_entries, const char** names, float* ms, int* n_entries) {
  if (!h || !img_dev || !names || !ms || !n_entries) { set_error("cvt_profile_step: null argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m;
  int r = check_batch(m, batch, "cvt_profile_step"); if (r) return r;
  if (training && !y_dev) { set_error("cvt_profile_step: training needs targets"); return CVT_ERR_INVALID; }
  // the profiled step must leave the model as it found it: Adam moments, BatchNorm moving statistics and the dropout counter are
  // saved here and restored below (the weights themselves do not move: the Adam launch runs with a zero step size)
  const size_t n_mom = (size_t)m->n_train * sizeof(float), n_st = (size_t)std::max<int64_t>(m->n_state, 4) * sizeof(float);
  float* save = nullptr;
  const uint32_t save_step = m->dropout_step;
  if (training) {
    API_CUDA(cudaMalloc(&save, 2 * n_mom + n_st));
    API_CUDA(cudaMemcpyAsync(save, m->Mo, n_mom, cudaMemcpyDeviceToDevice, m->stream));
    API_CUDA(cudaMemcpyAsync(save + m->n_train, m->Vo, n_mom, cudaMemcpyDeviceToDevice, m->stream));
    API_CUDA(cudaMemcpyAsync(save + 2 * m->n_train, m->St, n_st, cudaMemcpyDeviceToDevice, m->stream));
  }
  std::vector<ProfMark> marks;
  g_prof = &marks;
  int rc = CVT_OK;
  do {
    prof_mark(m->stream, "prep");          // weight-only kernels ahead of stage 1: K-major / TF32 weight copies, the folded q/k/v products
    if (forward_impl(m, img_dev, proc_dev, batch, training) != cudaSuccess) { rc = CVT_ERR_CUDA; break; }
    if (training) {
      if (launch_mse_loss(m->pred, y_dev, batch * m->cfg.num_classes, m->metrics, m->dpred, 1.f, 0, m->stream) != cudaSuccess) { rc = CVT_ERR_CUDA; break; }
      if (backward_impl(m, img_dev, proc_dev, batch, 1, -1, nullptr) != cudaSuccess) { rc = CVT_ERR_CUDA; break; }
      prof_mark(m->stream, "adam");
      if (adam_impl(m, 0.f, 1.f) != cudaSuccess) { rc = CVT_ERR_CUDA; break; }
      m->adam_t -= 1;   // lr = 0: weights untouched; keep the step counter where it was
      prof_mark(m->stream, "end");
    }
  } while (0);
  g_prof = nullptr;
  if (save) {
    cudaMemcpyAsync(m->Mo, save, n_mom, cudaMemcpyDeviceToDevice, m->stream);
    cudaMemcpyAsync(m->Vo, save + m->n_train, n_mom, cudaMemcpyDeviceToDevice, m->stream);
    cudaMemcpyAsync(m->St, save + 2 * m->n_train, n_st, cudaMemcpyDeviceToDevice, m->stream);
    m->dropout_step = save_step;
  }
  cudaStreamSynchronize(m->stream);
  if (save) cudaFree(save);
  int n = 0;
  m->prof_groups.clear();
  for (size_t i = 0; i + 1 < marks.size() && n < max_entries; ++i) {
    float t = 0.f; cudaEventElapsedTime(&t, marks[i].ev, marks[i + 1].ev);
    const char* nm = marks[i].name;
    if (!strcmp(nm, "end_forward") && training) nm = "loss";      // the separate loss kernel of the profiled step (the fused head emits it in cvt_train_step)
    else if (!strncmp(nm, "end", 3)) continue;
    names[n] = nm; ms[n] = t; ++n;
    m->prof_groups.push_back({nm, (int)(marks[i + 1].launches - marks[i].launches)});
  }
  for (auto& mk : marks) cudaEventDestroy(mk.ev);
  *n_entries = n;
  m->last_img = img_dev; m->last_proc = proc_dev;
  return rc;
}

// kernels each group of the last cvt_profile_step launched, in issue order (joins an ncu launch list to the groups: tools/ncu_traffic.py)
int cvt_profile_group_launches(cvt_model* h, int max_entries, const char** names, int* launches, int* n_entries) {
  if (!h || !names || !launches || !n_entries) { set_error("cvt_profile_group_launches: null argument"); return CVT_ERR_INVALID; }
  Model* m = &h->m;
  int n = 0;
  for (const auto& g : m->prof_groups) { if (n >= max_entries) break; names[n] = g.first; launches[n] = g.second; ++n; }
  *n_entries = n;
  return CVT_OK;
}

}  // extern "C"
