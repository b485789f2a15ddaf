// Attention of keras.layers.MultiHeadAttention as called at models/CvT(Par).py:185 (query=q, value=v, key=k):
//   S = Q K^T (Q already scaled by 1/sqrt(key_dim)), P = softmax(S), O = P V, heads interleaved in the feature axis.
// fp32 path: batched contractions that materialise P (kept for the backward pass).  The tensor-core path
// (attention_tc.cu) replaces it with a flash-style kernel when the model runs in TF32/BF16 precision.
#include "model.cuh"

namespace cvt {

cudaError_t attention_forward_tc(Model* m, int si, int B, cudaStream_t s);
cudaError_t attention_backward_tc(Model* m, int si, int B, const float* dO, float* dQKV, cudaStream_t s);
bool attention_tc_enabled(const Model* m, int si);

namespace {
constexpr int HD = 64;

GemmP batched(const Model* m, int si, int B) {
  const StageGeo& g = m->geo[si];
  GemmP p; p.batch = B * g.heads; p.bdiv = g.heads;
  return p;
}
}  // namespace

cudaError_t attention_forward(Model* m, int si, int B, cudaStream_t s) {
  if (attention_tc_enabled(m, si)) return attention_forward_tc(m, si, B, s);
  const StageGeo& g = m->geo[si]; StageBuf& a = m->buf[si];
  const int D = g.D, Tc = g.Tc;
  const long long RD = (long long)B * Tc * D, TT = (long long)Tc * Tc;
  const float* Q = a.qkv2; const float* K = Q + RD; const float* V = K + RD;
  GemmP p = batched(m, si, B);
  p.M = Tc; p.N = Tc; p.K = HD;
  p.A = Q; p.lda = D; p.a_bs0 = (long long)Tc * D; p.a_bs1 = HD;
  p.B = K; p.ldb = D; p.b_mode = B_NK; p.b_bs0 = (long long)Tc * D; p.b_bs1 = HD;
  p.C = a.scores; p.ldc = Tc; p.c_bs0 = g.heads * TT; p.c_bs1 = TT;
  CVT_CUDA(launch_gemm_simt(p, s));
  CVT_CUDA(launch_softmax_rows(a.scores, (long long)B * g.heads * Tc, Tc, s));
  GemmP q = batched(m, si, B);
  q.M = Tc; q.N = HD; q.K = Tc;
  q.A = a.scores; q.lda = Tc; q.a_bs0 = g.heads * TT; q.a_bs1 = TT;
  q.B = V; q.ldb = D; q.b_mode = B_KN; q.b_bs0 = (long long)Tc * D; q.b_bs1 = HD;
  q.C = a.attn_o; q.ldc = D; q.c_bs0 = (long long)Tc * D; q.c_bs1 = HD;
  CVT_CUDA(launch_gemm_simt(q, s));
  return cudaSuccess;
}

// dQKV = [dQ (w.r.t. the un-scaled query projection), dK, dV], each [B,Tc,D]
cudaError_t attention_backward(Model* m, int si, int B, const float* dO, float* dQKV, cudaStream_t s) {
  if (attention_tc_enabled(m, si)) return attention_backward_tc(m, si, B, dO, dQKV, s);
  const StageGeo& g = m->geo[si]; StageBuf& a = m->buf[si];
  const int D = g.D, Tc = g.Tc;
  const long long RD = (long long)B * Tc * D, TT = (long long)Tc * Tc, TD = (long long)Tc * D;
  const float* Q = a.qkv2; const float* K = Q + RD; const float* V = K + RD;
  float* dQ = dQKV; float* dK = dQ + RD; float* dV = dK + RD;
  float* dS = m->gS;
  {  // dP = dO V^T
    GemmP p = batched(m, si, B); p.M = Tc; p.N = Tc; p.K = HD;
    p.A = dO; p.lda = D; p.a_bs0 = TD; p.a_bs1 = HD;
    p.B = V; p.ldb = D; p.b_mode = B_NK; p.b_bs0 = TD; p.b_bs1 = HD;
    p.C = dS; p.ldc = Tc; p.c_bs0 = g.heads * TT; p.c_bs1 = TT;
    CVT_CUDA(launch_gemm_simt(p, s));
  }
  {  // dV = P^T dO
    GemmP p = batched(m, si, B); p.M = Tc; p.N = HD; p.K = Tc;
    p.A = a.scores; p.lda = Tc; p.a_mode = A_TRANS; p.a_bs0 = g.heads * TT; p.a_bs1 = TT;
    p.B = dO; p.ldb = D; p.b_mode = B_KN; p.b_bs0 = TD; p.b_bs1 = HD;
    p.C = dV; p.ldc = D; p.c_bs0 = TD; p.c_bs1 = HD;
    CVT_CUDA(launch_gemm_simt(p, s));
  }
  CVT_CUDA(launch_softmax_bwd_rows(dS, a.scores, (long long)B * g.heads * Tc, Tc, s));     // dS = P * (dP - rowsum(dP*P))
  {  // dQ_scaled = dS K ; gradient w.r.t. the projection output = 0.125 * dQ_scaled
    GemmP p = batched(m, si, B); p.M = Tc; p.N = HD; p.K = Tc;
    p.A = dS; p.lda = Tc; p.a_bs0 = g.heads * TT; p.a_bs1 = TT;
    p.B = K; p.ldb = D; p.b_mode = B_KN; p.b_bs0 = TD; p.b_bs1 = HD;
    p.C = dQ; p.ldc = D; p.c_bs0 = TD; p.c_bs1 = HD; p.alpha = 0.125f;
    CVT_CUDA(launch_gemm_simt(p, s));
  }
  {  // dK = dS^T Q_scaled
    GemmP p = batched(m, si, B); p.M = Tc; p.N = HD; p.K = Tc;
    p.A = dS; p.lda = Tc; p.a_mode = A_TRANS; p.a_bs0 = g.heads * TT; p.a_bs1 = TT;
    p.B = Q; p.ldb = D; p.b_mode = B_KN; p.b_bs0 = TD; p.b_bs1 = HD;
    p.C = dK; p.ldc = D; p.c_bs0 = TD; p.c_bs1 = HD;
    CVT_CUDA(launch_gemm_simt(p, s));
  }
  return cudaSuccess;
}

}  // namespace cvt
