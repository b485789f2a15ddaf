// Launchers for the memory-bound layers of the CvT block (everything that is not a contraction).
#pragma once
#include "common.cuh"

namespace cvt {

// LayerNormalization(epsilon) over the last axis, biased variance — models/CvT(Par).py:248,272,278,328.
// y[row*ldy + c] = (x[in_map(row)*ldx + c] - mu) * rstd * gamma[c] + beta[c]; mean/rstd saved per row.
cudaError_t launch_ln_fwd(const float* x, RowMap in_map, long long ldx, const float* gamma, const float* beta, float eps,
                          float* y, long long ldy, float* mean, float* rstd, int rows, int D, cudaStream_t s, int round_out = 0);
// dx[out_map(row)*ldo + c] = (dres ? dres[out_map(row)*ldo + c] : 0) + LNbwd(dy)[c];  dgamma/dbeta += (atomics)
cudaError_t launch_ln_bwd(const float* dy, long long lddy, const float* x, RowMap in_map, long long ldx,
                          const float* mean, const float* rstd, const float* gamma,
                          const float* dres, float* dx, RowMap out_map, long long ldo,
                          float* dgamma, float* dbeta, int rows, int D, cudaStream_t s,
                          const uint8_t* mask = nullptr, float mscale = 0.f, float* dxm = nullptr, int round_m = 0);
// mask != nullptr (identity out_map, ldo == D): also dxm = dx * (mask ? mscale : 0), the gradient through the dropout that follows in
// backward order, in the same pass

// vectorised instantiations (ln.cu); cudaErrorNotSupported -> caller uses the generic kernel
cudaError_t launch_ln_fwd_v4(const float* x, RowMap in_map, long long ldx, const float* gamma, const float* beta, float eps,
                             float* y, long long ldy, float* mean, float* rstd, int rows, int D, cudaStream_t s, int round_out = 0);
cudaError_t launch_ln_bwd_v4(const float* dy, long long lddy, const float* x, RowMap in_map, long long ldx, const float* mean,
                             const float* rstd, const float* gamma, const float* dres, float* dx, RowMap out_map, long long ldo,
                             float* dgamma, float* dbeta, int rows, int D, cudaStream_t s,
                             const uint8_t* mask = nullptr, float mscale = 0.f, float* dxm = nullptr, int round_m = 0);

// Projection.call for q, k, v in one pass over the normalised map — models/CvT(Par).py:102-112,157-170.
//   method[z]: 0 depthwise 3x3 (SAME, no bias), 1 3x3 average pool (SAME, valid-count divisor), 2 identity
//   xn [B,Tc,D] (cls row first when has_cls); out [3][B,Tc,D]; cls rows are copied through.
//   stats (nullable) [6][D]: per-channel sum (q,k,v) then sum of squares (q,k,v) over the B*H*W non-cls tokens.
struct DwProjP {
  const float* xn; float* out; long long out_bs; const float* w; int method[3];
  int B, H, W, D, Tc, has_cls; float* stats;
  int round_out = 0;   // store the maps rounded to TF32 (they only feed tensor-core GEMMs; BN statistics use the unrounded values)
  int cls_zero = 0;    // write zeros instead of the cls row (folded mode: the cls row never enters the tensor-core products)
  int nparts = 1;      // dwproj.cu: stats is [nparts][6][D] per-CTA partials (bn_finalize sums them); generic kernel: 1 (atomics)
};
cudaError_t launch_dwproj_fwd(const DwProjP& p, cudaStream_t s);

// BatchNormalization bookkeeping for the three branches: batch (training) or moving (inference) statistics ->
// per-channel scale/shift applied when the q/k/v GEMM loads its A operand; moving-stat update (momentum, Bessel).
struct BnFinalizeP {
  const float* stats; int nparts = 1; long long count; const float* gamma; const float* beta; float* moving_mean; float* moving_var;
  int training; int update_moving; float eps, momentum; float* scale; float* shift; float* save_mean; float* save_rstd; int D;
};
cudaError_t launch_bn_finalize(const BnFinalizeP& p, cudaStream_t s);

// BN backward, pass 1: dgamma[z][c] += sum dy*xhat, dbeta[z][c] += sum dy over the non-cls tokens.
cudaError_t launch_bn_bwd_reduce(const float* dy, const float* conv, long long bs, const float* save_mean, const float* save_rstd,
                                 float* dgamma, float* dbeta, int B, int T, int Tc, int has_cls, int D, cudaStream_t s);
// BN backward, pass 2 (in place on dy): dconv = gamma*rstd*(dy - dbeta/N - xhat*dgamma/N)  (training)
//                                       dconv = dy*scale                                   (inference statistics)
cudaError_t launch_bn_bwd_apply(float* dy, const float* conv, long long bs, const float* save_mean, const float* save_rstd,
                                const float* gamma, const float* scale, const float* dgamma, const float* dbeta, int training,
                                int B, int T, int Tc, int has_cls, int D, cudaStream_t s);
// transposed Projection: dxn = sum_z proj_z^T(dconv_z) (+ cls rows summed), ddw[z][9][D] += (atomics)
struct DwProjBwdP {
  const float* dconv; long long bs; const float* xn; const float* w; int method[3];
  int B, H, W, D, Tc, has_cls; float* dxn; float* ddw;
  float* ddw_part = nullptr;                                  // dwproj.cu: scratch [148][27][D] for the per-CTA weight-gradient partials
  const float* conv = nullptr; const float* coef = nullptr;   // register-window kernel only: BN backward applied on load (dwproj.cu)
};
cudaError_t launch_dwproj_bwd(const DwProjBwdP& p, cudaStream_t s);
// register sliding-window versions (dwproj.cu); cudaErrorNotSupported -> caller uses the generic kernels above
cudaError_t launch_dwproj_fwd2(const DwProjP& p, cudaStream_t s);
cudaError_t launch_dwproj_bwd2(const DwProjBwdP& p, cudaStream_t s);
// register column walkers, one channel per lane (dwcol.cu): D % 32 == 0; cudaErrorNotSupported -> caller uses the kernels above
cudaError_t launch_dwcol_fwd(const DwProjP& p, cudaStream_t s, int* stat_slices);
cudaError_t launch_dwcol_bwd(const DwProjBwdP& p, cudaStream_t s);
// coef [3][3][D] = (A, E, F) with dconv = A dn + E conv + F  (BatchNorm backward as an affine map, training or moving statistics)
cudaError_t launch_bn_bwd_coef(const float* gamma, const float* mean, const float* rstd, const float* scale, const float* dgamma,
                               const float* dbeta, long long count, int training, int D, float* coef, cudaStream_t s);

// folded q/k/v path (fold.cu)
cudaError_t launch_bn_fold(const float* Wc, const float* bc, const float* sc, const float* sh, float alpha0, int D, float* Wft, float* bf,
                           cudaStream_t s, int round = 1);
cudaError_t launch_cls_fix(const float* xn, long long bs, const float* Wc, const float* bc, float alpha0, int B, int Tc, int D, float* q,
                           cudaStream_t s);
cudaError_t launch_qkv_unfold(float* Z, const float* s_all, const float* xn, const float* dq, long long bs, int B, int T, int Tc, int has_cls,
                              int D, const float* Wc, const float* sc, const float* sh, const float* gamma, const float* mu, const float* rstd,
                              int training, float* dgamma, float* dbeta, float* coef, float* scratch /* [3][D][D] + [3][D] */, cudaStream_t s);

// softmax over the last axis, in place (fp32 attention path) and its backward dS = P*(dP - sum(dP*P)), in place on dP
cudaError_t launch_softmax_rows(float* S, long long rows, int cols, cudaStream_t s);
cudaError_t launch_softmax_bwd_rows(float* dP, const float* P, long long rows, int cols, cudaStream_t s);

// dropout keep-masks from Philox4x32-10 (same stream as oracle.dropout_mask)
cudaError_t launch_philox_mask(uint8_t* mask, long long n, float rate, uint64_t seed, uint32_t step, uint32_t site, cudaStream_t s,
                               const float* dyn = nullptr);
cudaError_t launch_set_step_params(float* dyn, float alpha, float gs, uint32_t step, cudaStream_t s);
// dst = src * mask * scale
cudaError_t launch_mask_scale(float* dst, const float* src, const uint8_t* mask, float scale, long long n, cudaStream_t s, int round = 0);
// dst[map(r)*ld + c] = src[c] for r < rows (cls token broadcast)
cudaError_t launch_fill_rows(float* dst, RowMap map, long long ld, const float* src, int rows, int D, cudaStream_t s);
// out[b*ldo + c] = mean_t x[(b*T+t)*D + c]  and its transpose dx[(b*T+t)*D + c] = dy[b*ldy + c] / T
cudaError_t launch_token_mean(const float* x, int B, int T, int D, float* out, long long ldo, cudaStream_t s);
cudaError_t launch_token_mean_bwd(const float* dy, long long lddy, int B, int T, int D, float* dx, cudaStream_t s);
// copy rows: dst[dmap(r)*ldd + c] = src[smap(r)*lds + c]
cudaError_t launch_copy_rows(float* dst, RowMap dmap, long long ldd, const float* src, RowMap smap, long long lds, int rows, int D, cudaStream_t s);

// input pipeline (dataset.cu): rows moved by index between a resident dataset and the staging buffers of a step (idx nullable = identity);
// float64 host rows converted after their copy; OpenCV's 8-bit bilinear resize + BGR2GRAY + /255 (models/CvT(Par).py:420-423)
cudaError_t launch_move_rows(float* dst_img, const int32_t* dst_idx, const float* src_img, const int32_t* src_idx, int rows, long long row_floats,
                             float* dst_p, const float* src_p, int P, float* dst_y, const float* src_y, int nc, cudaStream_t s);
cudaError_t launch_f64_to_f32(float* dst, const double* src, long long n, cudaStream_t s);
void resize_axis_table(int src, int dst, int clamp_weights, int32_t* s_out, int32_t* w0, int32_t* w1);      // host: cv::resize INTER_LINEAR tables of one axis
cudaError_t launch_preprocess_u8(const uint8_t* pix_dev, int sh, int sw, int sc, const int32_t* tab_dev, const float* lut_dev, int dh, int dw,
                                 long long count, float* out, cudaStream_t s);

// MSE + MAE + dL/dpred  (model.compile(loss='mean_squared_error', metrics=['mae']), models/CvT(Par).py:464-466)
// metrics[0] = mean((p-y)^2), metrics[1] = mean(|p-y|); dpred = 2 (p-y) / (B*nc) * grad_scale (nullable)
cudaError_t launch_mse_loss(const float* pred, const float* y, int n, float* metrics, float* dpred, float grad_scale, int accumulate, cudaStream_t s);

// Keras Adam on a flat arena: m,v update; theta -= alpha * m / (sqrt(v) + eps); alpha = lr*sqrt(1-b2^t)/(1-b1^t)
cudaError_t launch_adam(float* params, const float* grads, float* m, float* v, long long n, float alpha, float b1, float b2,
                        float eps, float grad_scale, cudaStream_t s, const float* dyn = nullptr);

// fused regression head + loss (head.cu): models/CvT(Par).py:325-350,464-466
struct HeadP {
  int B, D, F, PH, proc_dim, nc, cls_ln;
  float eps;
  const float* x2; long long x2_bs;          // last block's output [B, Tc, D]; the cls token is row 0 of every image (cls_ln head)
  const float* hn_g; const float* hn_b; float* h_mean; float* h_rstd;
  float* feat;                               // [B, F]: image feature (written here when cls_ln, else read) ++ process feature
  const float* proc; const float* w1; const float* b1; const float* w2; const float* b2; float* p1;
  const float* fw; const float* fb; float* pred;
  const float* y; float* dpred; float* metrics; float* partial; unsigned* ticket;     // y == nullptr: no loss
  // backward
  float* dfeat; float* dz2; float* dz1; float* dx2;
  float* g_w1; float* g_b1; float* g_w2; float* g_b2; float* g_fw; float* g_fb; float* g_hn_g; float* g_hn_b;
};
cudaError_t launch_head_fwd(const HeadP& p, cudaStream_t s);
cudaError_t launch_head_bwd(const HeadP& p, cudaStream_t s);

}  // namespace cvt
