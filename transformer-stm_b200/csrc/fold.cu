// Glue of the folded q/k/v path (TF32 mode): the two stacked linear maps of ConvAttention (proj_q/k/v Dense, then Keras
// MultiHeadAttention's query/key/value EinsumDense; models/CvT(Par).py:180-185) and the BatchNorm affine in front of
// them (Projection, :94) are applied as ONE D x D GEMM per token:
//     q_z = alpha_z ( BN(c_z) Wp_z Wm_z + bp_z Wm_z + bm_z ) = c_z Wf_z + bf_z,    Wf_z = alpha_z diag(sc_z) Wc_z,  Wc_z = Wp_z Wm_z
// The cls row bypasses BatchNorm (models/CvT(Par).py:150,168-170), so it is recomputed with the unscaled Wc_z by cls_fix_kernel.
// Backward recovers every individual gradient (Wp, bp, Wm, bm, BN gamma/beta) from Z_z = c_z^T dq_z and s_z = colsum(dq_z):
// see qkv_unfold_kernel and the multi-GEMM list built in cvt_api.cu.
#include "layers.cuh"

namespace cvt {
namespace {

// Wft[z][j][i] = alpha_z sc_z[i] Wc[z][i][j];   bf[z][j] = alpha_z (sum_i sh_z[i] Wc[z][i][j] + bc[z][j])
// grid (D/32 j-tiles, D/32 i-tiles, 3), block (32, 8)
__global__ void __launch_bounds__(256) bn_fold_kernel(const float* __restrict__ Wc, const float* __restrict__ bc, const float* __restrict__ sc,
                                                      const float* __restrict__ sh, float alpha0, int D, float* __restrict__ Wft,
                                                      float* __restrict__ bf, int round) {
  pdl_wait(); pdl_trigger();
  __shared__ float t[32][33];
  __shared__ float red[8][33];
  const int z = blockIdx.z, j0 = blockIdx.x * 32, i0 = blockIdx.y * 32, tx = threadIdx.x, ty = threadIdx.y;
  const float alpha = z == 0 ? alpha0 : 1.f;
  const float* W = Wc + (long long)z * D * D;
  for (int r = ty; r < 32; r += 8) {
    const int i = i0 + r;
    const float s = sc ? sc[z * D + i] : 1.f;
    const float wv = W[(long long)i * D + j0 + tx] * s * alpha;      // B operand of the folded q/k/v GEMM
    t[r][tx] = round ? round_tf32(wv) : wv;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) Wft[(long long)z * D * D + (long long)(j0 + r) * D + i0 + tx] = t[tx][r];
  if (blockIdx.y == 0) {
    float a = 0.f;
    if (sh) {
      for (int ib = ty; ib < D; ib += 8 * 16) {            // 16 independent weight loads in flight (the loop is pure latency otherwise)
        float wv[16], sv[16];
#pragma unroll
        for (int u = 0; u < 16; ++u) {
          const int i = ib + 8 * u; const bool ok = i < D;
          wv[u] = ok ? W[(long long)i * D + j0 + tx] : 0.f; sv[u] = ok ? sh[z * D + i] : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 16; ++u) a = fmaf(sv[u], wv[u], a);
      }
    }
    red[ty][tx] = a;
    __syncthreads();
    if (ty == 0) {
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) s += red[k][tx];
      bf[z * D + j0 + tx] = alpha * (s + bc[z * D + j0 + tx]);
    }
  }
}

// q[z][b, cls row, :] = alpha_z (xn[b, cls row, :] Wc[z] + bc[z]): a [B, D] x [D, D] product per branch, far too small for the tensor
// pipe and pure latency if done as one matvec per image.  grid (D/32 column tiles, B/16 image chunks, 3), block (32, 8): thread
// (tx, ty) owns output column j0 + tx and the ty-th eighth of the contraction; it loads its weights in batches of 16 independent
// loads, multiplies them into 16 images' cls rows held in shared memory, and the eight partial sums meet in shared memory.
constexpr int CF_B = 16, CF_U = 16;
__global__ void __launch_bounds__(256) cls_fix_kernel(const float* __restrict__ xn, long long bs, const float* __restrict__ Wc,
                                                      const float* __restrict__ bc, float alpha0, int B, int Tc, int D, float* __restrict__ q) {
  pdl_wait(); pdl_trigger();
  extern __shared__ float4 cf_smem[];
  float* sx = reinterpret_cast<float*>(cf_smem);          // [CF_B][D] cls rows; reused as [8][CF_B][32] partial sums
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * 32 + tx;
  const int j = blockIdx.x * 32 + tx, b0 = blockIdx.y * CF_B, z = blockIdx.z;
  const int nb = min(CF_B, B - b0);
  for (int e = tid; e < CF_B * (D >> 2); e += 256) {
    const int b = e / (D >> 2), c = e - b * (D >> 2);
    reinterpret_cast<float4*>(sx)[e] = b < nb ? *reinterpret_cast<const float4*>(xn + (long long)(b0 + b) * Tc * D + c * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const int per = D >> 3, i0 = ty * per;                  // D % 32 == 0: per is a multiple of 4
  const float* W = Wc + (long long)z * D * D + (long long)i0 * D + j;
  float acc[CF_B];
#pragma unroll
  for (int b = 0; b < CF_B; ++b) acc[b] = 0.f;
  __syncthreads();
  for (int u0 = 0; u0 < per; u0 += CF_U) {
    float wv[CF_U];
#pragma unroll
    for (int u = 0; u < CF_U; ++u) wv[u] = u0 + u < per ? __ldg(W + (long long)(u0 + u) * D) : 0.f;
#pragma unroll
    for (int b = 0; b < CF_B; ++b) {
      const float4* xr = reinterpret_cast<const float4*>(sx + b * D + i0 + u0);
#pragma unroll
      for (int u = 0; u < CF_U; u += 4) {
        if (u0 + u < per) {
          const float4 xv = xr[u >> 2];
          acc[b] = fmaf(xv.x, wv[u], fmaf(xv.y, wv[u + 1], fmaf(xv.z, wv[u + 2], fmaf(xv.w, wv[u + 3], acc[b]))));
        }
      }
    }
  }
  __syncthreads();
#pragma unroll
  for (int b = 0; b < CF_B; ++b) sx[(ty * CF_B + b) * 32 + tx] = acc[b];
  __syncthreads();
  const float alpha = z == 0 ? alpha0 : 1.f, bias = bc[z * D + j];
  for (int b = ty; b < nb; b += 8) {
    float a = bias;
#pragma unroll
    for (int k = 0; k < 8; ++k) a += sx[(k * CF_B + b) * 32 + tx];
    q[(long long)z * bs + (long long)(b0 + b) * Tc * D + j] = round_tf32(alpha * a);     // attention operand (TMA): rounded by its producer
  }
}

// Zcls[z][c][j] = sum_b xn[b, cls, c] dq[z][b, cls, j]  and  scls[z][j] = sum_b dq[z][b, cls, j]   (fp32; the cls row is the only row
// of the block that reaches the q/k/v GEMM without going through conv + BatchNorm).  grid (D/32 j, D/32 c, 3), block (32, 8)
__global__ void __launch_bounds__(256) cls_outer_kernel(const float* __restrict__ xn, const float* __restrict__ dq, long long bs, int B, int Tc,
                                                        int D, float* __restrict__ Zcls, float* __restrict__ scls) {
  pdl_wait(); pdl_trigger();
  __shared__ float sx[32][33], sg[32][33];
  const int z = blockIdx.z, j0 = blockIdx.x * 32, c0 = blockIdx.y * 32, tx = threadIdx.x, ty = threadIdx.y;
  float acc[4] = {0.f, 0.f, 0.f, 0.f}, ssum = 0.f;
  for (int b0 = 0; b0 < B; b0 += 32) {
    for (int r = ty; r < 32; r += 8) {
      const int b = b0 + r;
      const long long row = (long long)b * Tc * D;
      sx[r][tx] = b < B ? xn[row + c0 + tx] : 0.f;
      sg[r][tx] = b < B ? dq[(long long)z * bs + row + j0 + tx] : 0.f;
    }
    __syncthreads();
#pragma unroll 8
    for (int r = 0; r < 32; ++r) {
      const float g = sg[r][tx];
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[q] = fmaf(sx[r][ty + 8 * q], g, acc[q]);
      ssum += g;
    }
    __syncthreads();
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) Zcls[(long long)z * D * D + (long long)(c0 + ty + 8 * q) * D + j0 + tx] = acc[q];
  if (blockIdx.y == 0 && ty == 0) scls[z * D + j0 + tx] = ssum;
}

// After Z[z] = c_z^T dq_z over the spatial rows (the cls rows of the conv maps are kept at zero in this mode, so the tensor-core
// product never contains them and nothing has to be subtracted) and s[z] = colsum(dq_z) over all rows:
//   Z_sp = Z,  Z_cls / s_cls from cls_outer_kernel (nullptr without a cls token),  s_sp = s - s_cls
//   dbeta[c]  = sum_j s_sp[j] Wc[c][j]                      (= sum over spatial rows of dn, dn = dq Wc^T)
//   dgamma[c] = rstd[c] (sum_j Wc[c][j] Z_sp[c][j] - mu[c] dbeta[c])
//   coef = (A, E, F): dconv = A dn + E conv + F   (training: batch statistics; else A = scale, E = F = 0)
//   N[c][j] = sc[c] Z_sp[c][j] + sh[c] s_sp[j] + Z_cls[c][j]      (= BN(c)^T dq incl. the un-normalised cls row), in place on Z
// grid (D/8, 3), block 256: one warp per row c, lanes over the columns (coalesced)
__global__ void __launch_bounds__(256) qkv_unfold_kernel(float* __restrict__ Z, const float* __restrict__ s_all, const float* __restrict__ Zcls,
                                                         const float* __restrict__ scls, int D,
                                                         const float* __restrict__ Wc, const float* __restrict__ sc, const float* __restrict__ sh,
                                                         const float* __restrict__ gamma, const float* __restrict__ mu, const float* __restrict__ rstd,
                                                         float invN, int training, float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                         float* __restrict__ coef) {
  pdl_wait(); pdl_trigger();
  const int z = blockIdx.y, c = blockIdx.x * 8 + (threadIdx.x >> 5), l = threadIdx.x & 31;
  float* Zr = Z + (long long)z * D * D + (long long)c * D;
  const float* Wr = Wc + (long long)z * D * D + (long long)c * D;
  const float* Zc = Zcls ? Zcls + (long long)z * D * D + (long long)c * D : nullptr;
  const float scc = sc[z * D + c], shc = sh[z * D + c];
  float dot = 0.f, db = 0.f;
  for (int j = l; j < D; j += 32) {
    const float zc = Zc ? Zc[j] : 0.f;
    const float zsp = Zr[j], ssp = s_all[z * D + j] - (scls ? scls[z * D + j] : 0.f);
    const float w = Wr[j];
    dot = fmaf(w, zsp, dot); db = fmaf(ssp, w, db);
    Zr[j] = scc * zsp + shc * ssp + zc;
  }
  dot = warp_sum(dot); db = warp_sum(db);
  if (l == 0) {
    const int i = z * D + c;
    const float r = rstd[i], m = mu[i], g = gamma[i];
    const float dg = r * (dot - m * db);
    dgamma[i] += dg; dbeta[i] += db;
    float A, E, F;
    if (training) { A = g * r; E = -g * r * r * dg * invN; F = -g * r * db * invN + g * r * r * m * dg * invN; }
    else { A = scc; E = 0.f; F = 0.f; }
    coef[(z * 3 + 0) * D + c] = A; coef[(z * 3 + 1) * D + c] = E; coef[(z * 3 + 2) * D + c] = F;
  }
}

}  // namespace

cudaError_t launch_bn_fold(const float* Wc, const float* bc, const float* sc, const float* sh, float alpha0, int D, float* Wft, float* bf,
                           cudaStream_t s, int round) {
  if (D % 32 != 0) { set_error("bn_fold: D must be a multiple of 32"); return cudaErrorInvalidValue; }
  CVT_LAUNCH((bn_fold_kernel), dim3(D / 32, D / 32, 3), dim3(32, 8), 0, s, Wc, bc, sc, sh, alpha0, D, Wft, bf, round);
  return cudaSuccess;
}

cudaError_t launch_cls_fix(const float* xn, long long bs, const float* Wc, const float* bc, float alpha0, int B, int Tc, int D, float* q,
                           cudaStream_t s) {
  if (D % 32 != 0 || D > 1024) { set_error("cls_fix: D must be a multiple of 32, at most 1024"); return cudaErrorInvalidValue; }
  static bool configured = false;
  if (!configured) { CVT_CUDA(cudaFuncSetAttribute(cls_fix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CF_B * 1024 * 4)); configured = true; }
  const size_t smem = (size_t)CF_B * (D > 256 ? D : 256) * sizeof(float);      // cls rows, later the [8][CF_B][32] partial sums
  CVT_LAUNCH((cls_fix_kernel), dim3(D / 32, (B + CF_B - 1) / CF_B, 3), dim3(32, 8), smem, s, xn, bs, Wc, bc, alpha0, B, Tc, D, q);
  return cudaSuccess;
}

cudaError_t launch_qkv_unfold(float* Z, const float* s_all, const float* xn, const float* dq, long long bs, int B, int T, int Tc, int has_cls,
                              int D, const float* Wc, const float* sc, const float* sh, const float* gamma, const float* mu, const float* rstd,
                              int training, float* dgamma, float* dbeta, float* coef, float* scratch, cudaStream_t s) {
  if (D % 32 != 0) { set_error("qkv_unfold: D must be a multiple of 32"); return cudaErrorInvalidValue; }
  float* Zcls = nullptr; float* scls = nullptr;
  if (has_cls) {
    Zcls = scratch; scls = scratch + (long long)3 * D * D;
    CVT_LAUNCH((cls_outer_kernel), dim3(D / 32, D / 32, 3), dim3(32, 8), 0, s, xn, dq, bs, B, Tc, D, Zcls, scls);
  }
  CVT_LAUNCH((qkv_unfold_kernel), dim3(D / 8, 3), 256, 0, s, Z, s_all, Zcls, scls, D, Wc, sc, sh, gamma, mu, rstd,
                                                   1.f / (float)((long long)B * T), training, dgamma, dbeta, coef);
  return cudaSuccess;
}

}  // namespace cvt
