// Generic dense-contraction descriptor shared by the fp32 FFMA kernel (gemm_simt.cu) and the
// tcgen05/TMEM kernel (gemm_tc.cu).  One description covers every dense op of the CvT block:
//   forward  Dense / EinsumDense / Conv2D-as-implicit-GEMM   (models/CvT(Par).py:180-188,203-208,253-258)
//   backward dgrad (B read k-contiguous), wgrad (A read transposed, split-K + atomics), conv scatter.
#pragma once
#include "common.cuh"

namespace cvt {

enum AMode : int { A_PLAIN = 0,   // A(m,k) = A[tok(m)*lda + k]
                   A_TRANS = 1,   // A(m,k) = A[tok(k)*lda + m]            (wgrad: X^T)
                   A_CONV  = 2,   // A(m,k) = im2col(x)[m=(b,i,j), k=(kh,kw,ci)]   (Conv2D SAME forward)
                   A_CONVT = 3 }; // A(m,k) = im2col(x)[k=(b,i,j), m=(kh,kw,ci)]   (Conv2D wgrad)
enum BMode : int { B_KN = 0,      // B(k,n) = B[row(k)*ldb + n]
                   B_NK = 1 };    // B(k,n) = B[n*ldb + k]
enum Epi : uint32_t {
  EPI_BIAS = 1u, EPI_GELU = 2u, EPI_RESID = 4u, EPI_MASK = 8u, EPI_ATOMIC = 16u, EPI_SCATTER = 32u,
  EPI_GELU_BWD = 64u, EPI_RELU = 128u, EPI_RELU_BWD = 256u, EPI_ALPHA0 = 512u,
  EPI_ROUND = 2048u,      // store C rounded to TF32 (C only feeds tensor-core GEMMs); aux keeps full precision
  EPI_PHILOX = 4096u,     // with EPI_MASK: the keep-mask is drawn here (Philox4x32-10, same counters as philox_mask_kernel) and stored to mask_gen
  EPI_GELU_DSAVE = 8192u, // tcgen05/TMA kernel only.  With EPI_GELU: aux receives d out / d pre = gelu'(pre) (times the dropout keep factor when
                          // EPI_MASK) instead of pre; with EPI_GELU_BWD: aux holds that derivative, v *= aux (no erf, no mask load)
  EPI_KEEP_ROUND = 16384u, // orchestration only (never reaches a kernel): EPI_ROUND stays on in CVT_PREC_TF32X3 too, the result feeds the attention kernels
  EPI_BN_BWD = 1024u      // v = A[n] v + E[n] aux[idx] + F[n]  (BatchNorm backward as an affine map of (dn, conv); coef = [A|E|F][N])
};

// Conv2D(padding='same') geometry.  The input map of image b starts at x + b*bstride + boff (floats), which lets a
// [B,Tc,D] block output with a leading cls row be read as the next stage's [B,H,W,C] input without a copy.
struct ConvGeom { int hin, win, cin, ho, wo, k, s, pt, pl; long long bstride; int boff; int noround; };      // noround: im2col keeps full fp32 (CVT_PREC_TF32X3)

struct GemmP {
  int M = 0, N = 0, K = 0;
  int batch = 1, splitk = 1, bdiv = 1;          // z -> (z / bdiv, z % bdiv) selects the operand slices
  // A
  const float* A = nullptr; long long lda = 0, a_bs0 = 0, a_bs1 = 0;
  int a_mode = A_PLAIN; RowMap a_map{0, 0, 0};
  const float* a_scale = nullptr; const float* a_shift = nullptr; long long a_aff_bs = 0;   // per-feature affine (BN on load)
  int a_skip_period = 0;                       // tokens with tok % period == 0 (cls rows) skip the affine
  ConvGeom cv{0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  // B
  const float* B = nullptr; long long ldb = 0, b_bs0 = 0, b_bs1 = 0;
  int b_mode = B_KN; RowMap b_map{0, 0, 0};
  // C
  float* C = nullptr; long long ldc = 0, c_bs0 = 0, c_bs1 = 0; RowMap c_map{0, 0, 0};
  // epilogue:  v = (acc + bias[n]) * alpha ; gelu/relu ; *gelu'(aux) ; *mask*mask_scale ; + resid ; store|atomic|scatter
  const float* bias = nullptr; long long bias_bs = 0;
  float alpha = 1.f, alpha0 = 1.f;
  uint32_t epi = 0;
  const float* resid = nullptr; long long ldr = 0;
  const uint8_t* mask = nullptr; float mask_scale = 1.f;
  uint8_t* mask_gen = nullptr; uint32_t ph_k0 = 0, ph_k1 = 0, ph_site = 0, ph_step = 0, ph_thr = 0; const float* ph_dyn = nullptr;   // EPI_PHILOX
  const float* epi_coef = nullptr; long long epi_coef_bs = 0; int epi_skip_period = 0;   // EPI_BN_BWD (rows m % period == 0 are cls rows: identity)
  long long* trace = nullptr;                  // diagnostics only (CVT_GEMM_TRACE): clock64 timeline of CTA 0 of gemm_tma_kernel
  int dbg = 0;                                 // diagnostics only (cvt_debug_gemm): bit 0 swaps LBO/SBO of MN-major descriptors
  int x3 = 0;                                  // TMA kernel only: error-compensated product (hi.hi + lo.hi + hi.lo, operands split in shared memory)
  float* aux = nullptr;                        // same indexing as C: GELU pre-activation (write) / saved activation (read)
};

// fp32 FFMA path (always available, exact fp32 arithmetic)
cudaError_t launch_gemm_simt(const GemmP& p, cudaStream_t s);
// tcgen05 kind::tf32 path; returns cudaErrorNotSupported for shapes it does not take
cudaError_t launch_gemm_tc(const GemmP& p, cudaStream_t s);
bool gemm_tc_supports(const GemmP& p);

// persistent TMA-fed tcgen05 kernel (gemm_tma.cu): A_PLAIN x B_NK (both K-major), no split-K
cudaError_t launch_gemm_tma(const GemmP& p, cudaStream_t s);
bool gemm_tma_supports(const GemmP& p);
bool gemm_tma_conv_supports(const GemmP& p);      // a_mode == A_CONV: ConvEmbed as an implicit GEMM (4-D strided TMA boxes, no im2col)
// TMA-fed split-K weight-gradient kernel (gemm_wgrad.cu):  dW[z][Kin,N] += X[z][tokens,Kin]^T . dY[z][tokens,N],
// db[z][N] += column sums of dY (nullable).  Both operands are read MN-major straight from their row-major activations.
struct WgradP {
  const float* X = nullptr; long long ldx = 0, x_bs = 0; int Kin = 0;
  const float* dY = nullptr; long long lddy = 0, dy_bs = 0; int N = 0;
  int tokens = 0, batch = 1;
  float* dW = nullptr; long long dw_bs = 0;
  float* db = nullptr; long long db_bs = 0;
  int dbg = 0;
  int x3 = 0;                                  // error-compensated product: both operand tiles are split into high and low parts in shared memory
  // conv != 0: X is not a matrix but the im2col view of a Conv2D input (ConvEmbed weight gradient): X(token, feature) =
  // x_pad[b, s*i + kh, s*j + kw, c] with token = (b, i, j), feature = (kh, kw, c); boxes are strided TMA boxes of the NHWC map itself
  int conv = 0; ConvGeom cv{0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
};
cudaError_t launch_wgrad_tma(const WgradP& p, cudaStream_t s);
bool wgrad_tma_supports(const WgradP& p);
// many small fp32 products per launch (multigemm.cu): C = alpha op(A) op(B) [+ u (x) v] [+ 1 (x) v] [+ C]
struct MgDesc {
  const float* A; const float* B; float* C; int M, N, K; int lda, ldb, ldc; int ta, tb;   // ta: A(m,k) = A[k*lda + m]; tb: B(k,n) = B[n*ldb + k]
  float alpha; int accumulate; const float* u; const float* v; int u1; int tile0;
  int round;      // store round-to-nearest TF32 (the result is a tensor-core operand)
};
int mg_tiles(const MgDesc& d);
// tile_base: tile0 of descs_dev[0] when the launch covers a sub-range of a descriptor table (per-stage unfold)
cudaError_t launch_multigemm(const MgDesc* descs_dev, int ndesc, int total_tiles, cudaStream_t s, int tile_base = 0);
// dst[off + n*K + k] = src[off + k*N + n] for every registered dense kernel (forward GEMMs read their weights K-major)
struct TrDesc { long long off; int K, N, tile0; int copy; long long toff; int ldt; };   // transposed copy at dst[toff + n*ldt + k]   // copy: dst2[off + k*N + n] = round(src) (a TF32-rounded straight copy, dgrad operand)
cudaError_t launch_transpose_weights(const float* src, float* dst, float* dst2, const TrDesc* descs_dev, int ndesc, int total_tiles, cudaStream_t s, int round = 1);

// explicit im2col / col2im for ConvEmbed in TF32 mode (im2col.cu)
cudaError_t launch_im2col(const float* x, const ConvGeom& g, int M, int K, int ldc, float* col, cudaStream_t s);
cudaError_t launch_col2im(const float* dcol, const ConvGeom& g, int B, int ldc, float* dx, cudaStream_t s, const uint8_t* mask = nullptr,
                          float mscale = 0.f, float* dxm = nullptr, int round_m = 0);

// column sums (bias gradients): out[z][n] += sum_r X[z][map(r)*ld + n]
cudaError_t launch_colsum(const float* X, long long ld, RowMap map, int rows, int N, int batch, long long x_bs,
                          float* out, long long out_bs, cudaStream_t s);

}  // namespace cvt
