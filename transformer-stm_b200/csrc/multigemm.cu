// Small fp32 products on the weights themselves, many per launch ("multi-GEMM"), used to fold / unfold the stacked linear maps
// of ConvAttention (models/CvT(Par).py:180-188: Dense(dim) followed by Keras MultiHeadAttention's own EinsumDense):
//     forward   Wc = Wp . Wm,  bc = bp . Wm + bm            (one D x D GEMM per token instead of two)
//     backward  dWm = Wp^T . N + bp (x) s,  dWp = N . Wm^T,  dbp = s . Wm^T        (N = n^T dq, s = colsum dq)
// These are D x D x D (D <= 512) with fp32 FFMA accumulation: exact enough that folding does not cost parity, and tiny next to
// the token GEMMs.  One CTA computes one 64 x 64 tile of one product; the descriptor list lives in device memory.
#include "gemm.cuh"

namespace cvt {
namespace {

__global__ void __launch_bounds__(256) multigemm_kernel(const MgDesc* __restrict__ descs, int ndesc) {
  __shared__ float sA[16][65], sB[16][65];
  int d = 0;
  while (d + 1 < ndesc && (int)blockIdx.x >= descs[d + 1].tile0) ++d;
  const MgDesc e = descs[d];
  const int tile = blockIdx.x - e.tile0, tn = cdiv(e.N, 64);
  const int m0 = (tile / tn) * 64, n0 = (tile % tn) * 64;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;      // thread computes rows ty*4.., cols tx*4..
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < e.K; k0 += 16) {
    // stage A(m, k) and B(k, n) for 64 x 16 and 16 x 64
    for (int i = tid; i < 64 * 16; i += 256) {
      int m, k;
      if (e.ta) { m = i & 63; k = i >> 6; } else { k = i & 15; m = i >> 4; }       // contiguous index follows the memory layout
      const int gm = m0 + m, gk = k0 + k;
      float v = 0.f;
      if (gm < e.M && gk < e.K) v = e.ta ? e.A[(long long)gk * e.lda + gm] : e.A[(long long)gm * e.lda + gk];
      sA[k][m] = v;
    }
    for (int i = tid; i < 64 * 16; i += 256) {
      int n, k;
      if (e.tb) { k = i & 15; n = i >> 4; } else { n = i & 63; k = i >> 6; }
      const int gn = n0 + n, gk = k0 + k;
      float v = 0.f;
      if (gn < e.N && gk < e.K) v = e.tb ? e.B[(long long)gn * e.ldb + gk] : e.B[(long long)gk * e.ldb + gn];
      sB[k][n] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = sA[k][ty * 4 + i]; b[i] = sB[k][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= e.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= e.N) continue;
      float v = acc[i][j] * e.alpha;
      if (e.u) v += e.u[m] * e.v[n];
      if (e.u1) v += e.v[n];                 // u == vector of ones (bias-style term)
      float* c = e.C + (long long)m * e.ldc + n;
      v = e.accumulate ? *c + v : v;
      *c = e.round ? round_tf32(v) : v;
    }
  }
}

}  // namespace

int mg_tiles(const MgDesc& d) { return cdiv(d.M, 64) * cdiv(d.N, 64); }

cudaError_t launch_multigemm(const MgDesc* descs_dev, int ndesc, int total_tiles, cudaStream_t s) {
  if (ndesc < 1 || total_tiles < 1) return cudaSuccess;
  multigemm_kernel<<<total_tiles, 256, 0, s>>>(descs_dev, ndesc);
  CVT_LAUNCH_CHECK();
  return cudaSuccess;
}

}  // namespace cvt
