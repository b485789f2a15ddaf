// Small fp32 products on the weights themselves, many per launch ("multi-GEMM"), used to fold / unfold the stacked linear maps
// of ConvAttention (models/CvT(Par).py:180-188: Dense(dim) followed by Keras MultiHeadAttention's own EinsumDense):
//     forward   Wc = Wp . Wm,  bc = bp . Wm + bm            (one D x D GEMM per token instead of two)
//     backward  dWm = Wp^T . N + bp (x) s,  dWp = N . Wm^T,  dbp = s . Wm^T        (N = n^T dq, s = colsum dq)
// These are D x D x D (D <= 512) with fp32 FFMA accumulation: exact enough that folding does not cost parity, and tiny next to
// the token GEMMs.  One CTA computes one 64 x 64 tile of one product; the descriptor list lives in device memory.
#include "gemm.cuh"

namespace cvt {
namespace {

// 64 x 64 tile, K chunks of 32, 256 threads (4 x 4 outputs each).  Operand chunks are fetched as float4s along whichever index is
// contiguous in memory and prefetched into registers while the previous chunk is multiplied out of shared memory.
// Shapes: K % 8 == 0; a dimension read along its contiguous axis must be a multiple of 4 (all sizes here are 1 or multiples of 64).
__global__ void __launch_bounds__(256) multigemm_kernel(const MgDesc* __restrict__ descs, int ndesc, int tile_base) {
  pdl_wait(); pdl_trigger();
  __shared__ __align__(16) float sA[32][68], sB[32][68];
  const int bid = (int)blockIdx.x + tile_base;          // tile0 of the descriptors is absolute; a launch may cover a sub-range of the table
  const MgDesc e = descs[find_desc(descs, ndesc, bid)];
  const int tile = bid - e.tile0, tn = cdiv(e.N, 64);
  const int m0 = (tile / tn) * 64, n0 = (tile % tn) * 64;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;      // thread computes rows ty*4.., cols tx*4..
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  // fetch coordinates: "row-contiguous-in-k" operands (A plain, B transposed): row = tid / 4, k segment = (tid % 4) * 8
  //                    "k-major" operands (A transposed, B plain):            k = tid / 8,  mn segment = (tid % 8) * 8
  float4 ra[2], rb[2];
  auto fetch = [&](int k0) {
    if (!e.ta) {
      const int m = m0 + (tid >> 2), k = k0 + ((tid & 3) << 3);
      const bool ok = m < e.M && k < e.K;
      const float4* src = reinterpret_cast<const float4*>(e.A + (long long)(ok ? m : 0) * e.lda + (ok ? k : 0));
      ra[0] = ok ? src[0] : make_float4(0.f, 0.f, 0.f, 0.f); ra[1] = ok ? src[1] : make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      const int k = k0 + (tid >> 3), m = m0 + ((tid & 7) << 3);
      const bool ok = k < e.K && m < e.M;
      const float4* src = reinterpret_cast<const float4*>(e.A + (long long)(ok ? k : 0) * e.lda + (ok ? m : 0));
      ra[0] = ok ? src[0] : make_float4(0.f, 0.f, 0.f, 0.f); ra[1] = (ok && m + 4 < e.M) ? src[1] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (e.tb) {
      const int n = n0 + (tid >> 2), k = k0 + ((tid & 3) << 3);
      const bool ok = n < e.N && k < e.K;
      const float4* src = reinterpret_cast<const float4*>(e.B + (long long)(ok ? n : 0) * e.ldb + (ok ? k : 0));
      rb[0] = ok ? src[0] : make_float4(0.f, 0.f, 0.f, 0.f); rb[1] = ok ? src[1] : make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      const int k = k0 + (tid >> 3), n = n0 + ((tid & 7) << 3);
      const bool ok = k < e.K && n < e.N;
      const float4* src = reinterpret_cast<const float4*>(e.B + (long long)(ok ? k : 0) * e.ldb + (ok ? n : 0));
      rb[0] = ok ? src[0] : make_float4(0.f, 0.f, 0.f, 0.f); rb[1] = (ok && n + 4 < e.N) ? src[1] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  auto stash = [&]() {
    if (!e.ta) {
      const int m = tid >> 2, k = (tid & 3) << 3;
      const float v[8] = {ra[0].x, ra[0].y, ra[0].z, ra[0].w, ra[1].x, ra[1].y, ra[1].z, ra[1].w};
#pragma unroll
      for (int i = 0; i < 8; ++i) sA[k + i][m] = v[i];
    } else {
      const int k = tid >> 3, m = (tid & 7) << 3;
      *reinterpret_cast<float4*>(&sA[k][m]) = ra[0]; *reinterpret_cast<float4*>(&sA[k][m + 4]) = ra[1];
    }
    if (e.tb) {
      const int n = tid >> 2, k = (tid & 3) << 3;
      const float v[8] = {rb[0].x, rb[0].y, rb[0].z, rb[0].w, rb[1].x, rb[1].y, rb[1].z, rb[1].w};
#pragma unroll
      for (int i = 0; i < 8; ++i) sB[k + i][n] = v[i];
    } else {
      const int k = tid >> 3, n = (tid & 7) << 3;
      *reinterpret_cast<float4*>(&sB[k][n]) = rb[0]; *reinterpret_cast<float4*>(&sB[k][n + 4]) = rb[1];
    }
  };
  fetch(0);
  for (int k0 = 0; k0 < e.K; k0 += 32) {
    stash();
    __syncthreads();
    if (k0 + 32 < e.K) fetch(k0 + 32);
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      const float4 a4 = *reinterpret_cast<const float4*>(&sA[k][ty * 4]);
      const float4 b4 = *reinterpret_cast<const float4*>(&sB[k][tx * 4]);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w}, b[4] = {b4.x, b4.y, b4.z, b4.w};
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

cudaError_t launch_multigemm(const MgDesc* descs_dev, int ndesc, int total_tiles, cudaStream_t s, int tile_base) {
  if (ndesc < 1 || total_tiles < 1) return cudaSuccess;
  CVT_LAUNCH((multigemm_kernel), total_tiles, 256, 0, s, descs_dev, ndesc, tile_base);
  return cudaSuccess;
}

}  // namespace cvt
