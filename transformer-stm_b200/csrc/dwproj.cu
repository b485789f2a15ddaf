// Projection (models/CvT(Par).py:83-112,157-170) for q, k, v: depthwise 3x3 / 3x3 average pool / identity, forward and backward,
// as register sliding-window kernels.  A thread owns one float2 channel pair of one image column and walks down a strip of rows
// keeping the 3x3 neighbourhood in registers, so each step loads 3 new values per tensor (the column neighbours come from L1).
// Backward optionally applies the BatchNorm backward transform while it loads d(BN output) (coef != nullptr), which removes the
// separate apply pass:  dconv = A[c] * dn + E[c] * conv + F[c]   (A, E, F from bn_bwd_coef_kernel).
#include "layers.cuh"

namespace cvt {
namespace {

constexpr int TH = 8;   // rows per strip

__device__ __forceinline__ float2 f2_fma(float2 a, float2 b, float2 c) { return make_float2(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)); }
__device__ __forceinline__ float2 f2_zero() { return make_float2(0.f, 0.f); }

__device__ __forceinline__ float inv_count(int i, int j, int H, int W) {
  const int ch = min(i + 1, H - 1) - max(i - 1, 0) + 1, cw = min(j + 1, W - 1) - max(j - 1, 0) + 1;
  return 1.f / (float)(ch * cw);
}

// block = (D/2 channel pairs, CW columns, 3 branches): thread (c2, jl, z) produces branch z of column j for a strip of TH rows
__global__ void __launch_bounds__(384) dwproj_fwd2_kernel(const __grid_constant__ DwProjP p) {
  pdl_wait(); pdl_trigger();
  extern __shared__ __align__(16) float sstat[];          // [6][D] block partials of the BN statistics
  const int D = p.D, H = p.H, W = p.W;
  const int c = threadIdx.x * 2, jl = threadIdx.y, z = threadIdx.z;
  const int tid = (z * blockDim.y + jl) * blockDim.x + threadIdx.x, nthreads = blockDim.x * blockDim.y * 3;
  if (p.stats) {
    for (int i = tid; i < 6 * D; i += nthreads) sstat[i] = 0.f;
    __syncthreads();
  }
  const int method = p.method[z];
  const int ncx = cdiv(W, (int)blockDim.y), ncy = cdiv(H, TH), ntiles = ncx * ncy * p.B;
  float2 w[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    if (method == 0) w[t] = *reinterpret_cast<const float2*>(p.w + (z * 9 + t) * D + c);
    else if (method == 1) w[t] = make_float2(1.f, 1.f);
    else w[t] = (t == 4) ? make_float2(1.f, 1.f) : f2_zero();
  }
  float2 ssum = f2_zero(), ssq = f2_zero();
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int b = tile / (ncx * ncy), rem = tile - b * (ncx * ncy);
    const int j = (rem % ncx) * blockDim.y + jl, r0 = (rem / ncx) * TH;
    const long long base = ((long long)b * p.Tc + p.has_cls) * D + c;
    const float* xn = p.xn + base;
    float* out = p.out + z * p.out_bs + base;
    if (j >= W) continue;
    auto ld = [&](int ii, int jj) -> float2 {
      if (ii < 0 || ii >= H || jj < 0 || jj >= W) return f2_zero();
      return *reinterpret_cast<const float2*>(xn + ((long long)ii * W + jj) * D);
    };
    float2 win[3][3];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int q = 0; q < 3; ++q) win[a][q] = ld(r0 - 1 + a, j - 1 + q);
    const int rend = min(r0 + TH, H);
#pragma unroll 2
    for (int i = r0; i < rend; ++i) {
      float2 nx[3];
#pragma unroll
      for (int q = 0; q < 3; ++q) nx[q] = ld(i + 2, j - 1 + q);        // next step's bottom row, in flight during this step's math
      float2 y = f2_zero();
#pragma unroll
      for (int t = 0; t < 9; ++t) y = f2_fma(win[t / 3][t % 3], w[t], y);
      if (method == 1) { const float ic = inv_count(i, j, H, W); y.x *= ic; y.y *= ic; }
      *reinterpret_cast<float2*>(out + ((long long)i * W + j) * D) = p.round_out ? make_float2(round_tf32(y.x), round_tf32(y.y)) : y;
      ssum.x += y.x; ssum.y += y.y; ssq = f2_fma(y, y, ssq);
#pragma unroll
      for (int q = 0; q < 3; ++q) { win[0][q] = win[1][q]; win[1][q] = win[2][q]; win[2][q] = nx[q]; }
    }
  }
  // cls rows pass through untouched (models/CvT(Par).py:150,168-170)
  if (p.has_cls && jl == 0) {
    for (int b = blockIdx.x; b < p.B; b += gridDim.x) {
      const long long row = (long long)b * p.Tc * D + c;
      *reinterpret_cast<float2*>(p.out + z * p.out_bs + row) = p.cls_zero ? f2_zero() : *reinterpret_cast<const float2*>(p.xn + row);
    }
  }
  if (p.stats) {    // per-CTA partials, plain stores: bn_finalize sums the gridDim.x slices (no same-line atomics storm)
    atomicAdd(sstat + z * D + c, ssum.x); atomicAdd(sstat + z * D + c + 1, ssum.y);
    atomicAdd(sstat + (3 + z) * D + c, ssq.x); atomicAdd(sstat + (3 + z) * D + c + 1, ssq.y);
    __syncthreads();
    for (int i = tid; i < 6 * D; i += nthreads) p.stats[(long long)blockIdx.x * 6 * D + i] = sstat[i];
  }
}

// block = (D/2 channel pairs, CW columns, 6): threadIdx.z = role * 3 + z; role 0 threads compute branch z's contribution to the
// input gradient, role 1 threads accumulate branch z's depthwise-weight gradient.
__global__ void __launch_bounds__(768, 1) dwproj_bwd2_kernel(const __grid_constant__ DwProjBwdP p) {
  pdl_wait(); pdl_trigger();
  extern __shared__ __align__(16) float sdw[];            // [3][9][D] block partials of ddw, then [3][TH][nxy] float2 partial dx per branch
  const int D = p.D, H = p.H, W = p.W;
  const int c = threadIdx.x * 2, jl = threadIdx.y, z = threadIdx.z % 3, role = threadIdx.z / 3;
  const int nxy = blockDim.x * blockDim.y, txy = jl * blockDim.x + threadIdx.x;
  const int tid = threadIdx.z * nxy + txy, nthreads = nxy * 6;
  const int method = p.method[z];
  float2* dxp = reinterpret_cast<float2*>(sdw + 27 * D) + (z * TH) * nxy + txy;      // dxp[r * nxy]
  const int ncx = cdiv(W, (int)blockDim.y), ncy = cdiv(H, TH), ntiles = ncx * ncy * p.B;
  for (int i = tid; i < 27 * D; i += nthreads) sdw[i] = 0.f;
  __syncthreads();
  float2 cA = make_float2(1.f, 1.f), cE = f2_zero(), cF = f2_zero();
  const bool bn = p.coef != nullptr && method == 0;
  if (bn) {
    cA = *reinterpret_cast<const float2*>(p.coef + (z * 3 + 0) * D + c);
    cE = *reinterpret_cast<const float2*>(p.coef + (z * 3 + 1) * D + c);
    cF = *reinterpret_cast<const float2*>(p.coef + (z * 3 + 2) * D + c);
  }
  float2 aw[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) aw[t] = f2_zero();
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
  const int b = tile / (ncx * ncy), rem = tile - b * (ncx * ncy);
  const int j = (rem % ncx) * blockDim.y + jl, r0 = (rem / ncx) * TH;
  const long long base = ((long long)b * p.Tc + p.has_cls) * D + c;
  const float* xn = p.xn + base;
  const float* dn = p.dconv + z * p.bs + base;
  const float* cv = p.conv ? p.conv + z * p.bs + base : nullptr;
  const bool active = j < W;
  const int rend = min(r0 + TH, H);
  auto ldx = [&](int ii, int jj) -> float2 {
    if (ii < 0 || ii >= H || jj < 0 || jj >= W) return f2_zero();
    return *reinterpret_cast<const float2*>(xn + ((long long)ii * W + jj) * D);
  };
  auto ldd = [&](int ii, int jj) -> float2 {
    if (ii < 0 || ii >= H || jj < 0 || jj >= W) return f2_zero();
    const long long o = ((long long)ii * W + jj) * D;
    float2 v = *reinterpret_cast<const float2*>(dn + o);
    if (bn) {
      const float2 q = *reinterpret_cast<const float2*>(cv + o);
      v.x = fmaf(cA.x, v.x, fmaf(cE.x, q.x, cF.x)); v.y = fmaf(cA.y, v.y, fmaf(cE.y, q.y, cF.y));
    }
    if (method == 1) { const float ic = inv_count(ii, jj, H, W); v.x *= ic; v.y *= ic; }
    return v;
  };
  if (active && role == 0) {
    if (method == 2) {      // identity: dx = dconv
#pragma unroll 1
      for (int r = 0; r < TH; ++r) dxp[r * nxy] = (r0 + r < rend) ? ldd(r0 + r, j) : f2_zero();
    } else {
      float2 w[9];
#pragma unroll
      for (int t = 0; t < 9; ++t) w[t] = method == 0 ? *reinterpret_cast<const float2*>(p.w + (z * 9 + t) * D + c) : make_float2(1.f, 1.f);
      float2 dw[3][3];
#pragma unroll
      for (int a = 0; a < 3; ++a)
#pragma unroll
        for (int q = 0; q < 3; ++q) dw[a][q] = ldd(r0 - 1 + a, j - 1 + q);
#pragma unroll 2
      for (int r = 0; r < TH; ++r) {
        const int i = r0 + r;
        if (i >= rend) { dxp[r * nxy] = f2_zero(); continue; }
        float2 nd[3];
#pragma unroll
        for (int q = 0; q < 3; ++q) nd[q] = ldd(i + 2, j - 1 + q);
        float2 dx = f2_zero();
        // x[i,j] fed output (i-di, j-dj) through tap t  ->  window element [1-di][1-dj]
#pragma unroll
        for (int t = 0; t < 9; ++t) dx = f2_fma(dw[2 - t / 3][2 - t % 3], w[t], dx);
        dxp[r * nxy] = dx;
#pragma unroll
        for (int q = 0; q < 3; ++q) { dw[0][q] = dw[1][q]; dw[1][q] = dw[2][q]; dw[2][q] = nd[q]; }
      }
    }
  } else if (active && role == 1 && method == 0 && p.ddw) {
    float2 xw[3][3];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int q = 0; q < 3; ++q) xw[a][q] = ldx(r0 - 1 + a, j - 1 + q);
    float2 dc = ldd(r0, j);
#pragma unroll 2
    for (int i = r0; i < rend; ++i) {
      float2 nx[3];
#pragma unroll
      for (int q = 0; q < 3; ++q) nx[q] = ldx(i + 2, j - 1 + q);
      const float2 ndc = (i + 1 < rend) ? ldd(i + 1, j) : f2_zero();
      // output (i,j) read x[i+di, j+dj] through tap t
#pragma unroll
      for (int t = 0; t < 9; ++t) aw[t] = f2_fma(xw[t / 3][t % 3], dc, aw[t]);
      dc = ndc;
#pragma unroll
      for (int q = 0; q < 3; ++q) { xw[0][q] = xw[1][q]; xw[1][q] = xw[2][q]; xw[2][q] = nx[q]; }
    }
  }
  __syncthreads();
  if (active && threadIdx.z == 0) {
    const float2* d0 = reinterpret_cast<const float2*>(sdw + 27 * D) + txy;
#pragma unroll 1
    for (int r = 0; r < TH; ++r)
      if (r0 + r < rend) {
        const float2 a = d0[r * nxy], bb = d0[(TH + r) * nxy], cc = d0[(2 * TH + r) * nxy];
        *reinterpret_cast<float2*>(p.dxn + base + ((long long)(r0 + r) * W + j) * D) = make_float2(a.x + bb.x + cc.x, a.y + bb.y + cc.y);
      }
  }
  __syncthreads();        // dxp is rewritten by the next tile
  }
  if (role == 1 && method == 0 && p.ddw) {
#pragma unroll
    for (int t = 0; t < 9; ++t) { atomicAdd(sdw + (z * 9 + t) * D + c, aw[t].x); atomicAdd(sdw + (z * 9 + t) * D + c + 1, aw[t].y); }
  }
  __syncthreads();
  // per-CTA partial of the depthwise-weight gradient, plain stores: reduce_partials_kernel adds the gridDim.x slices into ddw
  if (p.ddw) for (int i = tid; i < 27 * D; i += nthreads) p.ddw_part[(long long)blockIdx.x * 27 * D + i] = sdw[i];
  if (p.has_cls && jl == 0 && threadIdx.z == 0) {   // the cls row bypassed conv and BN: plain sum
    for (int b = blockIdx.x; b < p.B; b += gridDim.x) {
      const long long row = (long long)b * p.Tc * D + c;
      float2 s = f2_zero();
#pragma unroll
      for (int zz = 0; zz < 3; ++zz) { const float2 v = *reinterpret_cast<const float2*>(p.dconv + zz * p.bs + row); s.x += v.x; s.y += v.y; }
      *reinterpret_cast<float2*>(p.dxn + row) = s;
    }
  }
}

// dst[i] += sum_k part[k][i];  block = (32 elements, 8 partial lanes)
__global__ void __launch_bounds__(256) reduce_partials_kernel(const float* __restrict__ part, int nparts, int n, float* __restrict__ dst) {
  pdl_wait(); pdl_trigger();
  __shared__ float red[8][33];
  const int i = blockIdx.x * 32 + threadIdx.x;
  float a = 0.f;
  if (i < n) for (int k = threadIdx.y; k < nparts; k += 8) a += part[(long long)k * n + i];
  red[threadIdx.y][threadIdx.x] = a;
  __syncthreads();
  if (threadIdx.y == 0 && i < n) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += red[k][threadIdx.x];
    dst[i] += t;
  }
}

// BatchNorm backward as an affine map of (dn, conv):  dconv = A dn + E conv + F   per (z, channel)
//   training : A = g r, E = -g r^2 dgamma / N, F = -g r dbeta / N + g r^2 mu dgamma / N
//   inference: A = scale, E = F = 0
__global__ void bn_bwd_coef_kernel(const float* __restrict__ gamma, const float* __restrict__ mean, const float* __restrict__ rstd,
                                   const float* __restrict__ scale, const float* __restrict__ dgamma, const float* __restrict__ dbeta,
                                   float invN, int training, int D, float* __restrict__ coef) {
  pdl_wait(); pdl_trigger();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;       // (z, c)
  if (i >= 3 * D) return;
  const int z = i / D, c = i - z * D;
  float A, E, F;
  if (training) {
    const float g = gamma[i], r = rstd[i], mu = mean[i], dg = dgamma[i] * invN, db = dbeta[i] * invN;
    A = g * r; E = -g * r * r * dg; F = -g * r * db + g * r * r * mu * dg;
  } else { A = scale[i]; E = 0.f; F = 0.f; }
  coef[(z * 3 + 0) * D + c] = A; coef[(z * 3 + 1) * D + c] = E; coef[(z * 3 + 2) * D + c] = F;
}

dim3 block_for(int D) {
  const int c2 = D / 2;
  int cw = 128 / c2; if (cw < 1) cw = 1;
  return dim3(c2, cw, 3);
}

}  // namespace

cudaError_t launch_dwproj_fwd2(const DwProjP& p, cudaStream_t s) {
  if ((p.D & 3) || p.D > 256) return cudaErrorNotSupported;
  const dim3 blk = block_for(p.D);
  const int ntiles = cdiv(p.W, blk.y) * cdiv(p.H, TH) * p.B;
  const int grid = ntiles < p.nparts ? ntiles : p.nparts;
  if (p.stats && grid < p.nparts) CVT_CUDA(cudaMemsetAsync(p.stats + (size_t)grid * 6 * p.D, 0, (size_t)(p.nparts - grid) * 6 * p.D * sizeof(float), s));
  const size_t smem = p.stats ? (size_t)6 * p.D * sizeof(float) : 0;
  CVT_LAUNCH((dwproj_fwd2_kernel), grid, blk, smem, s, p);
  return cudaSuccess;
}

cudaError_t launch_dwproj_bwd2(const DwProjBwdP& p, cudaStream_t s) {
  if ((p.D & 3) || p.D > 256) return cudaErrorNotSupported;
  dim3 blk = block_for(p.D); blk.z = 6;
  const int ntiles = cdiv(p.W, blk.y) * cdiv(p.H, TH) * p.B;
  const int grid = ntiles < kNumSMs ? ntiles : kNumSMs;
  if (p.ddw && !p.ddw_part) return cudaErrorInvalidValue;
  const size_t smem = (size_t)27 * p.D * sizeof(float) + (size_t)3 * TH * blk.x * blk.y * sizeof(float2);
  static bool configured = false;
  if (!configured) { CVT_CUDA(cudaFuncSetAttribute(dwproj_bwd2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024)); configured = true; }
  CVT_LAUNCH((dwproj_bwd2_kernel), grid, blk, smem, s, p);
  if (p.ddw) {
    CVT_LAUNCH((reduce_partials_kernel), cdiv(27 * p.D, 32), dim3(32, 8), 0, s, p.ddw_part, grid, 27 * p.D, p.ddw);
  }
  return cudaSuccess;
}

cudaError_t launch_reduce_partials(const float* part, int nparts, int n, float* dst, cudaStream_t s) {
  CVT_LAUNCH((reduce_partials_kernel), cdiv(n, 32), dim3(32, 8), 0, s, part, nparts, n, dst);
  return cudaSuccess;
}

cudaError_t launch_bn_bwd_coef(const float* gamma, const float* mean, const float* rstd, const float* scale, const float* dgamma,
                               const float* dbeta, long long count, int training, int D, float* coef, cudaStream_t s) {
  CVT_LAUNCH((bn_bwd_coef_kernel), cdiv(3 * D, 128), 128, 0, s, gamma, mean, rstd, scale, dgamma, dbeta, 1.f / (float)count, training, D, coef);
  return cudaSuccess;
}

}  // namespace cvt
