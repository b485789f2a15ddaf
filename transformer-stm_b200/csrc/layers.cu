// Memory-bound layers of the CvT block: LayerNorm, the q/k/v Projection (depthwise 3x3 + BatchNorm statistics,
// average pool, identity), softmax, dropout masks, loss and Adam.  All fp32; channel-contiguous (NHWC) accesses.
#include "layers.cuh"

namespace cvt {
namespace {

constexpr int LN_MAXV = 16;   // D <= 512

// ------------------------------------------------------------------------------------------------
// LayerNorm
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ln_fwd_kernel(const float* __restrict__ x, RowMap in_map, long long ldx,
                                                     const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                                                     float* __restrict__ y, long long ldy, float* __restrict__ mean,
                                                     float* __restrict__ rstd, int rows, int D) {
  pdl_wait(); pdl_trigger();
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const int n = D >> 5;
  const float invD = 1.f / (float)D;
  for (int row = warp; row < rows; row += nwarps) {
    const float* xr = x + in_map(row) * ldx;
    float v[LN_MAXV];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < LN_MAXV; ++i)
      if (i < n) { v[i] = xr[lane + 32 * i]; s += v[i]; }
    const float mu = warp_sum(s) * invD;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < LN_MAXV; ++i)
      if (i < n) { const float d = v[i] - mu; q += d * d; }
    const float var = warp_sum(q) * invD;
    const float r = 1.0f / sqrtf(var + eps);
    float* yr = y + (long long)row * ldy;
#pragma unroll
    for (int i = 0; i < LN_MAXV; ++i)
      if (i < n) { const int c = lane + 32 * i; yr[c] = (v[i] - mu) * r * gamma[c] + beta[c]; }
    if (lane == 0) {
      if (mean) mean[row] = mu;
      if (rstd) rstd[row] = r;
    }
  }
}

__global__ void __launch_bounds__(256) ln_bwd_kernel(const float* __restrict__ dy, long long lddy, const float* __restrict__ x,
                                                     RowMap in_map, long long ldx, const float* __restrict__ mean,
                                                     const float* __restrict__ rstd, const float* __restrict__ gamma,
                                                     const float* dres, float* dx, RowMap out_map,   // may alias (in-place add)
                                                     long long ldo, float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                     int rows, int D) {
  pdl_wait(); pdl_trigger();
  extern __shared__ float sm[];   // [2][D]
  float* sg = sm; float* sb = sm + D;
  for (int i = threadIdx.x; i < 2 * D; i += blockDim.x) sm[i] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const int n = D >> 5;
  const float invD = 1.f / (float)D;
  float ag[LN_MAXV], ab[LN_MAXV];
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i) { ag[i] = 0.f; ab[i] = 0.f; }
  for (int row = warp; row < rows; row += nwarps) {
    const float* xr = x + in_map(row) * ldx;
    const float* dyr = dy + (long long)row * lddy;
    const float mu = mean[row], r = rstd[row];
    float xh[LN_MAXV], g[LN_MAXV];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < LN_MAXV; ++i)
      if (i < n) {
        const int c = lane + 32 * i;
        const float d = dyr[c];
        xh[i] = (xr[c] - mu) * r;
        g[i] = d * gamma[c];
        s1 += g[i]; s2 += g[i] * xh[i];
        ag[i] += d * xh[i]; ab[i] += d;
      }
    const float m1 = warp_sum(s1) * invD, m2 = warp_sum(s2) * invD;
    const long long orow = out_map(row);
    float* dxr = dx + orow * ldo;
    const float* rr = dres ? dres + orow * ldo : nullptr;
#pragma unroll
    for (int i = 0; i < LN_MAXV; ++i)
      if (i < n) {
        const int c = lane + 32 * i;
        const float v = r * (g[i] - m1 - xh[i] * m2);
        dxr[c] = rr ? rr[c] + v : v;
      }
  }
#pragma unroll
  for (int i = 0; i < LN_MAXV; ++i)
    if (i < n) { atomicAdd(&sg[lane + 32 * i], ag[i]); atomicAdd(&sb[lane + 32 * i], ab[i]); }
  __syncthreads();
  for (int c = threadIdx.x; c < D; c += blockDim.x) { atomicAdd(dgamma + c, sg[c]); atomicAdd(dbeta + c, sb[c]); }
}

// ------------------------------------------------------------------------------------------------
// Projection (depthwise 3x3 / avg pool / identity) for q,k,v + BN statistics
// block = (D/4, PY) threads: x = float4 channel group, y = pixel lane; grid-stride over the B*H*W pixels.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 f4_fma(float4 a, float4 b, float4 c) {
  return make_float4(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y), fmaf(a.z, b.z, c.z), fmaf(a.w, b.w, c.w));
}
__device__ __forceinline__ float4 f4_add(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 f4_scale(float4 a, float s) { return make_float4(a.x * s, a.y * s, a.z * s, a.w * s); }
__device__ __forceinline__ float4 f4_zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }

__global__ void dwproj_fwd_kernel(const __grid_constant__ DwProjP p) {
  pdl_wait(); pdl_trigger();
  extern __shared__ __align__(16) float smw[];            // [3][9][D] weights, then (stats) [6][D] block partials
  const int D = p.D, C4 = D >> 2;
  const int c4 = threadIdx.x, py = threadIdx.y, PY = blockDim.y;
  const int tid = py * C4 + c4, nthreads = C4 * PY;
  float* sstat = smw + 27 * D;
  if (p.w) for (int i = tid; i < 27 * D; i += nthreads) smw[i] = p.w[i];
  if (p.stats) for (int i = tid; i < 6 * D; i += nthreads) sstat[i] = 0.f;
  __syncthreads();
  const long long npix = (long long)p.B * p.H * p.W;
  const int HW = p.H * p.W;
  float4 ssum[3], ssq[3];
#pragma unroll
  for (int z = 0; z < 3; ++z) { ssum[z] = f4_zero(); ssq[z] = f4_zero(); }
  for (long long pix = (long long)blockIdx.x * PY + py; pix < npix; pix += (long long)gridDim.x * PY) {
    const int b = (int)(pix / HW), r = (int)(pix - (long long)b * HW);
    const int i = r / p.W, j = r - i * p.W;
    const long long base = ((long long)b * p.Tc + p.has_cls) * D;      // first spatial token of image b
    float4 nb[9];
    int cnt = 0;
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const int ii = i + t / 3 - 1, jj = j + t % 3 - 1;
      if (ii >= 0 && ii < p.H && jj >= 0 && jj < p.W) {
        nb[t] = *reinterpret_cast<const float4*>(p.xn + base + ((long long)ii * p.W + jj) * D + c4 * 4);
        ++cnt;
      } else nb[t] = f4_zero();
    }
    const long long orow = ((long long)b * p.Tc + p.has_cls + r) * D + c4 * 4;
#pragma unroll
    for (int z = 0; z < 3; ++z) {
      float4 y;
      if (p.method[z] == 0) {
        y = f4_zero();
#pragma unroll
        for (int t = 0; t < 9; ++t) y = f4_fma(nb[t], *reinterpret_cast<const float4*>(smw + (z * 9 + t) * D + c4 * 4), y);
      } else if (p.method[z] == 1) {
        y = f4_zero();
#pragma unroll
        for (int t = 0; t < 9; ++t) y = f4_add(y, nb[t]);
        y = f4_scale(y, 1.f / (float)cnt);
      } else y = nb[4];
      *reinterpret_cast<float4*>(p.out + z * p.out_bs + orow) = y;
      ssum[z] = f4_add(ssum[z], y);
      ssq[z] = f4_fma(y, y, ssq[z]);
    }
  }
  // cls rows pass through untouched (models/CvT(Par).py:150,168-170)
  if (p.has_cls) {
    for (int b = blockIdx.x * PY + py; b < p.B; b += gridDim.x * PY) {
      const long long row = (long long)b * p.Tc * D + c4 * 4;
      const float4 v = p.cls_zero ? f4_zero() : *reinterpret_cast<const float4*>(p.xn + row);
#pragma unroll
      for (int z = 0; z < 3; ++z) *reinterpret_cast<float4*>(p.out + z * p.out_bs + row) = v;
    }
  }
  if (p.stats) {
#pragma unroll
    for (int z = 0; z < 3; ++z) {
      float* a = sstat + z * D + c4 * 4; float* q = sstat + (3 + z) * D + c4 * 4;
      atomicAdd(a + 0, ssum[z].x); atomicAdd(a + 1, ssum[z].y); atomicAdd(a + 2, ssum[z].z); atomicAdd(a + 3, ssum[z].w);
      atomicAdd(q + 0, ssq[z].x); atomicAdd(q + 1, ssq[z].y); atomicAdd(q + 2, ssq[z].z); atomicAdd(q + 3, ssq[z].w);
    }
    __syncthreads();
    for (int i = tid; i < 6 * D; i += nthreads) atomicAdd(p.stats + i, sstat[i]);
  }
}

// block = (32 channels, 32 partial lanes): the per-CTA partial slices of the statistics are summed 32-way in parallel
__global__ void __launch_bounds__(1024) bn_finalize_kernel(const __grid_constant__ BnFinalizeP p) {
  pdl_wait(); pdl_trigger();
  __shared__ double red[2][32][33];
  const int i = blockIdx.x * 32 + threadIdx.x;      // (z, c)
  const bool ok = i < 3 * p.D;
  double s1 = 0.0, s2 = 0.0;
  if (ok && p.training) {
#pragma unroll 4
    for (int k = threadIdx.y; k < p.nparts; k += 32) {
      s1 += (double)p.stats[(long long)k * 6 * p.D + i]; s2 += (double)p.stats[(long long)k * 6 * p.D + 3 * p.D + i];
    }
  }
  red[0][threadIdx.y][threadIdx.x] = s1; red[1][threadIdx.y][threadIdx.x] = s2;
  __syncthreads();
  if (threadIdx.y != 0 || !ok) return;
  float mu, var;
  if (p.training) {
    s1 = 0.0; s2 = 0.0;
    for (int k = 0; k < 32; ++k) { s1 += red[0][k][threadIdx.x]; s2 += red[1][k][threadIdx.x]; }
    const double n = (double)p.count;
    const double m = s1 / n;
    double v = s2 / n - m * m;
    if (v < 0.0) v = 0.0;
    mu = (float)m; var = (float)v;
    if (p.update_moving) {
      const float unb = (float)(v * (n / (n > 1.0 ? n - 1.0 : 1.0)));   // TF fused BN feeds the unbiased variance
      p.moving_mean[i] = p.momentum * p.moving_mean[i] + (1.f - p.momentum) * mu;
      p.moving_var[i] = p.momentum * p.moving_var[i] + (1.f - p.momentum) * unb;
    }
  } else { mu = p.moving_mean[i]; var = p.moving_var[i]; }
  const float r = 1.0f / sqrtf(var + p.eps);
  const float sc = p.gamma[i] * r;
  p.scale[i] = sc;
  p.shift[i] = p.beta[i] - mu * sc;
  p.save_mean[i] = mu;
  p.save_rstd[i] = r;
}

__global__ void bn_bwd_reduce_kernel(const float* __restrict__ dy, const float* __restrict__ conv, long long bs,
                                     const float* __restrict__ save_mean, const float* __restrict__ save_rstd,
                                     float* __restrict__ dgamma, float* __restrict__ dbeta, int B, int T, int Tc, int has_cls, int D) {
  pdl_wait(); pdl_trigger();
  extern __shared__ __align__(16) float sred[];             // [6][D]
  const int C4 = D >> 2, c4 = threadIdx.x, py = threadIdx.y, PY = blockDim.y;
  const int tid = py * C4 + c4, nthreads = C4 * PY;
  for (int i = tid; i < 6 * D; i += nthreads) sred[i] = 0.f;
  __syncthreads();
  float4 ag[3], ab[3], mu[3], rs[3];
#pragma unroll
  for (int z = 0; z < 3; ++z) {
    ag[z] = f4_zero(); ab[z] = f4_zero();
    mu[z] = *reinterpret_cast<const float4*>(save_mean + z * D + c4 * 4);
    rs[z] = *reinterpret_cast<const float4*>(save_rstd + z * D + c4 * 4);
  }
  const long long npix = (long long)B * T;
  for (long long pix = (long long)blockIdx.x * PY + py; pix < npix; pix += (long long)gridDim.x * PY) {
    const long long b = pix / T, r = pix - b * T;
    const long long off = (b * Tc + has_cls + r) * D + c4 * 4;
#pragma unroll
    for (int z = 0; z < 3; ++z) {
      const float4 d = *reinterpret_cast<const float4*>(dy + z * bs + off);
      const float4 c = *reinterpret_cast<const float4*>(conv + z * bs + off);
      const float4 xh = make_float4((c.x - mu[z].x) * rs[z].x, (c.y - mu[z].y) * rs[z].y, (c.z - mu[z].z) * rs[z].z, (c.w - mu[z].w) * rs[z].w);
      ag[z] = f4_fma(d, xh, ag[z]);
      ab[z] = f4_add(ab[z], d);
    }
  }
#pragma unroll
  for (int z = 0; z < 3; ++z) {
    float* a = sred + z * D + c4 * 4; float* q = sred + (3 + z) * D + c4 * 4;
    atomicAdd(a + 0, ag[z].x); atomicAdd(a + 1, ag[z].y); atomicAdd(a + 2, ag[z].z); atomicAdd(a + 3, ag[z].w);
    atomicAdd(q + 0, ab[z].x); atomicAdd(q + 1, ab[z].y); atomicAdd(q + 2, ab[z].z); atomicAdd(q + 3, ab[z].w);
  }
  __syncthreads();
  for (int i = tid; i < 3 * D; i += nthreads) { atomicAdd(dgamma + i, sred[i]); atomicAdd(dbeta + i, sred[3 * D + i]); }
}

__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(float* __restrict__ dy, const float* __restrict__ conv, long long bs,
                                                           const float* __restrict__ save_mean, const float* __restrict__ save_rstd,
                                                           const float* __restrict__ gamma, const float* __restrict__ scale,
                                                           const float* __restrict__ dgamma, const float* __restrict__ dbeta,
                                                           int training, int B, int T, int Tc, int has_cls, int D) {
  pdl_wait(); pdl_trigger();
  const int C4 = D >> 2;
  const long long per = (long long)B * T * C4;
  const long long total = 3 * per;
  const float invN = 1.f / (float)((long long)B * T);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int z = (int)(i / per);
    const long long rem = i - z * per;
    const long long pix = rem / C4; const int c = (int)(rem - pix * C4) * 4;
    const long long b = pix / T, r = pix - b * T;
    const long long off = z * bs + (b * Tc + has_cls + r) * D + c;
    float4 d = *reinterpret_cast<const float4*>(dy + off);
    if (training) {
      const float4 cv = *reinterpret_cast<const float4*>(conv + off);
      const float4 mu = *reinterpret_cast<const float4*>(save_mean + z * D + c);
      const float4 rs = *reinterpret_cast<const float4*>(save_rstd + z * D + c);
      const float4 g = *reinterpret_cast<const float4*>(gamma + z * D + c);
      const float4 dg = *reinterpret_cast<const float4*>(dgamma + z * D + c);
      const float4 db = *reinterpret_cast<const float4*>(dbeta + z * D + c);
      d.x = g.x * rs.x * (d.x - db.x * invN - (cv.x - mu.x) * rs.x * dg.x * invN);
      d.y = g.y * rs.y * (d.y - db.y * invN - (cv.y - mu.y) * rs.y * dg.y * invN);
      d.z = g.z * rs.z * (d.z - db.z * invN - (cv.z - mu.z) * rs.z * dg.z * invN);
      d.w = g.w * rs.w * (d.w - db.w * invN - (cv.w - mu.w) * rs.w * dg.w * invN);
    } else {
      const float4 sc = *reinterpret_cast<const float4*>(scale + z * D + c);
      d.x *= sc.x; d.y *= sc.y; d.z *= sc.z; d.w *= sc.w;
    }
    *reinterpret_cast<float4*>(dy + off) = d;
  }
}

__global__ void dwproj_bwd_kernel(const __grid_constant__ DwProjBwdP p) {
  pdl_wait(); pdl_trigger();
  extern __shared__ __align__(16) float smw[];            // [3][9][D] weights, then [9][D] block partial of ddw (per z)
  const int D = p.D, C4 = D >> 2;
  const int c4 = threadIdx.x, py = threadIdx.y, PY = blockDim.y;
  const int tid = py * C4 + c4, nthreads = C4 * PY;
  float* sdw = smw + 27 * D;
  if (p.w) for (int i = tid; i < 27 * D; i += nthreads) smw[i] = p.w[i];
  __syncthreads();
  const long long npix = (long long)p.B * p.H * p.W;
  const int HW = p.H * p.W;
  for (int z = 0; z < 3; ++z) {
    const int method = p.method[z];
    const float* dc = p.dconv + z * p.bs;
    float4 aw[9];
#pragma unroll
    for (int t = 0; t < 9; ++t) aw[t] = f4_zero();
    for (long long pix = (long long)blockIdx.x * PY + py; pix < npix; pix += (long long)gridDim.x * PY) {
      const int b = (int)(pix / HW), r = (int)(pix - (long long)b * HW);
      const int i = r / p.W, j = r - i * p.W;
      const long long base = ((long long)b * p.Tc + p.has_cls) * D + c4 * 4;
      const long long me = base + (long long)r * D;
      float4 dx = f4_zero();
      if (method == 2) {
        dx = *reinterpret_cast<const float4*>(dc + me);
      } else {
        const float4 dcc = *reinterpret_cast<const float4*>(dc + me);
#pragma unroll
        for (int t = 0; t < 9; ++t) {
          const int di = t / 3 - 1, dj = t % 3 - 1;
          // (a) input gradient: x[i,j] fed output (i-di, j-dj) through tap t
          const int oi = i - di, oj = j - dj;
          if (oi >= 0 && oi < p.H && oj >= 0 && oj < p.W) {
            const float4 g = *reinterpret_cast<const float4*>(dc + base + ((long long)oi * p.W + oj) * D);
            if (method == 0) dx = f4_fma(g, *reinterpret_cast<const float4*>(smw + (z * 9 + t) * D + c4 * 4), dx);
            else {
              const int ch = min(oi + 1, p.H - 1) - max(oi - 1, 0) + 1, cw = min(oj + 1, p.W - 1) - max(oj - 1, 0) + 1;
              dx = f4_add(dx, f4_scale(g, 1.f / (float)(ch * cw)));
            }
          }
          // (b) weight gradient: output (i,j) read x[i+di, j+dj] through tap t
          if (method == 0) {
            const int ii = i + di, jj = j + dj;
            if (ii >= 0 && ii < p.H && jj >= 0 && jj < p.W) {
              const float4 xv = *reinterpret_cast<const float4*>(p.xn + base + ((long long)ii * p.W + jj) * D);
              aw[t] = f4_fma(xv, dcc, aw[t]);
            }
          }
        }
      }
      float4* dst = reinterpret_cast<float4*>(p.dxn + me);
      *dst = (z == 0) ? dx : f4_add(*dst, dx);
    }
    if (method == 0) {
      __syncthreads();
      for (int i = tid; i < 9 * D; i += nthreads) sdw[i] = 0.f;
      __syncthreads();
#pragma unroll
      for (int t = 0; t < 9; ++t) {
        float* a = sdw + t * D + c4 * 4;
        atomicAdd(a + 0, aw[t].x); atomicAdd(a + 1, aw[t].y); atomicAdd(a + 2, aw[t].z); atomicAdd(a + 3, aw[t].w);
      }
      __syncthreads();
      for (int i = tid; i < 9 * D; i += nthreads) atomicAdd(p.ddw + z * 9 * D + i, sdw[i]);
    }
  }
  if (p.has_cls) {
    for (int b = blockIdx.x * PY + py; b < p.B; b += gridDim.x * PY) {
      const long long row = (long long)b * p.Tc * D + c4 * 4;
      float4 s = f4_zero();
#pragma unroll
      for (int z = 0; z < 3; ++z) s = f4_add(s, *reinterpret_cast<const float4*>(p.dconv + z * p.bs + row));
      *reinterpret_cast<float4*>(p.dxn + row) = s;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// softmax rows (fp32 attention path)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) softmax_rows_kernel(float* __restrict__ S, long long rows, int cols) {
  pdl_wait(); pdl_trigger();
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long row = warp; row < rows; row += nwarps) {
    float* s = S + row * cols;
    float mx = -INFINITY;
    for (int c = lane; c < cols; c += 32) mx = fmaxf(mx, s[c]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int c = lane; c < cols; c += 32) { const float e = expf(s[c] - mx); s[c] = e; sum += e; }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    for (int c = lane; c < cols; c += 32) s[c] *= inv;
  }
}
__global__ void __launch_bounds__(256) softmax_bwd_rows_kernel(float* __restrict__ dP, const float* __restrict__ P, long long rows, int cols) {
  pdl_wait(); pdl_trigger();
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long row = warp; row < rows; row += nwarps) {
    float* d = dP + row * cols; const float* pr = P + row * cols;
    float dot = 0.f;
    for (int c = lane; c < cols; c += 32) dot += d[c] * pr[c];
    dot = warp_sum(dot);
    for (int c = lane; c < cols; c += 32) d[c] = pr[c] * (d[c] - dot);
  }
}

// ------------------------------------------------------------------------------------------------
// small elementwise kernels
// ------------------------------------------------------------------------------------------------
// dyn (optional): per-step values kept in device memory so that a captured CUDA graph can be replayed: {alpha, grad_scale, dropout step}
__global__ void set_step_params_kernel(float* dyn, float alpha, float gs, uint32_t step) {
  pdl_wait(); pdl_trigger();
  dyn[0] = alpha; dyn[1] = gs; reinterpret_cast<uint32_t*>(dyn)[2] = step;
}
__global__ void __launch_bounds__(256) philox_mask_kernel(uint8_t* __restrict__ mask, long long n, uint32_t thr, uint32_t k0, uint32_t k1,
                                                          uint32_t step, uint32_t site, const float* __restrict__ dyn) {
  pdl_wait(); pdl_trigger();
  if (dyn) step = reinterpret_cast<const uint32_t*>(dyn)[2];
  const long long nq = (n + 3) >> 2;
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < nq; q += (long long)gridDim.x * blockDim.x) {
    uint32_t c[4] = {(uint32_t)q, site, step, 0u};
    philox4x32_10(c, k0, k1);
    const long long base = q << 2;
    if (base + 3 < n) {
      uchar4 m = make_uchar4(c[0] >= thr, c[1] >= thr, c[2] >= thr, c[3] >= thr);
      *reinterpret_cast<uchar4*>(mask + base) = m;
    } else {
      for (int i = 0; i < 4; ++i) if (base + i < n) mask[base + i] = c[i] >= thr;
    }
  }
}
__global__ void __launch_bounds__(256) mask_scale_kernel(float* __restrict__ dst, const float* __restrict__ src,
                                                         const uint8_t* __restrict__ mask, float scale, long long n, int round) {
  pdl_wait(); pdl_trigger();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float v = mask[i] ? src[i] * scale : 0.f;
    dst[i] = round ? round_tf32(v) : v;
  }
}
__global__ void __launch_bounds__(256) fill_rows_kernel(float* __restrict__ dst, RowMap map, long long ld, const float* __restrict__ src, int rows, int D) {
  pdl_wait(); pdl_trigger();
  const long long total = (long long)rows * D;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / D; const int c = (int)(i - r * D);
    dst[map(r) * ld + c] = src[c];
  }
}
__global__ void __launch_bounds__(256) copy_rows_kernel(float* __restrict__ dst, RowMap dmap, long long ldd, const float* __restrict__ src,
                                                        RowMap smap, long long lds, int rows, int D) {
  pdl_wait(); pdl_trigger();
  const long long total = (long long)rows * D;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / D; const int c = (int)(i - r * D);
    dst[dmap(r) * ldd + c] = src[smap(r) * lds + c];
  }
}
// 16-byte version (D, ldd, lds multiples of 4, aligned bases): the row-map divisions are paid once per float4
__global__ void __launch_bounds__(256) copy_rows_v4_kernel(float* __restrict__ dst, RowMap dmap, long long ldd, const float* __restrict__ src,
                                                           RowMap smap, long long lds, int rows, int D4) {
  pdl_wait(); pdl_trigger();
  const long long total = (long long)rows * D4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / D4), c = (int)(i - (long long)r * D4) << 2;
    *reinterpret_cast<float4*>(dst + dmap(r) * ldd + c) = *reinterpret_cast<const float4*>(src + smap(r) * lds + c);
  }
}
__global__ void __launch_bounds__(256) token_mean_kernel(const float* __restrict__ x, int B, int T, int D, float* __restrict__ out, long long ldo) {
  pdl_wait(); pdl_trigger();
  // block = one image, threads over channels; coalesced over c
  const int b = blockIdx.x;
  for (int c = threadIdx.x; c < D; c += blockDim.x) {
    float s = 0.f;
    for (int t = 0; t < T; ++t) s += x[((long long)b * T + t) * D + c];
    out[(long long)b * ldo + c] = s / (float)T;
  }
}
__global__ void __launch_bounds__(256) token_mean_bwd_kernel(const float* __restrict__ dy, long long lddy, int B, int T, int D, float* __restrict__ dx) {
  pdl_wait(); pdl_trigger();
  const long long total = (long long)B * T * D;
  const float inv = 1.f / (float)T;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / D; const int c = (int)(i - row * D);
    dx[i] = dy[(row / T) * lddy + c] * inv;
  }
}

__global__ void __launch_bounds__(256) mse_loss_kernel(const float* __restrict__ pred, const float* __restrict__ y, int n,
                                                       float* __restrict__ metrics, float* __restrict__ dpred, float grad_scale, int accumulate) {
  pdl_wait(); pdl_trigger();
  __shared__ float s1[8], s2[8];
  float a = 0.f, b = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float d = pred[i] - y[i];
    a += d * d; b += fabsf(d);
    if (dpred) dpred[i] = 2.f * d / (float)n * grad_scale;
  }
  a = warp_sum(a); b = warp_sum(b);
  if ((threadIdx.x & 31) == 0) { s1[threadIdx.x >> 5] = a; s2[threadIdx.x >> 5] = b; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float ta = 0.f, tb = 0.f;
    for (int i = 0; i < 8; ++i) { ta += s1[i]; tb += s2[i]; }
    if (accumulate) { metrics[0] += ta; metrics[1] += tb; }
    else { metrics[0] = ta / (float)n; metrics[1] = tb / (float)n; }
  }
}

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                   float* __restrict__ v, long long n4, float alpha, float b1, float b2, float eps, float gs,
                                                   const float* __restrict__ dyn) {
  pdl_wait(); pdl_trigger();
  if (dyn) { alpha = dyn[0]; gs = dyn[1]; }
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 pp = reinterpret_cast<float4*>(p)[i];
    float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    gg.x *= gs; gg.y *= gs; gg.z *= gs; gg.w *= gs;
    mm.x = b1 * mm.x + (1.f - b1) * gg.x; mm.y = b1 * mm.y + (1.f - b1) * gg.y;
    mm.z = b1 * mm.z + (1.f - b1) * gg.z; mm.w = b1 * mm.w + (1.f - b1) * gg.w;
    vv.x = b2 * vv.x + (1.f - b2) * gg.x * gg.x; vv.y = b2 * vv.y + (1.f - b2) * gg.y * gg.y;
    vv.z = b2 * vv.z + (1.f - b2) * gg.z * gg.z; vv.w = b2 * vv.w + (1.f - b2) * gg.w * gg.w;
    pp.x -= alpha * mm.x / (sqrtf(vv.x) + eps); pp.y -= alpha * mm.y / (sqrtf(vv.y) + eps);
    pp.z -= alpha * mm.z / (sqrtf(vv.z) + eps); pp.w -= alpha * mm.w / (sqrtf(vv.w) + eps);
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
}

inline int ew_blocks(long long n, int per_block = 256) {
  long long b = cdivl(n, per_block);
  const long long cap = (long long)kNumSMs * 8;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}
inline dim3 proj_block(int D) { const int c4 = D / 4; int py = 256 / c4; if (py < 1) py = 1; return dim3(c4, py); }
inline int proj_grid(long long npix, int py) {
  long long b = cdivl(npix, py);
  const long long cap = (long long)kNumSMs * 4;
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}
}  // namespace

cudaError_t launch_ln_fwd(const float* x, RowMap in_map, long long ldx, const float* gamma, const float* beta, float eps,
                          float* y, long long ldy, float* mean, float* rstd, int rows, int D, cudaStream_t s, int round_out) {
  if (rows <= 0) return cudaSuccess;
  if (D % 32 != 0 || D > 32 * LN_MAXV) { set_error("layer norm: D must be a multiple of 32 and <= 512"); return cudaErrorInvalidValue; }
  {
    const cudaError_t e = launch_ln_fwd_v4(x, in_map, ldx, gamma, beta, eps, y, ldy, mean, rstd, rows, D, s, round_out);
    if (e != cudaErrorNotSupported) return e;
  }
  CVT_LAUNCH((ln_fwd_kernel), ew_blocks(rows, 8), 256, 0, s, x, in_map, ldx, gamma, beta, eps, y, ldy, mean, rstd, rows, D);
  return cudaSuccess;
}

cudaError_t launch_ln_bwd(const float* dy, long long lddy, const float* x, RowMap in_map, long long ldx, const float* mean,
                          const float* rstd, const float* gamma, const float* dres, float* dx, RowMap out_map, long long ldo,
                          float* dgamma, float* dbeta, int rows, int D, cudaStream_t s,
                          const uint8_t* mask, float mscale, float* dxm, int round_m) {
  if (rows <= 0) return cudaSuccess;
  if (D % 32 != 0 || D > 32 * LN_MAXV) { set_error("layer norm: D must be a multiple of 32 and <= 512"); return cudaErrorInvalidValue; }
  if (mask && (out_map.period || ldo != D)) { set_error("layer norm backward: the masked output needs a dense [rows, D] dx"); return cudaErrorInvalidValue; }
  {
    const cudaError_t e = launch_ln_bwd_v4(dy, lddy, x, in_map, ldx, mean, rstd, gamma, dres, dx, out_map, ldo, dgamma, dbeta, rows, D, s,
                                           mask, mscale, dxm, round_m);
    if (e != cudaErrorNotSupported) return e;
  }
  int blocks = cdiv(rows, 8);
  if (blocks > kNumSMs * 2) blocks = kNumSMs * 2;
  CVT_LAUNCH((ln_bwd_kernel), blocks, 256, 2 * D * sizeof(float), s, dy, lddy, x, in_map, ldx, mean, rstd, gamma, dres, dx, out_map, ldo,
                                                           dgamma, dbeta, rows, D);
  if (mask) CVT_CUDA(launch_mask_scale(dxm, dx, mask, mscale, (long long)rows * D, s, round_m));
  return cudaSuccess;
}

cudaError_t launch_dwproj_fwd(const DwProjP& p, cudaStream_t s) {
  if (p.D % 4 != 0 || p.D > 1024) { set_error("projection: D must be a multiple of 4, <= 1024"); return cudaErrorInvalidValue; }
  const dim3 blk = proj_block(p.D);
  const size_t smem = (size_t)(27 + 6) * p.D * sizeof(float);
  if (smem > 48 * 1024) CVT_CUDA(cudaFuncSetAttribute(dwproj_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CVT_LAUNCH((dwproj_fwd_kernel), proj_grid((long long)p.B * p.H * p.W, blk.y), blk, smem, s, p);
  return cudaSuccess;
}

cudaError_t launch_bn_finalize(const BnFinalizeP& p, cudaStream_t s) {
  CVT_LAUNCH((bn_finalize_kernel), cdiv(3 * p.D, 32), dim3(32, 32), 0, s, p);
  return cudaSuccess;
}

cudaError_t launch_bn_bwd_reduce(const float* dy, const float* conv, long long bs, const float* save_mean, const float* save_rstd,
                                 float* dgamma, float* dbeta, int B, int T, int Tc, int has_cls, int D, cudaStream_t s) {
  const dim3 blk = proj_block(D);
  CVT_LAUNCH((bn_bwd_reduce_kernel), proj_grid((long long)B * T, blk.y), blk, 6 * D * sizeof(float), s, dy, conv, bs, save_mean, save_rstd,
                                                                                             dgamma, dbeta, B, T, Tc, has_cls, D);
  return cudaSuccess;
}

cudaError_t launch_bn_bwd_apply(float* dy, const float* conv, long long bs, const float* save_mean, const float* save_rstd,
                                const float* gamma, const float* scale, const float* dgamma, const float* dbeta, int training,
                                int B, int T, int Tc, int has_cls, int D, cudaStream_t s) {
  const long long total = 3ll * B * T * (D / 4);
  CVT_LAUNCH((bn_bwd_apply_kernel), ew_blocks(total), 256, 0, s, dy, conv, bs, save_mean, save_rstd, gamma, scale, dgamma, dbeta, training,
                                                       B, T, Tc, has_cls, D);
  return cudaSuccess;
}

cudaError_t launch_dwproj_bwd(const DwProjBwdP& p, cudaStream_t s) {
  const dim3 blk = proj_block(p.D);
  const size_t smem = (size_t)(27 + 9) * p.D * sizeof(float);
  if (smem > 48 * 1024) CVT_CUDA(cudaFuncSetAttribute(dwproj_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CVT_LAUNCH((dwproj_bwd_kernel), proj_grid((long long)p.B * p.H * p.W, blk.y), blk, smem, s, p);
  return cudaSuccess;
}

cudaError_t launch_softmax_rows(float* S, long long rows, int cols, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  CVT_LAUNCH((softmax_rows_kernel), ew_blocks(rows, 8), 256, 0, s, S, rows, cols);
  return cudaSuccess;
}
cudaError_t launch_softmax_bwd_rows(float* dP, const float* P, long long rows, int cols, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  CVT_LAUNCH((softmax_bwd_rows_kernel), ew_blocks(rows, 8), 256, 0, s, dP, P, rows, cols);
  return cudaSuccess;
}

cudaError_t launch_philox_mask(uint8_t* mask, long long n, float rate, uint64_t seed, uint32_t step, uint32_t site, cudaStream_t s, const float* dyn) {
  if (n <= 0) return cudaSuccess;
  double t = (double)rate * 4294967296.0;
  const uint32_t thr = t >= 4294967295.0 ? 0xFFFFFFFFu : (uint32_t)t;
  CVT_LAUNCH((philox_mask_kernel), ew_blocks((n + 3) / 4), 256, 0, s, mask, n, thr, (uint32_t)(seed & 0xFFFFFFFFu), (uint32_t)(seed >> 32), step, site, dyn);
  return cudaSuccess;
}
cudaError_t launch_mask_scale(float* dst, const float* src, const uint8_t* mask, float scale, long long n, cudaStream_t s, int round) {
  if (n <= 0) return cudaSuccess;
  CVT_LAUNCH((mask_scale_kernel), ew_blocks(n), 256, 0, s, dst, src, mask, scale, n, round);
  return cudaSuccess;
}
cudaError_t launch_fill_rows(float* dst, RowMap map, long long ld, const float* src, int rows, int D, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  CVT_LAUNCH((fill_rows_kernel), ew_blocks((long long)rows * D), 256, 0, s, dst, map, ld, src, rows, D);
  return cudaSuccess;
}
cudaError_t launch_copy_rows(float* dst, RowMap dmap, long long ldd, const float* src, RowMap smap, long long lds, int rows, int D, cudaStream_t s) {
  if (rows <= 0) return cudaSuccess;
  if (!(D & 3) && !(ldd & 3) && !(lds & 3) && !(reinterpret_cast<uintptr_t>(dst) & 15) && !(reinterpret_cast<uintptr_t>(src) & 15))
    CVT_LAUNCH((copy_rows_v4_kernel), ew_blocks((long long)rows * (D >> 2)), 256, 0, s, dst, dmap, ldd, src, smap, lds, rows, D >> 2);
  else
    CVT_LAUNCH((copy_rows_kernel), ew_blocks((long long)rows * D), 256, 0, s, dst, dmap, ldd, src, smap, lds, rows, D);
  return cudaSuccess;
}
cudaError_t launch_token_mean(const float* x, int B, int T, int D, float* out, long long ldo, cudaStream_t s) {
  CVT_LAUNCH((token_mean_kernel), B, 256, 0, s, x, B, T, D, out, ldo);
  return cudaSuccess;
}
cudaError_t launch_token_mean_bwd(const float* dy, long long lddy, int B, int T, int D, float* dx, cudaStream_t s) {
  CVT_LAUNCH((token_mean_bwd_kernel), ew_blocks((long long)B * T * D), 256, 0, s, dy, lddy, B, T, D, dx);
  return cudaSuccess;
}
cudaError_t launch_mse_loss(const float* pred, const float* y, int n, float* metrics, float* dpred, float grad_scale, int accumulate, cudaStream_t s) {
  CVT_LAUNCH((mse_loss_kernel), 1, 256, 0, s, pred, y, n, metrics, dpred, grad_scale, accumulate);
  return cudaSuccess;
}
cudaError_t launch_set_step_params(float* dyn, float alpha, float gs, uint32_t step, cudaStream_t s) {
  CVT_LAUNCH((set_step_params_kernel), 1, 1, 0, s, dyn, alpha, gs, step);
  return cudaSuccess;
}
cudaError_t launch_adam(float* params, const float* grads, float* m, float* v, long long n, float alpha, float b1, float b2,
                        float eps, float grad_scale, cudaStream_t s, const float* dyn) {
  if (n <= 0) return cudaSuccess;
  if (n % 4 != 0) { set_error("adam: arena length must be a multiple of 4"); return cudaErrorInvalidValue; }
  CVT_LAUNCH((adam_kernel), ew_blocks(n / 4), 256, 0, s, params, grads, m, v, n / 4, alpha, b1, b2, eps, grad_scale, dyn);
  return cudaSuccess;
}

}  // namespace cvt
