// Fused regression head of create_cvt_model (models/CvT(Par).py:325-350, CvT(Img).py:325-337) and the loss of model.compile
// (:464-466): LayerNormalization(cls token) | [precomputed image feature]  ++  Proc_Dense_1/2 (relu)  ->  concatenate ->
// Final_Dense (linear) -> mean squared error + mean absolute error, and its backward.  Everything here is a few hundred
// kilobytes of fp32 FFMA work per batch, so the design goal is launch count: one kernel forward (one CTA per image), two
// backward (per-image input gradients, then the batch reductions of every head parameter gradient).
#include "layers.cuh"

namespace cvt {
namespace {
constexpr int HT = 256;                                   // threads per CTA

__device__ __forceinline__ float block_sum(float v, float* red) {   // red: 8 floats of shared memory; result broadcast to all threads
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < HT / 32; ++i) t += red[i];
  return t;
}

// ---- forward: grid = B, block = 256.  Shared: feat[F] + p1[PH] ------------------------------------------------------------
__global__ void __launch_bounds__(HT) head_fwd_kernel(const HeadP p) {
  pdl_wait(); pdl_trigger();
  extern __shared__ float sh[];
  float* s_feat = sh; float* s_p1 = sh + p.F; float* s_part = s_p1 + p.PH;     // s_part: [8][PH] partial sums of Proc_Dense_2
  __shared__ float red[8];
  __shared__ int s_last;
  const int b = blockIdx.x, tid = threadIdx.x;
  const int D = p.D, PH = p.PH;
  float* feat = p.feat + (long long)b * p.F;

  if (p.cls_ln) {                                         // LayerNormalization(epsilon)(cls_tokens)  :328
    const float* x = p.x2 + (long long)b * p.x2_bs;
    float a = 0.f;
    for (int i = tid; i < D; i += HT) a += x[i];
    const float mean = block_sum(a, red) / (float)D;
    float v = 0.f;
    for (int i = tid; i < D; i += HT) { const float d = x[i] - mean; v += d * d; }
    const float rstd = rsqrtf(block_sum(v, red) / (float)D + p.eps);
    for (int i = tid; i < D; i += HT) { const float o = (x[i] - mean) * rstd * p.hn_g[i] + p.hn_b[i]; s_feat[i] = o; feat[i] = o; }
    if (tid == 0) { p.h_mean[b] = mean; p.h_rstd[b] = rstd; }
  } else {
    for (int i = tid; i < D; i += HT) s_feat[i] = feat[i];       // LN + token mean done by the caller  :334-337
  }
  if (p.proc_dim > 0) {                                   // Proc_Dense_1 / Proc_Dense_2 (relu)  :343-344
    const float* pr = p.proc + (long long)b * p.proc_dim;
    for (int i = tid; i < PH; i += HT) {
      float a = p.b1[i];
      for (int k = 0; k < p.proc_dim; ++k) a = fmaf(pr[k], p.w1[(long long)k * PH + i], a);
      a = fmaxf(a, 0.f);
      s_p1[i] = a; p.p1[(long long)b * PH + i] = a;
    }
    __syncthreads();
    if ((PH & 3) == 0 && PH <= 4 * HT) {
      // thread = (group of 4 consecutive outputs, slice of the contraction): many independent 16-byte weight loads in flight
      const int groups = PH >> 2, slices = HT / groups > 0 ? min(HT / groups, 8) : 1;
      const int grp = tid % groups, sl = tid / groups;
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      {
        // every thread runs the same number of rounds (idle ones load row 0 and multiply by zero) so that a CTA barrier can stand
        // between the 16 loads of a round and their uses: without it ptxas re-pipelines the batch only four loads deep
        const bool active = sl < slices;
        const int per = (PH + slices - 1) / slices, i0 = active ? sl * per : 0, i1 = active ? min(PH, i0 + per) : 0;
        const float4* w = reinterpret_cast<const float4*>(p.w2) + grp;
        for (int r = 0; r < per; r += 16) {
          const int i = i0 + r;
          float4 wv[16];
#pragma unroll
          for (int u = 0; u < 16; ++u) wv[u] = w[(long long)max(min(i + u, i1 - 1), 0) * groups];      // plain ld.global: ld.global.nc may sink below the barrier
          __syncthreads();
#pragma unroll
          for (int u = 0; u < 16; ++u) {
            const float a = i + u < i1 ? s_p1[i + u] : 0.f;
            acc.x = fmaf(a, wv[u].x, acc.x); acc.y = fmaf(a, wv[u].y, acc.y); acc.z = fmaf(a, wv[u].z, acc.z); acc.w = fmaf(a, wv[u].w, acc.w);
          }
        }
      }
      float4* part = reinterpret_cast<float4*>(s_part);
      if (sl < slices) part[sl * groups + grp] = acc;
      __syncthreads();
      for (int j = tid; j < PH; j += HT) {
        float a = p.b2[j];
        for (int q = 0; q < slices; ++q) a += s_part[q * PH + j];
        const float o = fmaxf(a, 0.f);
        s_feat[D + j] = o; feat[D + j] = o;               // concatenate([img, proc])  :347
      }
    } else {
      for (int j = tid; j < PH; j += HT) {
        float a = p.b2[j];
        for (int i = 0; i < PH; ++i) a = fmaf(s_p1[i], p.w2[(long long)i * PH + j], a);
        const float o = fmaxf(a, 0.f);
        s_feat[D + j] = o; feat[D + j] = o;
      }
    }
  }
  __syncthreads();
  // Final_Dense (linear)  :350, then the per-image terms of the loss
  float sq = 0.f, ab = 0.f;
  const float inv_n = 1.f / (float)(p.B * p.nc);
  for (int c = 0; c < p.nc; ++c) {
    float a = 0.f;
    for (int f = tid; f < p.F; f += HT) a = fmaf(s_feat[f], p.fw[(long long)f * p.nc + c], a);
    const float o = block_sum(a, red) + p.fb[c];
    if (tid == 0) {
      p.pred[(long long)b * p.nc + c] = o;
      if (p.y) {
        const float d = o - p.y[(long long)b * p.nc + c];
        sq += d * d; ab += fabsf(d);
        p.dpred[(long long)b * p.nc + c] = 2.f * d * inv_n;
      }
    }
  }
  if (!p.y) return;
  // loss = mean over (batch, classes): the last CTA to finish adds the per-image terms in image order (deterministic)
  if (tid == 0) {
    p.partial[2 * b] = sq; p.partial[2 * b + 1] = ab;
    __threadfence();
    s_last = atomicAdd(p.ticket, 1u) == (unsigned)(p.B - 1);
  }
  __syncthreads();
  if (s_last && tid == 0) {
    __threadfence();
    float ts = 0.f, ta = 0.f;
    for (int i = 0; i < p.B; ++i) { ts += __ldcg(p.partial + 2 * i); ta += __ldcg(p.partial + 2 * i + 1); }
    p.metrics[0] = ts * inv_n; p.metrics[1] = ta * inv_n;
    *p.ticket = 0u;
  }
}

// ---- backward A: per-image input gradients.  grid = B, block = 256.  Shared: dfeat[F] + dz2[PH] -------------------------------
__global__ void __launch_bounds__(HT) head_bwd_image_kernel(const HeadP p) {
  pdl_wait(); pdl_trigger();
  extern __shared__ float sh[];
  float* s_dfeat = sh; float* s_dz2 = sh + p.F;
  __shared__ float red[8];
  const int b = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int D = p.D, PH = p.PH, nc = p.nc;
  const float* dpred = p.dpred + (long long)b * nc;
  for (int f = tid; f < p.F; f += HT) {                  // d feat = d pred . Final_Dense^T
    float a = 0.f;
    for (int c = 0; c < nc; ++c) a = fmaf(dpred[c], p.fw[(long long)f * nc + c], a);
    s_dfeat[f] = a;
    p.dfeat[(long long)b * p.F + f] = a;
  }
  __syncthreads();
  if (p.proc_dim > 0) {
    const float* feat = p.feat + (long long)b * p.F + D;
    for (int j = tid; j < PH; j += HT) {                 // through relu of Proc_Dense_2
      const float g = feat[j] > 0.f ? s_dfeat[D + j] : 0.f;
      s_dz2[j] = g; p.dz2[(long long)b * PH + j] = g;
    }
    __syncthreads();
    // d p1 = dz2 . W2^T, through relu of Proc_Dense_1: a warp owns rows of W2
    if ((PH & 127) == 0 && (p.F & 3) == 0) {
      // eight rows at a time, every lane holding two 16-byte pieces of each: 16 independent loads in flight before the arithmetic
      const int nv = PH >> 7, per = PH >> 3;
      const float4* dz4 = reinterpret_cast<const float4*>(s_dz2);
      for (int r0 = warp * per; r0 < (warp + 1) * per; r0 += 8) {
        float a[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) a[r] = 0.f;
        for (int v0 = 0; v0 < nv; v0 += 2) {
          float4 wv[8][2];
#pragma unroll
          for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int v = 0; v < 2; ++v)
              wv[r][v] = (r0 + r < (warp + 1) * per && v0 + v < nv) ? __ldg(reinterpret_cast<const float4*>(p.w2 + (long long)(r0 + r) * PH) + (v0 + v) * 32 + lane)
                                                                    : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int v = 0; v < 2; ++v) {
            const float4 d = v0 + v < nv ? dz4[(v0 + v) * 32 + lane] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int r = 0; r < 8; ++r) a[r] = fmaf(d.x, wv[r][v].x, fmaf(d.y, wv[r][v].y, fmaf(d.z, wv[r][v].z, fmaf(d.w, wv[r][v].w, a[r]))));
          }
        }
        const int i = r0 + (lane & 7);
        const float p1v = i < (warp + 1) * per ? p.p1[(long long)b * PH + i] : 0.f;
        float mine = 0.f;
#pragma unroll
        for (int r = 0; r < 8; ++r) { const float t = warp_sum(a[r]); if ((lane & 7) == r) mine = t; }
        if (lane < 8 && i < (warp + 1) * per) p.dz1[(long long)b * PH + i] = p1v > 0.f ? mine : 0.f;
      }
    } else {
      for (int i = warp; i < PH; i += HT / 32) {
        const float* w = p.w2 + (long long)i * PH;
        float a = 0.f;
        for (int j = lane; j < PH; j += 32) a = fmaf(s_dz2[j], w[j], a);
        a = warp_sum(a);
        if (lane == 0) p.dz1[(long long)b * PH + i] = p.p1[(long long)b * PH + i] > 0.f ? a : 0.f;
      }
    }
  }
  if (p.cls_ln) {                                         // LayerNormalization backward on the cls row; every other row of dX2 is zero
    const float* x = p.x2 + (long long)b * p.x2_bs;
    float* dx = p.dx2 + (long long)b * p.x2_bs;
    const float mean = p.h_mean[b], rstd = p.h_rstd[b];
    float s1 = 0.f, s2 = 0.f;
    for (int i = tid; i < D; i += HT) { const float g = s_dfeat[i] * p.hn_g[i]; s1 += g; s2 += g * (x[i] - mean) * rstd; }
    s1 = block_sum(s1, red) / (float)D;
    s2 = block_sum(s2, red) / (float)D;
    for (int i = tid; i < D; i += HT) {
      const float xh = (x[i] - mean) * rstd;
      dx[i] = rstd * (s_dfeat[i] * p.hn_g[i] - s1 - xh * s2);
    }
  }
}

// ---- backward B: parameter gradients = reductions over the batch.  block = (32, 8) ------------------------------------------------
// blocks [0, tiles2): 32 x 32 tiles of dW2 = p1^T dz2;  then groups of blocks that each own 32 entries, thread (tx, ty) summing the
// images b = ty, ty + 8, ... of entry tx and the eight partial sums meeting in shared memory:
//   n0 blocks {db2, db1, dW1} (32 hidden units each) | n1 blocks dW_final (+ db_final in the first) | n2 blocks LayerNorm d gamma, d beta
// Every loop issues its loads for a batch of images before the arithmetic: the work is tiny, the cost is exposed memory latency.
__global__ void __launch_bounds__(HT) head_bwd_param_kernel(const HeadP p, int tiles2, int tn, int n0, int n1) {
  pdl_wait(); pdl_trigger();
  __shared__ float red[8][10][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5, tid = threadIdx.x;
  const int B = p.B, PH = p.PH, D = p.D, nc = p.nc;
  int job = blockIdx.x;
  if (job < tiles2) {
    const int i0 = (job / tn) * 32, j = (job % tn) * 32 + tx;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int b0 = 0; b0 < B; b0 += 16) {
      float g[16], pv[16][4];
#pragma unroll
      for (int u = 0; u < 16; ++u) {
        const int b = b0 + u;
        g[u] = (b < B && j < PH) ? p.dz2[(long long)b * PH + j] : 0.f;
#pragma unroll
        for (int r = 0; r < 4; ++r) { const int i = i0 + ty + 8 * r; pv[u][r] = (b < B && i < PH) ? p.p1[(long long)b * PH + i] : 0.f; }
      }
#pragma unroll
      for (int u = 0; u < 16; ++u)
#pragma unroll
        for (int r = 0; r < 4; ++r) acc[r] = fmaf(pv[u][r], g[u], acc[r]);
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) { const int i = i0 + ty + 8 * r; if (i < PH && j < PH) p.g_w2[(long long)i * PH + j] = acc[r]; }
    return;
  }
  job -= tiles2;
  if (job < n0) {                                        // db2, db1 and the first 8 rows of dW1 for hidden unit i
    const int i = job * 32 + tx, pd = p.proc_dim;
    float v[10];
#pragma unroll
    for (int k = 0; k < 10; ++k) v[k] = 0.f;
    for (int b0 = ty; b0 < B; b0 += 64) {
      float g2[8], g1[8], pr[8][8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int b = b0 + 8 * u; const bool ok = b < B && i < PH;
        g2[u] = ok ? p.dz2[(long long)b * PH + i] : 0.f;
        g1[u] = ok ? p.dz1[(long long)b * PH + i] : 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) pr[u][k] = (b < B && k < pd) ? p.proc[(long long)b * pd + k] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        v[0] += g2[u]; v[1] += g1[u];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[2 + k] = fmaf(pr[u][k], g1[u], v[2 + k]);
      }
    }
#pragma unroll
    for (int k = 0; k < 10; ++k) red[ty][k][tx] = v[k];
    __syncthreads();
    for (int k = ty; k < 10; k += 8) {
      float a = 0.f;
#pragma unroll
      for (int q = 0; q < 8; ++q) a += red[q][k][tx];
      if (i < PH) {
        if (k == 0) p.g_b2[i] = a;
        else if (k == 1) p.g_b1[i] = a;
        else if (k - 2 < pd) p.g_w1[(long long)(k - 2) * PH + i] = a;
      }
    }
    for (int k = 8 + ty; k < pd; k += 8) {               // rare: more than 8 process parameters
      float a = 0.f;
      if (i < PH) { for (int b = 0; b < B; ++b) a = fmaf(p.proc[(long long)b * pd + k], p.dz1[(long long)b * PH + i], a); p.g_w1[(long long)k * PH + i] = a; }
    }
    return;
  }
  job -= n0;
  if (job < n1) {                                        // dW_final[f][c] = sum_b feat[b][f] dpred[b][c]
    const int e = job * 32 + tx; const bool eok = e < p.F * nc;
    const int f = eok ? e / nc : 0, c = eok ? e - f * nc : 0;
    float a = 0.f;
    for (int b0 = ty; b0 < B; b0 += 64) {
      float x[8], g[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int b = b0 + 8 * u; const bool ok = eok && b < B;
        x[u] = ok ? p.feat[(long long)b * p.F + f] : 0.f; g[u] = ok ? p.dpred[(long long)b * nc + c] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) a = fmaf(x[u], g[u], a);
    }
    red[ty][0][tx] = a;
    __syncthreads();
    if (ty == 0 && eok) {
      float t = 0.f;
#pragma unroll
      for (int q = 0; q < 8; ++q) t += red[q][0][tx];
      p.g_fw[e] = t;
    }
    if (job == 0) {
      for (int c2 = 0; c2 < nc; ++c2) {                  // db_final[c] = sum_b dpred[b][c]
        float t = 0.f;
        for (int b = tid; b < B; b += HT) t += p.dpred[(long long)b * nc + c2];
        t = warp_sum(t);
        __syncthreads();
        if (tx == 0) red[ty][1][0] = t;
        __syncthreads();
        if (tid == 0) { float q = 0.f; for (int k = 0; k < 8; ++k) q += red[k][1][0]; p.g_fb[c2] = q; }
      }
    }
    return;
  }
  job -= n1;
  if (p.cls_ln) {                                        // LayerNormalization(cls): d gamma[i] = sum_b dfeat xhat, d beta[i] = sum_b dfeat
    const int i = job * 32 + tx;
    float dg = 0.f, db = 0.f;
    for (int b0 = ty; b0 < B; b0 += 64) {
      float g[8], x[8], mu[8], rs[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int b = b0 + 8 * u; const bool ok = b < B && i < D;
        g[u] = ok ? p.dfeat[(long long)b * p.F + i] : 0.f; x[u] = ok ? p.x2[(long long)b * p.x2_bs + i] : 0.f;
        mu[u] = ok ? p.h_mean[b] : 0.f; rs[u] = ok ? p.h_rstd[b] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) { dg = fmaf(g[u], (x[u] - mu[u]) * rs[u], dg); db += g[u]; }
    }
    red[ty][0][tx] = dg; red[ty][1][tx] = db;
    __syncthreads();
    if (ty < 2 && i < D) {
      float t = 0.f;
#pragma unroll
      for (int q = 0; q < 8; ++q) t += red[q][ty][tx];
      (ty == 0 ? p.g_hn_g : p.g_hn_b)[i] = t;
    }
  }
}
}  // namespace

cudaError_t launch_head_fwd(const HeadP& p, cudaStream_t s) {
  const size_t smem = (size_t)(p.F + 9 * p.PH) * sizeof(float);
  if (smem > 48 * 1024) { set_error("head: feature + hidden width too large for the fused head kernel"); return cudaErrorInvalidValue; }
  CVT_LAUNCH((head_fwd_kernel), p.B, HT, smem, s, p);
  return cudaSuccess;
}

cudaError_t launch_head_bwd(const HeadP& p, cudaStream_t s) {
  const size_t smem = (size_t)(p.F + p.PH) * sizeof(float);
  if (smem > 48 * 1024) { set_error("head: feature + hidden width too large for the fused head kernel"); return cudaErrorInvalidValue; }
  CVT_LAUNCH((head_bwd_image_kernel), p.B, HT, smem, s, p);
  const int tn = p.proc_dim > 0 ? (p.PH + 31) / 32 : 0, tiles2 = tn * tn;
  const int n0 = tn, n1 = (p.F * p.nc + 31) / 32, n2 = p.cls_ln ? (p.D + 31) / 32 : 0;
  CVT_LAUNCH((head_bwd_param_kernel), tiles2 + n0 + n1 + n2, HT, 0, s, p, tiles2, tn, n0, n1);
  return cudaSuccess;
}

}  // namespace cvt
