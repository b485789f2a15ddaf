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
      if (sl < slices) {
        const int per = (PH + slices - 1) / slices, i0 = sl * per, i1 = min(PH, i0 + per);
        const float4* w = reinterpret_cast<const float4*>(p.w2) + grp;
#pragma unroll 8
        for (int i = i0; i < i1; ++i) {
          const float4 wv = __ldg(w + (long long)i * groups);
          const float a = s_p1[i];
          acc.x = fmaf(a, wv.x, acc.x); acc.y = fmaf(a, wv.y, acc.y); acc.z = fmaf(a, wv.z, acc.z); acc.w = fmaf(a, wv.w, acc.w);
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
#pragma unroll 4
    for (int i = warp; i < PH; i += HT / 32) {           // d p1 = dz2 . W2^T, through relu of Proc_Dense_1 (one warp per row of W2)
      const float* w = p.w2 + (long long)i * PH;
      float a = 0.f;
      for (int j = lane; j < PH; j += 32) a = fmaf(s_dz2[j], w[j], a);
      a = warp_sum(a);
      if (lane == 0) p.dz1[(long long)b * PH + i] = p.p1[(long long)b * PH + i] > 0.f ? a : 0.f;
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
// blocks [0, tiles2): 32 x 32 tiles of dW2 = p1^T dz2;  then one block each for {db2, db1, dW1}, {dW_final, db_final}, {d gamma, d beta}
__global__ void __launch_bounds__(HT) head_bwd_param_kernel(const HeadP p, int tiles2, int tn) {
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5, tid = threadIdx.x;
  const int B = p.B, PH = p.PH, D = p.D, nc = p.nc;
  int job = blockIdx.x;
  if (job < tiles2) {
    const int i0 = (job / tn) * 32, j = (job % tn) * 32 + tx;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 8
    for (int b = 0; b < B; ++b) {
      const float g = j < PH ? p.dz2[(long long)b * PH + j] : 0.f;
#pragma unroll
      for (int r = 0; r < 4; ++r) { const int i = i0 + ty + 8 * r; if (i < PH) acc[r] = fmaf(p.p1[(long long)b * PH + i], g, acc[r]); }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) { const int i = i0 + ty + 8 * r; if (i < PH && j < PH) p.g_w2[(long long)i * PH + j] = acc[r]; }
    return;
  }
  job -= tiles2;
  if (p.proc_dim == 0) job += 1;                         // Img variant: no process-parameter branch
  if (job == 0) {
    for (int i = tid; i < PH; i += HT) {
      float s2 = 0.f, s1 = 0.f;
      float w[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) w[k] = 0.f;
#pragma unroll 4
      for (int b = 0; b < B; ++b) {
        s2 += p.dz2[(long long)b * PH + i];
        const float g = p.dz1[(long long)b * PH + i];
        s1 += g;
#pragma unroll
        for (int k = 0; k < 8; ++k) if (k < p.proc_dim) w[k] = fmaf(p.proc[(long long)b * p.proc_dim + k], g, w[k]);
      }
      p.g_b2[i] = s2; p.g_b1[i] = s1;
      for (int k = 0; k < p.proc_dim && k < 8; ++k) p.g_w1[(long long)k * PH + i] = w[k];
      for (int k = 8; k < p.proc_dim; ++k) {               // rare: more than 8 process parameters
        float a = 0.f;
        for (int b = 0; b < B; ++b) a = fmaf(p.proc[(long long)b * p.proc_dim + k], p.dz1[(long long)b * PH + i], a);
        p.g_w1[(long long)k * PH + i] = a;
      }
    }
  } else if (job == 1) {
    for (int e = tid; e < p.F * nc; e += HT) {
      const int f = e / nc, c = e - f * nc;
      float a = 0.f;
#pragma unroll 8
      for (int b = 0; b < B; ++b) a = fmaf(p.feat[(long long)b * p.F + f], p.dpred[(long long)b * nc + c], a);
      p.g_fw[e] = a;
    }
    for (int c = tid; c < nc; c += HT) {
      float a = 0.f;
      for (int b = 0; b < B; ++b) a += p.dpred[(long long)b * nc + c];
      p.g_fb[c] = a;
    }
  } else if (job == 2 && p.cls_ln) {
    for (int i = tid; i < D; i += HT) {
      float dg = 0.f, db = 0.f;
#pragma unroll 8
      for (int b = 0; b < B; ++b) {
        const float g = p.dfeat[(long long)b * p.F + i];
        dg = fmaf(g, (p.x2[(long long)b * p.x2_bs + i] - p.h_mean[b]) * p.h_rstd[b], dg);
        db += g;
      }
      p.g_hn_g[i] = dg; p.g_hn_b[i] = db;
    }
  }
}
}  // namespace

cudaError_t launch_head_fwd(const HeadP& p, cudaStream_t s) {
  const size_t smem = (size_t)(p.F + 9 * p.PH) * sizeof(float);
  if (smem > 48 * 1024) { set_error("head: feature + hidden width too large for the fused head kernel"); return cudaErrorInvalidValue; }
  head_fwd_kernel<<<p.B, HT, smem, s>>>(p);
  CVT_LAUNCH_CHECK();
  return cudaSuccess;
}

cudaError_t launch_head_bwd(const HeadP& p, cudaStream_t s) {
  const size_t smem = (size_t)(p.F + p.PH) * sizeof(float);
  if (smem > 48 * 1024) { set_error("head: feature + hidden width too large for the fused head kernel"); return cudaErrorInvalidValue; }
  head_bwd_image_kernel<<<p.B, HT, smem, s>>>(p);
  CVT_LAUNCH_CHECK();
  const int tn = p.proc_dim > 0 ? (p.PH + 31) / 32 : 0, tiles2 = tn * tn;
  const int misc = (p.proc_dim > 0 ? 1 : 0) + 1 + (p.cls_ln ? 1 : 0);
  head_bwd_param_kernel<<<tiles2 + misc, HT, 0, s>>>(p, tiles2, tn);
  CVT_LAUNCH_CHECK();
  return cudaSuccess;
}

}  // namespace cvt
