// Vectorised LayerNorm forward / backward (LayerNormalization(epsilon) over the last axis, models/CvT(Par).py:248,272,278,328).
// HBM-bound: a row of D floats is spread over LPR = min(32, D/16)... lanes as float4s, 32/LPR rows per warp pass and U passes in
// flight per warp, so every warp keeps several KB of loads outstanding.  gamma/beta gradients are accumulated in registers over
// all the rows a lane sees, reduced once per CTA through shared memory and added to global memory with one atomic per channel.
#include "layers.cuh"

namespace cvt {
namespace {

__device__ __forceinline__ float group_sum(float v, int lpr) {
  for (int o = lpr >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// VPL float4s per lane, LPR lanes per row (LPR * VPL * 4 == D)
template <int VPL, int LPR, int U>
__global__ void __launch_bounds__(256) ln_fwd_v4_kernel(const float* __restrict__ x, RowMap in_map, long long ldx,
                                                        const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                                                        float* __restrict__ y, long long ldy, float* __restrict__ mean,
                                                        float* __restrict__ rstd, int rows, int round_out) {
  pdl_wait(); pdl_trigger();
  constexpr int RPW = 32 / LPR, D = LPR * VPL * 4;
  const int lane = threadIdx.x & 31, sub = lane / LPR, li = lane % LPR;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
  float4 g[VPL], b[VPL];
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    g[v] = *reinterpret_cast<const float4*>(gamma + (li + v * LPR) * 4);
    b[v] = *reinterpret_cast<const float4*>(beta + (li + v * LPR) * 4);
  }
  const float invD = 1.f / (float)D;
  for (int base = warp * RPW * U; base < rows; base += nwarps * RPW * U) {
    float4 xv[U][VPL];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int row = base + u * RPW + sub;
      if (row < rows) {
        const float* xr = x + in_map(row) * ldx;
#pragma unroll
        for (int v = 0; v < VPL; ++v) xv[u][v] = *reinterpret_cast<const float4*>(xr + (li + v * LPR) * 4);
      } else {
#pragma unroll
        for (int v = 0; v < VPL; ++v) xv[u][v] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int row = base + u * RPW + sub;
      float s = 0.f;
#pragma unroll
      for (int v = 0; v < VPL; ++v) s += (xv[u][v].x + xv[u][v].y) + (xv[u][v].z + xv[u][v].w);
      const float mu = group_sum(s, LPR) * invD;
      float q = 0.f;
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        const float a = xv[u][v].x - mu, bb = xv[u][v].y - mu, c = xv[u][v].z - mu, d = xv[u][v].w - mu;
        q += (a * a + bb * bb) + (c * c + d * d);
      }
      const float r = 1.0f / sqrtf(group_sum(q, LPR) * invD + eps);
      if (row < rows) {
        float* yr = y + (long long)row * ldy;
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
          float4 o;
          o.x = (xv[u][v].x - mu) * r * g[v].x + b[v].x; o.y = (xv[u][v].y - mu) * r * g[v].y + b[v].y;
          o.z = (xv[u][v].z - mu) * r * g[v].z + b[v].z; o.w = (xv[u][v].w - mu) * r * g[v].w + b[v].w;
          if (round_out) o = make_float4(round_tf32(o.x), round_tf32(o.y), round_tf32(o.z), round_tf32(o.w));
          *reinterpret_cast<float4*>(yr + (li + v * LPR) * 4) = o;
        }
        if (li == 0) { if (mean) mean[row] = mu; if (rstd) rstd[row] = r; }
      }
    }
  }
}

template <int VPL, int LPR, int U>
__global__ void __launch_bounds__(256) ln_bwd_v4_kernel(const float* __restrict__ dy, long long lddy, const float* __restrict__ x,
                                                        RowMap in_map, long long ldx, const float* __restrict__ mean,
                                                        const float* __restrict__ rstd, const float* __restrict__ gamma,
                                                        const float* dres, float* dx, RowMap out_map, long long ldo,   // may alias
                                                        float* __restrict__ dgamma, float* __restrict__ dbeta, int rows,
                                                        const uint8_t* __restrict__ mask, float mscale, float* __restrict__ dxm, int round_m) {
  pdl_wait(); pdl_trigger();
  // mask != nullptr: a second output dxm = dx * (mask ? mscale : 0) (the gradient through the dropout that follows in backward order;
  // mask and dxm are indexed like dx), optionally rounded to TF32 for the tensor-core GEMMs that consume it
  constexpr int RPW = 32 / LPR, D = LPR * VPL * 4;
  __shared__ float red[2 * D];
  for (int i = threadIdx.x; i < 2 * D; i += blockDim.x) red[i] = 0.f;
  __syncthreads();
  const int lane = threadIdx.x & 31, sub = lane / LPR, li = lane % LPR;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
  float4 g[VPL], ag[VPL], ab[VPL];
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    g[v] = *reinterpret_cast<const float4*>(gamma + (li + v * LPR) * 4);
    ag[v] = make_float4(0.f, 0.f, 0.f, 0.f); ab[v] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const float invD = 1.f / (float)D;
  for (int base = warp * RPW * U; base < rows; base += nwarps * RPW * U) {
    float4 xv[U][VPL], dv[U][VPL], rv[U][VPL];
    uint32_t mk[U][VPL];
    float mu[U], rs[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int row = base + u * RPW + sub;
      const bool ok = row < rows;
      const float* xr = x + (ok ? in_map(row) : 0) * ldx;
      const float* dr = dy + (long long)(ok ? row : 0) * lddy;
      const float* rr = dres ? dres + (ok ? out_map(row) : 0) * ldo : nullptr;
      mu[u] = ok ? mean[row] : 0.f; rs[u] = ok ? rstd[row] : 0.f;
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        const int c = (li + v * LPR) * 4;
        xv[u][v] = ok ? *reinterpret_cast<const float4*>(xr + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        dv[u][v] = ok ? *reinterpret_cast<const float4*>(dr + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        rv[u][v] = (ok && rr) ? *reinterpret_cast<const float4*>(rr + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        mk[u][v] = (ok && mask) ? *reinterpret_cast<const uint32_t*>(mask + out_map(row) * ldo + c) : 0u;
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int row = base + u * RPW + sub;
      float s1 = 0.f, s2 = 0.f;
      float4 xh[VPL], gg[VPL];
#pragma unroll
      for (int v = 0; v < VPL; ++v) {
        xh[v] = make_float4((xv[u][v].x - mu[u]) * rs[u], (xv[u][v].y - mu[u]) * rs[u], (xv[u][v].z - mu[u]) * rs[u], (xv[u][v].w - mu[u]) * rs[u]);
        gg[v] = make_float4(dv[u][v].x * g[v].x, dv[u][v].y * g[v].y, dv[u][v].z * g[v].z, dv[u][v].w * g[v].w);
        s1 += (gg[v].x + gg[v].y) + (gg[v].z + gg[v].w);
        s2 += (gg[v].x * xh[v].x + gg[v].y * xh[v].y) + (gg[v].z * xh[v].z + gg[v].w * xh[v].w);
        ag[v].x += dv[u][v].x * xh[v].x; ag[v].y += dv[u][v].y * xh[v].y; ag[v].z += dv[u][v].z * xh[v].z; ag[v].w += dv[u][v].w * xh[v].w;
        ab[v].x += dv[u][v].x; ab[v].y += dv[u][v].y; ab[v].z += dv[u][v].z; ab[v].w += dv[u][v].w;
      }
      const float m1 = group_sum(s1, LPR) * invD, m2 = group_sum(s2, LPR) * invD;
      if (row < rows) {
        float* dxr = dx + out_map(row) * ldo;
#pragma unroll
        for (int v = 0; v < VPL; ++v) {
          float4 o;
          o.x = rv[u][v].x + rs[u] * (gg[v].x - m1 - xh[v].x * m2); o.y = rv[u][v].y + rs[u] * (gg[v].y - m1 - xh[v].y * m2);
          o.z = rv[u][v].z + rs[u] * (gg[v].z - m1 - xh[v].z * m2); o.w = rv[u][v].w + rs[u] * (gg[v].w - m1 - xh[v].w * m2);
          *reinterpret_cast<float4*>(dxr + (li + v * LPR) * 4) = o;
          if (mask) {
            float4 q;
            q.x = (mk[u][v] & 0xFFu) ? o.x * mscale : 0.f; q.y = (mk[u][v] & 0xFF00u) ? o.y * mscale : 0.f;
            q.z = (mk[u][v] & 0xFF0000u) ? o.z * mscale : 0.f; q.w = (mk[u][v] & 0xFF000000u) ? o.w * mscale : 0.f;
            if (round_m) { q.x = round_tf32(q.x); q.y = round_tf32(q.y); q.z = round_tf32(q.z); q.w = round_tf32(q.w); }
            *reinterpret_cast<float4*>(dxm + out_map(row) * ldo + (li + v * LPR) * 4) = q;
          }
        }
      }
    }
  }
  // lanes holding the same channels (different `sub`) fold through shuffles, then one shared-memory atomic per warp and channel
#pragma unroll
  for (int v = 0; v < VPL; ++v) {
    float a[8] = {ag[v].x, ag[v].y, ag[v].z, ag[v].w, ab[v].x, ab[v].y, ab[v].z, ab[v].w};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      for (int o = LPR; o < 32; o <<= 1) a[j] += __shfl_xor_sync(0xffffffffu, a[j], o);
    }
    if (sub == 0) {
      const int c = (li + v * LPR) * 4;
#pragma unroll
      for (int j = 0; j < 4; ++j) { atomicAdd(&red[c + j], a[j]); atomicAdd(&red[D + c + j], a[4 + j]); }
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < D; c += blockDim.x) { atomicAdd(dgamma + c, red[c]); atomicAdd(dbeta + c, red[D + c]); }
}

inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace

// returns cudaErrorNotSupported when the shape is not one of the vectorised instantiations (the caller falls back)
cudaError_t launch_ln_fwd_v4(const float* x, RowMap in_map, long long ldx, const float* gamma, const float* beta, float eps,
                             float* y, long long ldy, float* mean, float* rstd, int rows, int D, cudaStream_t s, int round_out) {
  if ((ldx & 3) || (ldy & 3) || !al16(x) || !al16(y) || !al16(gamma) || !al16(beta)) return cudaErrorNotSupported;
  auto grid = [&](int rpw_u) { long long w = cdivl(rows, rpw_u); long long b = cdivl(w, 8); return (int)(b < 1 ? 1 : (b > kNumSMs * 8 ? kNumSMs * 8 : b)); };
#define LN_FWD(VPL, LPR, U) CVT_LAUNCH((ln_fwd_v4_kernel<VPL, LPR, U>), grid((32 / LPR) * U), 256, 0, s, x, in_map, ldx, gamma, beta, eps, y, ldy, mean, rstd, rows, round_out)
  if (D == 64) LN_FWD(1, 16, 4);
  else if (D == 128) LN_FWD(1, 32, 4);
  else if (D == 192) LN_FWD(3, 16, 2);
  else if (D == 256) LN_FWD(2, 32, 4);
  else if (D == 384) LN_FWD(3, 32, 2);
  else if (D == 512) LN_FWD(4, 32, 2);
  else return cudaErrorNotSupported;
#undef LN_FWD
  return cudaSuccess;
}

cudaError_t launch_ln_bwd_v4(const float* dy, long long lddy, const float* x, RowMap in_map, long long ldx, const float* mean,
                             const float* rstd, const float* gamma, const float* dres, float* dx, RowMap out_map, long long ldo,
                             float* dgamma, float* dbeta, int rows, int D, cudaStream_t s, const uint8_t* mask, float mscale, float* dxm, int round_m) {
  if ((ldx & 3) || (lddy & 3) || (ldo & 3) || !al16(x) || !al16(dy) || !al16(dx) || !al16(gamma) || (dres && !al16(dres))) return cudaErrorNotSupported;
  if (mask && ((reinterpret_cast<uintptr_t>(mask) & 3) || !al16(dxm))) return cudaErrorNotSupported;
  auto grid = [&](int rpw_u) { long long w = cdivl(rows, rpw_u); long long b = cdivl(w, 8); return (int)(b < 1 ? 1 : (b > kNumSMs * 4 ? kNumSMs * 4 : b)); };
#define LN_BWD(VPL, LPR, U) CVT_LAUNCH((ln_bwd_v4_kernel<VPL, LPR, U>), grid((32 / LPR) * U), 256, 0, s, dy, lddy, x, in_map, ldx, mean, rstd, gamma, dres, dx, out_map, ldo, dgamma, dbeta, rows, mask, mscale, dxm, round_m)
  if (D == 64) LN_BWD(1, 16, 4);
  else if (D == 128) LN_BWD(1, 32, 4);
  else if (D == 192) LN_BWD(3, 16, 2);
  else if (D == 256) LN_BWD(2, 32, 2);
  else if (D == 384) LN_BWD(3, 32, 1);
  else if (D == 512) LN_BWD(4, 32, 1);
  else return cudaErrorNotSupported;
#undef LN_BWD
  return cudaSuccess;
}

}  // namespace cvt
