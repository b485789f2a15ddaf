// Coalesced-domain epilogue shared by the tcgen05 contraction kernels (gemm_tc.cu, gemm_tma.cu).
#pragma once
#include "gemm.cuh"

namespace cvt {
namespace {
// ---- epilogue of 4 consecutive output columns of one row (coalesced domain: a quarter-warp covers 128 contiguous bytes) ----------
// E = the compile-time epilogue kind (GELU | RESID | GELU_BWD | SCATTER | ATOMIC); bias / alpha / dropout mask stay run-time
template <uint32_t E>
__device__ __forceinline__ void epilogue_f4(const GemmP& p, float4 acc, float* Cp, float* aux, const float* bias, float alpha, int m, int n0,
                                            const float* coef = nullptr) {
  float v[4] = {acc.x, acc.y, acc.z, acc.w};
  if (E & EPI_SCATTER) {
    // C is the conv input gradient [B,hin,win,cin]; m = output pixel, n = (tap, c)
    const ConvGeom& g = p.cv;
    const int hw = g.ho * g.wo; const int b = m / hw; const int rr = m - b * hw; const int oi = rr / g.wo; const int oj = rr - oi * g.wo;
    float* img = Cp + (long long)b * g.bstride + g.boff;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + j;
      const int tap = n / g.cin, c = n - tap * g.cin; const int kh = tap / g.k, kw = tap - kh * g.k;
      const int yy = g.s * oi + kh - g.pt, xx = g.s * oj + kw - g.pl;
      if (yy >= 0 && yy < g.hin && xx >= 0 && xx < g.win) atomicAdd(img + ((long long)yy * g.win + xx) * g.cin + c, v[j] * alpha);
    }
    return;
  }
  const long long crow = p.c_map(m);
  const long long idx0 = crow * p.ldc + n0;
  if (bias) {
    const float4 b4 = *reinterpret_cast<const float4*>(bias + n0);
    v[0] += b4.x; v[1] += b4.y; v[2] += b4.z; v[3] += b4.w;
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) v[j] *= alpha;
  if (coef != nullptr && !(p.epi_skip_period && (m % p.epi_skip_period) == 0)) {     // EPI_BN_BWD
    const float4 a4 = *reinterpret_cast<const float4*>(coef + n0), e4 = *reinterpret_cast<const float4*>(coef + p.N + n0);
    const float4 f4 = *reinterpret_cast<const float4*>(coef + 2 * p.N + n0), c4 = *reinterpret_cast<const float4*>(aux + idx0);
    v[0] = fmaf(a4.x, v[0], fmaf(e4.x, c4.x, f4.x)); v[1] = fmaf(a4.y, v[1], fmaf(e4.y, c4.y, f4.y));
    v[2] = fmaf(a4.z, v[2], fmaf(e4.z, c4.z, f4.z)); v[3] = fmaf(a4.w, v[3], fmaf(e4.w, c4.w, f4.w));
  }
  if (E & EPI_GELU) {
    if (aux) *reinterpret_cast<float4*>(aux + idx0) = make_float4(v[0], v[1], v[2], v[3]);
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = gelu_erf(v[j]);
  }
  if (E & EPI_GELU_BWD) {
    const float4 a4 = *reinterpret_cast<const float4*>(aux + idx0);
    v[0] *= gelu_erf_grad(a4.x); v[1] *= gelu_erf_grad(a4.y); v[2] *= gelu_erf_grad(a4.z); v[3] *= gelu_erf_grad(a4.w);
  }
  if (p.epi & EPI_MASK) {
    const uint32_t mk = *reinterpret_cast<const uint32_t*>(p.mask + idx0);
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = ((mk >> (8 * j)) & 0xFFu) ? v[j] * p.mask_scale : 0.f;
  }
  if (E & EPI_RESID) {
    const float4 r4 = *reinterpret_cast<const float4*>(p.resid + crow * p.ldr + n0);
    v[0] += r4.x; v[1] += r4.y; v[2] += r4.z; v[3] += r4.w;
  }
  if (p.epi & EPI_ROUND) {
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = round_tf32(v[j]);
  }
  if (E & EPI_ATOMIC) {
#pragma unroll
    for (int j = 0; j < 4; ++j) atomicAdd(Cp + idx0 + j, v[j]);
  } else {
    *reinterpret_cast<float4*>(Cp + idx0) = make_float4(v[0], v[1], v[2], v[3]);
  }
}

}  // namespace
}  // namespace cvt
