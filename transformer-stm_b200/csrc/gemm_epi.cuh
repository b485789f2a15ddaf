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
    uint32_t mk;
    if (E & EPI_PHILOX) {         // (compile-time) Dropout(rate) mask of 4 consecutive elements = one Philox block (counter = element index / 4, site, step)
      uint32_t c[4] = {(uint32_t)(idx0 >> 2), p.ph_site, p.ph_dyn ? __ldg(reinterpret_cast<const uint32_t*>(p.ph_dyn) + 2) : p.ph_step, 0u};
      philox4x32_10(c, p.ph_k0, p.ph_k1);
      mk = (c[0] >= p.ph_thr ? 1u : 0u) | (c[1] >= p.ph_thr ? 0x100u : 0u) | (c[2] >= p.ph_thr ? 0x10000u : 0u) | (c[3] >= p.ph_thr ? 0x1000000u : 0u);
      *reinterpret_cast<uint32_t*>(p.mask_gen + idx0) = mk;         // kept for the backward pass
    } else {
      mk = *reinterpret_cast<const uint32_t*>(p.mask + idx0);
    }
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

// ---- the same epilogue split in two so that a warp can issue the global operand loads of several rows before it consumes any
// (gemm_tma.cu: the loads do not depend on the accumulator, so they overlap the wait for it and each other) ------------------------
struct EpiOps { float4 r4, a4; uint32_t mk; };
template <uint32_t E>
__device__ __forceinline__ EpiOps epilogue_load(const GemmP& p, const float* aux, int m, int n0, bool use_coef) {
  EpiOps o; o.r4 = make_float4(0.f, 0.f, 0.f, 0.f); o.a4 = o.r4; o.mk = 0u;
  const long long crow = p.c_map(m);
  const long long idx0 = crow * p.ldc + n0;
  if ((E & EPI_GELU_BWD) || use_coef) o.a4 = __ldg(reinterpret_cast<const float4*>(aux + idx0));
  if ((p.epi & EPI_MASK) && !(E & EPI_PHILOX)) o.mk = __ldg(reinterpret_cast<const uint32_t*>(p.mask + idx0));
  if (E & EPI_RESID) o.r4 = __ldg(reinterpret_cast<const float4*>(p.resid + crow * p.ldr + n0));
  return o;
}
template <uint32_t E>
__device__ __forceinline__ void epilogue_apply(const GemmP& p, float4 acc, float* Cp, float* aux, float4 b4, float alpha, int m, int n0,
                                               const float4* cf /*[3] or nullptr*/, const EpiOps& o) {
  float v[4] = {acc.x + b4.x, acc.y + b4.y, acc.z + b4.z, acc.w + b4.w};
  const long long crow = p.c_map(m);
  const long long idx0 = crow * p.ldc + n0;
#pragma unroll
  for (int j = 0; j < 4; ++j) v[j] *= alpha;
  if (cf != nullptr && !(p.epi_skip_period && (m % p.epi_skip_period) == 0)) {     // EPI_BN_BWD
    v[0] = fmaf(cf[0].x, v[0], fmaf(cf[1].x, o.a4.x, cf[2].x)); v[1] = fmaf(cf[0].y, v[1], fmaf(cf[1].y, o.a4.y, cf[2].y));
    v[2] = fmaf(cf[0].z, v[2], fmaf(cf[1].z, o.a4.z, cf[2].z)); v[3] = fmaf(cf[0].w, v[3], fmaf(cf[1].w, o.a4.w, cf[2].w));
  }
  if (E & EPI_GELU) {
    if (aux) *reinterpret_cast<float4*>(aux + idx0) = make_float4(v[0], v[1], v[2], v[3]);
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = gelu_erf(v[j]);
  }
  if (E & EPI_GELU_BWD) {
    v[0] *= gelu_erf_grad(o.a4.x); v[1] *= gelu_erf_grad(o.a4.y); v[2] *= gelu_erf_grad(o.a4.z); v[3] *= gelu_erf_grad(o.a4.w);
  }
  if (p.epi & EPI_MASK) {
    uint32_t mk = o.mk;
    if (E & EPI_PHILOX) {
      uint32_t c[4] = {(uint32_t)(idx0 >> 2), p.ph_site, p.ph_dyn ? __ldg(reinterpret_cast<const uint32_t*>(p.ph_dyn) + 2) : p.ph_step, 0u};
      philox4x32_10(c, p.ph_k0, p.ph_k1);
      mk = (c[0] >= p.ph_thr ? 1u : 0u) | (c[1] >= p.ph_thr ? 0x100u : 0u) | (c[2] >= p.ph_thr ? 0x10000u : 0u) | (c[3] >= p.ph_thr ? 0x1000000u : 0u);
      *reinterpret_cast<uint32_t*>(p.mask_gen + idx0) = mk;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = ((mk >> (8 * j)) & 0xFFu) ? v[j] * p.mask_scale : 0.f;
  }
  if (E & EPI_RESID) { v[0] += o.r4.x; v[1] += o.r4.y; v[2] += o.r4.z; v[3] += o.r4.w; }
  if (p.epi & EPI_ROUND) {
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = round_tf32(v[j]);
  }
  *reinterpret_cast<float4*>(Cp + idx0) = make_float4(v[0], v[1], v[2], v[3]);
}

}  // namespace
}  // namespace cvt
