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
struct EpiCoef { float4 a, e, f; };          // EPI_BN_BWD coefficients of 4 columns (registers, never indexed)
template <uint32_t E>
__device__ __forceinline__ EpiOps epilogue_load(const GemmP& p, const float* aux, int m, int n0, bool use_coef) {
  EpiOps o; o.r4 = make_float4(0.f, 0.f, 0.f, 0.f); o.a4 = o.r4; o.mk = 0u;
  const long long crow = p.c_map(m);
  const long long idx0 = crow * p.ldc + n0;
  if ((E & EPI_GELU_BWD) || ((E & EPI_BN_BWD) && use_coef)) o.a4 = __ldg(reinterpret_cast<const float4*>(aux + idx0));
  if ((p.epi & EPI_MASK) && !(E & EPI_PHILOX)) o.mk = __ldg(reinterpret_cast<const uint32_t*>(p.mask + idx0));
  if (E & EPI_RESID) o.r4 = __ldg(reinterpret_cast<const float4*>(p.resid + crow * p.ldr + n0));
  return o;
}
// the arithmetic of epilogue_f4 on 4 columns with pre-loaded operands and no global stores: v = result, pre = GELU pre-activation,
// mk = keep-mask bytes (drawn here when E has EPI_PHILOX)
template <uint32_t E>
__device__ __forceinline__ void epilogue_math(const GemmP& p, float4 acc, float4 b4, float alpha, int m, long long idx0, bool use_cf,
                                              const EpiCoef& cf, const EpiOps& o, float v[4], float pre[4], uint32_t& mk) {
  v[0] = (acc.x + b4.x) * alpha; v[1] = (acc.y + b4.y) * alpha; v[2] = (acc.z + b4.z) * alpha; v[3] = (acc.w + b4.w) * alpha;
  if ((E & EPI_BN_BWD) && use_cf && !(p.epi_skip_period && (m % p.epi_skip_period) == 0)) {
    v[0] = fmaf(cf.a.x, v[0], fmaf(cf.e.x, o.a4.x, cf.f.x)); v[1] = fmaf(cf.a.y, v[1], fmaf(cf.e.y, o.a4.y, cf.f.y));
    v[2] = fmaf(cf.a.z, v[2], fmaf(cf.e.z, o.a4.z, cf.f.z)); v[3] = fmaf(cf.a.w, v[3], fmaf(cf.e.w, o.a4.w, cf.f.w));
  }
  if (E & EPI_GELU) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (E & EPI_GELU_DSAVE) {                          // one erf serves the activation and its derivative
        const float x = v[j];
        gelu_and_grad(x, v[j], pre[j]);
      } else { pre[j] = v[j]; v[j] = gelu_erf(v[j]); }
    }
  }
  if (E & EPI_GELU_BWD) {
    if (E & EPI_GELU_DSAVE) { v[0] *= o.a4.x; v[1] *= o.a4.y; v[2] *= o.a4.z; v[3] *= o.a4.w; }
    else { v[0] *= gelu_erf_grad(o.a4.x); v[1] *= gelu_erf_grad(o.a4.y); v[2] *= gelu_erf_grad(o.a4.z); v[3] *= gelu_erf_grad(o.a4.w); }
  }
  mk = o.mk;
  if (p.epi & EPI_MASK) {
    if (E & EPI_PHILOX) {
      uint32_t c[4] = {(uint32_t)(idx0 >> 2), p.ph_site, p.ph_dyn ? __ldg(reinterpret_cast<const uint32_t*>(p.ph_dyn) + 2) : p.ph_step, 0u};
      philox4x32_10(c, p.ph_k0, p.ph_k1);
      mk = (c[0] >= p.ph_thr ? 1u : 0u) | (c[1] >= p.ph_thr ? 0x100u : 0u) | (c[2] >= p.ph_thr ? 0x10000u : 0u) | (c[3] >= p.ph_thr ? 0x1000000u : 0u);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const bool keep = ((mk >> (8 * j)) & 0xFFu) != 0u;
      v[j] = keep ? v[j] * p.mask_scale : 0.f;
      if ((E & EPI_GELU) && (E & EPI_GELU_DSAVE)) pre[j] = keep ? pre[j] * p.mask_scale : 0.f;
    }
  }
  if (E & EPI_RESID) { v[0] += o.r4.x; v[1] += o.r4.y; v[2] += o.r4.z; v[3] += o.r4.w; }
  if (p.epi & EPI_ROUND) {
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = round_tf32(v[j]);
  }
}
template <uint32_t E>
__device__ __forceinline__ void epilogue_apply(const GemmP& p, float4 acc, float* Cp, float* aux, float4 b4, float alpha, int m, int n0,
                                               bool use_cf, const EpiCoef& cf, const EpiOps& o) {
  const long long idx0 = p.c_map(m) * p.ldc + n0;
  float v[4], pre[4]; uint32_t mk;
  epilogue_math<E>(p, acc, b4, alpha, m, idx0, use_cf, cf, o, v, pre, mk);
  if ((E & EPI_GELU) && aux) *reinterpret_cast<float4*>(aux + idx0) = make_float4(pre[0], pre[1], pre[2], pre[3]);
  if ((E & EPI_PHILOX) && !(E & EPI_GELU_DSAVE) && (p.epi & EPI_MASK)) *reinterpret_cast<uint32_t*>(p.mask_gen + idx0) = mk;   // (DSAVE: the mask lives on inside aux)
  *reinterpret_cast<float4*>(Cp + idx0) = make_float4(v[0], v[1], v[2], v[3]);
}
// 8 consecutive columns of one row with 32-byte stores (st.global.v8.f32, sm_100): the per-warp store issue rate, not bytes, is what
// bounds this epilogue (tools/gemm_trace.py), so every store instruction carries twice the data
__device__ __forceinline__ void st_global_v8(float* ptr, const float a[4], const float b[4]) {
  asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(ptr), "f"(a[0]), "f"(a[1]), "f"(a[2]), "f"(a[3]), "f"(b[0]),
               "f"(b[1]), "f"(b[2]), "f"(b[3])
               : "memory");
}
template <uint32_t E>
__device__ __forceinline__ void epilogue_apply8(const GemmP& p, float4 acc0, float4 acc1, float* Cp, float* aux, float4 b0, float4 b1, float alpha,
                                                int m, int n0, bool use_cf, const EpiCoef& cf0, const EpiCoef& cf1, const EpiOps& o0, const EpiOps& o1) {
  const long long idx0 = p.c_map(m) * p.ldc + n0;
  float v0[4], v1[4], pre0[4], pre1[4]; uint32_t mk0, mk1;
  epilogue_math<E>(p, acc0, b0, alpha, m, idx0, use_cf, cf0, o0, v0, pre0, mk0);
  epilogue_math<E>(p, acc1, b1, alpha, m, idx0 + 4, use_cf, cf1, o1, v1, pre1, mk1);
  if ((E & EPI_GELU) && aux) st_global_v8(aux + idx0, pre0, pre1);
  if ((E & EPI_PHILOX) && !(E & EPI_GELU_DSAVE) && (p.epi & EPI_MASK)) *reinterpret_cast<uint2*>(p.mask_gen + idx0) = make_uint2(mk0, mk1);
  st_global_v8(Cp + idx0, v0, v1);
}

}  // namespace
}  // namespace cvt
