// Guarded operand fetchers shared by the FFMA and the tcgen05 contraction kernels: every access checks its bounds
// and alignment, so ragged shapes (Tc = 65, K = 49, num_classes = 1) go through the same code.
#pragma once
#include "gemm.cuh"

namespace cvt {
namespace fetch {

struct Frag { float v[4]; };

__device__ __forceinline__ bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// ---- A fetchers -------------------------------------------------------------------------------
// row-fixed mapping: thread owns row m, fetches k..k+3
template <int AM>
__device__ __forceinline__ Frag fetchA_row(const GemmP& p, const float* A, const float* sc, const float* sh,
                                           int m, int k, int kend, bool mvalid, long long tok, int cb, int ci_, int cj) {
  Frag f{{0.f, 0.f, 0.f, 0.f}};
  if (!mvalid || k >= kend) return f;
  if (AM == A_PLAIN) {
    const float* ptr = A + tok * p.lda + k;
    if (k + 3 < kend && aligned16(ptr)) {
      const float4 t = *reinterpret_cast<const float4*>(ptr);
      f.v[0] = t.x; f.v[1] = t.y; f.v[2] = t.z; f.v[3] = t.w;
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) if (k + i < kend) f.v[i] = ptr[i];
    }
    if (sc != nullptr && !(p.a_skip_period && (tok % p.a_skip_period) == 0)) {
#pragma unroll
      for (int i = 0; i < 4; ++i) if (k + i < kend) f.v[i] = f.v[i] * sc[k + i] + sh[k + i];
    }
  } else {  // A_CONV: m = (cb, ci_, cj) decoded by the caller
    const ConvGeom& g = p.cv;
    if ((g.cin & 3) == 0) {
      const int tap = k / g.cin, c = k - tap * g.cin;
      const int kh = tap / g.k, kw = tap - kh * g.k;
      const int yy = g.s * ci_ + kh - g.pt, xx = g.s * cj + kw - g.pl;
      if (yy >= 0 && yy < g.hin && xx >= 0 && xx < g.win) {
        const float4 t = *reinterpret_cast<const float4*>(A + (long long)cb * g.bstride + g.boff + ((long long)yy * g.win + xx) * g.cin + c);
        f.v[0] = t.x; f.v[1] = t.y; f.v[2] = t.z; f.v[3] = t.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int kk = k + i;
        if (kk < kend) {
          const int tap = kk / g.cin, c = kk - tap * g.cin;
          const int kh = tap / g.k, kw = tap - kh * g.k;
          const int yy = g.s * ci_ + kh - g.pt, xx = g.s * cj + kw - g.pl;
          if (yy >= 0 && yy < g.hin && xx >= 0 && xx < g.win)
            f.v[i] = A[(long long)cb * g.bstride + g.boff + ((long long)yy * g.win + xx) * g.cin + c];
        }
      }
    }
  }
  return f;
}

// k-fixed mapping: thread owns contraction index k (a token), fetches m0..m0+3
template <int AM>
__device__ __forceinline__ Frag fetchA_col(const GemmP& p, const float* A, const float* sc, const float* sh,
                                           int k, int kend, int m0) {
  Frag f{{0.f, 0.f, 0.f, 0.f}};
  if (k >= kend || m0 >= p.M) return f;
  if (AM == A_TRANS) {
    const long long tok = p.a_map(k);
    const float* ptr = A + tok * p.lda + m0;
    if (m0 + 3 < p.M && aligned16(ptr)) {
      const float4 t = *reinterpret_cast<const float4*>(ptr);
      f.v[0] = t.x; f.v[1] = t.y; f.v[2] = t.z; f.v[3] = t.w;
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) if (m0 + i < p.M) f.v[i] = ptr[i];
    }
    if (sc != nullptr && !(p.a_skip_period && (tok % p.a_skip_period) == 0)) {
#pragma unroll
      for (int i = 0; i < 4; ++i) if (m0 + i < p.M) f.v[i] = f.v[i] * sc[m0 + i] + sh[m0 + i];
    }
  } else {  // A_CONVT: k = output pixel (b,i,j), m = (tap, c)
    const ConvGeom& g = p.cv;
    const int hw = g.ho * g.wo;
    const int b = k / hw, r = k - b * hw;
    const int oi = r / g.wo, oj = r - oi * g.wo;
    if ((g.cin & 3) == 0) {
      const int tap = m0 / g.cin, c = m0 - tap * g.cin;
      const int kh = tap / g.k, kw = tap - kh * g.k;
      const int yy = g.s * oi + kh - g.pt, xx = g.s * oj + kw - g.pl;
      if (yy >= 0 && yy < g.hin && xx >= 0 && xx < g.win) {
        const float4 t = *reinterpret_cast<const float4*>(A + (long long)b * g.bstride + g.boff + ((long long)yy * g.win + xx) * g.cin + c);
        f.v[0] = t.x; f.v[1] = t.y; f.v[2] = t.z; f.v[3] = t.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int mm = m0 + i;
        if (mm < p.M) {
          const int tap = mm / g.cin, c = mm - tap * g.cin;
          const int kh = tap / g.k, kw = tap - kh * g.k;
          const int yy = g.s * oi + kh - g.pt, xx = g.s * oj + kw - g.pl;
          if (yy >= 0 && yy < g.hin && xx >= 0 && xx < g.win)
            f.v[i] = A[(long long)b * g.bstride + g.boff + ((long long)yy * g.win + xx) * g.cin + c];
        }
      }
    }
  }
  return f;
}

// ---- B fetchers -------------------------------------------------------------------------------
__device__ __forceinline__ Frag fetchB_kn(const GemmP& p, const float* B, int k, int kend, int n0) {
  Frag f{{0.f, 0.f, 0.f, 0.f}};
  if (k >= kend || n0 >= p.N) return f;
  const float* ptr = B + p.b_map(k) * p.ldb + n0;
  if (n0 + 3 < p.N && aligned16(ptr)) {
    const float4 t = *reinterpret_cast<const float4*>(ptr);
    f.v[0] = t.x; f.v[1] = t.y; f.v[2] = t.z; f.v[3] = t.w;
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) if (n0 + i < p.N) f.v[i] = ptr[i];
  }
  return f;
}
__device__ __forceinline__ Frag fetchB_nk(const GemmP& p, const float* B, int n, int k, int kend) {
  Frag f{{0.f, 0.f, 0.f, 0.f}};
  if (n >= p.N || k >= kend) return f;
  const float* ptr = B + (long long)n * p.ldb + k;
  if (k + 3 < kend && aligned16(ptr)) {
    const float4 t = *reinterpret_cast<const float4*>(ptr);
    f.v[0] = t.x; f.v[1] = t.y; f.v[2] = t.z; f.v[3] = t.w;
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) if (k + i < kend) f.v[i] = ptr[i];
  }
  return f;
}

}  // namespace fetch
}  // namespace cvt
