// ConvEmbed (Conv2D(D, k, s, 'same'), models/CvT(Par).py:203-208) as an explicit-im2col GEMM in TF32 mode: the patch matrix
//     col[(b,i,j)][(kh,kw,c)] = x[b, s i + kh - pt, s j + kw - pl, c]          (0 outside the image: TF 'SAME' padding)
// is written once (pre-rounded to TF32, row stride ldc = K rounded up to 4 floats so TMA can box it), then the forward product,
// the weight gradient (col^T dY) and the input gradient (dcol = dY W^T, gathered back by col2im) all run on the TMA-fed tcgen05
// kernels.  Both kernels here are pure data movement: coalesced along the channel axis, one pass, no atomics.
#include "gemm.cuh"

namespace cvt {
namespace {

// cin % 4 == 0: one thread per (row m, tap, float4 channel group)
template <int KK, int SS>
__global__ void __launch_bounds__(256) im2col_v4_kernel(const float* __restrict__ x, ConvGeom g, int M, int ldc, float* __restrict__ col) {
  pdl_wait(); pdl_trigger();
  if (KK) { g.k = KK; g.s = SS; }
  const int c4n = g.cin >> 2, per_row = g.k * g.k * c4n;
  const long long total = (long long)M * per_row;
  const int hw = g.ho * g.wo;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int m = (int)(idx / per_row), r = (int)(idx - (long long)m * per_row);
    const int tap = r / c4n, c = (r - tap * c4n) << 2;
    const int kh = tap / g.k, kw = tap - kh * g.k;
    const int b = m / hw, pix = m - b * hw, oi = pix / g.wo, oj = pix - oi * g.wo;
    const int yy = g.s * oi + kh - g.pt, xx = g.s * oj + kw - g.pl;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (yy >= 0 && yy < g.hin && xx >= 0 && xx < g.win) {
      v = *reinterpret_cast<const float4*>(x + (long long)b * g.bstride + g.boff + ((long long)yy * g.win + xx) * g.cin + c);
      if (!g.noround) v = make_float4(round_tf32(v.x), round_tf32(v.y), round_tf32(v.z), round_tf32(v.w));
    }
    *reinterpret_cast<float4*>(col + (long long)m * ldc + tap * g.cin + c) = v;
  }
}

// cin % 4 == 0 and k * k * cin / 4 <= 1024: a block walks a run of consecutive rows m; a thread keeps ONE (tap, float4 channel group) slot,
// decoded once, and the (image, output row, output column) of the walk advances by increments: no division in the loop
template <int KK, int SS>
__global__ void __launch_bounds__(1024) im2col_rows_kernel(const float* __restrict__ x, ConvGeom g, int M, int ldc, float* __restrict__ col, int rpb) {
  pdl_wait(); pdl_trigger();
  if (KK) { g.k = KK; g.s = SS; }
  const int c4n = g.cin >> 2, per_row = g.k * g.k * c4n, r = threadIdx.x;
  if (r >= per_row) return;
  const int tap = r / c4n, c = (r - tap * c4n) << 2, kh = tap / g.k, kw = tap - kh * g.k;
  const int dy = kh - g.pt, dx = kw - g.pl, koff = tap * g.cin + c;
  const int m0 = blockIdx.x * rpb, m1 = min(M, m0 + rpb), hw = g.ho * g.wo;
  int b = m0 / hw, pix = m0 - b * hw, oi = pix / g.wo, oj = pix - oi * g.wo;
#pragma unroll 4
  for (int m = m0; m < m1; ++m) {
    const int yy = g.s * oi + dy, xx = g.s * oj + dx;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (yy >= 0 && yy < g.hin && xx >= 0 && xx < g.win) {
      v = __ldg(reinterpret_cast<const float4*>(x + (long long)b * g.bstride + g.boff + ((long long)yy * g.win + xx) * g.cin + c));
      if (!g.noround) v = make_float4(round_tf32(v.x), round_tf32(v.y), round_tf32(v.z), round_tf32(v.w));
    }
    *reinterpret_cast<float4*>(col + (long long)m * ldc + koff) = v;
    if (++oj == g.wo) { oj = 0; if (++oi == g.ho) { oi = 0; ++b; } }
  }
}

// any cin (stage 1: cin = 1): one thread per (row m, padded column kk)
__global__ void __launch_bounds__(256) im2col_scalar_kernel(const float* __restrict__ x, ConvGeom g, int M, int K, int ldc, float* __restrict__ col) {
  pdl_wait(); pdl_trigger();
  const long long total = (long long)M * ldc;
  const int hw = g.ho * g.wo;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int m = (int)(idx / ldc), kk = (int)(idx - (long long)m * ldc);
    float v = 0.f;
    if (kk < K) {
      const int tap = kk / g.cin, c = kk - tap * g.cin;
      const int kh = tap / g.k, kw = tap - kh * g.k;
      const int b = m / hw, pix = m - b * hw, oi = pix / g.wo, oj = pix - oi * g.wo;
      const int yy = g.s * oi + kh - g.pt, xx = g.s * oj + kw - g.pl;
      if (yy >= 0 && yy < g.hin && xx >= 0 && xx < g.win)
        { v = x[(long long)b * g.bstride + g.boff + ((long long)yy * g.win + xx) * g.cin + c]; if (!g.noround) v = round_tf32(v); }
    }
    col[idx] = v;
  }
}

// any cin, ldc % 4 == 0: one thread per (row m, group of 4 padded columns): the index decode (runtime divisions) is paid once per 16-byte
// store instead of once per element, and consecutive threads write consecutive 16-byte pieces of the contiguous patch matrix
__global__ void __launch_bounds__(256) im2col_vec_kernel(const float* __restrict__ x, ConvGeom g, int M, int K, int ldc, float* __restrict__ col) {
  pdl_wait(); pdl_trigger();
  const int q4n = ldc >> 2;
  const long long total = (long long)M * q4n;
  const int hw = g.ho * g.wo;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int m = (int)(idx / q4n), kk = (int)(idx - (long long)m * q4n) << 2;
    const int b = m / hw, pix = m - b * hw, oi = pix / g.wo, oj = pix - oi * g.wo;
    int tap = kk / g.cin, c = kk - tap * g.cin, kh = tap / g.k, kw = tap - kh * g.k;
    const float* xb = x + (long long)b * g.bstride + g.boff;
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      v[e] = 0.f;
      const int yy = g.s * oi + kh - g.pt, xx = g.s * oj + kw - g.pl;
      if (kk + e < K && yy >= 0 && yy < g.hin && xx >= 0 && xx < g.win) { v[e] = __ldg(xb + ((long long)yy * g.win + xx) * g.cin + c); if (!g.noround) v[e] = round_tf32(v[e]); }
      if (++c == g.cin) { c = 0; if (++kw == g.k) { kw = 0; ++kh; } }
    }
    *reinterpret_cast<float4*>(col + idx * 4) = make_float4(v[0], v[1], v[2], v[3]);
  }
}

// transposed gather: dx[b, y, x, c] = sum over taps (kh, kw) with s | (y + pt - kh), s | (x + pl - kw) of dcol[(b,i,j)][(kh,kw,c)]
// one thread per (b, y, x, float4 channel group); cin % 4 == 0
// mask != nullptr: also dxm = dx * (mask ? mscale : 0) at the same offsets (the dropout at the end of the previous block's Mlp, next in
// backward order), optionally rounded to TF32
// KK / SS: compile-time kernel size and stride (0 = read them from g): the tap loops are full of `% s` and `/ s`
template <int KK, int SS>
__global__ void __launch_bounds__(256) col2im_v4_kernel(const float* __restrict__ dcol, ConvGeom g, int B, int ldc, float* __restrict__ dx,
                                                        const uint8_t* __restrict__ mask, float mscale, float* __restrict__ dxm, int round_m) {
  pdl_wait(); pdl_trigger();
  if (KK) { g.k = KK; g.s = SS; }
  const int c4n = g.cin >> 2;
  const long long total = (long long)B * g.hin * g.win * c4n;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(idx % c4n) << 2;
    long long r = idx / c4n;
    const int xx = (int)(r % g.win); r /= g.win;
    const int yy = (int)(r % g.hin); const int b = (int)(r / g.hin);
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int kh = 0; kh < (KK ? KK : g.k); ++kh) {
      const int ty = yy + g.pt - kh;
      if (ty < 0 || ty % (SS ? SS : g.s)) continue;
      const int oi = ty / (SS ? SS : g.s);
      if (oi >= g.ho) continue;
#pragma unroll
      for (int kw = 0; kw < (KK ? KK : g.k); ++kw) {
        const int tx = xx + g.pl - kw;
        if (tx < 0 || tx % (SS ? SS : g.s)) continue;
        const int oj = tx / (SS ? SS : g.s);
        if (oj >= g.wo) continue;
        const float4 v = *reinterpret_cast<const float4*>(dcol + ((long long)(b * g.ho + oi) * g.wo + oj) * ldc + (kh * g.k + kw) * g.cin + c);
        a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
      }
    }
    const long long off = (long long)b * g.bstride + g.boff + ((long long)yy * g.win + xx) * g.cin + c;
    *reinterpret_cast<float4*>(dx + off) = a;
    if (mask) {
      const uint32_t mk = *reinterpret_cast<const uint32_t*>(mask + off);
      float4 q;
      q.x = (mk & 0xFFu) ? a.x * mscale : 0.f; q.y = (mk & 0xFF00u) ? a.y * mscale : 0.f;
      q.z = (mk & 0xFF0000u) ? a.z * mscale : 0.f; q.w = (mk & 0xFF000000u) ? a.w * mscale : 0.f;
      if (round_m) { q.x = round_tf32(q.x); q.y = round_tf32(q.y); q.z = round_tf32(q.z); q.w = round_tf32(q.w); }
      *reinterpret_cast<float4*>(dxm + off) = q;
    }
  }
}

inline int ew_grid(long long n) { long long b = (n + 255) / 256; const long long cap = (long long)kNumSMs * 16; return (int)(b < 1 ? 1 : (b > cap ? cap : b)); }

}  // namespace

cudaError_t launch_im2col(const float* x, const ConvGeom& g, int M, int K, int ldc, float* col, cudaStream_t s) {
  if ((g.cin & 3) == 0 && (g.boff & 3) == 0 && (g.bstride & 3) == 0 && ldc == K)
  {
    const int per_row = g.k * g.k * (g.cin >> 2), threads = (per_row + 31) / 32 * 32;
    if (threads <= 1024) {
      const int rpb = M >= kNumSMs * 16 ? 4 : 1;           // rows per block (measured: 4 beats 1 and 16 on both 3x3 ConvEmbeds)
      const int grid = (M + rpb - 1) / rpb;
      if (g.k == 3 && g.s == 2) CVT_LAUNCH((im2col_rows_kernel<3, 2>), grid, threads, 0, s, x, g, M, ldc, col, rpb);
      else CVT_LAUNCH((im2col_rows_kernel<0, 0>), grid, threads, 0, s, x, g, M, ldc, col, rpb);
    } else
    if (g.k == 3 && g.s == 2) CVT_LAUNCH((im2col_v4_kernel<3, 2>), ew_grid((long long)M * (K >> 2)), 256, 0, s, x, g, M, ldc, col);
    else CVT_LAUNCH((im2col_v4_kernel<0, 0>), ew_grid((long long)M * (K >> 2)), 256, 0, s, x, g, M, ldc, col);
  }
  else if ((ldc & 3) == 0 && (reinterpret_cast<uintptr_t>(col) & 15) == 0)
    CVT_LAUNCH((im2col_vec_kernel), ew_grid((long long)M * (ldc >> 2)), 256, 0, s, x, g, M, K, ldc, col);
  else
    CVT_LAUNCH((im2col_scalar_kernel), ew_grid((long long)M * ldc), 256, 0, s, x, g, M, K, ldc, col);
  return cudaSuccess;
}

cudaError_t launch_col2im(const float* dcol, const ConvGeom& g, int B, int ldc, float* dx, cudaStream_t s, const uint8_t* mask, float mscale,
                          float* dxm, int round_m) {
  if ((g.cin & 3) || (g.boff & 3) || (g.bstride & 3)) { set_error("col2im: cin must be a multiple of 4"); return cudaErrorInvalidValue; }
  const int grid = ew_grid((long long)B * g.hin * g.win * (g.cin >> 2));
  if (g.k == 3 && g.s == 2) CVT_LAUNCH((col2im_v4_kernel<3, 2>), grid, 256, 0, s, dcol, g, B, ldc, dx, mask, mscale, dxm, round_m);     // ConvEmbed of stages 2, 3
  else CVT_LAUNCH((col2im_v4_kernel<0, 0>), grid, 256, 0, s, dcol, g, B, ldc, dx, mask, mscale, dxm, round_m);
  return cudaSuccess;
}

}  // namespace cvt
