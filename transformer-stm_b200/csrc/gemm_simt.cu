// fp32 FFMA contraction kernel: the exact-arithmetic path (CVT_PREC_FP32) and the reference the
// tcgen05 path is checked against on the device.  64x64x16 tiles, 256 threads, 4x4 register tile,
// register-prefetch double buffering.  Every operand access is guarded, so ragged shapes
// (Tc = 65, 85x87 maps, num_classes = 1) take the same kernel.
#include "gemm_fetch.cuh"

namespace cvt {

namespace {
constexpr int BM = 64, BN = 64, BK = 16, LDS = 68;

using namespace fetch;

template <int AM, int BMODE>
__global__ void __launch_bounds__(256) gemm_simt_kernel(const __grid_constant__ GemmP p) {
  pdl_wait(); pdl_trigger();
  __shared__ __align__(16) float As[2][BK][LDS];
  __shared__ __align__(16) float Bs[2][BK][LDS];
  constexpr bool A_ROWFIXED = (AM == A_PLAIN || AM == A_CONV);

  const int tid = threadIdx.x;
  const int zz = blockIdx.z;
  const int z = zz / p.splitk, ks = zz - z * p.splitk;
  const int zb = z / p.bdiv, zh = z - zb * p.bdiv;
  const float* A = p.A + zb * p.a_bs0 + zh * p.a_bs1;
  const float* B = p.B + zb * p.b_bs0 + zh * p.b_bs1;
  const float* sc = p.a_scale ? p.a_scale + z * p.a_aff_bs : nullptr;
  const float* sh = p.a_shift ? p.a_shift + z * p.a_aff_bs : nullptr;

  const int m_blk = blockIdx.y * BM, n_blk = blockIdx.x * BN;
  // split-K range (multiple of BK per split)
  const int ktiles = cdiv(p.K, BK);
  const int tiles_per = cdiv(ktiles, p.splitk);
  const int kbeg = ks * tiles_per * BK;
  const int kend = min(p.K, kbeg + tiles_per * BK);
  if (kbeg >= kend && p.splitk > 1) return;
  const int nk = cdiv(max(kend - kbeg, 0), BK);

  // row-fixed thread coordinates
  const int rf_row = tid >> 2, rf_kq = (tid & 3) << 2;
  // k-fixed thread coordinates
  const int kf_k = tid >> 4, kf_q = (tid & 15) << 2;

  // per-thread constants for the A row (row-fixed modes)
  const int am = m_blk + rf_row;
  const bool amvalid = am < p.M;
  long long atok = 0; int cb = 0, ci_ = 0, cj = 0;
  if (A_ROWFIXED && amvalid) {
    if (AM == A_PLAIN) atok = p.a_map(am);
    else { const int hw = p.cv.ho * p.cv.wo; cb = am / hw; const int r = am - cb * hw; ci_ = r / p.cv.wo; cj = r - ci_ * p.cv.wo; }
  }

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  auto fetchA = [&](int k0) -> Frag {
    if (A_ROWFIXED) return fetchA_row<AM>(p, A, sc, sh, am, k0 + rf_kq, kend, amvalid, atok, cb, ci_, cj);
    return fetchA_col<AM>(p, A, sc, sh, k0 + kf_k, kend, m_blk + kf_q);
  };
  auto fetchB = [&](int k0) -> Frag {
    if (BMODE == B_KN) return fetchB_kn(p, B, k0 + kf_k, kend, n_blk + kf_q);
    return fetchB_nk(p, B, n_blk + rf_row, k0 + rf_kq, kend);
  };
  auto storeA = [&](int buf, const Frag& f) {
    if (A_ROWFIXED) {
#pragma unroll
      for (int i = 0; i < 4; ++i) As[buf][rf_kq + i][rf_row] = f.v[i];
    } else {
      *reinterpret_cast<float4*>(&As[buf][kf_k][kf_q]) = make_float4(f.v[0], f.v[1], f.v[2], f.v[3]);
    }
  };
  auto storeB = [&](int buf, const Frag& f) {
    if (BMODE == B_KN) {
      *reinterpret_cast<float4*>(&Bs[buf][kf_k][kf_q]) = make_float4(f.v[0], f.v[1], f.v[2], f.v[3]);
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) Bs[buf][rf_kq + i][rf_row] = f.v[i];
    }
  };

  const int ty = tid >> 4, tx = tid & 15;
  if (nk > 0) {
    Frag fa = fetchA(kbeg), fb = fetchB(kbeg);
    storeA(0, fa); storeB(0, fb);
  }
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    Frag fa, fb;
    const bool more = kt + 1 < nk;
    if (more) { fa = fetchA(kbeg + (kt + 1) * BK); fb = fetchB(kbeg + (kt + 1) * BK); }
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&As[buf][k][ty << 2]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[buf][k][tx << 2]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    if (more) { storeA(buf ^ 1, fa); storeB(buf ^ 1, fb); }
    __syncthreads();
  }

  // ---- epilogue ----
  float* C = p.C + zb * p.c_bs0 + zh * p.c_bs1;
  float* aux = p.aux ? p.aux + zb * p.c_bs0 + zh * p.c_bs1 : nullptr;
  const float* bias = (p.epi & EPI_BIAS) ? p.bias + z * p.bias_bs : nullptr;
  const float alpha = ((p.epi & EPI_ALPHA0) && z == 0) ? p.alpha0 : p.alpha;
  const int n0 = n_blk + (tx << 2);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m_blk + (ty << 2) + i;
    if (m >= p.M) continue;
    if (p.epi & EPI_SCATTER) {
      // C is the conv input gradient [B,hin,win,cin]; m = output pixel, n = (tap, c)
      const ConvGeom& g = p.cv;
      const int hw = g.ho * g.wo;
      const int b = m / hw, r = m - b * hw;
      const int oi = r / g.wo, oj = r - oi * g.wo;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int n = n0 + j;
        if (n >= p.N) continue;
        const int tap = n / g.cin, c = n - tap * g.cin;
        const int kh = tap / g.k, kw = tap - kh * g.k;
        const int yy = g.s * oi + kh - g.pt, xx = g.s * oj + kw - g.pl;
        if (yy >= 0 && yy < g.hin && xx >= 0 && xx < g.win)
          atomicAdd(C + (long long)b * g.bstride + g.boff + ((long long)yy * g.win + xx) * g.cin + c, acc[i][j] * alpha);
      }
      continue;
    }
    const long long crow = p.c_map(m);
    float v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + j;
      float x = acc[i][j];
      if (n < p.N) {
        const long long idx = crow * p.ldc + n;
        if (bias) x += bias[n];
        x *= alpha;
        if (p.epi & EPI_GELU) { if (aux) aux[idx] = x; x = gelu_erf(x); }
        if (p.epi & EPI_RELU) x = fmaxf(x, 0.f);
        if (p.epi & EPI_GELU_BWD) x *= gelu_erf_grad(aux[idx]);
        if (p.epi & EPI_RELU_BWD) x = aux[idx] > 0.f ? x : 0.f;
        if (p.epi & EPI_MASK) x *= p.mask[idx] ? p.mask_scale : 0.f;
        if (p.epi & EPI_RESID) x += p.resid[crow * p.ldr + n];
      }
      v[j] = x;
    }
    float* dst = C + crow * p.ldc + n0;
    if (p.epi & EPI_ATOMIC) {
#pragma unroll
      for (int j = 0; j < 4; ++j) if (n0 + j < p.N) atomicAdd(dst + j, v[j]);
    } else if (n0 + 3 < p.N && aligned16(dst)) {
      *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j) if (n0 + j < p.N) dst[j] = v[j];
    }
  }
}

template <int AM, int BMODE>
cudaError_t launch_one(const GemmP& p, cudaStream_t s) {
  dim3 grid(cdiv(p.N, BN), cdiv(p.M, BM), p.batch * p.splitk);
  CVT_LAUNCH((gemm_simt_kernel<AM, BMODE>), grid, 256, 0, s, p);
  return cudaSuccess;
}

// ---- column sums -------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) colsum_kernel(const float* __restrict__ X, long long ld, RowMap map, int rows, int N,
                                                     long long x_bs, float* __restrict__ out, long long out_bs, int rows_per_chunk) {
  pdl_wait(); pdl_trigger();
  __shared__ float red[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int n = blockIdx.x * 32 + tx;
  const int z = blockIdx.z;
  const float* Xz = X + z * x_bs;
  const int r0 = blockIdx.y * rows_per_chunk, r1 = min(rows, r0 + rows_per_chunk);
  float s = 0.f;
  if (n < N)
    for (int r = r0 + ty; r < r1; r += 8) s += Xz[map(r) * ld + n];
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && n < N) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][tx];
    atomicAdd(out + z * out_bs + n, t);
  }
}
}  // namespace

cudaError_t launch_gemm_simt(const GemmP& p, cudaStream_t s) {
  if (p.M <= 0 || p.N <= 0 || p.batch <= 0) return cudaSuccess;
  if (p.splitk > 1 && !(p.epi & (EPI_ATOMIC | EPI_SCATTER))) {
    set_error("gemm_simt: split-K needs an atomic epilogue");
    return cudaErrorInvalidValue;
  }
  switch (p.a_mode * 2 + p.b_mode) {
    case A_PLAIN * 2 + B_KN: return launch_one<A_PLAIN, B_KN>(p, s);
    case A_PLAIN * 2 + B_NK: return launch_one<A_PLAIN, B_NK>(p, s);
    case A_TRANS * 2 + B_KN: return launch_one<A_TRANS, B_KN>(p, s);
    case A_TRANS * 2 + B_NK: return launch_one<A_TRANS, B_NK>(p, s);
    case A_CONV * 2 + B_KN: return launch_one<A_CONV, B_KN>(p, s);
    case A_CONV * 2 + B_NK: return launch_one<A_CONV, B_NK>(p, s);
    case A_CONVT * 2 + B_KN: return launch_one<A_CONVT, B_KN>(p, s);
    case A_CONVT * 2 + B_NK: return launch_one<A_CONVT, B_NK>(p, s);
  }
  set_error("gemm_simt: bad operand modes");
  return cudaErrorInvalidValue;
}

cudaError_t launch_colsum(const float* X, long long ld, RowMap map, int rows, int N, int batch, long long x_bs,
                          float* out, long long out_bs, cudaStream_t s) {
  if (rows <= 0 || N <= 0) return cudaSuccess;
  const int colblocks = cdiv(N, 32);
  int chunks = max(1, min(cdiv(rows, 64), (kNumSMs * 4) / max(1, colblocks * batch)));
  const int rows_per_chunk = cdiv(rows, chunks);
  chunks = cdiv(rows, rows_per_chunk);
  dim3 grid(colblocks, chunks, batch);
  CVT_LAUNCH((colsum_kernel), grid, 256, 0, s, X, ld, map, rows, N, x_bs, out, out_bs, rows_per_chunk);
  return cudaSuccess;
}

}  // namespace cvt
