// tcgen05 / TMEM contraction kernel (CVT_PREC_TF32): the tensor-core twin of gemm_simt.cu, same GemmP description.
//
//   tile       128 (M) x BN (64|128|256) x 32 (K, one 128-byte swizzle line of fp32), UMMA 128xBNx8 kind::tf32
//   operands   staged by all 128 threads global -> registers -> 128B-swizzled shared memory, so the A-side
//              prologues (im2col gather, BatchNorm affine, cls-row remap) cost nothing extra.  No transposes:
//              an operand that is contiguous along K is written K-major, one contiguous along M/N is written
//              MN-major, and the UMMA descriptors carry the major-ness.
//   pipeline   S smem stages; thread 0 issues the MMAs and commits them to the stage's mbarrier (frees the slot);
//              register prefetch of the next K-tile overlaps the tensor core.
//   epilogue   accumulator TMEM -> registers (tcgen05.ld 32x32b) -> per-warp smem transpose -> coalesced
//              bias/GELU/mask/residual/atomic/scatter epilogue (identical arithmetic to the FFMA kernel).
#include "gemm_fetch.cuh"
#include "model.cuh"
#include "tc_common.cuh"

namespace cvt {
using namespace fetch;
using namespace tc;

namespace {
constexpr int TBM = 128, TBK = 32, A_TILE_BYTES = TBM * TBK * 4;

template <int BN> struct Cfg {
  static constexpr int S = BN == 64 ? 3 : 2;
  static constexpr int B_TILE_BYTES = BN * TBK * 4;
  static constexpr int SMEM = 1024 + S * (A_TILE_BYTES + B_TILE_BYTES) + 4 * 32 * 33 * 4 + 16 * 8 + 16;
};

template <int AM, int BMODE, int BN>
__global__ void __launch_bounds__(128) gemm_tc_kernel(const __grid_constant__ GemmP p) {
  using C = Cfg<BN>;
  constexpr int S = C::S;
  constexpr bool A_ROWFIXED = (AM == A_PLAIN || AM == A_CONV);     // K-contiguous A  -> K-major
  constexpr bool B_KMAJOR = (BMODE == B_NK);
  constexpr int NB_FRAGS = BN / 16;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sA = smem;
  uint8_t* sB = sA + S * A_TILE_BYTES;
  float* sT = reinterpret_cast<float*>(sB + S * C::B_TILE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(sT + 4 * 32 * 33);      // [S] stage-free barriers, [S] accumulator-ready
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int zz = blockIdx.z;
  const int z = zz / p.splitk, ks = zz - z * p.splitk;
  const int zb = z / p.bdiv, zh = z - zb * p.bdiv;
  const float* A = p.A + zb * p.a_bs0 + zh * p.a_bs1;
  const float* B = p.B + zb * p.b_bs0 + zh * p.b_bs1;
  const float* sc = p.a_scale ? p.a_scale + z * p.a_aff_bs : nullptr;
  const float* sh = p.a_shift ? p.a_shift + z * p.a_aff_bs : nullptr;
  const int m_blk = blockIdx.y * TBM, n_blk = blockIdx.x * BN;
  const int ktiles = cdiv(p.K, TBK);
  const int tiles_per = cdiv(ktiles, p.splitk);
  const int kbeg = ks * tiles_per * TBK;
  const int kend = min(p.K, kbeg + tiles_per * TBK);
  if (kbeg >= kend) return;                                           // empty split (whole CTA exits together)
  const int nk = cdiv(kend - kbeg, TBK);

  if (warp == 0) tmem_alloc(tmem_slot, BN);
  if (tid == 32) {
    for (int i = 0; i <= S; ++i) mbar_init(&bars[i], 1);
    mbar_fence_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // ---- per-thread operand coordinates ----
  // A, K-major:  chunk = tid & 7 (16 B along K), rows (tid >> 3) + 16 i
  // A, MN-major: m4 = tid & 31 (float4 along M),  k rows (tid >> 5) + 4 i
  long long a_tok[8]; int a_cb[8], a_cij[8]; uint32_t a_valid = 0;
  if (A_ROWFIXED) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int m = m_blk + (tid >> 3) + 16 * i;
      a_tok[i] = 0; a_cb[i] = 0; a_cij[i] = 0;
      if (m < p.M) {
        a_valid |= 1u << i;
        if (AM == A_PLAIN) a_tok[i] = p.a_map(m);
        else {
          const int hw = p.cv.ho * p.cv.wo; const int b = m / hw; const int r = m - b * hw; const int oi = r / p.cv.wo;
          a_cb[i] = b; a_cij[i] = (oi << 16) | (r - oi * p.cv.wo);
        }
      }
    }
  }
  Frag ra[8], rb[NB_FRAGS];
  auto load_tile = [&](int k0) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (A_ROWFIXED)
        ra[i] = fetchA_row<AM>(p, A, sc, sh, m_blk + (tid >> 3) + 16 * i, k0 + ((tid & 7) << 2), kend, (a_valid >> i) & 1, a_tok[i],
                               a_cb[i], a_cij[i] >> 16, a_cij[i] & 0xFFFF);
      else
        ra[i] = fetchA_col<AM>(p, A, sc, sh, k0 + (tid >> 5) + 4 * i, kend, m_blk + ((tid & 31) << 2));
    }
#pragma unroll
    for (int i = 0; i < NB_FRAGS; ++i) {
      const int idx = tid + 128 * i;
      if (B_KMAJOR) rb[i] = fetchB_nk(p, B, n_blk + (idx >> 3), k0 + ((idx & 7) << 2), kend);
      else rb[i] = fetchB_kn(p, B, k0 + idx / (BN / 4), kend, n_blk + ((idx % (BN / 4)) << 2));
    }
  };
  auto st4 = [](uint8_t* base, uint32_t off, const Frag& f) {
    *reinterpret_cast<float4*>(base + off) = make_float4(to_tf32(f.v[0]), to_tf32(f.v[1]), to_tf32(f.v[2]), to_tf32(f.v[3]));
  };
  auto store_tile = [&](int s) {
    uint8_t* a = sA + s * A_TILE_BYTES; uint8_t* b = sB + s * C::B_TILE_BYTES;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (A_ROWFIXED) st4(a, sw128_kmajor_off((tid >> 3) + 16 * i, tid & 7), ra[i]);
      else st4(a, sw128_mnmajor_off((tid >> 5) + 4 * i, tid & 31, TBM / 32), ra[i]);
    }
#pragma unroll
    for (int i = 0; i < NB_FRAGS; ++i) {
      const int idx = tid + 128 * i;
      if (B_KMAJOR) st4(b, sw128_kmajor_off(idx >> 3, idx & 7), rb[i]);
      else st4(b, sw128_mnmajor_off(idx / (BN / 4), idx % (BN / 4), BN / 32), rb[i]);
    }
  };

  constexpr uint32_t idesc = make_idesc(kFmtTF32, TBM, BN, A_ROWFIXED ? 0u : 1u, B_KMAJOR ? 0u : 1u);
  load_tile(kbeg);
  for (int kt = 0; kt < nk; ++kt) {
    const int s = kt % S, use = kt / S;
    if (use > 0) mbar_wait(&bars[s], (use - 1) & 1);               // the MMAs that read this slot have completed
    store_tile(s);
    fence_proxy_async_smem();
    __syncthreads();
    if (kt + 1 < nk) load_tile(kbeg + (kt + 1) * TBK);              // global loads in flight while the tensor core works
    if (tid == 0) {
      tc_fence_after();
      const uint32_t a0 = smem_u32(sA + s * A_TILE_BYTES), b0 = smem_u32(sB + s * C::B_TILE_BYTES);
#pragma unroll
      for (int j = 0; j < TBK / 8; ++j) {
        const uint64_t ad = A_ROWFIXED ? make_smem_desc(a0 + j * 32, 16, 1024) : make_smem_desc(a0 + j * (TBM / 32) * 1024, (p.dbg & 1) ? (TBM / 32) * 512 : 512, (p.dbg & 1) ? 512 : (TBM / 32) * 512, kLayoutSw128Base32);
        const uint64_t bd = B_KMAJOR ? make_smem_desc(b0 + j * 32, 16, 1024) : make_smem_desc(b0 + j * (BN / 32) * 1024, (p.dbg & 1) ? (BN / 32) * 512 : 512, (p.dbg & 1) ? 512 : (BN / 32) * 512, kLayoutSw128Base32);
        mma_tf32_ss(tmem_base, ad, bd, idesc, (kt > 0 || j > 0) ? 1u : 0u);
      }
      mma_commit(&bars[s]);
      if (kt == nk - 1) mma_commit(&bars[S]);
    }
  }
  mbar_wait(&bars[S], 0);
  tc_fence_after();

  // ---- epilogue ----
  float* Cp = p.C + zb * p.c_bs0 + zh * p.c_bs1;
  float* aux = p.aux ? p.aux + zb * p.c_bs0 + zh * p.c_bs1 : nullptr;
  const float* bias = (p.epi & EPI_BIAS) ? p.bias + z * p.bias_bs : nullptr;
  const float alpha = ((p.epi & EPI_ALPHA0) && z == 0) ? p.alpha0 : p.alpha;
  float* tb = sT + warp * 32 * 33;
  const uint32_t trow = tmem_base + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
  for (int c0 = 0; c0 < BN; c0 += 32) {
    if (n_blk + c0 >= p.N) break;
    float v[32];
    tmem_ld32(trow + c0, v);
#pragma unroll
    for (int j = 0; j < 32; ++j) tb[lane * 33 + j] = v[j];
    __syncwarp();
    const int n = n_blk + c0 + lane;
    const bool nvalid = n < p.N;
    const float bn = (bias && nvalid) ? bias[n] : 0.f;
    int s_tap = 0, s_c = 0, s_kh = 0, s_kw = 0;
    if ((p.epi & EPI_SCATTER) && nvalid) { s_tap = n / p.cv.cin; s_c = n - s_tap * p.cv.cin; s_kh = s_tap / p.cv.k; s_kw = s_tap - s_kh * p.cv.k; }
#pragma unroll 4
    for (int r = 0; r < 32; ++r) {
      const int m = m_blk + warp * 32 + r;
      if (m >= p.M || !nvalid) continue;
      float x = tb[r * 33 + lane];
      if (p.epi & EPI_SCATTER) {
        const ConvGeom& g = p.cv;
        const int hw = g.ho * g.wo; const int b = m / hw; const int rr = m - b * hw; const int oi = rr / g.wo; const int oj = rr - oi * g.wo;
        const int yy = g.s * oi + s_kh - g.pt, xx = g.s * oj + s_kw - g.pl;
        if (yy >= 0 && yy < g.hin && xx >= 0 && xx < g.win)
          atomicAdd(Cp + (long long)b * g.bstride + g.boff + ((long long)yy * g.win + xx) * g.cin + s_c, x * alpha);
        continue;
      }
      const long long crow = p.c_map(m);
      const long long idx = crow * p.ldc + n;
      x = (x + bn) * alpha;
      if (p.epi & EPI_GELU) { if (aux) aux[idx] = x; x = gelu_erf(x); }
      if (p.epi & EPI_RELU) x = fmaxf(x, 0.f);
      if (p.epi & EPI_GELU_BWD) x *= gelu_erf_grad(aux[idx]);
      if (p.epi & EPI_RELU_BWD) x = aux[idx] > 0.f ? x : 0.f;
      if (p.epi & EPI_MASK) x *= p.mask[idx] ? p.mask_scale : 0.f;
      if (p.epi & EPI_RESID) x += p.resid[crow * p.ldr + n];
      if (p.epi & EPI_ATOMIC) atomicAdd(Cp + idx, x); else Cp[idx] = x;
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, BN);
}

template <int AM, int BMODE, int BN>
cudaError_t launch_cfg(const GemmP& p, cudaStream_t s) {
  static bool configured = false;
  if (!configured) {
    CVT_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<AM, BMODE, BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM));
    configured = true;
  }
  dim3 grid(cdiv(p.N, BN), cdiv(p.M, TBM), p.batch * p.splitk);
  gemm_tc_kernel<AM, BMODE, BN><<<grid, 128, Cfg<BN>::SMEM, s>>>(p);
  CVT_LAUNCH_CHECK();
  return cudaSuccess;
}

template <int AM, int BMODE>
cudaError_t launch_bn(const GemmP& p, int bn, cudaStream_t s) {
  if (bn == 256) return launch_cfg<AM, BMODE, 256>(p, s);
  if (bn == 128) return launch_cfg<AM, BMODE, 128>(p, s);
  return launch_cfg<AM, BMODE, 64>(p, s);
}

int pick_bn(const GemmP& p) {
  int bn = (p.N % 256 == 0) ? 256 : ((p.N % 128 == 0) ? 128 : 64);
  const long long mt = (long long)cdiv(p.M, TBM) * p.batch * p.splitk;
  while (bn > 64 && mt * cdiv(p.N, bn) < 2 * kNumSMs) bn >>= 1;
  return bn;
}
}  // namespace

bool gemm_tc_supports(const GemmP& p) {
  if (p.N % 64 != 0 || p.M < 1 || p.K < 1) return false;
  if (p.splitk > 1 && !(p.epi & (EPI_ATOMIC | EPI_SCATTER))) return false;
  return true;
}

cudaError_t launch_gemm_tc(const GemmP& p, cudaStream_t s) {
  if (!gemm_tc_supports(p)) { set_error("gemm_tc: unsupported shape"); return cudaErrorNotSupported; }
  const int bn = pick_bn(p);
  switch (p.a_mode * 2 + p.b_mode) {
    case A_PLAIN * 2 + B_KN: return launch_bn<A_PLAIN, B_KN>(p, bn, s);
    case A_PLAIN * 2 + B_NK: return launch_bn<A_PLAIN, B_NK>(p, bn, s);
    case A_TRANS * 2 + B_KN: return launch_bn<A_TRANS, B_KN>(p, bn, s);
    case A_TRANS * 2 + B_NK: return launch_bn<A_TRANS, B_NK>(p, bn, s);
    case A_CONV * 2 + B_KN: return launch_bn<A_CONV, B_KN>(p, bn, s);
    case A_CONV * 2 + B_NK: return launch_bn<A_CONV, B_NK>(p, bn, s);
    case A_CONVT * 2 + B_KN: return launch_bn<A_CONVT, B_KN>(p, bn, s);
    case A_CONVT * 2 + B_NK: return launch_bn<A_CONVT, B_NK>(p, bn, s);
  }
  set_error("gemm_tc: bad operand modes");
  return cudaErrorInvalidValue;
}

}  // namespace cvt

// diagnostics: one raw contraction through either kernel (tests/test_gpu_gemm.py)
extern "C" CVT_API int cvt_debug_gemm(int a_mode, int b_mode, int M, int N, int K, int splitk, int use_tc, const float* A, const float* B,
                                      float* C) {
  using namespace cvt;
  if (a_mode > A_TRANS || a_mode < 0 || b_mode < 0 || b_mode > B_NK || M < 1 || N < 1 || K < 1 || splitk < 1) { set_error("cvt_debug_gemm: bad argument"); return CVT_ERR_INVALID; }
  float *dA = nullptr, *dB = nullptr, *dC = nullptr;
  const size_t na = (size_t)M * K, nb = (size_t)K * N, nc = (size_t)M * N;
  auto fail = [&](cudaError_t e) { set_error(std::string("cvt_debug_gemm: ") + cudaGetErrorString(e)); cudaFree(dA); cudaFree(dB); cudaFree(dC); return (int)CVT_ERR_CUDA; };
  cudaError_t e;
  if ((e = cudaMalloc(&dA, na * 4)) != cudaSuccess || (e = cudaMalloc(&dB, nb * 4)) != cudaSuccess || (e = cudaMalloc(&dC, nc * 4)) != cudaSuccess) return fail(e);
  if ((e = cudaMemcpy(dA, A, na * 4, cudaMemcpyHostToDevice)) != cudaSuccess || (e = cudaMemcpy(dB, B, nb * 4, cudaMemcpyHostToDevice)) != cudaSuccess ||
      (e = cudaMemset(dC, 0, nc * 4)) != cudaSuccess) return fail(e);
  GemmP p; p.M = M; p.N = N; p.K = K; p.splitk = splitk;
  p.A = dA; p.a_mode = a_mode; p.lda = a_mode == A_PLAIN ? K : M;
  p.B = dB; p.b_mode = b_mode; p.ldb = b_mode == B_KN ? N : K;
  p.C = dC; p.ldc = N; if (splitk > 1) p.epi = EPI_ATOMIC;
  p.dbg = use_tc >> 1;
  if (use_tc && !gemm_tc_supports(p)) { cudaFree(dA); cudaFree(dB); cudaFree(dC); set_error("cvt_debug_gemm: shape not supported by the tcgen05 kernel"); return CVT_ERR_INVALID; }
  e = use_tc ? launch_gemm_tc(p, 0) : launch_gemm_simt(p, 0);
  if (e != cudaSuccess) return fail(e);
  if ((e = cudaDeviceSynchronize()) != cudaSuccess) return fail(e);
  if ((e = cudaMemcpy(C, dC, nc * 4, cudaMemcpyDeviceToHost)) != cudaSuccess) return fail(e);
  cudaFree(dA); cudaFree(dB); cudaFree(dC);
  return CVT_OK;
}

