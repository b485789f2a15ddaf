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
constexpr int TBM = 128, TBK = 32, A_TILE_BYTES = TBM * TBK * 4, NT = 256, STAGES = 3;

template <int BN> struct Cfg {
  static constexpr int B_TILE_BYTES = BN * TBK * 4;
  static constexpr int SMEM = 1024 + STAGES * (A_TILE_BYTES + B_TILE_BYTES) + 16 * 8 + 16;
};

// ---- epilogue of one accumulator row segment: 32 consecutive output columns held by one thread ------------------------------
__device__ __forceinline__ void epilogue_row32(const GemmP& p, float v[32], float* Cp, float* aux, const float* bias, float alpha,
                                               int m, int n0) {
  if (p.epi & EPI_SCATTER) {
    // C is the conv input gradient [B,hin,win,cin]; m = output pixel, n = (tap, c)
    const ConvGeom& g = p.cv;
    const int hw = g.ho * g.wo; const int b = m / hw; const int rr = m - b * hw; const int oi = rr / g.wo; const int oj = rr - oi * g.wo;
    float* img = Cp + (long long)b * g.bstride + g.boff;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
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
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 b4 = *reinterpret_cast<const float4*>(bias + n0 + j);
      v[j] += b4.x; v[j + 1] += b4.y; v[j + 2] += b4.z; v[j + 3] += b4.w;
    }
  }
  if (alpha != 1.f) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] *= alpha;
  }
  if (p.epi & EPI_GELU) {
    if (aux) {
#pragma unroll
      for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(aux + idx0 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
    }
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = gelu_erf(v[j]);
  }
  if (p.epi & EPI_RELU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
  }
  if (p.epi & (EPI_GELU_BWD | EPI_RELU_BWD)) {
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 a4 = *reinterpret_cast<const float4*>(aux + idx0 + j);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
      for (int t = 0; t < 4; ++t) v[j + t] = (p.epi & EPI_GELU_BWD) ? v[j + t] * gelu_erf_grad(a[t]) : (a[t] > 0.f ? v[j + t] : 0.f);
    }
  }
  if (p.epi & EPI_MASK) {
#pragma unroll
    for (int j = 0; j < 32; j += 16) {
      const uint4 mk = *reinterpret_cast<const uint4*>(p.mask + idx0 + j);
      const uint32_t w[4] = {mk.x, mk.y, mk.z, mk.w};
#pragma unroll
      for (int t = 0; t < 16; ++t) v[j + t] = ((w[t >> 2] >> ((t & 3) * 8)) & 0xFFu) ? v[j + t] * p.mask_scale : 0.f;
    }
  }
  if (p.epi & EPI_RESID) {
    const float* rp = p.resid + crow * p.ldr + n0;
#pragma unroll
    for (int j = 0; j < 32; j += 4) {
      const float4 r4 = *reinterpret_cast<const float4*>(rp + j);
      v[j] += r4.x; v[j + 1] += r4.y; v[j + 2] += r4.z; v[j + 3] += r4.w;
    }
  }
  if (p.epi & EPI_ATOMIC) {
#pragma unroll
    for (int j = 0; j < 32; ++j) atomicAdd(Cp + idx0 + j, v[j]);
  } else {
#pragma unroll
    for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(Cp + idx0 + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
  }
}

template <int AM, int BMODE, int BN>
__global__ void __launch_bounds__(NT, 2) gemm_tc_kernel(const __grid_constant__ GemmP p) {
  using C = Cfg<BN>;
  constexpr bool A_ROWFIXED = (AM == A_PLAIN || AM == A_CONV);     // K-contiguous A  -> K-major
  constexpr bool B_KMAJOR = (BMODE == B_NK);
  constexpr int NA_FRAGS = TBM * 8 / NT;       // 4
  constexpr int NB_FRAGS = BN * 8 / NT;        // 2 | 4
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
  uint8_t* sA = smem;
  uint8_t* sB = sA + STAGES * A_TILE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + STAGES * C::B_TILE_BYTES);      // [STAGES] stage-free, [STAGES] accumulator-ready
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
    for (int i = 0; i <= STAGES; ++i) mbar_init(&bars[i], 1);
    mbar_fence_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // ---- per-thread operand coordinates ----
  // A, K-major:  chunk = tid & 7 (16 B along K), rows (tid >> 3) + 32 i
  // A, MN-major: m4 = tid & 31 (float4 along M),  k rows (tid >> 5) + 8 i
  long long a_tok[NA_FRAGS]; int a_cb[NA_FRAGS], a_cij[NA_FRAGS]; uint32_t a_valid = 0;
  if (A_ROWFIXED) {
#pragma unroll
    for (int i = 0; i < NA_FRAGS; ++i) {
      const int m = m_blk + (tid >> 3) + 32 * i;
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
  Frag ra[NA_FRAGS], rb[NB_FRAGS];
  auto load_tile = [&](int k0) {
#pragma unroll
    for (int i = 0; i < NA_FRAGS; ++i) {
      if (A_ROWFIXED)
        ra[i] = fetchA_row<AM>(p, A, sc, sh, m_blk + (tid >> 3) + 32 * i, k0 + ((tid & 7) << 2), kend, (a_valid >> i) & 1, a_tok[i],
                               a_cb[i], a_cij[i] >> 16, a_cij[i] & 0xFFFF);
      else
        ra[i] = fetchA_col<AM>(p, A, sc, sh, k0 + (tid >> 5) + 8 * i, kend, m_blk + ((tid & 31) << 2));
    }
#pragma unroll
    for (int i = 0; i < NB_FRAGS; ++i) {
      const int idx = tid + NT * i;
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
    for (int i = 0; i < NA_FRAGS; ++i) {
      if (A_ROWFIXED) st4(a, sw128_kmajor_off((tid >> 3) + 32 * i, tid & 7), ra[i]);
      else st4(a, sw128_mnmajor_off((tid >> 5) + 8 * i, tid & 31, TBM / 32), ra[i]);
    }
#pragma unroll
    for (int i = 0; i < NB_FRAGS; ++i) {
      const int idx = tid + NT * i;
      if (B_KMAJOR) st4(b, sw128_kmajor_off(idx >> 3, idx & 7), rb[i]);
      else st4(b, sw128_mnmajor_off(idx / (BN / 4), idx % (BN / 4), BN / 32), rb[i]);
    }
  };

  constexpr uint32_t idesc = make_idesc(kFmtTF32, TBM, BN, A_ROWFIXED ? 0u : 1u, B_KMAJOR ? 0u : 1u);
  load_tile(kbeg);
  for (int kt = 0; kt < nk; ++kt) {
    const int s = kt % STAGES, use = kt / STAGES;
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
        const uint64_t ad = A_ROWFIXED ? make_smem_desc(a0 + j * 32, 16, 1024)
                                       : make_smem_desc(a0 + j * (TBM / 32) * 1024, 512, (TBM / 32) * 512, kLayoutSw128Base32);
        const uint64_t bd = B_KMAJOR ? make_smem_desc(b0 + j * 32, 16, 1024)
                                     : make_smem_desc(b0 + j * (BN / 32) * 1024, 512, (BN / 32) * 512, kLayoutSw128Base32);
        mma_tf32_ss(tmem_base, ad, bd, idesc, (kt > 0 || j > 0) ? 1u : 0u);
      }
      mma_commit(&bars[s]);
      if (kt == nk - 1) mma_commit(&bars[STAGES]);
    }
  }
  mbar_wait(&bars[STAGES], 0);
  tc_fence_after();

  // ---- epilogue: warp w reads TMEM lane quadrant w & 3, column half w >> 2; one accumulator row per thread ----
  float* Cp = p.C + zb * p.c_bs0 + zh * p.c_bs1;
  float* aux = p.aux ? p.aux + zb * p.c_bs0 + zh * p.c_bs1 : nullptr;
  const float* bias = (p.epi & EPI_BIAS) ? p.bias + z * p.bias_bs : nullptr;
  const float alpha = ((p.epi & EPI_ALPHA0) && z == 0) ? p.alpha0 : p.alpha;
  const int quad = warp & 3, half = warp >> 2;
  const uint32_t trow = tmem_base + ((uint32_t)(quad * 32) << 16);
  const int m = m_blk + quad * 32 + lane;
#pragma unroll 1
  for (int c0 = half * (BN / 2); c0 < (half + 1) * (BN / 2); c0 += 32) {
    float v[32];
    tmem_ld32(trow + c0, v);                                         // warp-collective: issued by every lane, rows beyond M included
    if (m < p.M) epilogue_row32(p, v, Cp, aux, bias, alpha, m, n_blk + c0);
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
  gemm_tc_kernel<AM, BMODE, BN><<<grid, NT, Cfg<BN>::SMEM, s>>>(p);
  CVT_LAUNCH_CHECK();
  return cudaSuccess;
}

template <int AM, int BMODE>
cudaError_t launch_bn(const GemmP& p, int bn, cudaStream_t s) {
  if (bn == 128) return launch_cfg<AM, BMODE, 128>(p, s);
  return launch_cfg<AM, BMODE, 64>(p, s);
}

int pick_bn(const GemmP& p) {
  if (p.N % 128 != 0) return 64;
  const long long ctas128 = (long long)cdiv(p.M, TBM) * p.batch * p.splitk * (p.N / 128);
  return ctas128 >= 2 * kNumSMs ? 128 : 64;
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

